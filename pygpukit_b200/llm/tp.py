"""Tensor parallelism for the decode path (net-new: the reference is single-GPU,
docs/scheduler.md:358).

One process per GPU.  Megatron-style sharding decided at weight-creation time:
  column-parallel  q/k/v (heads split), gate/up (intermediate split)   -> no communication
  row-parallel     o_proj (K = local heads * hd), down_proj (K = I/tp) -> partial [M, H] sums,
                   ONE all-reduce (bf16, sum) each on the compute stream (graph-capturable)
  attention        local heads, KV cache sharded by KV head (replicated over world / n_kv ranks
                   when world > n_kv, e.g. Qwen3-30B-A3B at TP8)        -> no communication
  lm_head          column-parallel over the vocabulary; greedy = all-gather of the per-rank
                   (max logit, argmax) pairs (8 bytes per rank), not of the logits
  MoE              experts sharded E/tp per rank with a replicated router -> same all-reduce

The collective itself is NCCL over NVLink/NVSwitch through the C ABI (pgk_nccl_*); the unique id
travels over whatever host-side channel the launcher has (torch.distributed store under torchrun,
or a file).  The shard arithmetic below is pure Python and is what the world_size-2 gloo tests
exercise on CPU.
"""
from __future__ import annotations

import ctypes as C
import os
import time

import numpy as np


def shard_range(n: int, world: int, rank: int, multiple: int = 1) -> tuple[int, int]:
    """[lo, hi) of rank's contiguous shard of n items; n must divide into world * multiple."""
    if n % (world * multiple):
        raise ValueError(f"cannot shard {n} over {world} ranks in multiples of {multiple}")
    per = n // world
    return rank * per, (rank + 1) * per


def shard_plan(config, world: int, rank: int) -> dict:
    """Row/column ranges of every sharded tensor of one layer for `rank` (global index space)."""
    hd = config.head_dim
    q0, q1 = shard_range(config.num_heads, world, rank)
    if world > config.num_kv_heads:      # world / num_kv_heads ranks replicate one KV head
        if world % config.num_kv_heads:
            raise ValueError(f"cannot replicate {config.num_kv_heads} KV heads over {world} ranks")
        kv0 = rank // (world // config.num_kv_heads)
        kv1 = kv0 + 1
    else:
        kv0, kv1 = shard_range(config.num_kv_heads, world, rank)
    if config.vocab_size % world != 0:
        # equal lm_head shards only: the logits all-gather and the (local argmax, global index) exchange both assume
        # V / world rows per rank -- pad the vocabulary (rows of -inf logits) rather than giving one rank the remainder
        raise ValueError(f"vocab_size {config.vocab_size} is not a multiple of the tensor-parallel world size {world}")
    plan = {
        "q_rows": (q0 * hd, q1 * hd), "k_rows": (kv0 * hd, kv1 * hd), "v_rows": (kv0 * hd, kv1 * hd),
        "o_cols": (q0 * hd, q1 * hd),
        "vocab_rows": shard_range(config.vocab_size, world, rank),
    }
    if config.is_moe:
        plan["experts"] = shard_range(config.num_experts, world, rank)
    else:
        # FP8 128-blocks / NVF4 32-blocks must not straddle ranks
        plan["inter"] = shard_range(config.intermediate_size, world, rank, multiple=128)
    return plan


def combine_greedy(pairs: np.ndarray, vocab_offsets: list[int]) -> int:
    """pairs [world, 2] = (max logit as f32, local argmax as int32 bits): global greedy token with
    np.argmax's tie rule (first maximum wins = lowest global index)."""
    vals = pairs[:, 0].astype(np.float32)
    idx = pairs[:, 1].view(np.int32) if pairs.dtype == np.float32 else pairs[:, 1].astype(np.int64)
    best = int(np.argmax(vals))           # first max over ranks; ranks are in vocabulary order
    return int(vocab_offsets[best] + int(idx[best]))


_FUSED_AR = __import__("os").environ.get("PGK_TP_FUSED_AR", "1") != "0"

class TensorParallel:
    """NCCL communicator of this process (rank <-> GPU)."""

    def __init__(self, rank: int, world: int):
        self.rank, self.world = rank, world
        self._ready = False
        self._p2p = False
        self._p2p_max_bytes = 0
        self._ll = os.environ.get("PGK_P2P_LL", "1") != "0"      # LL form of the decode-row all-reduce (A/B switch)

    def greedy_token(self, logits_local, token_buf, vocab_offset: int) -> bool:
        """token_buf[0]: local argmax index -> global greedy token over all ranks' shards (16 bytes per rank over
        NVLink).  False when the peer buffers are not set up (the caller gathers the logits instead)."""
        if not (self._p2p and self._ll):
            return False
        from .._lib import api
        from ..core.stream import current_stream_handle
        api.p2p_ll_argmax_bf16(logits_local.device_ptr, token_buf.device_ptr, int(vocab_offset), current_stream_handle())
        return True

    @classmethod
    def from_env(cls) -> "TensorParallel":
        return cls(int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")))

    def init(self, exchange=None, id_file: str | None = None, allgather=None) -> "TensorParallel":
        """exchange(bytes|None) -> bytes: broadcast of rank 0's 128-byte id (e.g. over
        torch.distributed); or id_file: a path on a shared filesystem.  allgather(bytes) -> list[bytes]
        (one entry per rank, in rank order) additionally sets up the NVLink peer-memory all-reduce
        (pgk_p2p_*), which then replaces NCCL for the row-parallel layers (PGK_P2P=0 keeps NCCL)."""
        from .._lib import api
        if self.world == 1 or self._ready:
            self._ready = True
            return self
        if allgather is not None and os.environ.get("PGK_P2P", "1") != "0":
            from .._lib import load
            hb = int(load().pgk_p2p_handle_bytes())
            mine = (C.c_char * hb)()
            api.p2p_create(self.rank, self.world, mine)
            handles = allgather(bytes(mine))
            blob = (C.c_char * (hb * self.world)).from_buffer_copy(b"".join(handles))
            api.p2p_open(blob)
            self._p2p = True
            self._p2p_max_bytes = int(load().pgk_p2p_max_bytes())   # the kernel's own limit (one region per source rank)
        buf = (C.c_char * 128)()
        if self.rank == 0:
            api.nccl_unique_id(buf)
        if exchange is not None:
            raw = exchange(bytes(buf) if self.rank == 0 else None)
        elif id_file is not None:
            if self.rank == 0:
                with open(id_file + ".tmp", "wb") as f:
                    f.write(bytes(buf))
                os.replace(id_file + ".tmp", id_file)
            t0 = time.time()
            while not os.path.exists(id_file):
                if time.time() - t0 > 120:
                    raise RuntimeError("TensorParallel.init: timed out waiting for the NCCL id file")
                time.sleep(0.05)
            with open(id_file, "rb") as f:
                raw = f.read()
        else:
            raise ValueError("TensorParallel.init needs exchange= or id_file= when world > 1")
        idbuf = (C.c_char * 128).from_buffer_copy(raw)
        api.nccl_init(idbuf, self.rank, self.world)
        self._ready = True
        return self

    def allreduce(self, x) -> None:
        """In-place bf16 sum over ranks on the current stream."""
        if self.world == 1:
            return
        from .._lib import api
        from ..core.stream import current_stream_handle
        api.nccl_allreduce_bf16(x.device_ptr, x.size, current_stream_handle())

    def fuses_allreduce(self, linear, M: int) -> bool:
        """True when a row-parallel layer can run its GEMV and the all-reduce after it as ONE launch
        (`linear.forward_allreduce`): one token, FP8 / NVF4 weights without bias, peer buffers with the LL regions open.
        PGK_TP_FUSED_AR=0 keeps the two launches (A/B timing, parity tests)."""
        from .layers.linear import LinearFP8, LinearNVF4
        if not (_FUSED_AR and self.world > 1 and self._p2p and self._ll and M == 1):
            return False
        if not isinstance(linear, (LinearFP8, LinearNVF4)) or linear.bias is not None:
            return False
        if linear.out_features > 8192 or linear.in_features % 32 or linear.in_features * 2 + 1024 > 200 * 1024:
            return False
        return not isinstance(linear, LinearNVF4) or linear.out_features % 16 == 0

    def allreduce_add(self, part, hidden) -> None:
        """hidden <- hidden + allreduce(part): the tail of a row-parallel layer.  One peer-memory kernel over
        NVLink when the p2p buffers are set up, else ncclAllReduce followed by the in-place add."""
        if self.world == 1:
            return
        from .._lib import api
        from ..core.stream import current_stream_handle
        if self._p2p and part.size % 8 == 0 and part.size <= 8192 and self._ll:
            # one decode row: LL words (payload + exchange number in one 8-byte store), one NVLink hop, no fence
            api.p2p_ll_allreduce_add_bf16(part.device_ptr, hidden.device_ptr, hidden.device_ptr, part.size,
                                          current_stream_handle())
            return
        if self._p2p and part.size % 8 == 0 and part.nbytes <= self._p2p_max_bytes:
            api.p2p_allreduce_add_bf16(part.device_ptr, hidden.device_ptr, hidden.device_ptr, part.size,
                                       current_stream_handle())
            return
        from ..ops.elementwise import add_inplace
        self.allreduce(part)
        add_inplace(hidden, part)

    def allreduce_rows(self, x) -> None:
        """x <- bf16(sum over ranks of x), in place, no residual (the MoE layers exchange their un-combined
        down-projection rows with it: one non-zero contributor per row, so the sum is exact)."""
        if self.world == 1:
            return
        from .._lib import api
        from ..core.stream import current_stream_handle
        if self._p2p and x.size % 8 == 0 and x.nbytes <= self._p2p_max_bytes:
            api.p2p_allreduce_add_bf16(x.device_ptr, None, x.device_ptr, x.size, current_stream_handle())
            return
        self.allreduce(x)

    def allgather_bytes(self, send, recv, nbytes: int) -> None:
        from .._lib import api
        from ..core.stream import current_stream_handle
        api.nccl_allgather(send.device_ptr, recv.device_ptr, nbytes, current_stream_handle())

    def destroy(self) -> None:
        if self._ready and self.world > 1:
            from .._lib import api
            if self._p2p:
                api.p2p_destroy()
                self._p2p = False
            api.nccl_destroy()
        self._ready = False


def torch_allgather():
    """allgather() over an initialised torch.distributed process group (harness plumbing)."""
    import torch.distributed as dist

    def ag(payload: bytes):
        out = [None] * dist.get_world_size()
        dist.all_gather_object(out, payload)
        return out
    return ag


def torch_exchange():
    """exchange() over an initialised torch.distributed process group (harness plumbing)."""
    import torch.distributed as dist

    def ex(payload):
        obj = [payload]
        dist.broadcast_object_list(obj, src=0)
        return obj[0]
    return ex


__all__ = ["shard_range", "shard_plan", "combine_greedy", "TensorParallel", "torch_exchange", "torch_allgather"]
