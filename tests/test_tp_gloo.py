"""Host-side tensor-parallel logic on CPU: two processes over `gloo` (no GPU, no NCCL).

Covers what pygpukit_b200/llm/tp.py decides on the host for N > 1: the shard plan
partitions every sharded tensor exactly once, the NCCL-id exchange over torch.distributed delivers rank 0's
payload, a row-parallel layer's partial sums all-reduce to the unsharded product, and the 8-byte-per-rank
greedy combine picks np.argmax's token (first maximum wins)."""
import os
import socket

import numpy as np
import pytest

torch = pytest.importorskip("torch")
import torch.distributed as dist  # noqa: E402
import torch.multiprocessing as mp  # noqa: E402

from pygpukit_b200.llm.config import named_config  # noqa: E402
from pygpukit_b200.llm.tp import combine_greedy, shard_plan, shard_range, torch_exchange  # noqa: E402

WORLD = 2


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, port: int, out_dir: str) -> None:
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(WORLD))
    dist.init_process_group("gloo", rank=rank, world_size=WORLD)
    try:
        # 1. the id exchange used by TensorParallel.init(exchange=...)
        payload = bytes(range(128)) if rank == 0 else None
        got = torch_exchange()(payload)
        assert got == bytes(range(128))

        # 2. row-parallel layer: every rank multiplies its K-shard, one all-reduce(sum) gives the full product
        cfg, _ = named_config("llama3-70b", max_seq=64)
        plan = shard_plan(cfg, WORLD, rank)
        rng = np.random.default_rng(7)                      # same stream on both ranks
        K, N = 1024, 256
        a = rng.standard_normal((2, K)).astype(np.float32)
        w = rng.standard_normal((N, K)).astype(np.float32)
        k0, k1 = shard_range(K, WORLD, rank, multiple=128)
        part = torch.from_numpy(a[:, k0:k1] @ w[:, k0:k1].T)
        dist.all_reduce(part)
        np.testing.assert_allclose(part.numpy(), a @ w.T, rtol=2e-5, atol=2e-4)

        # 3. column-parallel lm_head + greedy: all-gather of (max, argmax) pairs, ties included
        V = 1000
        logits = rng.standard_normal(V).astype(np.float32)
        logits[[137, 640, 911]] = logits.max() + 1.0        # a three-way tie across both shards
        v0, v1 = shard_range(V, WORLD, rank)
        local = logits[v0:v1]
        pair = torch.tensor([float(local.max()), float(np.argmax(local))], dtype=torch.float64)
        pairs = [torch.zeros(2, dtype=torch.float64) for _ in range(WORLD)]
        dist.all_gather(pairs, pair)
        arr = np.stack([p.numpy() for p in pairs])
        tok = combine_greedy(arr, [shard_range(V, WORLD, r)[0] for r in range(WORLD)])
        assert tok == int(np.argmax(logits)) == 137

        # 4. the plan itself is written out for the parent to check the partition
        np.save(os.path.join(out_dir, f"plan{rank}.npy"), np.array(
            [plan["q_rows"], plan["k_rows"], plan["v_rows"], plan["o_cols"], plan["inter"], plan["vocab_rows"]]))
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo(tmp_path):
    mp.spawn(_worker, args=(_free_port(), str(tmp_path)), nprocs=WORLD, join=True)
    p0, p1 = (np.load(tmp_path / f"plan{r}.npy") for r in range(WORLD))
    cfg, _ = named_config("llama3-70b", max_seq=64)
    hd = cfg.head_dim
    totals = [cfg.num_heads * hd, cfg.num_kv_heads * hd, cfg.num_kv_heads * hd, cfg.num_heads * hd,
              cfg.intermediate_size, cfg.vocab_size]
    for (lo0, hi0), (lo1, hi1), n in zip(p0, p1, totals):
        assert lo0 == 0 and hi0 == lo1 and hi1 == n          # contiguous, disjoint, complete


@pytest.mark.parametrize("name,world", [("llama3-70b", 2), ("llama3-70b", 4), ("llama3-70b", 8),
                                        ("qwen3-30b-a3b", 2), ("qwen3-30b-a3b", 4), ("qwen3-30b-a3b", 8),
                                        ("qwen3-8b", 2), ("qwen3-8b", 8)])
def test_shard_plan_partitions(name, world):
    """Every rank's ranges tile the global index space; FP8 128-blocks / NVF4 32-blocks never straddle ranks."""
    cfg, _ = named_config(name, max_seq=64)
    plans = [shard_plan(cfg, world, r) for r in range(world)]
    keys = ["q_rows", "k_rows", "v_rows", "o_cols", "vocab_rows"] + (["experts"] if cfg.is_moe else ["inter"])
    hd = cfg.head_dim
    replicated = world > cfg.num_kv_heads          # e.g. Qwen3-30B-A3B (4 KV heads) at TP8
    for k in keys:
        if replicated and k in ("k_rows", "v_rows"):
            rep = world // cfg.num_kv_heads        # `rep` consecutive ranks hold the same whole KV head
            for r, p in enumerate(plans):
                assert p[k] == ((r // rep) * hd, (r // rep + 1) * hd)
            continue
        edge = 0
        for p in plans:
            lo, hi = p[k]
            assert lo == edge and hi > lo
            edge = hi
        total = {"q_rows": cfg.num_heads * cfg.head_dim, "o_cols": cfg.num_heads * cfg.head_dim,
                 "k_rows": cfg.num_kv_heads * cfg.head_dim, "v_rows": cfg.num_kv_heads * cfg.head_dim,
                 "vocab_rows": cfg.vocab_size, "experts": getattr(cfg, "num_experts", 0),
                 "inter": cfg.intermediate_size}[k]
        assert edge == total
        if k in ("q_rows", "k_rows", "v_rows", "o_cols", "inter"):
            assert all(p[k][0] % 128 == 0 for p in plans)


def test_shard_range_rejects_uneven():
    with pytest.raises(ValueError):
        shard_range(100, 8, 0)
    with pytest.raises(ValueError):
        shard_range(1024, 4, 0, multiple=512)


def test_shard_plan_rejects_vocabulary_remainder():
    """Equal lm_head shards only: the logits all-gather and the 16-byte (best logit, global index) exchange both assume
    V / world rows per rank, so a vocabulary that does not divide is refused at plan time."""
    cfg, _ = named_config("small-qwen3", max_seq=64)
    cfg.vocab_size = 2049
    with pytest.raises(ValueError, match="vocab_size"):
        shard_plan(cfg, 2, 0)


def test_fuses_allreduce_needs_peer_buffers_and_one_row():
    """TensorParallel.fuses_allreduce (GEMV + all-reduce in one launch) is a host decision: never without the
    peer-memory LL regions, never for more than one row, never for BF16 weights or a layer with a bias."""
    from pygpukit_b200.llm.layers.linear import LinearFP8, LinearNVF4
    from pygpukit_b200.llm.tp import TensorParallel

    class _FP8(LinearFP8):                      # host-only stand-ins: the decision reads shapes and the bias, nothing else
        def __init__(self, n, k, bias=None):
            self.in_features, self.out_features, self.bias = k, n, bias

    class _NVF4(LinearNVF4):
        def __init__(self, n, k, bias=None):
            self.in_features, self.out_features, self.bias = k, n, bias

    tp = TensorParallel(0, 2)
    tp._p2p, tp._ll = False, True
    assert not tp.fuses_allreduce(_FP8(4096, 2048), 1)          # NCCL fallback: two launches
    tp._p2p = True
    assert tp.fuses_allreduce(_FP8(4096, 2048), 1) and tp.fuses_allreduce(_NVF4(8192, 4096), 1)
    assert not tp.fuses_allreduce(_FP8(4096, 2048), 2)          # batch rows go through the slab all-reduce
    assert not tp.fuses_allreduce(_FP8(16384, 2048), 1)         # one LL region holds 8192 words per rank
    assert not tp.fuses_allreduce(_FP8(4096, 2048, bias=object()), 1)
    assert not tp.fuses_allreduce(_NVF4(4104, 4096), 1)         # NVF4 columns in 16s
    assert not tp.fuses_allreduce(object(), 1)                  # BF16 / unknown layers
    tp._ll = False
    assert not tp.fuses_allreduce(_FP8(4096, 2048), 1)
    assert not TensorParallel(0, 1).fuses_allreduce(_FP8(4096, 2048), 1)
