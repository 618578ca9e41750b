"""CausalTransformerModel (mirror of pygpukit/llm/models/causal.py:78-1493, decode path only).

Same constructor (config, embed_tokens, blocks, final_norm, lm_head, position_embed, spec) and the
reference's entry points on this path: `__call__`, `get_logits`, `generate`,
`_decode_step_fixed_cache`, `_decode_step_fixed_cache_batch`, `snapshot/restore` free.  Every
forward runs on the fixed KV cache; the KV-concat path of the reference (`past_key_values` tuples
growing by realloc, causal.py:245-256 of attention.py) is not reproduced -- `past_key_values` is
the number of cached positions.

Fast path (net-new orchestration, same math): `decode_step_device` runs one whole token on the
device -- token id, position and context length live in int32 device buffers, greedy sampling is
fused into the lm_head GEMV and feeds the next step's token buffer -- so the step captures into ONE
CUDA graph (`init_decode_graph`) that replays for every position with no host round trip.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from ..._lib import api
from ...core.array import GPUArray
from ...core.dtypes import bfloat16, int32
from ...core.factory import from_numpy
from ...core.stream import CudaGraph, Stream, current_stream_handle as _st, use_stream
from ..buffers import DecodeBuffers
from ..config import ModelSpec, TransformerConfig
from ..layers.block import TransformerBlock
from ..layers.linear import can_fuse_norm
from ..layers.norm import Norm
from ..decode.accept import accept_jacobi, accept_speculative
from ..sampling import sample_token


def _to_float32_logits(logits_np: np.ndarray) -> np.ndarray:
    """bf16-as-uint16 -> f32 (causal.py:62-71)."""
    if logits_np.dtype == np.uint16:
        return (logits_np.astype(np.uint32) << 16).view(np.float32)
    return logits_np.astype(np.float32)


class CausalTransformerModel:
    def __init__(self, config: TransformerConfig, embed_tokens: GPUArray, blocks: list[TransformerBlock],
                 final_norm: Norm, lm_head: GPUArray | None = None, position_embed: GPUArray | None = None,
                 spec: ModelSpec | None = None, comm=None):
        self.config = config
        self.embed_tokens = embed_tokens
        self.blocks = blocks
        self.final_norm = final_norm
        self._lm_head = lm_head
        self.position_embed = position_embed
        self.spec = spec
        self.comm = comm
        self._buffers: DecodeBuffers | None = None
        self._max_seq = 0
        self._graph: CudaGraph | None = None
        self._graph_stream: Stream | None = None
        self._gathered_logits: GPUArray | None = None
        # one-launch step kernel (llm/step_program.py), built on first use when the model fits it.  Opt-in: PGK_STEP=1 or
        # `use_step_program = True`.  Measured on B200 (DESIGN.md "step program"): its ring streams the BF16 lm_head at the
        # HBM peak, but an op boundary inside the kernel (device-side grid barrier + re-staging the activation vector)
        # costs ~7 us against ~3.3 us for a programmatic-dependent-launch boundary, and the CUDA-core dequantisation of the
        # FP8 / NVF4 rows runs at about the HBM rate, so it cannot catch up -- the per-op launches stay the default.
        env = os.environ.get("PGK_STEP", "0")
        self.use_step_program = env not in ("0", "", "auto")
        self._step = None
        self._step_tried = False
        self.step_unsupported_reason: str | None = None

    # ------------------------------------------------------------------ setup
    @property
    def lm_head(self) -> GPUArray | None:
        return self._lm_head

    @property
    def _head(self) -> GPUArray:
        return self._lm_head if self._lm_head is not None else self.embed_tokens

    @property
    def tp(self) -> int:
        return self.comm.world if self.comm is not None else 1

    def init_decode(self, max_seq_len: int, max_batch_size: int = 1, num_sequences: int = 1) -> DecodeBuffers:
        """Allocate the fixed KV caches (all layers), the shared RoPE tables and the step buffers.
        num_sequences = S > 1: S independent sequences decode together (one row each per step; continuous batching,
        SURVEY.md section 8f-2): every layer holds S cache regions, position / context are arrays of S."""
        S = max(1, int(num_sequences))
        if S > 1:
            a0 = self.blocks[0].attn
            need = 2 * len(self.blocks) * S * a0.num_kv_heads * max_seq_len * a0.head_dim * 2
            free, total = C.c_size_t(0), C.c_size_t(0)
            api.mem_info(C.byref(free), C.byref(total))
            if need > free.value - (2 << 30):
                raise MemoryError(f"{S} sequences x {max_seq_len} positions need {need / 2**30:.1f} GiB of KV cache, "
                                  f"{free.value / 2**30:.1f} GiB free")
        tables = None
        for blk in self.blocks:
            blk.attn.init_fixed_cache(max_seq_len, "bfloat16", rope_tables=tables, num_sequences=S)
            if tables is None and self.config.use_rope:
                tables = (blk.attn._cos_dev, blk.attn._sin_dev)
        v_local = self._head.shape[0]
        self._buffers = DecodeBuffers.allocate(self.config, "bfloat16", vocab_size=v_local,
                                               max_batch_size=max(max_batch_size, S), num_sequences=S)
        self._max_seq = max_seq_len
        self._num_sequences = S
        if self.tp > 1:
            self._gathered_logits = GPUArray((self.tp * v_local,), bfloat16)
        self._graph = None
        self._step, self._step_tried = None, False
        return self._buffers

    def select_sequence(self, s: int) -> None:
        """The single-sequence entry points (prefill, __call__, forward_fixed_cache) work on sequence `s`."""
        for blk in self.blocks:
            blk.attn.select_sequence(s)

    def step_program(self):
        """The one-launch decode step of this model, or None when the model does not fit it (reason kept in
        `step_unsupported_reason`) or it is switched off."""
        if not self.use_step_program:
            return None
        if not self._step_tried:
            from ..step_program import StepProgram, Unsupported
            self._step_tried = True
            try:
                self._step = StepProgram.from_model(self)
                self.fused_exchange = self.tp > 1        # the row-parallel all-reduces happen inside the step kernel
            except Unsupported as e:
                self._step, self.step_unsupported_reason = None, str(e)
            except RuntimeError as e:    # the library refused an op shape (message says which)
                if "step_create failed" not in str(e):
                    raise
                self._step, self.step_unsupported_reason = None, str(e)
        return self._step

    @property
    def buffers(self) -> DecodeBuffers:
        assert self._buffers is not None, "Call init_decode first"
        return self._buffers

    # ------------------------------------------------------------------ pieces
    def _embed(self, token_ids: list[int] | None, M: int, start_position: int, b: DecodeBuffers) -> GPUArray:
        hidden = b.hidden if b.hidden.shape[0] == M else b.hidden.slice_rows(M)
        if token_ids is not None:
            ids = np.asarray(token_ids, np.int32)
            api.memcpy_h2d(b.token_buf.device_ptr, ids.ctypes.data_as(C.c_void_p), ids.nbytes, _st())
            api.stream_sync(_st()) if _st() is not None else api.device_sync()   # ids is pageable
        self._embed_rows(hidden, M, start_position, None, b)
        return hidden

    def _embed_rows(self, hidden: GPUArray, M: int, position: int, pos_buf: GPUArray | None, b: DecodeBuffers) -> None:
        """hidden[m] = wte[token_buf[m]] (+ wpe[position + m] for learned positions, GPT-2), one launch."""
        H = self.config.hidden_size
        if self.position_embed is None:
            api.embedding_lookup_ptr(self.embed_tokens.device_ptr, hidden.device_ptr, H, b.token_buf.device_ptr, M, 2, _st())
            return
        if pos_buf is None and position + M > self.position_embed.shape[0]:
            raise ValueError(f"position {position + M - 1} exceeds the learned position table ({self.position_embed.shape[0]})")
        api.embedding_lookup_pos(self.embed_tokens.device_ptr, self.position_embed.device_ptr, hidden.device_ptr, H,
                                 b.token_buf.device_ptr, M, int(position), None if pos_buf is None else pos_buf.device_ptr,
                                 2, _st())

    def _layers(self, hidden: GPUArray, position: int, context_len: int, b: DecodeBuffers,
                pos_buf: GPUArray | None = None, ctx_buf: GPUArray | None = None) -> GPUArray:
        for blk in self.blocks:
            blk.decode(hidden, position, context_len, b, pos_buf, ctx_buf)
        return hidden

    def _final(self, hidden: GPUArray, b: DecodeBuffers) -> GPUArray:
        M = hidden.shape[0]
        out = b.norm_out if b.norm_out.shape[0] == M else b.norm_out.slice_rows(M)
        return self.final_norm(hidden, out=out)

    # ------------------------------------------------------------------ reference entry points
    def _decode_step_fixed_cache(self, token_id: int, position: int, context_len: int) -> GPUArray:
        """One token through all blocks + final norm -> hidden [1, H] (causal.py:799-840)."""
        return self._decode_step_fixed_cache_batch([token_id], position, context_len)

    def _decode_step_fixed_cache_batch(self, token_ids: list[int], start_position: int, context_len: int) -> GPUArray:
        """M consecutive tokens of one sequence -> hidden [M, H] (causal.py:842-891)."""
        b = self.buffers
        M = len(token_ids)
        if M > b.max_batch_size:
            raise ValueError(f"batch of {M} tokens exceeds max_batch_size {b.max_batch_size} (init_decode)")
        if context_len > self._max_seq:
            raise ValueError(f"context_len {context_len} exceeds the fixed cache ({self._max_seq})")
        hidden = self._embed(token_ids, M, start_position, b)
        self._layers(hidden, start_position, context_len, b)
        return self._final(hidden, b)

    def get_logits(self, hidden: GPUArray, *, out: GPUArray | None = None) -> GPUArray:
        """hidden [M, H] @ lm_head[V, H]^T -> logits [M, V] (causal.py:167-177).  Under TP every rank
        returns the logits of its vocabulary shard."""
        M = hidden.shape[0]
        V = self._head.shape[0]
        if out is None:
            out = GPUArray((M, V), bfloat16)
        api.gemv_bf16(hidden.device_ptr, self._head.device_ptr, out.device_ptr, M, self.config.hidden_size, V, _st())
        return out

    def _full_logits_host(self, hidden_row: GPUArray) -> np.ndarray:
        """Logits [V_global] of one final-normed row on the host; under TP the per-rank vocabulary shards are
        all-gathered first (ranks are in vocabulary order)."""
        logits = self.get_logits(hidden_row)
        if self.tp == 1:
            return logits.to_numpy()[0]
        V = self._head.shape[0]
        self.comm.allgather_bytes(logits, self._gathered_logits, V * 2)
        return self._gathered_logits.to_numpy()

    def _greedy_first_token(self, last: GPUArray, b: DecodeBuffers) -> None:
        """argmax(lm_head(last)) -> b.token_buf on the device (sharded lm_head: gather, then argmax)."""
        V, H = self._head.shape[0], self.config.hidden_size
        if self.tp == 1:
            api.gemv_bf16_argmax(last.device_ptr, self._head.device_ptr, b.logits.device_ptr, b.token_buf.device_ptr,
                                 H, V, _st())
            return
        if self._ll_greedy():
            api.gemv_bf16_argmax(last.device_ptr, self._head.device_ptr, b.logits.device_ptr, b.token_buf.device_ptr, H, V, _st())
            self.comm.greedy_token(b.logits, b.token_buf, self._vocab_offset())
            return
        api.gemv_bf16(last.device_ptr, self._head.device_ptr, b.logits.device_ptr, 1, H, V, _st())
        self.comm.allgather_bytes(b.logits, self._gathered_logits, V * 2)
        api.argmax(self._gathered_logits.device_ptr, b.token_buf.device_ptr, self.tp * V, 2, _st())

    def _vocab_offset(self) -> int:
        return self.comm.rank * (self.config.vocab_size // self.tp) if self.tp > 1 else 0

    def _ll_greedy(self) -> bool:
        """Sharded lm_head: exchange (best logit, global index) pairs over NVLink (16 bytes per rank) instead of
        gathering the logits -- needs the peer-memory buffers."""
        return self.tp > 1 and getattr(self.comm, "_p2p", False) and getattr(self.comm, "_ll", False)

    def __call__(self, input_ids: list[int], position_ids: list[int] | None = None, past_key_values=None,
                 use_cache: bool = False):
        """hidden states of `input_ids` (causal.py:108-160); past_key_values = cached positions."""
        start = position_ids[0] if position_ids else int(past_key_values or 0)
        hidden = None
        slab = self.buffers.max_batch_size
        for i in range(0, len(input_ids), slab):
            ids = input_ids[i:i + slab]
            hidden = self._decode_step_fixed_cache_batch(ids, start + i, start + i + len(ids))
        return hidden, (start + len(input_ids) if use_cache else None)

    # ------------------------------------------------------------------ device-resident step
    def decode_step_device(self) -> None:
        """One greedy token entirely on the device: reads token_buf / position_buf / context_buf,
        writes the sampled token to next_token AND token_buf, then advances position / context."""
        b = self.buffers
        if getattr(self, "_num_sequences", 1) > 1:
            return self._decode_step_device_multi()
        step = self.step_program()
        if step is not None:             # ONE persistent kernel: embedding .. argmax .. position advance
            step.run(_st())
            return
        hidden = b.hidden.slice_rows(1) if b.hidden.shape[0] != 1 else b.hidden
        self._embed_rows(hidden, 1, 0, b.position_buf, b)
        self._layers(hidden, 0, self._max_seq, b, b.position_buf, b.context_buf)
        V, H = self._head.shape[0], self.config.hidden_size
        fn = self.final_norm
        ll = self._ll_greedy()
        if can_fuse_norm(fn, H):        # final RMSNorm rides inside the lm_head launch
            g, e = fn.weight.device_ptr, float(fn.eps)
            if self.tp == 1 or ll:
                api.rmsnorm_gemv_bf16_argmax(hidden.device_ptr, g, e, self._head.device_ptr, b.logits.device_ptr,
                                             b.token_buf.device_ptr, H, V, _st())
            else:
                api.rmsnorm_gemv_bf16(hidden.device_ptr, g, e, self._head.device_ptr, b.logits.device_ptr, 1, H, V,
                                      None, None, _st())
        else:
            x = self._final(hidden, b)
            if self.tp == 1 or ll:
                api.gemv_bf16_argmax(x.device_ptr, self._head.device_ptr, b.logits.device_ptr, b.token_buf.device_ptr,
                                     H, V, _st())
            else:
                api.gemv_bf16(x.device_ptr, self._head.device_ptr, b.logits.device_ptr, 1, H, V, _st())
        if ll:
            self.comm.greedy_token(b.logits, b.token_buf, self._vocab_offset())
        elif self.tp > 1:
            self.comm.allgather_bytes(b.logits, self._gathered_logits, V * 2)
            api.argmax(self._gathered_logits.device_ptr, b.token_buf.device_ptr, self.tp * V, 2, _st())
        api.memcpy_d2d(b.next_token.device_ptr, b.token_buf.device_ptr, 4, _st())
        api.decode_advance(b.position_buf.device_ptr, b.context_buf.device_ptr, 1, _st())

    def _decode_step_device_multi(self) -> None:
        """One greedy token for each of the S independent sequences, entirely on the device: S rows through every
        linear layer (the weights are read once for all rows), ONE attention launch per layer over the S cache regions,
        per-row argmax, positions / contexts advanced."""
        b = self.buffers
        S = self._num_sequences
        if self.tp > 1:
            raise ValueError("multi-sequence decode runs on unsharded models")
        hidden = b.hidden if b.hidden.shape[0] == S else b.hidden.slice_rows(S)
        H, V = self.config.hidden_size, self._head.shape[0]
        api.embedding_lookup_ptr(self.embed_tokens.device_ptr, hidden.device_ptr, H, b.token_buf.device_ptr, S, 2, _st())
        for blk in self.blocks:
            blk.decode(hidden, 0, self._max_seq, b, b.position_buf, b.context_buf, multi=True)
        logits = b.logits if b.logits.shape[0] == S else b.logits.slice_rows(S)
        fn = self.final_norm
        if can_fuse_norm(fn, H):
            api.rmsnorm_gemv_bf16(hidden.device_ptr, fn.weight.device_ptr, float(fn.eps), self._head.device_ptr,
                                  logits.device_ptr, S, H, V, None, None, _st())
        else:
            x = self._final(hidden, b)
            api.gemv_bf16(x.device_ptr, self._head.device_ptr, logits.device_ptr, S, H, V, _st())
        api.argmax_rows(logits.device_ptr, b.next_token.device_ptr, b.token_buf.device_ptr, S, V, 2, _st())
        api.decode_advance_n(b.position_buf.device_ptr, b.context_buf.device_ptr, 1, S, _st())

    def set_decode_state(self, token_id, position) -> None:
        """position_buf = position, context_buf = position + 1, token_buf = token_id (one small copy).  With S
        sequences token_id is a list of S tokens and position an int (the same for all) or a list of S."""
        b = self.buffers
        S = getattr(self, "_num_sequences", 1)
        toks = [int(t) for t in token_id] if isinstance(token_id, (list, tuple, np.ndarray)) else [int(token_id)]
        pos = [int(p) for p in position] if isinstance(position, (list, tuple, np.ndarray)) else [int(position)] * S
        if len(toks) != S or len(pos) != S:
            raise ValueError(f"decode state needs {S} tokens / positions")
        st = np.array(pos + [p + 1 for p in pos] + toks, np.int32)
        api.memcpy_h2d(b.state_buf.device_ptr, st.ctypes.data_as(C.c_void_p), st.nbytes, _st())
        api.stream_sync(_st()) if _st() is not None else api.device_sync()

    def _pinned_state(self):
        """Pinned host staging: int32[3 S] step inputs + int32[S] results (allocated once per S)."""
        S = getattr(self, "_num_sequences", 1)
        if getattr(self, "_pin", None) is None or getattr(self, "_pin_S", 0) != S:
            h = C.c_void_p()
            api.host_alloc(C.byref(h), 16 * S)
            self._pin, self._pin_S = h, S
            self._pin_np = np.ctypeslib.as_array(C.cast(h, C.POINTER(C.c_int32)), shape=(4 * S,))
        return self._pin, self._pin_np

    def decode_step_host(self, token_id, position: int):
        """One greedy decode step with the HOST in the loop (what a serving loop does): token ids and positions go
        host -> device from pinned memory (12 bytes per sequence), the step graph replays, the sampled tokens come back
        device -> host (4 bytes per sequence).  Returns the next token (a list with S sequences)."""
        assert self._graph is not None, "Call init_decode_graph first"
        S = getattr(self, "_num_sequences", 1)
        pin, arr = self._pinned_state()
        toks = list(token_id) if S > 1 else [token_id]
        arr[0:S] = position
        arr[S:2 * S] = position + 1
        arr[2 * S:3 * S] = toks
        b = self.buffers
        sh = self._graph_stream.handle
        api.memcpy_h2d(b.state_buf.device_ptr, pin, 12 * S, sh)
        self._graph.replay(self._graph_stream)
        api.memcpy_d2h(pin.value + 12 * S, b.next_token.device_ptr, 4 * S, sh)
        api.stream_sync(sh)
        return [int(t) for t in arr[3 * S:4 * S]] if S > 1 else int(arr[3])

    def init_decode_graph(self) -> None:
        """Capture decode_step_device into one CUDA graph (replaces the reference's per-layer
        pre/post-SDPA graphs with syncs in between, decode/m1_graph.py:248-457)."""
        self._graph_stream = Stream()
        with use_stream(self._graph_stream):
            S = getattr(self, "_num_sequences", 1)
            self.set_decode_state([0] * S if S > 1 else 0, 0)
            self.decode_step_device()            # warm-up outside capture (workspaces, attributes)
            self._graph_stream.synchronize()
            g = CudaGraph()
            with g.capture(self._graph_stream):
                self.decode_step_device()
            self._graph = g

    def decode_graph_replay(self) -> None:
        assert self._graph is not None, "Call init_decode_graph first"
        self._graph.replay(self._graph_stream)

    # ------------------------------------------------------------------ generation
    def prefill(self, input_ids: list[int]) -> GPUArray:
        """Run the prompt through the fixed cache; returns the final-norm hidden of the last slab."""
        hidden, _ = self(list(input_ids), use_cache=True)
        return hidden

    def generate(self, input_ids: list[int], max_new_tokens: int = 20, temperature: float = 1.0, top_k: int = 50,
                 top_p: float = 0.9, eos_token_id: int | None = None, use_cache: bool = True,
                 gpu_sampling: bool = False, use_graph: bool = True) -> list[int]:
        """Autoregressive generation (causal.py:179-255).  temperature == 0 takes the device-resident
        greedy path (one graph replay per token, tokens read back once at the end unless an EOS id
        must be checked); otherwise `sample_token` on the host, or on the device with gpu_sampling=True."""
        tokens = list(input_ids)
        if self._buffers is None or self._max_seq < len(tokens) + max_new_tokens:
            self.init_decode(max(self.config.max_position_embeddings, len(tokens) + max_new_tokens),
                             max_batch_size=min(64, max(1, len(tokens))))
        if max_new_tokens <= 0:
            return tokens
        b = self.buffers
        greedy_dev = temperature == 0
        if greedy_dev and use_graph and self._graph is None:
            self.init_decode_graph()             # before prefill: capture clobbers cache slot 0 / step state
        hidden = self.prefill(tokens)
        last = hidden.row(hidden.shape[0] - 1).view((1, hidden.shape[1]))
        if greedy_dev:
            self._greedy_first_token(last, b)
            first = int(b.token_buf.to_numpy()[0])
            tokens.append(first)
            if (eos_token_id is not None and first == eos_token_id) or max_new_tokens == 1:
                return tokens
            n = max_new_tokens - 1
            self.set_decode_state(first, len(tokens) - 1)
            history = GPUArray((n,), int32)
            stream = self._graph_stream if (use_graph and self._graph is not None) else None
            api.device_sync()
            with use_stream(stream):
                for i in range(n):
                    if stream is not None:
                        self.decode_graph_replay()
                    else:
                        self.decode_step_device()
                    api.memcpy_d2d(history.device_ptr + 4 * i, b.next_token.device_ptr, 4, _st())
                    if eos_token_id is not None and int(b.next_token.to_numpy()[0]) == eos_token_id:
                        n = i + 1
                        break
                out = history.to_numpy()[:n]
            tokens.extend(int(v) for v in out)
            return tokens
        # sampling path: on the device when asked (gpu_sampling: the logits never leave the GPU, 4 bytes come back per
        # token; ops.sampling.sample_token_gpu, same semantics as sample_token with the generator of set_sampling_seed),
        # else the reference's host path (full logits D2H + NumPy, models/causal.py:62-71)
        on_device = gpu_sampling and self.tp == 1
        for step in range(max_new_tokens):
            if on_device:
                from ...ops.sampling import sample_token_gpu
                row = None if b.logits is None else (b.logits if b.logits.shape[0] == 1 else b.logits.slice_rows(1))
                nxt = sample_token_gpu(self.get_logits(last, out=row), temperature, top_k, top_p)
            else:
                logits = _to_float32_logits(self._full_logits_host(last))
                nxt = sample_token(logits, temperature, top_k, top_p)
            tokens.append(nxt)
            if (eos_token_id is not None and nxt == eos_token_id) or step == max_new_tokens - 1:
                break
            last = self._decode_step_fixed_cache(nxt, len(tokens) - 1, len(tokens))
        return tokens

    def snapshot_kv_cache(self) -> list[tuple[np.ndarray, np.ndarray]]:
        """Every layer's K / V cache on the host (causal.py:1003-1015): BF16 bit patterns, [sequences x n_kv, max_seq,
        head_dim].  The speculative / Jacobi paths here do not need it (lookahead pointers); kept for callers that do."""
        return [(blk.attn._k_cache.to_numpy().copy(), blk.attn._v_cache.to_numpy().copy()) for blk in self.blocks]

    def restore_kv_cache(self, snapshot: list[tuple[np.ndarray, np.ndarray]]) -> None:
        """Copy a snapshot back INTO the existing cache arrays (causal.py:1017-1034): device pointers are unchanged, so
        a captured decode graph stays valid."""
        if len(snapshot) != len(self.blocks):
            raise ValueError(f"snapshot holds {len(snapshot)} layers, the model has {len(self.blocks)}")
        for blk, (k_np, v_np) in zip(self.blocks, snapshot):
            blk.attn._k_cache.copy_from_numpy(np.ascontiguousarray(k_np, np.uint16))
            blk.attn._v_cache.copy_from_numpy(np.ascontiguousarray(v_np, np.uint16))

    def generate_stream(self, input_ids: list[int], max_new_tokens: int = 20, temperature: float = 1.0, top_k: int = 50,
                        top_p: float = 0.9, eos_token_id: int | None = None, gpu_sampling: bool = False):
        """generate() as a generator (causal.py:257-322): yields each new token as soon as it exists.  temperature == 0
        and gpu_sampling keep the logits on the device (4 bytes per token come back); otherwise `sample_token` runs on
        the host, as upstream."""
        from ...ops.sampling import sample_token_gpu
        tokens = list(input_ids)
        if max_new_tokens <= 0:
            return
        if self._buffers is None or self._max_seq < len(tokens) + max_new_tokens:
            self.init_decode(max(self.config.max_position_embeddings, len(tokens) + max_new_tokens),
                             max_batch_size=min(64, max(1, len(tokens))))
        b = self.buffers
        on_device = (gpu_sampling or temperature == 0) and self.tp == 1
        hidden = self.prefill(tokens)
        last = hidden.row(hidden.shape[0] - 1).view((1, hidden.shape[1]))
        for step in range(max_new_tokens):
            if on_device:
                row = None if b.logits is None else (b.logits if b.logits.shape[0] == 1 else b.logits.slice_rows(1))
                nxt = sample_token_gpu(self.get_logits(last, out=row), temperature, top_k, top_p)
            else:
                nxt = sample_token(_to_float32_logits(self._full_logits_host(last)), temperature, top_k, top_p)
            tokens.append(nxt)
            yield nxt
            if (eos_token_id is not None and nxt == eos_token_id) or step == max_new_tokens - 1:
                return
            last = self._decode_step_fixed_cache(nxt, len(tokens) - 1, len(tokens))

    # ------------------------------------------------------------------ self-speculative decode
    def set_lookahead_confirmed_pos(self, pos: int) -> None:
        """Tokens 0 .. pos-1 are committed in every layer's cache (causal.py:1204-1211)."""
        for blk in self.blocks:
            blk.attn.set_confirmed_pos(pos)

    def reset_lookahead_all(self) -> None:
        for blk in self.blocks:
            blk.attn.reset_lookahead()

    def commit_lookahead_all(self, n_accepted: int) -> None:
        for blk in self.blocks:
            blk.attn.commit_lookahead(n_accepted)

    def get_lookahead_confirmed_pos(self) -> int:
        return self.blocks[0].attn.get_confirmed_pos()

    def _draft_forward_early_layers(self, token_id: int, position: int, context_len: int, num_draft_layers: int) -> GPUArray:
        """The first `num_draft_layers` blocks as the draft model (causal.py:1036-1083): hidden [1, H] BEFORE the final
        norm.  It writes the early layers' K/V at `position`; the verification pass overwrites those rows."""
        b = self.buffers
        hidden = self._embed([token_id], 1, position, b)
        for blk in self.blocks[:max(0, min(num_draft_layers, len(self.blocks)))]:
            blk.decode(hidden, position, context_len, b)
        return hidden

    def _draft_get_logits(self, hidden: GPUArray) -> GPUArray:
        """final norm + lm_head on the draft's hidden state (causal.py:1085-1094)."""
        return self.get_logits(self._final(hidden, self.buffers))

    def decode_step_self_speculative_lookahead(self, token_id: int, max_draft_tokens: int = 4,
                                               draft_layers: int = 8) -> tuple[list[int], dict]:
        """One self-speculative step on the fixed cache (causal.py:1096-1198): the early layers draft
        `max_draft_tokens` greedy tokens, the full model verifies them in ONE batched pass, tokens are accepted up to
        the first disagreement (the full model's token replaces the rejected draft).  Call
        `set_lookahead_confirmed_pos(prefill_len)` after the prefill.  Greedy argmaxes run on the device (4 bytes per
        token come back, not the logits).  The reference re-runs the accepted tokens one by one after a partial accept;
        that is not needed here: the verification pass has already written, for every layer, the K/V rows of exactly
        the inputs that were accepted (input i sits at position confirmed + i), and rows past the committed position
        are never read."""
        if self.tp != 1:
            raise NotImplementedError("self-speculative decode runs on one GPU (the draft needs the full lm_head)")
        b = self.buffers
        n_draft = int(max_draft_tokens)
        if not 1 <= n_draft <= b.max_batch_size:
            raise ValueError(f"max_draft_tokens must be in [1, max_batch_size = {b.max_batch_size}]")
        confirmed = self.get_lookahead_confirmed_pos()
        if confirmed + n_draft > self._max_seq:
            raise ValueError(f"speculating {n_draft} tokens from position {confirmed} exceeds the fixed cache ({self._max_seq})")
        V = self._head.shape[0]
        tok = GPUArray((max(n_draft, 1),), int32)
        # 1. draft with the early layers
        self.reset_lookahead_all()
        draft, cur = [], int(token_id)
        for i in range(n_draft):
            hidden = self._draft_forward_early_layers(cur, confirmed + i, confirmed + i + 1, draft_layers)
            logits = self._draft_get_logits(hidden)
            api.argmax_rows(logits.device_ptr, tok.device_ptr, None, 1, V, 2, _st())
            cur = int(tok.to_numpy()[0])
            draft.append(cur)
        # 2. verify with the full model in one batched pass
        self.reset_lookahead_all()
        verify_input = [int(token_id)] + draft[:-1]
        hidden = self._decode_step_fixed_cache_batch(verify_input, confirmed, confirmed + len(verify_input))
        logits = self.get_logits(hidden)
        api.argmax_rows(logits.device_ptr, tok.device_ptr, None, len(verify_input), V, 2, _st())
        target = [int(t) for t in tok.to_numpy()[:len(verify_input)]]
        # 3. accept up to the first disagreement
        accepted = accept_speculative(draft, target)
        self.commit_lookahead_all(len(accepted))
        stats = {"draft_count": len(draft), "accepted_count": sum(1 for a, d in zip(accepted, draft) if a == d)}
        return accepted, stats

    # ------------------------------------------------------------------ Jacobi decode
    def _argmax_tokens(self, hidden: GPUArray, scratch: GPUArray) -> list[int]:
        """greedy token of every row of final-normed hidden states: lm_head + row argmax on the device"""
        M = hidden.shape[0]
        logits = self.get_logits(hidden)
        api.argmax_rows(logits.device_ptr, scratch.device_ptr, None, M, self._head.shape[0], 2, _st())
        return [int(t) for t in scratch.to_numpy()[:M]]

    def _init_jacobi_guess_lookahead(self, last_token: int, n_tokens: int, strategy: str) -> list[int]:
        """Initial guess of the next n_tokens (causal.py:1305-1355): "repeat" the last token, the "ngram" cache of
        earlier steps, or a sequential "greedy" decode on the lookahead rows (exact, but as slow as plain decoding)."""
        if strategy == "repeat":
            return [int(last_token)] * n_tokens
        if strategy == "ngram":
            cached = getattr(self, "_ngram_cache", {}).get(int(last_token))
            return list(cached[:n_tokens]) if cached is not None and len(cached) >= n_tokens else [int(last_token)] * n_tokens
        if strategy == "greedy":
            confirmed = self.get_lookahead_confirmed_pos()
            scratch = GPUArray((1,), int32)
            guess, cur = [], int(last_token)
            for i in range(n_tokens):
                cur = self._argmax_tokens(self._decode_step_fixed_cache(cur, confirmed + i, confirmed + i + 1), scratch)[0]
                guess.append(cur)
            self.reset_lookahead_all()
            return guess
        raise ValueError(f"Unknown init strategy: {strategy}")

    def decode_step_jacobi_lookahead(self, token_id: int, n_tokens: int = 8, max_iter: int = 3,
                                     init_strategy: str = "repeat") -> tuple[list[int], dict]:
        """One Jacobi step on the fixed cache (causal.py:1357-1476): guess n_tokens future tokens, run the batched
        forward pass on [token, guess[:-1]], replace the guess by the row argmaxes, repeat up to max_iter times or until
        nothing changes.  Output token i is the model's greedy token whenever the inputs before it were right, so the
        accepted tokens are the prefix on which the last two iterations agree PLUS the first token after it (computed
        from inputs that are all confirmed; the reference stops one token earlier).  No re-run after a partial accept:
        row confirmed + i of every layer's cache already holds the K/V of input i, which is an accepted token."""
        if self.tp != 1:
            raise NotImplementedError("Jacobi decode runs on one GPU (row argmaxes over the full vocabulary)")
        b = self.buffers
        n = int(n_tokens)
        if not 1 <= n <= b.max_batch_size:
            raise ValueError(f"n_tokens must be in [1, max_batch_size = {b.max_batch_size}]")
        confirmed = self.get_lookahead_confirmed_pos()
        if confirmed + n > self._max_seq:
            raise ValueError(f"decoding {n} tokens from position {confirmed} exceeds the fixed cache ({self._max_seq})")
        guess = self._init_jacobi_guess_lookahead(token_id, n, init_strategy)
        scratch = GPUArray((n,), int32)
        prev, converged, iterations = None, False, 0
        for _ in range(max(1, int(max_iter))):
            iterations += 1
            self.reset_lookahead_all()
            hidden = self._decode_step_fixed_cache_batch([int(token_id)] + guess[:-1], confirmed, confirmed + n)
            new = self._argmax_tokens(hidden, scratch)
            prev, guess = guess, new
            if new == prev:
                converged = True
                break
        accepted = accept_jacobi(prev, guess)      # guess[i] is right if prev[0 .. i-1] were: keep it, stop after a change
        self.commit_lookahead_all(len(accepted))
        if not hasattr(self, "_ngram_cache"):
            self._ngram_cache = {}
        self._ngram_cache[int(token_id)] = list(accepted)
        return accepted, {"iterations": iterations, "converged": converged, "accepted_count": len(accepted)}

    # ------------------------------------------------------------------ accounting
    def weight_bytes_per_token(self) -> int:
        """Algorithmic weight bytes one decode token streams on this rank."""
        return sum(blk.weight_bytes for blk in self.blocks) + self._head.nbytes + self.final_norm.weight.nbytes

    def kv_bytes_per_token(self, context_len: int) -> int:
        a = self.blocks[0].attn
        return len(self.blocks) * 2 * a.num_kv_heads * context_len * a.head_dim * 2


GPT2Model = CausalTransformerModel
LlamaModel = CausalTransformerModel
QwenModel = CausalTransformerModel   # named by BASELINE.json's north_star; absent upstream

__all__ = ["CausalTransformerModel", "GPT2Model", "LlamaModel", "QwenModel"]
_ = from_numpy
