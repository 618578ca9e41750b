"""Attention layer of the decode path (mirror of pygpukit/llm/layers/attention.py:43-461).

Same constructor, attributes (`q_proj/k_proj/v_proj/o_proj`, `q_norm/k_norm`, `_k_cache/_v_cache`,
`_max_cache_len`, lookahead bookkeeping) and methods (`init_fixed_cache`, `forward_fixed_cache`,
`forward_fixed_cache_batch`) as the reference.  What changes underneath:
  * q/k/v weights are concatenated row-wise into ONE projection for every weight format (rows are
    independent, so results equal the three separate GEMVs the reference runs for FP8,
    attention.py:315-319);
  * QK-norm + RoPE + KV-cache write + attention are ONE kernel for a single new token
    (`pgk_rope_sdpa_decode`; `pgk_qk_norm_rope_kv_write` + `pgk_sdpa_decode` for M > 1), reading f32
    cos/sin tables that were uploaded ONCE (the reference slices them on the host and does an H2D
    copy per layer per step, attention.py:338-347);
  * the KV cache is stored UN-expanded [num_kv_heads, max_seq, head_dim]; `sdpa_decode` serves the
    G query heads of a group from one read of K/V (the reference stores and reads the GQA-expanded
    [num_heads, max_seq, head_dim] copy, attention.py:135);
  * position / context length may live in int32 device buffers (`pos_buf`, `ctx_buf`) so one CUDA
    graph replays for every step.
"""
from __future__ import annotations

import os
from typing import TYPE_CHECKING

import numpy as np

from ..._lib import api
from ...core.array import GPUArray
from ...core.dtypes import bfloat16, float32, uint8
from ...core.factory import from_numpy, zeros
from ...core.stream import current_stream_handle as _st
from .linear import LinearBF16, LinearFP8, LinearNVF4
from .norm import Norm
from .rope import precompute_freqs_cis

if TYPE_CHECKING:
    from ..config import TransformerConfig

_LINEARS = (LinearBF16, LinearFP8, LinearNVF4)
# PGK_FUSED_BRACKET=0 keeps the two-launch bracket (qk_norm_rope_kv_write + sdpa_decode) for A/B timing
_FUSED_BRACKET = os.environ.get("PGK_FUSED_BRACKET", "1") != "0"


def _concat_rows(arrs: list[GPUArray]) -> GPUArray:
    """Row-wise concatenation of 2-D device arrays (init-time; device-to-device copies)."""
    cols, dt = arrs[0].shape[1], arrs[0].dtype
    out = GPUArray((sum(a.shape[0] for a in arrs), cols), dt)
    off = 0
    for a in arrs:
        api.memcpy_d2d(out.device_ptr + off, a.device_ptr, a.nbytes, _st())
        off += a.nbytes
    return out


def fuse_qkv(q, k, v):
    """One projection whose output row is [q | k | v]; None when the formats cannot be fused."""
    if all(isinstance(p, LinearBF16) for p in (q, k, v)):
        bias = None
        if any(p.bias is not None for p in (q, k, v)):
            parts = [p.bias.to_numpy() if p.bias is not None else np.zeros(p.out_features, np.uint16) for p in (q, k, v)]
            bias = from_numpy(np.concatenate(parts))
        return LinearBF16(_concat_rows([q.weight, k.weight, v.weight]), bias)
    if all(isinstance(p, LinearFP8) for p in (q, k, v)) and all(p.out_features % 128 == 0 for p in (q, k)) \
            and all(p.bias is None for p in (q, k, v)):
        return LinearFP8(_concat_rows([q.weight_fp8, k.weight_fp8, v.weight_fp8]),
                         _concat_rows([q.scale_inv, k.scale_inv, v.scale_inv]))
    if all(isinstance(p, LinearNVF4) for p in (q, k, v)) and all(p.bias is None for p in (q, k, v)):
        K = q.in_features
        d = np.concatenate([p.data.to_numpy().reshape(K // 2, p.out_features) for p in (q, k, v)], axis=1)
        s = np.concatenate([p.scale.to_numpy().reshape(-1, p.out_features) for p in (q, k, v)], axis=1)
        return LinearNVF4(from_numpy(np.ascontiguousarray(d)), from_numpy(np.ascontiguousarray(s)), K, d.shape[1])
    return None


class Attention:
    def __init__(self, q_proj, k_proj, v_proj, o_proj, config: "TransformerConfig",
                 q_bias: GPUArray | None = None, k_bias: GPUArray | None = None, v_bias: GPUArray | None = None,
                 o_bias: GPUArray | None = None, q_norm: Norm | None = None, k_norm: Norm | None = None,
                 qkv_proj=None):
        def wrap(p, b):
            return p if (p is None or isinstance(p, _LINEARS)) else LinearBF16(p, b)
        self.q_proj, self.k_proj, self.v_proj = wrap(q_proj, q_bias), wrap(k_proj, k_bias), wrap(v_proj, v_bias)
        self.o_proj = wrap(o_proj, o_bias)
        self.q_norm, self.k_norm = q_norm, k_norm
        self.config = config
        self.head_dim = config.head_dim
        # local (per tensor-parallel rank) head counts
        tp = max(1, config.tp_size)
        assert config.num_kv_heads is not None
        # tp > num_kv_heads: tp / num_kv_heads ranks share (replicate) one KV head, query heads still split
        if config.num_heads % tp or (config.num_kv_heads % tp and tp % config.num_kv_heads):
            raise ValueError(f"num_heads {config.num_heads} / num_kv_heads {config.num_kv_heads} not shardable over tp_size {tp}")
        self.num_heads = config.num_heads // tp
        self.num_kv_heads = max(1, config.num_kv_heads // tp)
        self.num_kv_groups = self.num_heads // self.num_kv_heads
        self.q_dim = self.num_heads * self.head_dim
        self.k_dim = self.num_kv_heads * self.head_dim
        self.v_dim = self.k_dim
        self.qkv_proj = qkv_proj if qkv_proj is not None else fuse_qkv(self.q_proj, self.k_proj, self.v_proj)
        if self.qkv_proj is None:
            raise ValueError("Attention: q/k/v projections must share one weight format")

        self._cos: np.ndarray | None = None
        self._sin: np.ndarray | None = None
        self._cos_dev: GPUArray | None = None
        self._sin_dev: GPUArray | None = None
        if config.use_rope:
            self._cos, self._sin = precompute_freqs_cis(self.head_dim, config.max_position_embeddings, config.rope_theta)
        self._k_cache: GPUArray | None = None
        self._v_cache: GPUArray | None = None
        self._max_cache_len = 0
        self._confirmed_pos = 0
        self._logical_pos = 0

    # ---- fixed KV cache
    def init_fixed_cache(self, max_seq_len: int, dtype: str = "bfloat16", rope_tables=None, num_sequences: int = 1) -> None:
        """Allocate the fixed-length KV cache (attention.py:128-146) -- un-expanded layout.  num_sequences > 1:
        one region per independent sequence, [S * n_kv, max_seq, hd] (multi-sequence decode, net-new on this path;
        the reference's paged / continuous-batching kernels are not wired to its Python layers)."""
        if dtype != "bfloat16":
            raise ValueError("the B200 decode path keeps the KV cache in bfloat16")
        shape = (num_sequences * self.num_kv_heads, max_seq_len, self.head_dim)
        self._k_cache, self._v_cache = zeros(shape, bfloat16), zeros(shape, bfloat16)
        self._num_sequences, self._cur_seq = num_sequences, 0
        self._max_cache_len = max_seq_len
        self._confirmed_pos = self._logical_pos = 0
        if self.config.use_rope:
            if rope_tables is not None:           # shared across layers (same head_dim / theta)
                self._cos_dev, self._sin_dev = rope_tables
            else:
                # the kernels index the tables by position and validate position < max_seq_len only, so the tables
                # always cover the whole cache (a cache longer than max_position_embeddings extends them with the
                # same float32 recipe, llm/layers/rope.py:13-24)
                if max_seq_len > self._cos.shape[0]:
                    self._cos, self._sin = precompute_freqs_cis(self.head_dim, max_seq_len, self.config.rope_theta)
                self._cos_dev = from_numpy(np.ascontiguousarray(self._cos[:max_seq_len], np.float32))
                self._sin_dev = from_numpy(np.ascontiguousarray(self._sin[:max_seq_len], np.float32))

    @property
    def _seq_stride(self) -> int:
        """elements between the cache regions of consecutive sequences"""
        return self.num_kv_heads * self._max_cache_len * self.head_dim

    def select_sequence(self, s: int) -> None:
        """Single-sequence entry points (prefill, forward_fixed_cache) work on sequence `s`'s cache region."""
        if not 0 <= s < getattr(self, "_num_sequences", 1):
            raise ValueError(f"sequence {s} out of range")
        self._cur_seq = s

    def _kv_ptrs(self) -> tuple[int, int]:
        off = getattr(self, "_cur_seq", 0) * self._seq_stride * 2
        return self._k_cache.device_ptr + off, self._v_cache.device_ptr + off

    def attend_multi(self, x: GPUArray, pos_buf: GPUArray, ctx_buf: GPUArray, *, qkv_buf: GPUArray | None = None,
                     attn_buf: GPUArray | None = None, norm=None) -> GPUArray:
        """Row s of x [S, hidden] is the new token of independent sequence s (its own cache region, position
        pos_buf[s], context ctx_buf[s]); returns [S, q_dim].  One qkv projection for all rows, ONE attention launch."""
        S = x.shape[0]
        if S != getattr(self, "_num_sequences", 1):
            raise ValueError(f"{S} rows for {getattr(self, '_num_sequences', 1)} sequences (init_decode num_sequences)")
        if self.head_dim != 128 or not self.config.use_rope:
            raise ValueError("multi-sequence decode needs head_dim 128 and RoPE")
        H, Hkv, hd = self.num_heads, self.num_kv_heads, self.head_dim
        qkv = self.qkv_proj.forward(x, out=qkv_buf, norm=norm)
        if attn_buf is None:
            attn_buf = GPUArray((S, H * hd), bfloat16)
        api.rope_sdpa_decode_batch(qkv.device_ptr, None if self.q_norm is None else self.q_norm.weight.device_ptr,
                                   None if self.k_norm is None else self.k_norm.weight.device_ptr,
                                   float(self.q_norm.eps if self.q_norm is not None else 0.0),
                                   self._cos_dev.device_ptr, self._sin_dev.device_ptr, self._k_cache.device_ptr,
                                   self._v_cache.device_ptr, attn_buf.device_ptr, H, Hkv, hd, self._max_cache_len, S,
                                   pos_buf.device_ptr, ctx_buf.device_ptr, self._seq_stride, 0.0, _st())
        return attn_buf

    def set_confirmed_pos(self, pos: int) -> None:
        assert 0 <= pos <= self._max_cache_len, f"Invalid pos {pos}"
        self._confirmed_pos = self._logical_pos = pos

    def reset_lookahead(self) -> None:
        self._logical_pos = self._confirmed_pos

    def commit_lookahead(self, n_accepted: int) -> None:
        new_pos = self._confirmed_pos + n_accepted
        assert new_pos <= self._max_cache_len, f"Commit exceeds cache: {new_pos}"
        self._confirmed_pos = self._logical_pos = new_pos

    def get_confirmed_pos(self) -> int:
        return self._confirmed_pos

    # ---- decode
    def attend(self, x: GPUArray, position: int, context_len: int, *, qkv_buf: GPUArray | None = None,
               attn_buf: GPUArray | None = None, pos_buf: GPUArray | None = None,
               ctx_buf: GPUArray | None = None, norm=None) -> GPUArray:
        """qkv projection -> QK-norm/RoPE/KV-write -> split-KV attention; returns [M, q_dim]
        (the input of o_proj).  M rows = M consecutive positions of one sequence."""
        assert self._k_cache is not None, "Call init_fixed_cache first"
        M = x.shape[0]
        H, Hkv, hd = self.num_heads, self.num_kv_heads, self.head_dim
        width = (H + 2 * Hkv) * hd
        if qkv_buf is not None and qkv_buf.shape[0] != M:
            qkv_buf = qkv_buf.slice_rows(M)
        qkv = self.qkv_proj.forward(x, out=qkv_buf, norm=norm)   # norm: pre-attention RMSNorm fused into the launch
        if attn_buf is None:
            attn_buf = GPUArray((M, H * hd), bfloat16)
        elif attn_buf.shape[0] != M:
            attn_buf = attn_buf.slice_rows(M)
        if M == 1 and hd == 128 and self.config.use_rope and _FUSED_BRACKET:
            # one launch: QK-norm + RoPE + KV write + split-KV attention (pgk_rope_sdpa_decode)
            kp, vp = self._kv_ptrs()
            api.rope_sdpa_decode(qkv.device_ptr, None if self.q_norm is None else self.q_norm.weight.device_ptr,
                                 None if self.k_norm is None else self.k_norm.weight.device_ptr,
                                 float(self.q_norm.eps if self.q_norm is not None else 0.0),
                                 self._cos_dev.device_ptr, self._sin_dev.device_ptr, kp,
                                 vp, attn_buf.device_ptr, H, Hkv, hd, self._max_cache_len,
                                 int(position), None if pos_buf is None else pos_buf.device_ptr, int(context_len),
                                 None if ctx_buf is None else ctx_buf.device_ptr, 0.0, _st())
            return attn_buf
        # QK-norm / RoPE / K,V cache write in one launch; without RoPE (GPT-2) the tables are NULL and the launch
        # only writes the K/V rows into the cache
        rope = self.config.use_rope
        kp, vp = self._kv_ptrs()
        api.qk_norm_rope_kv_write(qkv.device_ptr, None if self.q_norm is None else self.q_norm.weight.device_ptr,
                                  None if self.k_norm is None else self.k_norm.weight.device_ptr,
                                  float(self.q_norm.eps if self.q_norm is not None else 0.0),
                                  self._cos_dev.device_ptr if rope else None, self._sin_dev.device_ptr if rope else None,
                                  kp, vp, M, H, Hkv, hd, self._max_cache_len,
                                  int(position), None if pos_buf is None else pos_buf.device_ptr, _st())
        # Q read in place from the qkv rows (head stride hd, token stride `width`); out [M, H*hd]
        api.sdpa_decode(qkv.device_ptr, kp, vp, attn_buf.device_ptr, H, Hkv,
                        M, hd, self._max_cache_len, int(context_len), None if ctx_buf is None else ctx_buf.device_ptr,
                        0.0, hd, width, hd, H * hd, 2, _st())
        return attn_buf

    def forward_fixed_cache(self, x: GPUArray, position: int, context_len: int, *, out: GPUArray | None = None,
                            residual: GPUArray | None = None, **bufs) -> GPUArray:
        """x [1, hidden] -> [1, hidden] (attention.py:279-370)."""
        assert x.shape[0] == 1, "forward_fixed_cache expects single token"
        a = self.attend(x, position, context_len, **bufs)
        return self.o_proj.forward(a, out=out, residual=residual)

    def forward_fixed_cache_batch(self, x: GPUArray, start_position: int, context_len: int, *,
                                  out: GPUArray | None = None, residual: GPUArray | None = None, **bufs) -> GPUArray:
        """x [M, hidden]: M consecutive tokens of one sequence, causal inside the batch
        (attention.py:372-461)."""
        a = self.attend(x, start_position, context_len, **bufs)
        return self.o_proj.forward(a, out=out, residual=residual)

    def __call__(self, x: GPUArray, position_ids=None, past_kv=None, use_cache: bool = False):
        """Prefill/generic entry (attention.py:173-277) expressed on the fixed cache: positions must be
        consecutive; `past_kv` is the number of cached positions."""
        start = position_ids[0] if position_ids else 0
        out = self.forward_fixed_cache_batch(x, start, start + x.shape[0])
        return out, (start + x.shape[0] if use_cache else None)

    @property
    def weight_bytes(self) -> int:
        return self.qkv_proj.weight_bytes + self.o_proj.weight_bytes


__all__ = ["Attention", "fuse_qkv"]
_ = (float32, uint8)
