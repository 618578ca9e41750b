"""ctypes wrapper of the CPU oracle (oracle.c) -- TEST INFRASTRUCTURE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  NumPy in, NumPy out; bf16 tensors are uint16 arrays.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "liboracle.so")
_lib = None


def build() -> str:
    src = os.path.join(_HERE, "oracle.c")
    if not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.run(["make", "-C", _HERE], check=True, capture_output=True)
    return _LIB_PATH


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        # libgomp's default spin-wait burns the cgroup CPU quota between parallel regions (a 1-row
        # rmsnorm took 55 ms instead of 0.04 ms in the build container); read at libgomp start-up
        os.environ.setdefault("OMP_WAIT_POLICY", "PASSIVE")
        _lib = C.CDLL(build())
        _lib.orc_num_threads.restype = C.c_int
        _lib.orc_argmax_bf16.restype = C.c_int
        _lib.orc_argmax_f32.restype = C.c_int
    return _lib


def _p(a: np.ndarray):
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.c_void_p)


def _c(a, dt):
    return np.ascontiguousarray(a, dtype=dt)


def num_threads() -> int:
    return int(lib().orc_num_threads())


def set_num_threads(n: int) -> int:
    """OpenMP team size for every later call (bench.py: all host cores, whatever OMP_NUM_THREADS says)."""
    lib().orc_set_num_threads(int(n))
    return num_threads()


def fp8_table() -> np.ndarray:
    t = np.empty(256, np.float32)
    lib().orc_fp8_e4m3_table(_p(t))
    return t


def ue4m3_table() -> np.ndarray:
    t = np.empty(256, np.float32)
    lib().orc_ue4m3_table(_p(t))
    return t


def e2m1_table() -> np.ndarray:
    t = np.empty(16, np.float32)
    lib().orc_e2m1_table(_p(t))
    return t


def gemv_fp8_bf16(A, B, S) -> np.ndarray:
    """A [M,K] or [K] bf16 bits, B [N,K] u8, S [N/128,K/128] bf16 bits -> C [M,N] / [N]."""
    A2 = _c(np.atleast_2d(A), np.uint16)
    B, S = _c(B, np.uint8), _c(S, np.uint16)
    M, K = A2.shape
    N = B.shape[0]
    Cc = np.empty((M, N), np.uint16)
    lib().orc_gemv_fp8_bf16(_p(A2), _p(B), _p(S), _p(Cc), M, K, N)
    return Cc if np.ndim(A) == 2 else Cc[0]


def dequant_fp8_block(B, S) -> np.ndarray:
    B, S = _c(B, np.uint8), _c(S, np.uint16)
    N, K = B.shape
    W = np.empty((N, K), np.float32)
    lib().orc_dequant_fp8_block(_p(B), _p(S), _p(W), N, K)
    return W


def w8a16_gemm(A, B_kn, S) -> np.ndarray:
    A, B_kn, S = _c(A, np.uint16), _c(B_kn, np.uint8), _c(S, np.uint16)
    M, K = A.shape
    N = B_kn.shape[1]
    Cc = np.empty((M, N), np.uint16)
    lib().orc_w8a16_gemm(_p(A), _p(B_kn), _p(S), _p(Cc), M, N, K)
    return Cc


def grouped_gemm_fp8_bf16(A, B_enk, S, ids) -> np.ndarray:
    A, B_enk, S, ids = _c(A, np.uint16), _c(B_enk, np.uint8), _c(S, np.uint16), _c(ids, np.int32)
    M, K = A.shape
    N = B_enk.shape[1]
    Cc = np.empty((M, N), np.uint16)
    lib().orc_grouped_gemm_fp8_bf16(_p(A), _p(B_enk), _p(S), _p(Cc), _p(ids), M, N, K)
    return Cc


def gemv_bf16(A, B) -> np.ndarray:
    A2 = _c(np.atleast_2d(A), np.uint16)
    B = _c(B, np.uint16)
    M, K = A2.shape
    N = B.shape[0]
    Cc = np.empty((M, N), np.uint16)
    lib().orc_gemv_bf16(_p(A2), _p(B), _p(Cc), M, K, N)
    return Cc if np.ndim(A) == 2 else Cc[0]


def gemv_nvf4_bf16(A, Bd, Bs, alpha: float = 1.0) -> np.ndarray:
    A2 = _c(np.atleast_2d(A), np.uint16)
    Bd, Bs = _c(Bd, np.uint8), _c(Bs, np.uint8)
    M, K = A2.shape
    N = Bd.shape[1]
    Cc = np.empty((M, N), np.uint16)
    for m in range(M):
        lib().orc_gemv_nvf4_bf16(_p(A2[m]), _p(Bd), _p(Bs), _p(Cc[m]), K, N, C.c_float(alpha))
    return Cc if np.ndim(A) == 2 else Cc[0]


def dequant_nvf4(Bd, Bs, K: int) -> np.ndarray:
    Bd, Bs = _c(Bd, np.uint8), _c(Bs, np.uint8)
    N = Bd.shape[1]
    W = np.empty((K, N), np.float32)
    lib().orc_dequant_nvf4(_p(Bd), _p(Bs), _p(W), K, N)
    return W


def quantize_bf16_to_nvf4(W_kn):
    W_kn = _c(W_kn, np.uint16)
    K, N = W_kn.shape
    data = np.empty((K // 2, N), np.uint8)
    scale = np.empty(((K + 31) // 32, N), np.uint8)
    lib().orc_quantize_bf16_to_nvf4(_p(W_kn), _p(data), _p(scale), K, N, 0)
    return data, scale


def rmsnorm_bf16(x, gamma, eps: float) -> np.ndarray:
    x, gamma = _c(x, np.uint16), _c(gamma, np.uint16)
    y = np.empty_like(x)
    lib().orc_rmsnorm_bf16(_p(x), _p(gamma), _p(y), x.shape[0], x.shape[1], C.c_float(eps))
    return y


def rmsnorm_f32(x, gamma, eps: float) -> np.ndarray:
    x, gamma = _c(x, np.float32), _c(gamma, np.float32)
    y = np.empty_like(x)
    lib().orc_rmsnorm_f32(_p(x), _p(gamma), _p(y), x.shape[0], x.shape[1], C.c_float(eps))
    return y


def layernorm_f32(x, gamma, beta, eps: float) -> np.ndarray:
    x, gamma, beta = _c(x, np.float32), _c(gamma, np.float32), _c(beta, np.float32)
    y = np.empty_like(x)
    lib().orc_layernorm_f32(_p(x), _p(gamma), _p(beta), _p(y), x.shape[0], x.shape[1], C.c_float(eps))
    return y


def rope_bf16_f32table(q, k, cos, sin):
    q, k = _c(q, np.uint16).copy(), _c(k, np.uint16).copy()
    cos, sin = _c(cos, np.float32), _c(sin, np.float32)
    lib().orc_rope_bf16_f32table(_p(q), _p(k), _p(cos), _p(sin), q.shape[0], q.shape[1], k.shape[1], q.shape[2])
    return q, k


def rope_f32(q, k, cos, sin):
    q, k = _c(q, np.float32).copy(), _c(k, np.float32).copy()
    cos, sin = _c(cos, np.float32), _c(sin, np.float32)
    lib().orc_rope_f32(_p(q), _p(k), _p(cos), _p(sin), q.shape[0], q.shape[1], k.shape[1], q.shape[2])
    return q, k


def kv_cache_update_gqa(new_kv, cache, num_heads: int, position: int) -> None:
    """new_kv [seq, Hkv, hd] -> cache [H, max_seq, hd] in place (seq rows from `position`)."""
    new_kv = np.ascontiguousarray(new_kv)
    seq, hkv, hd = new_kv.shape
    lib().orc_kv_cache_prefill_gqa(_p(new_kv), _p(cache), num_heads, hkv, hd, cache.shape[1], position, seq,
                                   new_kv.itemsize)


_KIND = {np.dtype(np.float32): 0, np.dtype(np.float16): 1, np.dtype(np.uint16): 2}


def sdpa_causal_fixed_cache(Q, K, V, context_len: int, scale: float = 0.0) -> np.ndarray:
    """Q [H,q_len,hd], K/V [H,max_seq,hd] (same dtype; uint16 = bf16) -> out [H,q_len,hd]."""
    Q, K, V = np.ascontiguousarray(Q), np.ascontiguousarray(K), np.ascontiguousarray(V)
    out = np.empty_like(Q)
    H, ql, hd = Q.shape
    lib().orc_sdpa_causal_fixed_cache(_p(Q), _p(K), _p(V), _p(out), H, ql, context_len, K.shape[1], hd,
                                      C.c_float(scale), _KIND[Q.dtype])
    return out


def flash_decoding(Q, K, V, context_len: int, scale: float = 0.0) -> np.ndarray:
    Q, K, V = np.ascontiguousarray(Q), np.ascontiguousarray(K), np.ascontiguousarray(V)
    out = np.empty_like(Q)
    H, hd = Q.shape[0], Q.shape[-1]
    lib().orc_flash_decoding(_p(Q), _p(K), _p(V), _p(out), H, context_len, K.shape[1], hd, C.c_float(scale),
                             _KIND[Q.dtype])
    return out


def silu_bf16(x) -> np.ndarray:
    x = _c(x, np.uint16)
    y = np.empty_like(x)
    lib().orc_silu_bf16(_p(x), _p(y), C.c_size_t(x.size))
    return y


def gelu_f32(x) -> np.ndarray:
    x = _c(x, np.float32)
    y = np.empty_like(x)
    lib().orc_gelu_f32(_p(x), _p(y), C.c_size_t(x.size))
    return y


def add_inplace_bf16(a, b) -> np.ndarray:
    a, b = _c(a, np.uint16).copy(), _c(b, np.uint16)
    lib().orc_add_inplace_bf16(_p(a), _p(b), C.c_size_t(a.size))
    return a


def mul_inplace_bf16(a, b) -> np.ndarray:
    a, b = _c(a, np.uint16).copy(), _c(b, np.uint16)
    lib().orc_mul_inplace_bf16(_p(a), _p(b), C.c_size_t(a.size))
    return a


def add_bf16(a, b) -> np.ndarray:
    a, b = _c(a, np.uint16), _c(b, np.uint16)
    c = np.empty_like(a)
    lib().orc_add_bf16(_p(a), _p(b), _p(c), C.c_size_t(a.size))
    return c


def bias_add_bf16(x, bias) -> np.ndarray:
    x, bias = _c(x, np.uint16).copy(), _c(bias, np.uint16)
    lib().orc_bias_add_inplace_bf16(_p(x), _p(bias), x.shape[0], x.shape[1])
    return x


def argmax_bf16(x) -> int:
    x = _c(x, np.uint16).ravel()
    return int(lib().orc_argmax_bf16(_p(x), x.size))


def argmax_f32(x) -> int:
    x = _c(x, np.float32).ravel()
    return int(lib().orc_argmax_f32(_p(x), x.size))


def moe_topk_bf16(logits, k: int):
    logits = _c(logits, np.uint16)
    T, E = logits.shape
    vals = np.empty((T, k), np.uint16)
    idx = np.empty((T, k), np.int32)
    lib().orc_moe_topk_bf16(_p(logits), _p(vals), _p(idx), T, E, k)
    return vals, idx


def moe_softmax_topk_bf16(values) -> np.ndarray:
    values = _c(values, np.uint16).copy()
    lib().orc_moe_softmax_topk_bf16(_p(values), values.shape[0], values.shape[1])
    return values


def moe_compute_permutation(expert_indices, num_experts: int):
    ei = _c(expert_indices, np.int32)
    T, k = ei.shape
    counts = np.empty(num_experts, np.int32)
    offsets = np.empty(num_experts + 1, np.int32)
    perm = np.empty(T * k, np.int32)
    rev = np.empty(T * k, np.int32)
    lib().orc_moe_compute_permutation(_p(ei), _p(counts), _p(offsets), _p(perm), _p(rev), T, num_experts, k)
    return counts, offsets, perm, rev


def moe_expand_expert_offsets(offsets, m_total: int) -> np.ndarray:
    offsets = _c(offsets, np.int32)
    ids = np.empty(m_total, np.int32)
    lib().orc_moe_expand_expert_offsets(_p(offsets), _p(ids), offsets.size - 1, m_total)
    return ids


def moe_gather_bf16(hidden, perm, k: int) -> np.ndarray:
    hidden, perm = _c(hidden, np.uint16), _c(perm, np.int32)
    T, H = hidden.shape
    out = np.empty((T * k, H), np.uint16)
    lib().orc_moe_gather_bf16(_p(hidden), _p(perm), _p(out), T, H, k)
    return out


def moe_scatter_bf16(expert_out, weights, rev, T: int, k: int) -> np.ndarray:
    expert_out, weights, rev = _c(expert_out, np.uint16), _c(weights, np.uint16), _c(rev, np.int32)
    H = expert_out.shape[1]
    out = np.empty((T, H), np.uint16)
    lib().orc_moe_scatter_bf16(_p(expert_out), _p(weights), _p(rev), _p(out), T, H, k)
    return out
