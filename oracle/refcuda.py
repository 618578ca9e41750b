"""ctypes wrapper of oracle/_ref/libref_cuda.so -- the REFERENCE's own CUDA kernels compiled in
place by oracle/build_ref.sh (TEST INFRASTRUCTURE; used to pin the CPU oracle and as the
"reference kernel on the same B200" number).  Operates on pygpukit_b200 GPUArrays (device
pointers); every call synchronises."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
PATH = os.path.join(_HERE, "_ref", "libref_cuda.so")
_lib = None


def available() -> bool:
    return os.path.exists(PATH)


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        _lib = C.CDLL(PATH)
        _lib.ref_error_string.restype = C.c_char_p
        _lib.ref_flash_decoding_workspace.restype = C.c_size_t
    return _lib


def _chk(rc: int, what: str) -> None:
    if rc != 0:
        raise RuntimeError(f"{what} failed: {lib().ref_error_string(rc).decode()}")


def _p(a):
    return C.c_void_p(a.device_ptr)


def set_sync(on: bool) -> None:
    """False: launches stay asynchronous on the default stream (event timing); True: every call synchronises."""
    lib().ref_set_sync(1 if on else 0)


def gemv_fp8(A, B, S, Cout, K, N):
    _chk(lib().ref_gemv_fp8(_p(A), _p(B), _p(S), _p(Cout), K, N), "ref_gemv_fp8")


def gemv_fp8_batched(A, B, S, Cout, K, N, M):
    _chk(lib().ref_gemv_fp8_batched(_p(A), _p(B), _p(S), _p(Cout), K, N, M), "ref_gemv_fp8_batched")


def gemv_bf16(A, B, Cout, K, N):
    _chk(lib().ref_gemv_bf16(_p(A), _p(B), _p(Cout), K, N), "ref_gemv_bf16")


def gemv_nvf4(A, Bd, Bs, Cout, K, N, alpha=1.0):
    _chk(lib().ref_gemv_nvf4(_p(A), _p(Bd), _p(Bs), _p(Cout), K, N, C.c_float(alpha)), "ref_gemv_nvf4")


def quantize_nvf4(inp, data, scale, K, N):
    _chk(lib().ref_quantize_nvf4(_p(inp), _p(data), _p(scale), K, N), "ref_quantize_nvf4")


def w8a16_gemm(A, B_kn, S, Cout, M, N, K):
    _chk(lib().ref_w8a16_gemm(_p(A), _p(B_kn), _p(S), _p(Cout), M, N, K), "ref_w8a16_gemm")


def grouped_gemm(A, B, S, Cout, ids, M, N, K):
    _chk(lib().ref_grouped_gemm(_p(A), _p(B), _p(S), _p(Cout), _p(ids), M, N, K), "ref_grouped_gemm")


def rmsnorm_bf16(x, g, y, rows, features, eps):
    _chk(lib().ref_rmsnorm_bf16(_p(x), _p(g), _p(y), rows, features, C.c_float(eps)), "ref_rmsnorm_bf16")


def rope_bf16_f32table(q, k, cos, sin, seq, hq, hk, hd):
    _chk(lib().ref_rope_bf16_f32table(_p(q), _p(k), _p(cos), _p(sin), seq, hq, hk, hd), "ref_rope")


def kv_update_gqa_bf16(nkv, cache, H, Hkv, hd, max_seq, pos):
    _chk(lib().ref_kv_update_gqa_bf16(_p(nkv), _p(cache), H, Hkv, hd, max_seq, pos), "ref_kv_update")


def sdpa(Q, K, V, out, H, q_len, kv_len, kv_stride, hd, scale=0.0, f16=False):
    fn = lib().ref_sdpa_f16 if f16 else lib().ref_sdpa_bf16
    _chk(fn(_p(Q), _p(K), _p(V), _p(out), H, q_len, kv_len, kv_stride, hd, C.c_float(scale)), "ref_sdpa")


def flash_decoding_workspace(H, hd, kv_len) -> int:
    return int(lib().ref_flash_decoding_workspace(H, hd, kv_len))


def flash_decoding_f16(Q, K, V, out, ws, H, hd, kv_len, kv_stride):
    _chk(lib().ref_flash_decoding_f16(_p(Q), _p(K), _p(V), _p(out), _p(ws), H, hd, kv_len, kv_stride), "ref_flash_decoding")


def silu_bf16(x, y, n):
    _chk(lib().ref_silu_bf16(_p(x), _p(y), C.c_size_t(n)), "ref_silu")


def add_inplace_bf16(a, b, n):
    _chk(lib().ref_add_inplace_bf16(_p(a), _p(b), C.c_size_t(n)), "ref_add_inplace")


def mul_inplace_bf16(a, b, n):
    _chk(lib().ref_mul_inplace_bf16(_p(a), _p(b), C.c_size_t(n)), "ref_mul_inplace")


def moe_topk_bf16(logits, values, idx, T, E, k):
    _chk(lib().ref_moe_topk_bf16(_p(logits), _p(values), _p(idx), T, E, k), "ref_moe_topk")


def moe_softmax_topk_bf16(values, T, k):
    _chk(lib().ref_moe_softmax_topk_bf16(_p(values), T, k), "ref_moe_softmax_topk")
