"""Matmul family of the decode path (mirror of pygpukit/ops/matmul/*.py).

Same names, positional/keyword signatures (`*, out=None`), ValueError on bad shapes/dtypes
and return conventions (returns `out` when given) as the reference; every function dispatches
into libpgk_b200.so.  There is no CPU branch and no Triton/cuBLASLt dispatch on this path.
"""
from __future__ import annotations

from .._lib import api
from ..core.array import GPUArray
from ..core.dtypes import bfloat16, float16, float32, int32, uint8
from ._common import _out, _st


# ------------------------------------------------------------------ BF16 GEMV
def gemv_bf16(a: GPUArray, b: GPUArray, *, out: GPUArray | None = None) -> GPUArray:
    """BF16 GEMV: C[N] = A[K] @ B[N,K]^T (ops/matmul/gemv.py:15-70)."""
    if a.ndim != 1:
        raise ValueError(f"gemv_bf16 requires 1D input vector, got {a.ndim}D")
    if b.ndim != 2:
        raise ValueError(f"gemv_bf16 requires 2D weight matrix, got {b.ndim}D")
    if a.dtype != bfloat16 or b.dtype != bfloat16:
        raise ValueError("gemv_bf16 requires bfloat16 inputs")
    K, N = a.shape[0], b.shape[0]
    if b.shape[1] != K:
        raise ValueError(f"gemv_bf16 dimension mismatch: A[{K}] vs B[{N}, {b.shape[1]}]")
    out = _out(out, (N,), bfloat16)
    api.gemv_bf16(a.device_ptr, b.device_ptr, out.device_ptr, 1, K, N, _st())
    return out


gemv_bf16_bf16_sm120 = gemv_bf16


def gemv_bf16_batched(a: GPUArray, b: GPUArray, *, out: GPUArray | None = None) -> GPUArray:
    """C[M,N] = A[M,K] @ B[N,K]^T (weights read once for all rows)."""
    if a.ndim != 2 or b.ndim != 2:
        raise ValueError("gemv_bf16_batched requires 2D inputs")
    if a.dtype != bfloat16 or b.dtype != bfloat16:
        raise ValueError("gemv_bf16_batched requires bfloat16 inputs")
    M, K = a.shape
    N = b.shape[0]
    if b.shape[1] != K:
        raise ValueError(f"gemv_bf16_batched dimension mismatch: A[{M},{K}] vs B[{N},{b.shape[1]}]")
    out = _out(out, (M, N), bfloat16)
    api.gemv_bf16(a.device_ptr, b.device_ptr, out.device_ptr, M, K, N, _st())
    return out


# ------------------------------------------------------------------ W8A16 (FP8 weights)
def _check_fp8(name: str, a: GPUArray, b_nk: GPUArray, b_scale: GPUArray) -> None:
    if a.dtype != bfloat16:
        raise ValueError(f"{name} requires bfloat16 activation, got {a.dtype}")
    if b_nk.dtype != uint8:
        raise ValueError(f"{name} requires uint8 (FP8) weights, got {b_nk.dtype}")
    if b_scale.dtype != bfloat16:
        raise ValueError(f"{name} requires bfloat16 scale, got {b_scale.dtype}")


def gemv_fp8_bf16(a: GPUArray, b_nk: GPUArray, b_scale: GPUArray, *, out: GPUArray | None = None) -> GPUArray:
    """W8A16 GEMV: C[N] = A[K] @ dequant(B[N,K])^T (ops/matmul/gemv.py:73-131)."""
    if a.ndim != 1:
        raise ValueError(f"gemv_fp8_bf16 requires 1D input vector, got {a.ndim}D")
    if b_nk.ndim != 2:
        raise ValueError(f"gemv_fp8_bf16 requires 2D weight matrix, got {b_nk.ndim}D")
    _check_fp8("gemv_fp8_bf16", a, b_nk, b_scale)
    K, N = a.shape[0], b_nk.shape[0]
    if b_nk.shape[1] != K:
        raise ValueError(f"gemv_fp8_bf16 dimension mismatch: A[{K}] vs B[{N}, {b_nk.shape[1]}]")
    out = _out(out, (N,), bfloat16)
    api.gemv_fp8_bf16(a.device_ptr, b_nk.device_ptr, b_scale.device_ptr, out.device_ptr, 1, K, N, _st())
    return out


gemv_fp8_bf16_sm120 = gemv_fp8_bf16


def gemv_fp8_bf16_batched(a: GPUArray, b_nk: GPUArray, b_scale: GPUArray, *, out: GPUArray | None = None) -> GPUArray:
    """Batched W8A16 GEMV: C[M,N] = A[M,K] @ dequant(B[N,K])^T (ops/matmul/gemv.py:134-195)."""
    if a.ndim != 2:
        raise ValueError(f"gemv_fp8_bf16_batched requires 2D input matrix, got {a.ndim}D")
    if b_nk.ndim != 2:
        raise ValueError(f"gemv_fp8_bf16_batched requires 2D weight matrix, got {b_nk.ndim}D")
    _check_fp8("gemv_fp8_bf16_batched", a, b_nk, b_scale)
    M, K = a.shape
    N = b_nk.shape[0]
    if b_nk.shape[1] != K:
        raise ValueError(f"gemv_fp8_bf16_batched dimension mismatch: A[{M},{K}] vs B[{N},{b_nk.shape[1]}]")
    out = _out(out, (M, N), bfloat16)
    api.gemv_fp8_bf16(a.device_ptr, b_nk.device_ptr, b_scale.device_ptr, out.device_ptr, M, K, N, _st())
    return out


gemv_fp8_bf16_batched_sm120 = gemv_fp8_bf16_batched


def fp8_get_sizes(K: int, N: int) -> tuple[int, int, int]:
    """(scale_K, scale_N, scale_size) for 128x128 block scales (ops/matmul/fp8.py)."""
    sk, sn = (K + 127) // 128, (N + 127) // 128
    return sk, sn, sk * sn


def w8a16_gemm_init_lut() -> None:
    """No-op: the B200 kernels decode E4M3 with F2FP, there is no constant-memory LUT to load."""


def w8a16_gemm_sm120(a: GPUArray, b_fp8: GPUArray, b_scale: GPUArray, *, out: GPUArray | None = None) -> GPUArray:
    """W8A16 GEMM, reference layout B[K,N], scale [K/128, N/128] (ops/matmul/w8a16.py:38-119)."""
    if a.ndim != 2 or b_fp8.ndim != 2:
        raise ValueError("w8a16_gemm requires 2D inputs")
    if a.dtype != bfloat16:
        raise ValueError(f"w8a16_gemm requires bfloat16 activation, got {a.dtype}")
    if b_fp8.dtype != uint8:
        raise ValueError(f"w8a16_gemm requires uint8 (FP8) weights, got {b_fp8.dtype}")
    if b_scale.dtype != bfloat16:
        raise ValueError(f"w8a16_gemm requires bfloat16 scale, got {b_scale.dtype}")
    M, K = a.shape
    if b_fp8.shape[0] != K:
        raise ValueError(f"w8a16_gemm dimension mismatch: A[{M},{K}] vs B[{b_fp8.shape[0]},{b_fp8.shape[1]}]")
    N = b_fp8.shape[1]
    out = _out(out, (M, N), bfloat16)
    api.w8a16_gemm(a.device_ptr, b_fp8.device_ptr, b_scale.device_ptr, out.device_ptr, M, K, N, _st())
    return out


def grouped_gemm_init_lut() -> None:
    """No-op (see w8a16_gemm_init_lut)."""


def grouped_gemm_fp8_bf16(a: GPUArray, b_stacked: GPUArray, b_scale: GPUArray, row_expert_ids: GPUArray,
                          *, out: GPUArray | None = None) -> GPUArray:
    """MoE expert GEMM: row m of A[M,K] uses expert row_expert_ids[m] of B[E,N,K]
    (ops/matmul/grouped.py:36-131)."""
    if a.ndim != 2 or b_stacked.ndim != 3:
        raise ValueError("grouped_gemm_fp8_bf16 requires A[M,K] and B[E,N,K]")
    if a.dtype != bfloat16 or b_stacked.dtype != uint8 or b_scale.dtype != bfloat16:
        raise ValueError("grouped_gemm_fp8_bf16 requires bfloat16 A, uint8 B, bfloat16 scale")
    if row_expert_ids.dtype != int32:
        raise ValueError("grouped_gemm_fp8_bf16 requires int32 row_expert_ids")
    M, K = a.shape
    _, N, K2 = b_stacked.shape
    if K2 != K:
        raise ValueError(f"grouped_gemm_fp8_bf16 dimension mismatch: A[{M},{K}] vs B[..,{N},{K2}]")
    if row_expert_ids.size != M:
        raise ValueError(f"row_expert_ids size {row_expert_ids.size} must equal M={M}")
    out = _out(out, (M, N), bfloat16)
    api.moe_grouped_gemv_fp8(a.device_ptr, b_stacked.device_ptr, b_scale.device_ptr, out.device_ptr,
                             row_expert_ids.device_ptr, M, K, N, _st())
    return out


# ------------------------------------------------------------------ NVF4 (W4A16)
def nvf4_get_sizes(K: int, N: int) -> tuple[int, int]:
    """(data_size, scale_size) in bytes (ops/matmul/nvf4.py:14-26)."""
    return (K // 2) * N, ((K + 31) // 32) * N


gemv_nvf4_get_sizes = nvf4_get_sizes


def quantize_bf16_to_nvf4(input: GPUArray, out_data: GPUArray, out_scale: GPUArray) -> None:
    """BF16 [K,N] -> NVF4 data [K/2,N] + UE4M3 scale [K/32,N] (ops/matmul/nvf4.py:32-77)."""
    if input.ndim != 2:
        raise ValueError(f"quantize_bf16_to_nvf4 requires 2D input, got {input.ndim}D")
    if input.dtype != bfloat16:
        raise ValueError(f"quantize_bf16_to_nvf4 requires bfloat16 input, got {input.dtype}")
    K, N = input.shape
    dsz, ssz = nvf4_get_sizes(K, N)
    if out_data.dtype != uint8 or out_scale.dtype != uint8:
        raise ValueError("quantize_bf16_to_nvf4 requires uint8 output buffers")
    if out_data.size < dsz or out_scale.size < ssz:
        raise ValueError(f"output buffers too small: need data {dsz} / scale {ssz} bytes")
    api.quantize_bf16_to_nvf4(input.device_ptr, out_data.device_ptr, out_scale.device_ptr, K, N, _st())


def gemv_nvf4_bf16(a: GPUArray, b_data: GPUArray, b_scale: GPUArray, *, out: GPUArray | None = None,
                   alpha: float = 1.0) -> GPUArray:
    """NVF4 GEMV: C[N] = alpha * A[K] @ dequant(B[K,N]) (ops/matmul/nvf4.py:132-194)."""
    if a.ndim != 1:
        raise ValueError(f"gemv_nvf4_bf16 requires 1D input vector, got {a.ndim}D")
    if a.dtype != bfloat16:
        raise ValueError(f"gemv_nvf4_bf16 requires bfloat16 input, got {a.dtype}")
    if b_data.dtype != uint8 or b_scale.dtype != uint8:
        raise ValueError("gemv_nvf4_bf16 requires uint8 data and scale")
    K = a.shape[0]
    if b_data.ndim == 2:
        if b_data.shape[0] != K // 2:
            raise ValueError(f"gemv_nvf4_bf16 dimension mismatch: A[{K}] vs B_data{b_data.shape}")
        N = b_data.shape[1]
    else:
        if out is None:
            raise ValueError("gemv_nvf4_bf16 with flat weight buffers needs out= to infer N")
        N = out.shape[0]
    out = _out(out, (N,), bfloat16)
    api.gemv_nvf4_bf16(a.device_ptr, b_data.device_ptr, b_scale.device_ptr, out.device_ptr, 1, K, N,
                       float(alpha), _st())
    return out


gemv_nvf4_bf16_sm120 = gemv_nvf4_bf16


def gemv_nvf4_bf16_batched(a: GPUArray, b_data: GPUArray, b_scale: GPUArray, *, out: GPUArray | None = None,
                           alpha: float = 1.0) -> GPUArray:
    """C[M,N] = alpha * A[M,K] @ dequant(B[K,N]); B_data [K/2,N] (net-new: the reference has no batched NVF4 GEMV)."""
    if a.ndim != 2 or b_data.ndim != 2:
        raise ValueError("gemv_nvf4_bf16_batched requires A[M,K] and B_data[K/2,N]")
    if a.dtype != bfloat16 or b_data.dtype != uint8 or b_scale.dtype != uint8:
        raise ValueError("gemv_nvf4_bf16_batched requires bfloat16 A and uint8 data/scale")
    M, K = a.shape
    if b_data.shape[0] != K // 2:
        raise ValueError(f"gemv_nvf4_bf16_batched dimension mismatch: A[{M},{K}] vs B_data{b_data.shape}")
    N = b_data.shape[1]
    out = _out(out, (M, N), bfloat16)
    api.gemv_nvf4_bf16(a.device_ptr, b_data.device_ptr, b_scale.device_ptr, out.device_ptr, M, K, N,
                       float(alpha), _st())
    return out


# availability probes (ops/matmul/availability.py): this build targets sm_100a only
def gemv_nvf4_available() -> bool:
    return True


def fp8_available() -> bool:
    return True


nvf4_bf16_sm120_available = gemv_nvf4_available
w8a16_gemm_available = fp8_available
gemv_fp8_available = fp8_available


# ------------------------------------------------------------------ dense matmul
def matmul(a: GPUArray, b: GPUArray, *, out: GPUArray | None = None, use_tf32: bool | None = None) -> GPUArray:
    """C[M,N] = A[M,K] @ B[K,N] for the decode path's dense BF16 call (lm_head:
    models/causal.py:167-177 `matmul(hidden, lm_head.T)`).  B[K,N] is N-contiguous, which is the
    wrong order for weight streaming; the llm layers keep lm_head as [V,H] and call
    gemv_bf16(_batched) instead.  This entry point transposes B once per call (API parity, not
    a hot path) and is restricted to BF16."""
    if a.ndim != 2 or b.ndim != 2:
        raise ValueError("matmul requires 2D arrays")
    if a.shape[1] != b.shape[0]:
        raise ValueError(f"matmul dimension mismatch: {a.shape} @ {b.shape}")
    if a.dtype != b.dtype:
        raise ValueError("matmul requires arrays of same dtype")
    if a.dtype != bfloat16:
        raise ValueError(f"matmul on the B200 decode path supports bfloat16 only, got {a.dtype} "
                         "(the reference's cuBLASLt/CUTLASS dispatch is out of scope)")
    M, K = a.shape
    N = b.shape[1]
    from .tensor import transpose
    bt = transpose(b)
    out = _out(out, (M, N), bfloat16)
    api.gemv_bf16(a.device_ptr, bt.device_ptr, out.device_ptr, M, K, N, _st())
    return out


__all__ = [n for n in dir() if not n.startswith("_") and n not in ("annotations", "api", "GPUArray")]
_ = (float16, float32)
