"""Neural-network ops of the decode path (mirror of pygpukit/ops/nn/{norm,rope,attention,
activation,linear,fused}.py)."""
from __future__ import annotations

from .._lib import api
from ..core.array import GPUArray
from ..core.dtypes import bfloat16, float16, float32, float_code, int32
from ._common import _out, _st, _validate_float_dtype


# ------------------------------------------------------------------ norms
def rmsnorm(input: GPUArray, gamma: GPUArray, eps: float = 1e-5, *, out: GPUArray | None = None) -> GPUArray:
    """x / sqrt(mean(x^2) + eps) * gamma (ops/nn/norm.py:118-218)."""
    _validate_float_dtype(input, "rmsnorm")
    if input.ndim != 2:
        raise ValueError(f"rmsnorm expects 2D input [batch, features], got {input.ndim}D")
    if gamma.ndim != 1:
        raise ValueError("rmsnorm expects 1D gamma")
    if input.dtype != gamma.dtype:
        raise ValueError("rmsnorm: all inputs must have same dtype")
    features = input.shape[1]
    if gamma.shape[0] != features:
        raise ValueError(f"rmsnorm: gamma size {gamma.shape[0]} must match features {features}")
    if out is not None:
        if out.shape != input.shape:
            raise ValueError(f"out shape {out.shape} does not match input shape {input.shape}")
        if out.dtype != input.dtype:
            raise ValueError(f"out dtype {out.dtype} does not match input dtype {input.dtype}")
    else:
        out = GPUArray(input.shape, input.dtype)
    api.rmsnorm(input.device_ptr, gamma.device_ptr, out.device_ptr, input.shape[0], features, float(eps),
                float_code(input.dtype), _st())
    return out


def layernorm(input: GPUArray, gamma: GPUArray, beta: GPUArray, eps: float = 1e-5, *,
              out: GPUArray | None = None) -> GPUArray:
    """(x - mean) / sqrt(var + eps) * gamma + beta (ops/nn/norm.py:16-115)."""
    _validate_float_dtype(input, "layernorm")
    if input.ndim != 2:
        raise ValueError(f"layernorm expects 2D input [batch, features], got {input.ndim}D")
    if gamma.ndim != 1 or beta.ndim != 1:
        raise ValueError("layernorm expects 1D gamma and beta")
    if input.dtype != gamma.dtype or input.dtype != beta.dtype:
        raise ValueError("layernorm: all inputs must have same dtype")
    features = input.shape[1]
    if gamma.shape[0] != features or beta.shape[0] != features:
        raise ValueError(f"layernorm: gamma/beta size must match features {features}")
    out = _out(out, input.shape, input.dtype)
    api.layernorm(input.device_ptr, gamma.device_ptr, beta.device_ptr, out.device_ptr, input.shape[0], features,
                  float(eps), float_code(input.dtype), _st())
    return out


def rmsnorm_residual(input: GPUArray, residual: GPUArray, gamma: GPUArray, eps: float = 1e-5, *,
                     out: GPUArray | None = None) -> GPUArray:
    """rmsnorm(input + residual) * gamma (ops/nn/fused.py:16-73)."""
    _validate_float_dtype(input, "rmsnorm_residual")
    if input.ndim != 2 or residual.shape != input.shape:
        raise ValueError("rmsnorm_residual expects 2D input and residual of the same shape")
    if gamma.ndim != 1 or gamma.shape[0] != input.shape[1]:
        raise ValueError("rmsnorm_residual: gamma must be 1D of size features")
    if input.dtype != residual.dtype or input.dtype != gamma.dtype:
        raise ValueError("rmsnorm_residual: all inputs must have same dtype")
    out = _out(out, input.shape, input.dtype)
    api.rmsnorm_residual(input.device_ptr, residual.device_ptr, gamma.device_ptr, out.device_ptr, input.shape[0],
                         input.shape[1], float(eps), float_code(input.dtype), _st())
    return out


# ------------------------------------------------------------------ RoPE
def _rope_check(q: GPUArray, k: GPUArray, cos: GPUArray, sin: GPUArray, name: str) -> None:
    _validate_float_dtype(q, name)
    if q.ndim != 3 or k.ndim != 3:
        raise ValueError(f"{name} expects 3D q, k [seq_len, n_heads, head_dim]")
    if cos.ndim != 2 or sin.ndim != 2:
        raise ValueError(f"{name} expects 2D cos, sin [seq_len, head_dim]")
    if q.dtype != k.dtype:
        raise ValueError(f"{name}: q and k must have same dtype")
    if q.shape[0] != k.shape[0] or q.shape[2] != k.shape[2]:
        raise ValueError(f"{name}: q {q.shape} and k {k.shape} must share seq_len and head_dim")
    if cos.shape != sin.shape or cos.shape[0] < q.shape[0] or cos.shape[1] != q.shape[2]:
        raise ValueError(f"{name}: cos/sin shape {cos.shape} does not match [seq_len, head_dim]")


def rope_inplace(q: GPUArray, k: GPUArray, cos: GPUArray, sin: GPUArray) -> None:
    """NeoX half-split RoPE in place; tables in q's dtype (ops/nn/rope.py:16-106)."""
    _rope_check(q, k, cos, sin, "rope_inplace")
    if cos.dtype != q.dtype or sin.dtype != q.dtype:
        raise ValueError("rope_inplace: cos/sin must have the dtype of q (use rope_inplace_f32table for f32 tables)")
    dt = float_code(q.dtype)
    api.rope_inplace(q.device_ptr, k.device_ptr, cos.device_ptr, sin.device_ptr, q.shape[0], q.shape[1], k.shape[1],
                     q.shape[2], dt, dt if dt != 0 else 0, _st())


def rope_inplace_f32table(q: GPUArray, k: GPUArray, cos: GPUArray, sin: GPUArray) -> None:
    """RoPE with f32 cos/sin tables on bf16/f16 q, k (ops/nn/rope.py:109-133)."""
    _rope_check(q, k, cos, sin, "rope_inplace_f32table")
    if cos.dtype != float32 or sin.dtype != float32:
        raise ValueError("rope_inplace_f32table requires float32 cos/sin")
    api.rope_inplace(q.device_ptr, k.device_ptr, cos.device_ptr, sin.device_ptr, q.shape[0], q.shape[1], k.shape[1],
                     q.shape[2], float_code(q.dtype), 0, _st())


# ------------------------------------------------------------------ attention
def _sdpa_check(Q: GPUArray, K: GPUArray, V: GPUArray, out: GPUArray, name: str) -> None:
    _validate_float_dtype(Q, name)
    if Q.ndim != 3 or K.ndim != 3 or V.ndim != 3 or out.ndim != 3:
        raise ValueError(f"{name} expects 3D Q, K, V, out [n_heads, len, head_dim]")
    if K.shape != V.shape:
        raise ValueError(f"{name}: K {K.shape} and V {V.shape} must have the same shape")
    if Q.dtype != K.dtype or Q.dtype != V.dtype or Q.dtype != out.dtype:
        raise ValueError(f"{name}: Q, K, V, out must have same dtype")
    if out.shape != Q.shape:
        raise ValueError(f"{name}: out shape {out.shape} must match Q shape {Q.shape}")
    if Q.shape[2] != K.shape[2]:
        raise ValueError(f"{name}: head_dim mismatch Q {Q.shape[2]} vs K {K.shape[2]}")
    if Q.shape[0] % K.shape[0] != 0:
        raise ValueError(f"{name}: n_heads {Q.shape[0]} must be a multiple of KV heads {K.shape[0]}")


def sdpa_causal_fixed_cache(Q: GPUArray, K: GPUArray, V: GPUArray, out: GPUArray, context_len: int,
                            scale: float = 0.0) -> None:
    """Causal SDPA over the first `context_len` rows of a fixed KV cache
    (ops/nn/attention.py:159-192).  K/V [n_heads, max_seq_len, head_dim] as in the reference, or
    the un-expanded [n_kv_heads, max_seq_len, head_dim] (GQA handled in the kernel)."""
    _sdpa_check(Q, K, V, out, "sdpa_causal_fixed_cache")
    if context_len <= 0 or context_len > K.shape[1]:
        raise ValueError(f"sdpa_causal_fixed_cache: context_len {context_len} out of range (max_seq_len {K.shape[1]})")
    H, qlen, hd = Q.shape
    api.sdpa_decode(Q.device_ptr, K.device_ptr, V.device_ptr, out.device_ptr, H, K.shape[0], qlen, hd, K.shape[1],
                    int(context_len), None, float(scale), qlen * hd, hd, qlen * hd, hd, float_code(Q.dtype), _st())


def sdpa_causal_fixed_cache_ptr(Q: GPUArray, K: GPUArray, V: GPUArray, out: GPUArray, context_len_buf: GPUArray,
                                max_kv_len: int, scale: float = 0.0) -> None:
    """Same, context_len read from an int32 device buffer (ops/nn/attention.py:195-235)."""
    _sdpa_check(Q, K, V, out, "sdpa_causal_fixed_cache_ptr")
    if context_len_buf.dtype != int32:
        raise ValueError("sdpa_causal_fixed_cache_ptr: context_len_buf must be int32")
    if max_kv_len <= 0 or max_kv_len > K.shape[1]:
        raise ValueError(f"sdpa_causal_fixed_cache_ptr: max_kv_len {max_kv_len} out of range (max_seq_len {K.shape[1]})")
    H, qlen, hd = Q.shape
    api.sdpa_decode(Q.device_ptr, K.device_ptr, V.device_ptr, out.device_ptr, H, K.shape[0], qlen, hd, K.shape[1],
                    int(max_kv_len), context_len_buf.device_ptr, float(scale), qlen * hd, hd, qlen * hd, hd,
                    float_code(Q.dtype), _st())


def sdpa_causal(Q: GPUArray, K: GPUArray, V: GPUArray, scale: float = 0.0, *, out: GPUArray | None = None) -> GPUArray:
    """Causal SDPA, Q [H, q_len, hd], K/V [H or H_kv, kv_len, hd] (ops/nn/attention.py:16-157).
    Decode-shaped (small q_len) workloads only: prefill attention is out of scope of this path."""
    if out is None:
        out = GPUArray(Q.shape, Q.dtype)
    sdpa_causal_fixed_cache(Q, K, V, out, K.shape[1], scale)
    return out


# ------------------------------------------------------------------ activations / bias
def _unary(name: str, fn, a: GPUArray, out: GPUArray | None) -> GPUArray:
    _validate_float_dtype(a, name)
    out = _out(out, a.shape, a.dtype)
    fn(a.device_ptr, out.device_ptr, a.size, float_code(a.dtype), _st())
    return out


def silu(a: GPUArray, *, out: GPUArray | None = None) -> GPUArray:
    """x * sigmoid(x) (ops/nn/activation.py:61-108)."""
    return _unary("silu", api.silu, a, out)


def gelu(a: GPUArray, *, out: GPUArray | None = None) -> GPUArray:
    """tanh-approximation GELU (ops/nn/activation.py:16-58)."""
    return _unary("gelu", api.gelu, a, out)


def swiglu(gate: GPUArray, up: GPUArray, *, out: GPUArray | None = None) -> GPUArray:
    """silu(gate) * up with one rounding (ops/nn/fused.py:76-125)."""
    _validate_float_dtype(gate, "swiglu")
    if gate.shape != up.shape or gate.dtype != up.dtype:
        raise ValueError("swiglu: gate and up must have same shape and dtype")
    out = _out(out, gate.shape, gate.dtype)
    api.swiglu(gate.device_ptr, up.device_ptr, out.device_ptr, gate.size, float_code(gate.dtype), _st())
    return out


def bias_add_inplace(output: GPUArray, bias: GPUArray) -> None:
    """output[batch, features] += bias[features] (ops/nn/linear.py:14-69)."""
    _validate_float_dtype(output, "bias_add_inplace")
    if output.ndim != 2:
        raise ValueError(f"bias_add_inplace expects 2D output [batch, features], got {output.ndim}D")
    if bias.ndim != 1:
        raise ValueError(f"bias_add_inplace expects 1D bias [features], got {bias.ndim}D")
    if output.dtype != bias.dtype:
        raise ValueError("bias_add_inplace: output and bias must have same dtype")
    if output.shape[1] != bias.shape[0]:
        raise ValueError(f"bias_add_inplace: bias size {bias.shape[0]} must match features {output.shape[1]}")
    api.bias_add_inplace(output.device_ptr, bias.device_ptr, output.shape[0], output.shape[1],
                         float_code(output.dtype), _st())


__all__ = ["rmsnorm", "layernorm", "rmsnorm_residual", "rope_inplace", "rope_inplace_f32table",
           "sdpa_causal", "sdpa_causal_fixed_cache", "sdpa_causal_fixed_cache_ptr", "silu", "gelu", "swiglu",
           "bias_add_inplace"]
_ = (bfloat16, float16)
