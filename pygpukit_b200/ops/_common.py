from __future__ import annotations

from ..core.array import GPUArray
from ..core.dtypes import bfloat16, float16, float32
from ..core.stream import current_stream_handle as _st


def _validate_float_dtype(a: GPUArray, op: str) -> None:
    """Same check as pygpukit/ops/_common.py (float dtypes only)."""
    if a.dtype not in (float32, float16, bfloat16):
        raise ValueError(f"{op} requires float dtype, got {a.dtype}")


def _validate_same_shape(a: GPUArray, b: GPUArray, op: str) -> None:
    if a.shape != b.shape:
        raise ValueError(f"{op} requires arrays of same shape, got {a.shape} and {b.shape}")


def _validate_same_dtype(a: GPUArray, b: GPUArray, op: str) -> None:
    if a.dtype != b.dtype:
        raise ValueError(f"{op} requires arrays of same dtype, got {a.dtype} and {b.dtype}")


def _out(out: GPUArray | None, shape, dtype, what: str = "out") -> GPUArray:
    if out is None:
        return GPUArray(shape, dtype)
    if out.shape != tuple(shape):
        raise ValueError(f"{what} shape {out.shape} does not match expected {tuple(shape)}")
    if out.dtype != dtype:
        raise ValueError(f"{what} dtype {out.dtype} must be {dtype}")
    return out


__all__ = ["_st", "_validate_float_dtype", "_validate_same_shape", "_validate_same_dtype", "_out"]
