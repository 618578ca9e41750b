"""Layout helpers used on the path (mirror of pygpukit/ops/tensor.py)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from .._lib import api
from ..core.array import GPUArray
from ..core.dtypes import bfloat16, float16, float32, float_code
from ._common import _out, _st, _validate_float_dtype


def transpose_3d_021(input: GPUArray, *, out: GPUArray | None = None) -> GPUArray | None:
    """[d0, d1, d2] -> [d1, d0, d2] (ops/tensor.py:127-188)."""
    if input.ndim != 3:
        raise ValueError(f"transpose_3d_021 expects 3D input, got {input.ndim}D")
    d0, d1, d2 = input.shape
    given = out is not None
    out = _out(out, (d1, d0, d2), input.dtype)
    api.transpose_021(input.device_ptr, out.device_ptr, d0, d1, d2, input.itemsize, _st())
    return None if given else out


def transpose(a: GPUArray) -> GPUArray:
    """2D transpose (copy)."""
    if a.ndim != 2:
        raise ValueError(f"transpose expects 2D input, got {a.ndim}D")
    out = GPUArray((a.shape[1], a.shape[0]), a.dtype)
    api.transpose_021(a.device_ptr, out.device_ptr, a.shape[0], a.shape[1], 1, a.itemsize, _st())
    return out


def reshape_copy(input: GPUArray, new_shape=None, *, out: GPUArray | None = None) -> GPUArray | None:
    """Contiguous reshape with copy (ops/tensor.py:389-482)."""
    _validate_float_dtype(input, "reshape_copy")
    if out is not None:
        target = out.shape
    elif new_shape is not None:
        target = tuple(new_shape)
    else:
        raise ValueError("reshape_copy: either new_shape or out must be provided")
    n = 1
    for d in target:
        n *= d
    if n != input.size:
        raise ValueError(f"reshape_copy: total size mismatch ({input.size} vs {n})")
    given = out is not None
    if out is None:
        out = GPUArray(target, input.dtype)
    api.memcpy_d2d(out.device_ptr, input.device_ptr, input.nbytes, _st())
    return None if given else out


def _cast(src: GPUArray, dst_dtype) -> GPUArray:
    out = GPUArray(src.shape, dst_dtype)
    api.cast(src.device_ptr, out.device_ptr, src.size, float_code(src.dtype), float_code(dst_dtype), _st())
    return out


def cast_f32_to_bf16(src: GPUArray) -> GPUArray:
    return _cast(src, bfloat16)


def cast_f32_to_f16(src: GPUArray) -> GPUArray:
    return _cast(src, float16)


def cast_bf16_to_f32(src: GPUArray) -> GPUArray:
    return _cast(src, float32)


def cast_f16_to_f32(src: GPUArray) -> GPUArray:
    return _cast(src, float32)


__all__ = ["transpose_3d_021", "transpose", "reshape_copy", "cast_f32_to_bf16", "cast_f32_to_f16",
           "cast_bf16_to_f32", "cast_f16_to_f32"]
_ = (C, np)
