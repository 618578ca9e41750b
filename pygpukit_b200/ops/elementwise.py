"""Elementwise glue (mirror of pygpukit/ops/elementwise.py:18-252)."""
from __future__ import annotations

from .._lib import api
from ..core.array import GPUArray
from ..core.dtypes import float_code
from ._common import _out, _st, _validate_float_dtype, _validate_same_dtype, _validate_same_shape


def _binary(name: str, fn, a: GPUArray, b: GPUArray, out: GPUArray | None) -> GPUArray:
    _validate_same_shape(a, b, name)
    _validate_same_dtype(a, b, name)
    _validate_float_dtype(a, name)
    out = _out(out, a.shape, a.dtype)
    fn(a.device_ptr, b.device_ptr, out.device_ptr, a.size, float_code(a.dtype), _st())
    return out


def add(a: GPUArray, b: GPUArray, *, out: GPUArray | None = None) -> GPUArray:
    return _binary("add", api.add, a, b, out)


def mul(a: GPUArray, b: GPUArray, *, out: GPUArray | None = None) -> GPUArray:
    return _binary("mul", api.mul, a, b, out)


def add_inplace(a: GPUArray, b: GPUArray) -> None:
    """a += b, native-precision add (elementwise_kernels.cuh:370-410)."""
    if a.size != b.size:
        raise ValueError(f"add_inplace requires arrays of same size, got {a.shape} and {b.shape}")
    _validate_same_dtype(a, b, "add_inplace")
    api.add_inplace(a.device_ptr, b.device_ptr, a.size, float_code(a.dtype, "add_inplace"), _st())


def mul_inplace(a: GPUArray, b: GPUArray) -> None:
    if a.size != b.size:
        raise ValueError(f"mul_inplace requires arrays of same size, got {a.shape} and {b.shape}")
    _validate_same_dtype(a, b, "mul_inplace")
    api.mul_inplace(a.device_ptr, b.device_ptr, a.size, float_code(a.dtype, "mul_inplace"), _st())


def copy_to(src: GPUArray, dst: GPUArray) -> None:
    """GPU-to-GPU copy (ops/elementwise.py:237-251)."""
    if src.nbytes != dst.nbytes:
        raise ValueError(f"copy_to requires same size, got {src.shape}/{src.dtype} and {dst.shape}/{dst.dtype}")
    api.memcpy_d2d(dst.device_ptr, src.device_ptr, src.nbytes, _st())


__all__ = ["add", "mul", "add_inplace", "mul_inplace", "copy_to"]
