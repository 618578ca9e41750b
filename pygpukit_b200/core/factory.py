"""Factory functions (mirror of pygpukit/core/factory.py:16-198)."""
from __future__ import annotations

import numpy as np

from .._lib import api
from .array import GPUArray
from .dtypes import DataType, float32
from .stream import current_stream_handle


def _dtype(dtype) -> DataType:
    if isinstance(dtype, DataType):
        return dtype
    if isinstance(dtype, str):
        return DataType.from_string(dtype)
    return DataType.from_numpy_dtype(dtype)


def _shape(shape) -> tuple[int, ...]:
    if isinstance(shape, int):
        shape = (shape,)
    shape = tuple(int(s) for s in shape)
    if any(s < 0 for s in shape):
        raise ValueError(f"negative dimensions are not allowed: {shape}")
    return shape


def empty(shape, dtype=float32) -> GPUArray:
    return GPUArray(_shape(shape), _dtype(dtype))


def zeros(shape, dtype=float32) -> GPUArray:
    a = GPUArray(_shape(shape), _dtype(dtype))
    if a.nbytes:
        api.memset(a.device_ptr, 0, a.nbytes, current_stream_handle())
    return a


def ones(shape, dtype=float32) -> GPUArray:
    dt = _dtype(dtype)
    if dt.name == "bfloat16":
        host = np.full(_shape(shape), 0x3F80, dtype=np.uint16)
    else:
        host = np.ones(_shape(shape), dtype=dt.to_numpy_dtype())
    return from_numpy(host) if dt.name != "bfloat16" else from_numpy(host)


def from_numpy(array: np.ndarray) -> GPUArray:
    """Host -> device copy; uint16 arrays are BF16 storage (core/factory.py:168-198)."""
    a = np.ascontiguousarray(array)
    out = GPUArray(a.shape, DataType.from_numpy_dtype(a.dtype))
    out.copy_from_numpy(a)
    return out
