"""GPUArray: shape + dtype + raw device pointer (mirror of pygpukit/core/array.py:17-742).

Contiguous row-major only; BF16 <-> NumPy uint16; `narrow/view/slice_rows` are zero-copy
views that keep the owner alive (core/array.py:401-548).  The owner frees its allocation in
__del__ (native/core/memory.cpp:29-36).
"""
from __future__ import annotations

import ctypes as C
from typing import Any

import numpy as np

from .._lib import api
from . import dtypes as _dt
from .dtypes import DataType
from .stream import current_stream_handle


class GPUArray:
    __slots__ = ("_ptr", "_shape", "_dtype", "_owner", "_source_ref", "__weakref__")

    def __init__(self, shape, dtype: DataType, ptr: int | None = None, source: "GPUArray | None" = None) -> None:
        self._shape = tuple(int(s) for s in shape)
        self._dtype = dtype
        self._source_ref = source
        if ptr is None:
            h = C.c_void_p()
            api.malloc(C.byref(h), max(1, self.nbytes))
            self._ptr = int(h.value)
            self._owner = True
        else:
            self._ptr = int(ptr)
            self._owner = False

    # ---- attribute surface of the reference
    @property
    def shape(self) -> tuple[int, ...]:
        return self._shape

    @property
    def dtype(self) -> DataType:
        return self._dtype

    @property
    def size(self) -> int:
        n = 1
        for s in self._shape:
            n *= s
        return n

    @property
    def ndim(self) -> int:
        return len(self._shape)

    @property
    def itemsize(self) -> int:
        return self._dtype.itemsize

    @property
    def nbytes(self) -> int:
        return self.size * self._dtype.itemsize

    @property
    def device_ptr(self) -> int:
        return self._ptr

    @property
    def on_gpu(self) -> bool:
        return True

    def _get_native(self) -> "GPUArray":
        return self

    def is_contiguous(self) -> bool:
        return True

    def contiguous(self) -> "GPUArray":
        return self

    # ---- host transfer
    def to_numpy(self) -> np.ndarray:
        out = np.empty(self._shape, dtype=self._dtype.to_numpy_dtype())
        if self.nbytes:
            st = current_stream_handle()
            api.memcpy_d2h(out.ctypes.data_as(C.c_void_p), self._ptr, self.nbytes, st)
            if st is None:
                api.device_sync()
            else:
                api.stream_sync(st)
        return out

    def copy_from_numpy(self, array: np.ndarray) -> None:
        a = np.ascontiguousarray(array, dtype=self._dtype.to_numpy_dtype())
        if a.size != self.size:
            raise ValueError(f"size mismatch: array {a.shape} vs GPUArray {self._shape}")
        if self.nbytes:
            st = current_stream_handle()
            api.memcpy_h2d(self._ptr, a.ctypes.data_as(C.c_void_p), self.nbytes, st)
            # pageable source: make sure the copy has consumed it before `a` can die
            if st is None:
                api.device_sync()
            else:
                api.stream_sync(st)

    def clone(self) -> "GPUArray":
        out = GPUArray(self._shape, self._dtype)
        if self.nbytes:
            api.memcpy_d2d(out._ptr, self._ptr, self.nbytes, current_stream_handle())
        return out

    def astype(self, dtype: DataType) -> "GPUArray":
        if dtype == self._dtype:
            return self.clone()
        out = GPUArray(self._shape, dtype)
        api.cast(self._ptr, out._ptr, self.size, _dt.float_code(self._dtype, "astype"),
                 _dt.float_code(dtype, "astype"), current_stream_handle())
        return out

    # ---- zero-copy views
    def _view(self, offset_elems: int, shape) -> "GPUArray":
        owner = self._source_ref if (self._source_ref is not None and not self._owner) else self
        return GPUArray(shape, self._dtype, ptr=self._ptr + offset_elems * self._dtype.itemsize, source=owner)

    def narrow(self, offset: int, length: int) -> "GPUArray":
        if self.ndim == 2:
            shape = (self._shape[0], length)
        elif self.ndim == 1:
            shape = (length,)
        else:
            raise ValueError(f"narrow() only supports 1D or 2D arrays, got {self.ndim}D")
        return self._view(offset, shape)

    def view(self, new_shape) -> "GPUArray":
        n = 1
        for d in new_shape:
            n *= d
        if n != self.size:
            raise ValueError(f"Cannot view array of size {self.size} as shape {tuple(new_shape)} (size {n})")
        return self._view(0, new_shape)

    def reshape(self, *shape) -> "GPUArray":
        if len(shape) == 1 and isinstance(shape[0], (tuple, list)):
            shape = tuple(shape[0])
        shape = list(shape)
        if -1 in shape:
            known = 1
            for d in shape:
                if d != -1:
                    known *= d
            shape[shape.index(-1)] = self.size // max(1, known)
        return self.view(tuple(shape))

    def slice_rows(self, num_rows: int) -> "GPUArray":
        if self.ndim != 2:
            raise ValueError(f"slice_rows() requires 2D array, got {self.ndim}D")
        if num_rows > self._shape[0]:
            raise ValueError(f"num_rows ({num_rows}) exceeds batch dimension ({self._shape[0]})")
        return self._view(0, (num_rows, self._shape[1]))

    def row(self, index: int) -> "GPUArray":
        """View of one leading-axis slice (convenience for caches/buffers)."""
        inner = 1
        for d in self._shape[1:]:
            inner *= d
        return self._view(index * inner, self._shape[1:])

    # ---- operators used on the path
    def __add__(self, other: "GPUArray") -> "GPUArray":
        from ..ops.elementwise import add
        return add(self, other)

    def __mul__(self, other: "GPUArray") -> "GPUArray":
        from ..ops.elementwise import mul
        return mul(self, other)

    def __matmul__(self, other: "GPUArray") -> "GPUArray":
        from ..ops.matmul import matmul
        return matmul(self, other)

    @property
    def T(self) -> "GPUArray":
        from ..ops.tensor import transpose
        return transpose(self)

    def __repr__(self) -> str:
        return f"GPUArray(shape={self._shape}, dtype={self._dtype.name})"

    __str__ = __repr__

    def __del__(self) -> None:
        if getattr(self, "_owner", False) and getattr(self, "_ptr", 0):
            try:
                api.free(self._ptr)
            except Exception:
                pass
            self._ptr = 0
