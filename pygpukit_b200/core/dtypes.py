"""Data types (mirror of pygpukit/core/dtypes.py:10-123; BF16 is uint16 storage in NumPy)."""
from __future__ import annotations

from dataclasses import dataclass
from enum import Enum
from typing import Any

import numpy as np


class DataTypeKind(Enum):
    FLOAT64 = "float64"
    FLOAT32 = "float32"
    FLOAT16 = "float16"
    BFLOAT16 = "bfloat16"
    INT64 = "int64"
    INT32 = "int32"
    INT16 = "int16"
    INT8 = "int8"
    UINT8 = "uint8"
    INT4 = "int4"


_NP = {"float64": np.float64, "float32": np.float32, "float16": np.float16, "bfloat16": np.uint16,
       "int64": np.int64, "int32": np.int32, "int16": np.int16, "int8": np.int8, "uint8": np.uint8,
       "int4": np.uint8}


@dataclass(frozen=True)
class DataType:
    kind: DataTypeKind
    itemsize: int
    name: str

    def __str__(self) -> str:
        return self.name

    def __repr__(self) -> str:
        return f"DataType({self.name})"

    def to_numpy_dtype(self) -> Any:
        return np.dtype(_NP[self.name])

    @staticmethod
    def from_numpy_dtype(dtype: Any) -> "DataType":
        name = np.dtype(dtype).name
        if name == "uint16":
            return bfloat16
        if name in _BY_NAME and name not in ("bfloat16", "int4"):
            return _BY_NAME[name]
        raise ValueError(f"Unsupported dtype: {dtype}")

    @staticmethod
    def from_string(name: str) -> "DataType":
        if name not in _BY_NAME:
            raise ValueError(f"Unsupported dtype string: {name}")
        return _BY_NAME[name]


float64 = DataType(DataTypeKind.FLOAT64, 8, "float64")
float32 = DataType(DataTypeKind.FLOAT32, 4, "float32")
float16 = DataType(DataTypeKind.FLOAT16, 2, "float16")
bfloat16 = DataType(DataTypeKind.BFLOAT16, 2, "bfloat16")
int64 = DataType(DataTypeKind.INT64, 8, "int64")
int32 = DataType(DataTypeKind.INT32, 4, "int32")
int16 = DataType(DataTypeKind.INT16, 2, "int16")
int8 = DataType(DataTypeKind.INT8, 1, "int8")
uint8 = DataType(DataTypeKind.UINT8, 1, "uint8")
int4 = DataType(DataTypeKind.INT4, 1, "int4")
_BY_NAME = {d.name: d for d in (float64, float32, float16, bfloat16, int64, int32, int16, int8, uint8, int4)}

# dtype codes of the C ABI (include/pgk_b200.h: 0 = f32, 1 = f16, 2 = bf16)
_CODE = {"float32": 0, "float16": 1, "bfloat16": 2}


def float_code(dt: DataType, op: str = "op") -> int:
    if dt.name not in _CODE:
        raise ValueError(f"{op} requires float32/float16/bfloat16, got {dt}")
    return _CODE[dt.name]


def f32_to_bf16_bits(x: np.ndarray) -> np.ndarray:
    """float32 -> bf16 bits, round-to-nearest-even (core/array.py:386-395 bit trick)."""
    u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32)
    return ((u + 0x7FFF + ((u >> 16) & 1)) >> 16).astype(np.uint16)


def bf16_bits_to_f32(b: np.ndarray) -> np.ndarray:
    return (np.ascontiguousarray(b, dtype=np.uint16).astype(np.uint32) << 16).view(np.float32)
