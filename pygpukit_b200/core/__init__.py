from .array import GPUArray
from .dtypes import (DataType, DataTypeKind, bfloat16, float16, float32, float64, int4, int8, int16,
                     int32, int64, uint8)
from .factory import empty, from_numpy, ones, zeros
from .stream import CudaEvent, CudaGraph, Stream, default_stream, event_elapsed_ms, event_elapsed_us, synchronize

__all__ = ["GPUArray", "DataType", "DataTypeKind", "bfloat16", "float16", "float32", "float64", "int4",
           "int8", "int16", "int32", "int64", "uint8", "empty", "from_numpy", "ones", "zeros", "CudaEvent",
           "CudaGraph", "Stream", "default_stream", "event_elapsed_ms", "event_elapsed_us", "synchronize"]
