"""pygpukit_b200 -- B200-native (sm_100a) implementation of PyGPUkit's LLM decode hot path.

Drop-in for `import pygpukit as gk` on that path: GPUArray, from_numpy/zeros/empty/ones, dtypes,
`ops.*` (gemv_*, matmul, rmsnorm, rope_*, sdpa_*, kv_cache_*, ...) and `llm.*` layers / models /
decode strategies.  All compute goes through libpgk_b200.so (C ABI, include/pgk_b200.h);
there is no CPU fallback.
"""
from . import _lib
from .core import (CudaEvent, CudaGraph, DataType, GPUArray, Stream, bfloat16, default_stream, empty,
                   event_elapsed_ms, event_elapsed_us, float16, float32, float64, from_numpy, int4, int8,
                   int16, int32, int64, ones, synchronize, uint8, zeros)

__version__ = "0.1.0"


def is_cuda_available() -> bool:
    import ctypes as C
    try:
        n = C.c_int()
        return _lib.load().pgk_device_count(C.byref(n)) == 0 and n.value > 0
    except Exception:
        return False


def get_device_info(device: int = 0):
    import ctypes as C
    p = _lib.DeviceProps()
    _lib.api.device_props(device, C.byref(p))
    return {"name": p.name.decode(), "sm_count": p.sm_count, "compute_capability": (p.cc_major, p.cc_minor),
            "total_memory": p.total_mem, "l2_bytes": p.l2_bytes, "max_smem_per_block": p.max_smem_per_block}


from . import ops  # noqa: E402
from .ops import (add, bias_add_inplace, gelu, layernorm, matmul, mul, rmsnorm, silu, transpose)  # noqa: E402,F401


def __getattr__(name):
    if name == "llm":
        import importlib
        return importlib.import_module(".llm", __name__)
    raise AttributeError(name)
