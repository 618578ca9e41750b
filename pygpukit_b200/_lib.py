"""ctypes binding of libpgk_b200.so (the C ABI in include/pgk_b200.h).

The product path has NO CPU fallback: if the shared library is missing or cannot be loaded
this module raises, and every op raises RuntimeError on a non-zero return code with the
library's last error string ("<fn> failed: <cuda error>", the reference's RuntimeError text).
Replaces pygpukit/_native_loader.py:106-169 + the pybind11 module native/bindings/module.cpp.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PGK_B200_LIB", os.path.join(_HERE, "libpgk_b200.so"))

_lib = None

vp, i32, i64, f32, sz, cl = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_size_t, C.c_long


class DeviceProps(C.Structure):
    _fields_ = [("name", C.c_char * 128), ("sm_count", i32), ("cc_major", i32), ("cc_minor", i32),
                ("total_mem", sz), ("l2_bytes", sz), ("max_smem_per_block", i32)]


# name -> argtypes (every function returns int unless listed in _RESTYPE)
_SIGS = {
    "pgk_version": [],
    "pgk_device_count": [C.POINTER(i32)], "pgk_device_set": [i32], "pgk_device_get": [C.POINTER(i32)],
    "pgk_device_props": [i32, C.POINTER(DeviceProps)], "pgk_device_sync": [],
    "pgk_mem_info": [C.POINTER(sz), C.POINTER(sz)],
    "pgk_malloc": [C.POINTER(vp), sz], "pgk_free": [vp],
    "pgk_host_alloc": [C.POINTER(vp), sz], "pgk_host_free": [vp],
    "pgk_memset": [vp, i32, sz, vp],
    "pgk_memcpy_h2d": [vp, vp, sz, vp], "pgk_memcpy_d2h": [vp, vp, sz, vp], "pgk_memcpy_d2d": [vp, vp, sz, vp],
    "pgk_stream_create": [C.POINTER(vp)], "pgk_stream_destroy": [vp], "pgk_stream_sync": [vp],
    "pgk_event_create": [C.POINTER(vp)], "pgk_event_destroy": [vp], "pgk_event_record": [vp, vp],
    "pgk_event_sync": [vp], "pgk_event_elapsed_ms": [vp, vp, C.POINTER(f32)],
    "pgk_graph_begin": [vp], "pgk_graph_end": [vp, C.POINTER(vp)], "pgk_graph_launch": [vp, vp],
    "pgk_graph_destroy": [vp],
    "pgk_set_pdl": [i32], "pgk_flush_l2": [vp],
    "pgk_gemv_fp8_bf16": [vp, vp, vp, vp, i32, i32, i32, vp],
    "pgk_gemv_fp8_bf16_ex": [vp, vp, vp, vp, i32, i32, i32, vp, vp, vp],
    "pgk_gemv_fp8_swiglu_bf16": [vp, vp, vp, vp, vp, vp, i32, i32, i32, vp],
    "pgk_w8a16_gemm": [vp, vp, vp, vp, i32, i32, i32, vp],
    "pgk_gemv_nvf4_bf16": [vp, vp, vp, vp, i32, i32, i32, f32, vp],
    "pgk_gemv_nvf4_bf16_ex": [vp, vp, vp, vp, i32, i32, i32, f32, vp, vp, vp],
    "pgk_gemv_nvf4_swiglu_bf16": [vp, vp, vp, vp, i32, i32, i32, vp],
    "pgk_quantize_bf16_to_nvf4": [vp, vp, vp, i32, i32, vp],
    "pgk_gemm_nvf4_bf16_tc": [vp, vp, vp, vp, i32, i32, i32, f32, vp, vp, vp],
    "pgk_swiglu_interleaved8_bf16": [vp, vp, i32, i32, vp],
    "pgk_gemv_bf16": [vp, vp, vp, i32, i32, i32, vp],
    "pgk_gemv_bf16_ex": [vp, vp, vp, i32, i32, i32, vp, vp, vp],
    "pgk_gemv_bf16_argmax": [vp, vp, vp, vp, i32, i32, vp],
    "pgk_rmsnorm_gemv_fp8_bf16": [vp, vp, f32, vp, vp, vp, i32, i32, i32, vp, vp, vp],
    "pgk_rmsnorm_gemv_fp8_swiglu_bf16": [vp, vp, f32, vp, vp, vp, vp, vp, i32, i32, i32, vp],
    "pgk_rmsnorm_gemv_nvf4_bf16": [vp, vp, f32, vp, vp, vp, i32, i32, i32, f32, vp, vp, vp],
    "pgk_rmsnorm_gemv_nvf4_swiglu_bf16": [vp, vp, f32, vp, vp, vp, i32, i32, i32, vp],
    "pgk_rmsnorm_gemv_bf16": [vp, vp, f32, vp, vp, i32, i32, i32, vp, vp, vp],
    "pgk_rmsnorm_gemv_bf16_argmax": [vp, vp, f32, vp, vp, vp, i32, i32, vp],
    "pgk_rmsnorm": [vp, vp, vp, i32, i32, f32, i32, vp],
    "pgk_layernorm": [vp, vp, vp, vp, i32, i32, f32, i32, vp],
    "pgk_rmsnorm_residual": [vp, vp, vp, vp, i32, i32, f32, i32, vp],
    "pgk_rope_inplace": [vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp],
    "pgk_silu": [vp, vp, sz, i32, vp], "pgk_gelu": [vp, vp, sz, i32, vp],
    "pgk_swiglu": [vp, vp, vp, sz, i32, vp], "pgk_silu_mul": [vp, vp, vp, sz, i32, vp],
    "pgk_add": [vp, vp, vp, sz, i32, vp], "pgk_mul": [vp, vp, vp, sz, i32, vp],
    "pgk_add_inplace": [vp, vp, sz, i32, vp], "pgk_mul_inplace": [vp, vp, sz, i32, vp],
    "pgk_bias_add_inplace": [vp, vp, i32, i32, i32, vp],
    "pgk_cast": [vp, vp, sz, i32, i32, vp],
    "pgk_embedding_lookup": [vp, vp, i32, i32, i32, vp],
    "pgk_embedding_lookup_ptr": [vp, vp, i32, vp, i32, i32, vp],
    "pgk_embedding_lookup_pos": [vp, vp, vp, i32, vp, i32, i32, vp, i32, vp],
    "pgk_argmax": [vp, vp, i32, i32, vp],
    "pgk_transpose_021": [vp, vp, i32, i32, i32, i32, vp],
    "pgk_kv_cache_update": [vp, vp, i32, i32, i32, i32, i32, i32, vp, i32, vp],
    "pgk_qk_norm_rope_kv_write": [vp, vp, vp, f32, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp, vp],
    "pgk_gemm_fp8_bf16_tc": [vp, vp, vp, vp, i32, i32, i32, vp, vp, vp],
    "pgk_gemm_fp8_bf16_f8mma": [vp, vp, vp, vp, i32, i32, i32, vp, vp, vp],
    "pgk_rope_sdpa_decode": [vp, vp, vp, f32, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp, i32, vp, f32, vp],
    "pgk_sdpa_decode": [vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp, f32, cl, cl, cl, cl, i32, vp],
    "pgk_moe_topk": [vp, vp, vp, i32, i32, i32, i32, vp],
    "pgk_moe_softmax_topk": [vp, i32, i32, i32, vp],
    "pgk_moe_compute_permutation": [vp, vp, vp, vp, vp, i32, i32, i32, vp],
    "pgk_moe_expand_expert_offsets": [vp, vp, i32, i32, vp],
    "pgk_moe_gather": [vp, vp, vp, i32, i32, i32, vp],
    "pgk_moe_scatter": [vp, vp, vp, vp, i32, i32, i32, vp],
    "pgk_moe_grouped_gemv_fp8": [vp, vp, vp, vp, vp, i32, i32, i32, vp],
    "pgk_moe_grouped_gemv_fp8_ep": [vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp],
    "pgk_moe_grouped_gemv_nvf4": [vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp],
    "pgk_moe_grouped_gemv_fp8_ex": [vp, i32, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp],
    "pgk_moe_grouped_gemv_nvf4_ex": [vp, i32, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp],
    "pgk_moe_route_t1": [vp, vp, vp, vp, i32, i32, vp],
    "pgk_moe_gateup_fp8_t1": [vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp],
    "pgk_moe_gateup_nvf4_t1": [vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp],
    "pgk_moe_down_combine_fp8_t1": [vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp],
    "pgk_gemv_fp8_allreduce_bf16": [vp, vp, vp, vp, vp, i32, i32, vp],
    "pgk_gemv_nvf4_allreduce_bf16": [vp, vp, vp, vp, vp, i32, i32, f32, vp],
    "pgk_gemv_nvf4_bf16_noalloc": [vp, vp, vp, vp, i32, i32, i32, f32, vp],
    "pgk_moe_combine_t1": [vp, vp, vp, vp, i32, i32, vp],
    "pgk_sample_token_bf16": [vp, i32, f32, i32, f32, f32, vp, vp, vp],
    "pgk_fill_normal_bf16": [vp, sz, C.c_uint32, f32, vp],
    "pgk_quantize_bf16_to_fp8_block": [vp, vp, vp, i32, i32, vp],
    "pgk_decode_advance": [vp, vp, i32, vp], "pgk_decode_advance_n": [vp, vp, i32, i32, vp],
    "pgk_rope_sdpa_decode_batch": [vp, vp, vp, f32, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp, vp, cl, f32, vp],
    "pgk_argmax_rows": [vp, vp, vp, i32, i32, i32, vp],
    "pgk_step_create": [vp, i32, vp, vp, vp, vp, i32, i32, C.POINTER(vp)], "pgk_step_run": [vp, vp],
    "pgk_step_error": [vp, C.POINTER(i32)], "pgk_step_info": [vp, C.POINTER(i32), C.POINTER(i32), C.POINTER(i32), C.POINTER(i32)],
    "pgk_step_destroy": [vp], "pgk_step_trace": [vp, vp, i32], "pgk_step_set_vocab_offset": [vp, i32],
    "pgk_nvf4_pack_rows": [vp, vp, vp, i32, i32, i32, i32, i32, i32, vp],
    "pgk_nccl_unique_id": [vp], "pgk_nccl_init": [vp, i32, i32],
    "pgk_nccl_allreduce_bf16": [vp, sz, vp], "pgk_nccl_allgather": [vp, vp, sz, vp],
    "pgk_nccl_destroy": [],
    "pgk_p2p_create": [i32, i32, vp], "pgk_p2p_open": [vp], "pgk_p2p_allreduce_add_bf16": [vp, vp, vp, sz, vp],
    "pgk_p2p_ll_allreduce_add_bf16": [vp, vp, vp, sz, vp], "pgk_p2p_ll_argmax_bf16": [vp, vp, i32, vp],
    "pgk_p2p_error": [vp], "pgk_p2p_destroy": [],
}
_RESTYPE = {"pgk_last_error": C.c_char_p, "pgk_launch_count": i64, "pgk_launch_count_reset": None,
            "pgk_sdpa_workspace_bytes": sz, "pgk_p2p_handle_bytes": sz, "pgk_p2p_ready": i32,
            "pgk_p2p_max_bytes": sz}
_EXTRA = {"pgk_last_error": [], "pgk_launch_count": [], "pgk_launch_count_reset": [],
          "pgk_sdpa_workspace_bytes": [i32, i32, i32, i32], "pgk_p2p_handle_bytes": [], "pgk_p2p_ready": [], "pgk_p2p_max_bytes": []}


def load() -> C.CDLL:
    """Load the shared library (once). Raises RuntimeError when it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"pygpukit_b200: native library {LIB_PATH} not found. Build it with "
            "`python -m pygpukit_b200.build` (needs nvcc); there is no CPU fallback on this path.")
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    for name, args in _SIGS.items():
        fn = getattr(lib, name)
        fn.argtypes = args
        fn.restype = i32
    for name, args in _EXTRA.items():
        fn = getattr(lib, name)
        fn.argtypes = args
        fn.restype = _RESTYPE[name]
    _lib = lib
    return lib


def declared_symbols() -> list[str]:
    return sorted(list(_SIGS) + list(_EXTRA))


def last_error() -> str:
    return load().pgk_last_error().decode("utf-8", "replace")


def check(rc: int) -> None:
    """Raise RuntimeError (the reference's exception type for native failures) on rc != 0."""
    if rc != 0:
        raise RuntimeError(last_error() or f"libpgk_b200 call failed with code {rc}")


class _Api:
    """lib.<fn>(...) with automatic error checking: `api.gemv_fp8_bf16(...)`."""

    def __getattr__(self, name: str):
        fn = getattr(load(), "pgk_" + name)
        if fn.restype is i32 and name != "version":
            def call(*a, _fn=fn):
                check(_fn(*a))
            call.__name__ = name
            setattr(self, name, call)
            return call
        setattr(self, name, fn)
        return fn


api = _Api()
