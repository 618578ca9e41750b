"""CPU: the C-ABI library loads and exports every symbol include/pgk_b200.h declares
(no compute calls without a GPU), and the Python boundary validates like the reference."""
import os
import re

import pytest

import pygpukit_b200 as gk
from pygpukit_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    hdr = open(os.path.join(ROOT, "include", "pgk_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(pgk_[a-z0-9_]+)\s*\(", hdr)))


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    names = _declared()
    assert len(names) >= 70
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing


def test_python_binding_covers_header():
    assert sorted(_lib.declared_symbols()) == _declared()


def test_version_and_error_string():
    lib = _lib.load()
    assert lib.pgk_version() == 100
    assert isinstance(_lib.last_error(), str)


def test_missing_library_fails_loudly(monkeypatch):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libpgk_b200.so")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        _lib.load()


def test_reference_api_surface():
    # names pinned by the reference (ops/matmul/__init__.py:92-155, tests/test_llm_unified.py:198-239)
    for n in ("gemv_bf16", "gemv_bf16_bf16_sm120", "gemv_fp8_bf16", "gemv_fp8_bf16_sm120",
              "gemv_fp8_bf16_batched", "gemv_fp8_bf16_batched_sm120", "w8a16_gemm_sm120", "gemv_nvf4_bf16",
              "gemv_nvf4_bf16_sm120", "quantize_bf16_to_nvf4", "nvf4_get_sizes", "grouped_gemm_fp8_bf16", "matmul",
              "rmsnorm", "layernorm", "rope_inplace", "rope_inplace_f32table", "sdpa_causal",
              "sdpa_causal_fixed_cache", "sdpa_causal_fixed_cache_ptr", "silu", "gelu", "bias_add_inplace",
              "embedding_lookup", "embedding_lookup_ptr", "kv_cache_update_gqa", "kv_cache_update_gqa_ptr",
              "kv_cache_prefill_gqa", "add_inplace", "mul_inplace", "copy_to", "transpose_3d_021", "reshape_copy",
              "sample_greedy", "rmsnorm_residual", "swiglu"):
        assert callable(getattr(gk.ops, n)), n
    assert gk.ops.nvf4_get_sizes(64, 48) == (32 * 48, 2 * 48)
    for n in ("GPUArray", "from_numpy", "zeros", "empty", "ones", "bfloat16", "float16", "float32", "uint8", "int32"):
        assert hasattr(gk, n)
    assert gk.bfloat16.to_numpy_dtype().name == "uint16"
