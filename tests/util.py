"""Shared synthetic-data helpers for the tests (seeded)."""
import numpy as np

from pygpukit_b200.core.dtypes import bf16_bits_to_f32, f32_to_bf16_bits
from pygpukit_b200.llm.quant import quantize_fp8_e4m3_block

bf16 = f32_to_bf16_bits
f32 = bf16_bits_to_f32


def rand_bf16(rng, shape, std=1.0):
    return bf16((rng.standard_normal(shape) * std).astype(np.float32))


def rand_fp8_weights(rng, N, K, std=0.02):
    w = (rng.standard_normal((N, K)) * std).astype(np.float32)
    return quantize_fp8_e4m3_block(w)


def rel_err(out_bits, ref_bits):
    """max|out-ref| / max|ref| on bf16 bit arrays (the reference tests' metric,
    tests/test_gemv_correctness.py:105-149)."""
    o, r = f32(np.asarray(out_bits)).astype(np.float64), f32(np.asarray(ref_bits)).astype(np.float64)
    return float(np.abs(o - r).max() / max(np.abs(r).max(), 1e-30))


def ulp_diff_bf16(a_bits, b_bits):
    """max distance in bf16 ulps between two bit arrays (sign-magnitude -> ordered ints)."""
    def key(x):
        x = np.asarray(x).astype(np.int32)
        return np.where(x & 0x8000, -(x & 0x7FFF), x & 0x7FFF)
    return int(np.abs(key(a_bits) - key(b_bits)).max())


def f32acc_mismatch(out_bits, ref_bits, tol=2e-3):
    """Number of elements that break the f32-accumulate bar for a BF16-rounded output: |out - ref| must be within
    `tol` of max|ref| (the reference tests' max-norm metric) unless the two values are adjacent BF16 numbers -- a
    different f32 summation order may flip the final rounding of an element, and one flip of a large element is
    already 2^-8 of that element (up to 4e-3 of max|ref|), which says nothing about the f32 accumulation."""
    o, r = f32(np.asarray(out_bits)).astype(np.float64), f32(np.asarray(ref_bits)).astype(np.float64)
    def key(x):
        x = np.asarray(x).astype(np.int32)
        return np.where(x & 0x8000, -(x & 0x7FFF), x & 0x7FFF)
    far = np.abs(o - r) > tol * max(np.abs(r).max(), 1e-30)
    adjacent = np.abs(key(out_bits) - key(ref_bits)) <= 1
    return int(np.count_nonzero(far & ~adjacent))
