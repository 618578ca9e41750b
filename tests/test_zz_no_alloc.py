"""pgk_gemv_nvf4_bf16_noalloc -- the allocation-free variant of pgk_gemv_nvf4_bf16 (static device scratch, no cudaMalloc;
no counterpart in the reference).  Collected last on purpose (file name):
a failure here must not hide the tests of the path itself.

The new entry point is called through the C ABI BEFORE the existing one, so that it cannot lean on the workspace the
existing entry point allocates, on two calls in a row that share the static scratch (different shapes: the second call
sees the first call's partial sums and must not be disturbed by them), and compared with pgk_gemv_nvf4_bf16 at the
f32-accumulate bar of tests/test_gpu_gemv.py; the oracle is checked as well."""
import numpy as np
import pytest

import pygpukit_b200 as gk
from oracle import oracle as orc
from pygpukit_b200._lib import api, last_error, load
from tests.util import f32acc_mismatch, rand_bf16, rel_err

pytestmark = pytest.mark.gpu

# (M, K, N): split-K over many CTAs / a ragged last column group / unsplit, several rows
SHAPES = [(1, 8192, 1024), (1, 2048, 1552), (3, 1024, 4096)]


def _case(seed, M, K, N):
    rng = np.random.default_rng(seed)
    W = rand_bf16(rng, (K, N), 0.03)
    data, scale = orc.quantize_bf16_to_nvf4(W)
    A = rand_bf16(rng, (M, K))
    return A, data, scale


def _noalloc(A, data, scale, alpha=1.0):
    M, K = A.shape
    N = data.shape[1]
    gA, gD, gS = gk.from_numpy(A), gk.from_numpy(data), gk.from_numpy(scale)
    out = gk.empty((M, N), gk.bfloat16)
    api.gemv_nvf4_bf16_noalloc(gA.device_ptr, gD.device_ptr, gS.device_ptr, out.device_ptr, M, K, N, alpha, None)
    return out.to_numpy()


def _existing(A, data, scale, alpha=1.0):
    M, K = A.shape
    N = data.shape[1]
    gA, gD, gS = gk.from_numpy(A), gk.from_numpy(data), gk.from_numpy(scale)
    out = gk.empty((M, N), gk.bfloat16)
    api.gemv_nvf4_bf16(gA.device_ptr, gD.device_ptr, gS.device_ptr, out.device_ptr, M, K, N, alpha, None)
    return out.to_numpy()


def test_noalloc_matches_existing_entry_point_on_consecutive_calls():
    cases = [_case(900 + i, *shp) for i, shp in enumerate(SHAPES)]
    # two rounds over the shapes, back to back on the shared scratch, before the existing entry point has run at all
    got = [[_noalloc(*c) for c in cases] for _ in range(2)]
    for (A, data, scale), first, second in zip(cases, got[0], got[1]):
        assert np.array_equal(first, second)                       # same inputs, same bits, whatever the scratch held
        ref = _existing(A, data, scale)
        assert f32acc_mismatch(first, ref) == 0                    # 2e-3 of max|ref|, or adjacent BF16 values
        assert rel_err(first, orc.gemv_nvf4_bf16(A, data, scale)) < 1e-2
    # and once more after the existing entry point has allocated and used its own workspace
    A, data, scale = cases[0]
    assert np.array_equal(_noalloc(A, data, scale), got[0][0])
    assert f32acc_mismatch(_noalloc(A, data, scale, alpha=0.5), _existing(A, data, scale, alpha=0.5)) == 0


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    lib = load()
    buf = gk.zeros((1, 64), gk.bfloat16)            # never touched: the call must fail before any launch
    p = buf.device_ptr
    assert lib.pgk_gemv_nvf4_bf16_noalloc(p, p, p, p, 1, 64, 524288 + 16, 1.0, None) != 0
    assert "exceeds the static scratch" in last_error()
    assert lib.pgk_gemv_nvf4_bf16_noalloc(p, p, p, p, 1, 48, 64, 1.0, None) != 0      # K % 32
    assert lib.pgk_gemv_nvf4_bf16_noalloc(p, p, p, p, 1, 131072, 64, 1.0, None) != 0  # activation row beyond shared memory
    assert "noalloc" in last_error()
