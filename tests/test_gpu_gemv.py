"""GPU parity: weight-only-quantised GEMV family through the C ABI vs the CPU oracle and vs the
reference's own CUDA kernels (oracle/_ref/libref_cuda.so) on identical seeded inputs.
Tolerance (north_star): 2e-3 relative for f32-accumulate GEMV outputs (metric of the reference's
own tests: max|d|/max|ref|); dequantisation / quantisation bit-exact."""
import numpy as np
import pytest

import pygpukit_b200 as gk
from oracle import oracle as orc
from oracle import refcuda
from pygpukit_b200 import ops
from pygpukit_b200._lib import api
from tests.util import bf16, f32, f32acc_mismatch, rand_bf16, rand_fp8_weights, rel_err, ulp_diff_bf16

pytestmark = pytest.mark.gpu
TOL = 2e-3

# (K, N): reference LLM_CONFIGS (src/pygpukit/benchmark/gemv.py:11-19) scaled to finish on the CPU
# oracle in seconds + layer shapes of the configs + ragged cases
FP8_SHAPES = [(4096, 4096), (4096, 1024), (2048, 8192), (14336, 512), (12288, 256), (768, 2048), (2048, 768),
              (128, 128), (160, 96), (4100, 77), (64, 8)]


@pytest.mark.parametrize("K,N", FP8_SHAPES)
@pytest.mark.parametrize("M", [1, 2, 3, 4, 8])
def test_gemv_fp8_vs_oracle(K, N, M):
    if M > 1 and K * N > 4096 * 4096:
        pytest.skip("oracle time")
    rng = np.random.default_rng(1234 + K + N + M)
    codes, scale = rand_fp8_weights(rng, N, K)
    A = rand_bf16(rng, (M, K))
    ref = orc.gemv_fp8_bf16(A, codes, scale)
    gB, gS = gk.from_numpy(codes), gk.from_numpy(scale)
    if M == 1:
        out = ops.gemv_fp8_bf16(gk.from_numpy(A[0]), gB, gS).to_numpy()[None]
    else:
        out = ops.gemv_fp8_bf16_batched(gk.from_numpy(A), gB, gS).to_numpy()
    if M >= 5 and K % 128 == 0 and N % 128 == 0:
        # tcgen05 path (pgk_gemm_fp8_bf16_tc): exact BF16 weights x (hi + lo) BF16 split of activation*scale, f32
        # accumulation in TMEM.  The result is f32-accumulate class, but its ~2^-17 residual flips a BF16 output
        # rounding now and then, and ONE flip of a large element is already 4e-3..8e-3 of max|ref|: the bar is
        # <= 1 BF16 ulp on every significant element, few flips, and the BF16 bar on the max-norm error.
        big = np.abs(f32(ref)) > np.abs(f32(ref)).max() / 64
        assert ulp_diff_bf16(out[big], ref[big]) <= 1
        assert rel_err(out, ref) < 1e-2
        assert np.mean(out != ref) < 0.02
        return
    assert rel_err(out, ref) < TOL
    assert ulp_diff_bf16(out, ref) <= 64          # near-zero outputs may differ by many ulps; most by <= 1
    assert np.mean(out != ref) < 0.05             # f32 reordering flips few bf16 roundings


@pytest.mark.parametrize("K,N", [(4096, 4096), (1024, 256), (12288, 1024), (128, 128), (2048, 768), (256, 6144)])
@pytest.mark.parametrize("M", [1, 5, 8, 13, 16, 17, 32, 50, 64])
def test_gemm_fp8_tc_vs_oracle(K, N, M):
    """Decode batch (1..)5..64 on tcgen05 (pgk_gemm_fp8_bf16_tc, reached through gemv_fp8_bf16_batched) vs the
    oracle's FP8 GEMV (f32 weights x bf16 activations).  The block scale is folded into the activations and
    rounded to BF16 (the reference's M >= 16 path rounds weight*scale to BF16 instead, w8a16_gemm.cu:104-339):
    BF16-operand tolerance 1e-2 (north_star); and the slab GEMV (PGK_TC=0 semantics, M rows in slabs of 4)
    must agree with it to the same bar."""
    if K * N * M > 4096 * 4096 * 16 * 2:
        pytest.skip("oracle time")
    from pygpukit_b200._lib import api
    rng = np.random.default_rng(77 + K + N + M)
    codes, scale = rand_fp8_weights(rng, N, K)
    A = rand_bf16(rng, (M, K))
    ref = orc.gemv_fp8_bf16(A, codes, scale)
    gA, gB, gS = gk.from_numpy(A), gk.from_numpy(codes), gk.from_numpy(scale)
    out = gk.zeros((M, N), gk.bfloat16)
    api.gemm_fp8_bf16_tc(gA.device_ptr, gB.device_ptr, gS.device_ptr, out.device_ptr, M, K, N, None, None, None)
    got = out.to_numpy()
    assert rel_err(got, ref) < 1e-2
    # typical error is far below the bar: the mean relative deviation over the outputs stays under 2e-3
    d = np.abs(f32(got).astype(np.float64) - f32(ref).astype(np.float64))
    assert d.mean() / np.abs(f32(ref).astype(np.float64)).mean() < 4e-3
    # the public batched entry dispatches to the same kernel from the measured crossover (M >= 5) up; below it
    # the CUDA-core GEMV answers, to the same bar
    out2 = ops.gemv_fp8_bf16_batched(gA, gB, gS).to_numpy()
    if M >= 5:
        assert np.array_equal(out2, got)
    else:
        assert rel_err(out2, ref) < 1e-2 and rel_err(out2, got) < 1e-2
    # epilogues: bias + residual with the reference's rounding points
    bias, res = rand_bf16(rng, (N,), 0.1), rand_bf16(rng, (M, N), 0.5)
    gb, gr = gk.from_numpy(bias), gk.from_numpy(res)
    out3 = gk.zeros((M, N), gk.bfloat16)
    api.gemm_fp8_bf16_tc(gA.device_ptr, gB.device_ptr, gS.device_ptr, out3.device_ptr, M, K, N, gb.device_ptr,
                         gr.device_ptr, None)
    exp = bf16(f32(bf16(f32(got) + f32(bias)[None, :])) + f32(res))
    assert ulp_diff_bf16(out3.to_numpy(), exp) <= 1


def test_gemm_fp8_tc_deterministic_and_linear():
    """Full-size property checks (no oracle): bit-identical across launches (split-K folds in split order), and
    rows are independent: a batch equals its two halves computed separately, bit for bit."""
    from pygpukit_b200._lib import api
    rng = np.random.default_rng(5)
    M, K, N = 64, 4096, 12288
    codes, scale = rand_fp8_weights(rng, N, K)
    A = rand_bf16(rng, (M, K))
    gA, gB, gS = gk.from_numpy(A), gk.from_numpy(codes), gk.from_numpy(scale)
    o1, o2 = gk.zeros((M, N), gk.bfloat16), gk.zeros((M, N), gk.bfloat16)
    for o in (o1, o2):
        api.gemm_fp8_bf16_tc(gA.device_ptr, gB.device_ptr, gS.device_ptr, o.device_ptr, M, K, N, None, None, None)
    assert np.array_equal(o1.to_numpy(), o2.to_numpy())
    h = gk.zeros((32, N), gk.bfloat16)
    gA2 = gk.from_numpy(np.ascontiguousarray(A[32:]))
    api.gemm_fp8_bf16_tc(gA2.device_ptr, gB.device_ptr, gS.device_ptr, h.device_ptr, 32, K, N, None, None, None)
    assert np.array_equal(h.to_numpy(), o1.to_numpy()[32:])


@pytest.mark.parametrize("K,N", [(4096, 4096), (4096, 1024), (2048, 8192), (128, 128)])
def test_oracle_vs_reference_cuda_gemv_fp8(K, N):
    """Pins the oracle: the reference's own kernel on the same inputs must agree BIT FOR BIT with
    the C restatement (same accumulation order)."""
    if not refcuda.available():
        pytest.skip("libref_cuda.so not built")
    rng = np.random.default_rng(99 + K + N)
    codes, scale = rand_fp8_weights(rng, N, K)
    A = rand_bf16(rng, (4, K))
    gA, gB, gS = gk.from_numpy(A), gk.from_numpy(codes), gk.from_numpy(scale)
    out1 = gk.empty((N,), gk.bfloat16)
    refcuda.gemv_fp8(gA, gB, gS, out1, K, N)
    outM = gk.empty((4, N), gk.bfloat16)
    refcuda.gemv_fp8_batched(gA, gB, gS, outM, K, N, 4)
    ref = orc.gemv_fp8_bf16(A, codes, scale)
    assert np.array_equal(out1.to_numpy(), ref[0])
    assert np.array_equal(outM.to_numpy(), ref)
    ours = ops.gemv_fp8_bf16_batched(gA, gB, gS).to_numpy()
    assert rel_err(ours, outM.to_numpy()) < TOL


def test_gemv_fp8_large_k_reference_cannot_launch():
    """K=28672 (Llama-3-70B down_proj): the reference kernel needs 56 KB of dynamic smem without
    opting in (README.md:451,462); ours runs."""
    K, N = 28672, 256
    rng = np.random.default_rng(5)
    codes, scale = rand_fp8_weights(rng, N, K)
    A = rand_bf16(rng, (1, K))
    out = ops.gemv_fp8_bf16(gk.from_numpy(A[0]), gk.from_numpy(codes), gk.from_numpy(scale)).to_numpy()
    assert rel_err(out, orc.gemv_fp8_bf16(A[0], codes, scale)) < TOL
    if refcuda.available():
        with pytest.raises(RuntimeError):
            refcuda.gemv_fp8(gk.from_numpy(A[0]), gk.from_numpy(codes), gk.from_numpy(scale),
                             gk.empty((N,), gk.bfloat16), K, N)


def test_gemv_fp8_epilogues():
    from pygpukit_b200._lib import api
    rng = np.random.default_rng(11)
    K, N, M = 1024, 512, 2
    codes, scale = rand_fp8_weights(rng, N, K)
    codes2, scale2 = rand_fp8_weights(rng, N, K)
    A = rand_bf16(rng, (M, K))
    res = rand_bf16(rng, (M, N), 0.5)
    bias = rand_bf16(rng, (N,), 0.1)
    gA, gB, gS = gk.from_numpy(A), gk.from_numpy(codes), gk.from_numpy(scale)
    gB2, gS2 = gk.from_numpy(codes2), gk.from_numpy(scale2)
    out = gk.empty((M, N), gk.bfloat16)
    y = orc.gemv_fp8_bf16(A, codes, scale)
    # bias then residual, reference rounding points
    gbias, gres = gk.from_numpy(bias), gk.from_numpy(res)      # keep the owners alive across the launch
    api.gemv_fp8_bf16_ex(gA.device_ptr, gB.device_ptr, gS.device_ptr, out.device_ptr, M, K, N,
                         gbias.device_ptr, gres.device_ptr, None)
    exp = orc.add_inplace_bf16(orc.bias_add_bf16(y, bias), res)
    assert rel_err(out.to_numpy(), exp) < TOL
    # gate/up + silu*mul
    api.gemv_fp8_swiglu_bf16(gA.device_ptr, gB.device_ptr, gS.device_ptr, gB2.device_ptr, gS2.device_ptr,
                             out.device_ptr, M, K, N, None)
    u = orc.gemv_fp8_bf16(A, codes2, scale2)
    exp = orc.mul_inplace_bf16(orc.silu_bf16(y), u)
    assert rel_err(out.to_numpy(), exp) < 4e-3


def test_gemv_fp8_linearity_full_size():
    """Full benchmark size (no oracle): GEMV(a1 + a2) ~ GEMV(a1) + GEMV(a2), and a unit vector
    returns one dequantised column exactly."""
    K, N = 4096, 14336
    rng = np.random.default_rng(21)
    codes = rng.integers(0, 0x7E, size=(N, K), dtype=np.uint8) | (rng.integers(0, 2, size=(N, K), dtype=np.uint8) << 7)
    scale = bf16(rng.uniform(1e-3, 1e-2, (N // 128, K // 128)).astype(np.float32))
    gB, gS = gk.from_numpy(codes), gk.from_numpy(scale)
    e = np.zeros(K, np.float32)
    e[1234] = 1.0
    col = ops.gemv_fp8_bf16(gk.from_numpy(bf16(e)), gB, gS).to_numpy()
    from pygpukit_b200.llm.quant import fp8_e4m3_table
    exp = bf16(fp8_e4m3_table()[codes[:, 1234]] * f32(scale)[:, 1234 // 128].repeat(128))
    assert np.array_equal(f32(col), f32(exp))        # value equality: a -0.0 weight gives +0.0 (0 + 1*-0)
    a1, a2 = rand_bf16(rng, (K,)), rand_bf16(rng, (K,))
    s = bf16(f32(a1) + f32(a2))
    y1 = f32(ops.gemv_fp8_bf16(gk.from_numpy(a1), gB, gS).to_numpy())
    y2 = f32(ops.gemv_fp8_bf16(gk.from_numpy(a2), gB, gS).to_numpy())
    ys = f32(ops.gemv_fp8_bf16(gk.from_numpy(s), gB, gS).to_numpy())
    assert np.abs(ys - (y1 + y2)).max() / np.abs(ys).max() < 2e-2


@pytest.mark.parametrize("K,N", [(4096, 1024), (1024, 4096), (2048, 50000), (4096, 151936 // 8), (100, 37)])
@pytest.mark.parametrize("M", [1, 2, 4])
def test_gemv_bf16_vs_oracle(K, N, M):
    rng = np.random.default_rng(77 + K + N + M)
    B = rand_bf16(rng, (N, K), 0.02)
    A = rand_bf16(rng, (M, K))
    ref = orc.gemv_bf16(A, B)
    if M == 1:
        out = ops.gemv_bf16(gk.from_numpy(A[0]), gk.from_numpy(B)).to_numpy()[None]
    else:
        out = ops.gemv_bf16_batched(gk.from_numpy(A), gk.from_numpy(B)).to_numpy()
    assert rel_err(out, ref) < TOL
    if refcuda.available() and M == 1 and K % 8 == 0:
        o2 = gk.empty((N,), gk.bfloat16)
        refcuda.gemv_bf16(gk.from_numpy(A[0]), gk.from_numpy(B), o2, K, N)
        assert np.array_equal(o2.to_numpy(), ref[0])       # oracle pinned on the reference kernel


def test_gemv_bf16_argmax_first_max_wins():
    from pygpukit_b200._lib import api
    rng = np.random.default_rng(5)
    K, N = 1024, 40000
    B = rand_bf16(rng, (N, K), 0.02)
    B[31000] = B[12345]                       # exact tie between rows 12345 and 31000
    A = f32(B[12345]) * 40.0                   # makes that row the maximum
    A = bf16(A.astype(np.float32))
    gA, gB = gk.from_numpy(A), gk.from_numpy(B)
    logits = gk.empty((N,), gk.bfloat16)
    tok = gk.zeros((1,), gk.int32)
    for _ in range(3):                         # scratch must re-arm between launches
        api.gemv_bf16_argmax(gA.device_ptr, gB.device_ptr, logits.device_ptr, tok.device_ptr, K, N, None)
        lg = logits.to_numpy()
        assert int(tok.to_numpy()[0]) == orc.argmax_bf16(lg) == 12345
    assert rel_err(lg, orc.gemv_bf16(A, B)) < TOL
    assert ops.sample_greedy(logits) == 12345


NVF4_SHAPES = [(8192, 1024), (2048, 768), (768, 2048), (4096, 512), (1024, 28672 // 4), (64, 48), (96, 40), (34, 7)]


def _nvf4_weights(rng, K, N):
    W = rand_bf16(rng, (K, N), 0.02)
    return orc.quantize_bf16_to_nvf4(W) if K % 2 == 0 else None, W


@pytest.mark.parametrize("K,N", NVF4_SHAPES)
@pytest.mark.parametrize("M", [1, 2, 4])
def test_gemv_nvf4_vs_oracle(K, N, M):
    rng = np.random.default_rng(4321 + K + N + M)
    (data, scale), _ = _nvf4_weights(rng, K, N)
    A = rand_bf16(rng, (M, K))
    ref = orc.gemv_nvf4_bf16(A, data, scale, 0.75)
    gD, gS = gk.from_numpy(data), gk.from_numpy(scale)
    if M == 1:
        out = ops.gemv_nvf4_bf16(gk.from_numpy(A[0]), gD, gS, alpha=0.75).to_numpy()[None]
    else:
        out = ops.gemv_nvf4_bf16_batched(gk.from_numpy(A), gD, gS, alpha=0.75).to_numpy()
    assert rel_err(out, ref) < TOL
    if refcuda.available() and M == 1:
        o2 = gk.empty((N,), gk.bfloat16)
        refcuda.gemv_nvf4(gk.from_numpy(A[0]), gD, gS, o2, K, N, 0.75)
        assert np.array_equal(o2.to_numpy(), ref[0])       # oracle pinned on the reference kernel


@pytest.mark.parametrize("K,N", [(8192, 1024), (2048, 768), (768, 2048), (1024, 7168), (4096, 512), (96, 40)])
@pytest.mark.parametrize("M", [5, 8, 16, 32, 64])
def test_gemv_nvf4_batched_vs_oracle(K, N, M):
    """NVF4 decode batch 5..64 (BASELINE.json configs[4]: batch 1-32) vs the oracle's serial-K restatement of
    gemv_nvf4_bf16_kernel (nvf4_kernels.cu:79-200), row by row.  128-aligned shapes go to the tcgen05 GEMM
    (E2M1 x UE4M3 is exact in BF16, the activations are BF16: exact products, f32 accumulation in TMEM -- the
    f32-accumulate bar 2e-3 applies); the others stay on the CUDA-core GEMV in slabs of 4 rows."""
    rng = np.random.default_rng(977 + K + N + M)
    (data, scale), _ = _nvf4_weights(rng, K, N)
    A = rand_bf16(rng, (M, K))
    ref = orc.gemv_nvf4_bf16(A, data, scale, 1.0)
    out = ops.gemv_nvf4_bf16_batched(gk.from_numpy(A), gk.from_numpy(data), gk.from_numpy(scale)).to_numpy()
    assert f32acc_mismatch(out, ref, TOL) == 0    # 2e-3 of max|ref|, or a flipped final BF16 rounding (adjacent values)
    assert rel_err(out, ref) < 1e-2
    assert np.mean(out != ref) < 0.05             # f32 reordering flips few bf16 roundings
    assert ulp_diff_bf16(out[np.abs(f32(ref)) > np.abs(f32(ref)).max() / 64], ref[np.abs(f32(ref)) > np.abs(f32(ref)).max() / 64]) <= 1


@pytest.mark.parametrize("M", [1, 2, 4, 8, 33])
@pytest.mark.parametrize("K,N", [(1024, 256), (4096, 512), (160, 96)])
def test_gemv_fp8_codes_7f_ff_follow_the_reference_table(K, N, M):
    """E4M3 codes 0x7F / 0xFF are +-480 in every CUDA table of the reference (fp8.cuh:89,121) -- the hardware
    conversion the streaming kernels use would give NaN.  A weight tensor that holds them must produce the
    reference's values from the SAME entry point (pgk_gemv_fp8_bf16*), for every batch size: the library scans each
    weight tensor once and serves such tensors with the exact table-semantics kernel."""
    rng = np.random.default_rng(7 + K + N + M)
    codes, scale = rand_fp8_weights(rng, N, K)
    hot = rng.integers(0, codes.size, 64)
    codes.reshape(-1)[hot[:32]] = 0x7F
    codes.reshape(-1)[hot[32:]] = 0xFF
    A = rand_bf16(rng, (M, K))
    ref = orc.gemv_fp8_bf16(A, codes, scale)
    assert np.isfinite(f32(ref)).all()
    gB, gS = gk.from_numpy(codes), gk.from_numpy(scale)
    if M == 1:
        out = ops.gemv_fp8_bf16(gk.from_numpy(A[0]), gB, gS).to_numpy()[None]
    else:
        out = ops.gemv_fp8_bf16_batched(gk.from_numpy(A), gB, gS).to_numpy()
    assert np.isfinite(f32(out)).all()
    assert rel_err(out, ref) < TOL
    # the same shapes without those codes still take the streaming path and agree as well
    codes2 = codes.copy()
    codes2.reshape(-1)[hot] = 0x30
    out2 = ops.gemv_fp8_bf16_batched(gk.from_numpy(A), gk.from_numpy(codes2), gS).to_numpy()
    assert rel_err(out2, orc.gemv_fp8_bf16(A, codes2, scale)) < (1e-2 if M >= 5 else TOL)
    if refcuda.available() and M == 1 and K % 128 == 0:
        o2 = gk.empty((N,), gk.bfloat16)
        refcuda.gemv_fp8(gk.from_numpy(A[0]), gB, gS, o2, K, N)
        assert np.array_equal(o2.to_numpy(), ref[0])       # the oracle's +-480 is the reference kernel's


@pytest.mark.parametrize("K,N", [(8192, 1024), (4096, 4096), (64, 48), (96, 40), (2048, 768)])
def test_quantize_nvf4_bit_exact(K, N):
    rng = np.random.default_rng(8 + K + N)
    W = rand_bf16(rng, (K, N), 0.02)
    W[:32, 0] = 0                                            # an all-zero block (scale = 1.0 branch)
    gW = gk.from_numpy(W)
    d, s = gk.empty((K // 2, N), gk.uint8), gk.empty(((K + 31) // 32, N), gk.uint8)
    ops.quantize_bf16_to_nvf4(gW, d, s)
    od, os_ = orc.quantize_bf16_to_nvf4(W)
    if refcuda.available():
        rd, rs = gk.empty((K // 2, N), gk.uint8), gk.empty(((K + 31) // 32, N), gk.uint8)
        refcuda.quantize_nvf4(gW, rd, rs, K, N)
        assert np.array_equal(s.to_numpy(), rs.to_numpy())
        assert np.array_equal(d.to_numpy(), rd.to_numpy())   # bit-exact vs the reference kernel
        # the IEEE oracle may differ from the fast-math reference only on threshold ties
        assert np.array_equal(os_, rs.to_numpy())
        assert np.mean(od != rd.to_numpy()) < 1e-4
    else:
        assert np.array_equal(s.to_numpy(), os_)
        assert np.mean(d.to_numpy() != od) < 1e-4


def test_gemv_nvf4_swiglu():
    from pygpukit_b200._lib import api
    rng = np.random.default_rng(31)
    K, I, M = 1024, 768, 2
    Wg, Wu = rand_bf16(rng, (K, I), 0.03), rand_bf16(rng, (K, I), 0.03)
    (dg, sg), (du, su) = orc.quantize_bf16_to_nvf4(Wg), orc.quantize_bf16_to_nvf4(Wu)

    def interleave(g, u):                                   # groups of 8 columns: g0..7,u0..7,g8..15,...
        r = g.shape[0]
        return np.stack([g.reshape(r, I // 8, 8), u.reshape(r, I // 8, 8)], axis=2).reshape(r, 2 * I)
    d, s = interleave(dg, du), interleave(sg, su)
    A = rand_bf16(rng, (M, K))
    out = gk.empty((M, I), gk.bfloat16)
    gA, gd, gs = gk.from_numpy(A), gk.from_numpy(d), gk.from_numpy(s)
    api.gemv_nvf4_swiglu_bf16(gA.device_ptr, gd.device_ptr, gs.device_ptr, out.device_ptr, M, K, 2 * I, None)
    g = orc.gemv_nvf4_bf16(A, dg, sg)
    u = orc.gemv_nvf4_bf16(A, du, su)
    exp = orc.mul_inplace_bf16(orc.silu_bf16(g), u)
    assert rel_err(out.to_numpy(), exp) < 4e-3


@pytest.mark.parametrize("M,K,N", [(1, 256, 384), (4, 512, 256), (16, 256, 256), (33, 384, 200)])
def test_w8a16_gemm_kn_layout(M, K, N):
    rng = np.random.default_rng(M + K + N)
    codes_nk, scale_nk = rand_fp8_weights(rng, N, K)
    B_kn = np.ascontiguousarray(codes_nk.T)
    S_kn = np.ascontiguousarray(scale_nk.T)
    A = rand_bf16(rng, (M, K))
    out = ops.w8a16_gemm_sm120(gk.from_numpy(A), gk.from_numpy(B_kn), gk.from_numpy(S_kn)).to_numpy()
    ref = orc.w8a16_gemm(A, B_kn, S_kn)
    assert rel_err(out, ref) < (1e-2 if M >= 16 else TOL)   # M>=16: the reference rounds weights to bf16 first
    if refcuda.available() and K % 128 == 0 and N % 128 == 0:
        o2 = gk.empty((M, N), gk.bfloat16)
        refcuda.w8a16_gemm(gk.from_numpy(A), gk.from_numpy(B_kn), gk.from_numpy(S_kn), o2, M, N, K)
        assert rel_err(o2.to_numpy(), ref) < (1e-2 if M >= 16 else 1e-6)


def test_validation_errors_match_reference():
    a = gk.zeros((64,), gk.bfloat16)
    b = gk.zeros((32, 64), gk.uint8)
    s = gk.zeros((1, 1), gk.bfloat16)
    with pytest.raises(ValueError, match="requires 1D input vector"):
        ops.gemv_fp8_bf16(gk.zeros((1, 64), gk.bfloat16), b, s)
    with pytest.raises(ValueError, match="requires uint8"):
        ops.gemv_fp8_bf16(a, gk.zeros((32, 64), gk.bfloat16), s)
    with pytest.raises(ValueError, match="dimension mismatch"):
        ops.gemv_fp8_bf16(gk.zeros((32,), gk.bfloat16), b, s)
    with pytest.raises(ValueError, match="out shape"):
        ops.gemv_fp8_bf16(a, b, s, out=gk.zeros((31,), gk.bfloat16))
    out = gk.zeros((32,), gk.bfloat16)
    assert ops.gemv_fp8_bf16(a, b, s, out=out) is out


@pytest.mark.parametrize("K,N", [(1024, 256), (8192, 1024), (3584, 512), (128, 128), (2048, 768)])
@pytest.mark.parametrize("M", [1, 5, 16, 17, 33, 64])
def test_gemm_nvf4_tc_vs_oracle(K, N, M):
    """The tcgen05 NVF4 GEMM called directly (pgk_gemm_nvf4_bf16_tc): nibble x scale is exact in BF16, activations are
    BF16, products exact, F32 accumulation in TMEM -> the f32-accumulate bar against the oracle's serial-K restatement of
    gemv_nvf4_bf16_kernel (nvf4_kernels.cu:79-200), with alpha and the residual epilogue; deterministic across calls."""
    rng = np.random.default_rng(4242 + K + N + M)
    (data, scale), _ = _nvf4_weights(rng, K, N)
    A = rand_bf16(rng, (M, K))
    res = rand_bf16(rng, (M, N))
    ref = orc.add_inplace_bf16(orc.gemv_nvf4_bf16(A, data, scale, 0.5), res)
    gA, gD, gS, gR = gk.from_numpy(A), gk.from_numpy(data), gk.from_numpy(scale), gk.from_numpy(res)
    outs = []
    for _ in range(2):
        out = gk.zeros((M, N), gk.bfloat16)
        api.gemm_nvf4_bf16_tc(gA.device_ptr, gD.device_ptr, gS.device_ptr, out.device_ptr, M, K, N, 0.5, None, gR.device_ptr, None)
        outs.append(out.to_numpy())
    assert np.array_equal(outs[0], outs[1])
    assert f32acc_mismatch(outs[0], ref, TOL) == 0
    assert rel_err(outs[0], ref) < 1e-2


def test_swiglu_interleaved8_matches_oracle():
    rng = np.random.default_rng(5)
    M, I = 7, 1024
    gu = rand_bf16(rng, (M, 2 * I))
    out, ggu = gk.zeros((M, I), gk.bfloat16), gk.from_numpy(gu)
    api.swiglu_interleaved8_bf16(ggu.device_ptr, out.device_ptr, M, I, None)
    v = gu.reshape(M, I // 8, 2, 8)
    g, u = v[:, :, 0, :].reshape(M, I), v[:, :, 1, :].reshape(M, I)
    ref = orc.mul_inplace_bf16(orc.silu_bf16(np.ascontiguousarray(g)), np.ascontiguousarray(u))
    assert ulp_diff_bf16(out.to_numpy(), ref) <= 1          # __expf on both sides
    assert np.mean(out.to_numpy() != ref) < 0.01


@pytest.mark.parametrize("K,N", [(4096, 4096), (1024, 256), (12288, 1024), (128, 128), (2048, 768), (256, 6144)])
@pytest.mark.parametrize("M", [1, 4, 5, 8, 13, 16, 17, 32, 50, 64])
def test_gemm_fp8_f8mma_vs_oracle(K, N, M):
    """pgk_gemm_fp8_bf16_f8mma: raw E4M3 tiles by tensor-map TMA into tcgen05.mma.kind::f8f6f4, activations as two exact
    E4M3 terms, block scales applied when the partial products are promoted out of TMEM -- against the oracle's FP8 GEMV
    (fp8_opt.cuh:133-212 restated).  FP32-accumulate class: <= 1 BF16 ulp on every significant element, the f32-accumulate
    bar on the rest, few flipped roundings; deterministic; bias / residual epilogues."""
    if K * N * M > 4096 * 4096 * 16 * 2:
        pytest.skip("oracle time")
    rng = np.random.default_rng(99 + K + N + M)
    codes, scale = rand_fp8_weights(rng, N, K)
    A = rand_bf16(rng, (M, K))
    ref = orc.gemv_fp8_bf16(A, codes, scale)
    gA, gB, gS = gk.from_numpy(A), gk.from_numpy(codes), gk.from_numpy(scale)
    outs = []
    for _ in range(2):
        out = gk.zeros((M, N), gk.bfloat16)
        api.gemm_fp8_bf16_f8mma(gA.device_ptr, gB.device_ptr, gS.device_ptr, out.device_ptr, M, K, N, None, None, None)
        outs.append(out.to_numpy())
    assert np.array_equal(outs[0], outs[1])
    out = outs[0]
    assert f32acc_mismatch(out, ref, TOL) == 0
    big = np.abs(f32(ref)) > np.abs(f32(ref)).max() / 64
    assert ulp_diff_bf16(out[big], ref[big]) <= 1
    assert np.mean(out != ref) < 0.03
    # residual epilogue: bf16(acc) + residual with one rounding
    res = rand_bf16(rng, (M, N))
    out2, gR = gk.zeros((M, N), gk.bfloat16), gk.from_numpy(res)
    api.gemm_fp8_bf16_f8mma(gA.device_ptr, gB.device_ptr, gS.device_ptr, out2.device_ptr, M, K, N, None, gR.device_ptr, None)
    assert np.array_equal(out2.to_numpy(), orc.add_inplace_bf16(out, res))


def test_gemm_fp8_f8mma_wide_dynamic_range():
    """Activation rows whose elements span 20 binades (the two-term split keeps elements within 2^-10 of the row maximum
    exact and bounds the absolute error of the rest by 2^-18 of the maximum)."""
    rng = np.random.default_rng(7)
    K, N, M = 2048, 512, 6
    codes, scale = rand_fp8_weights(rng, N, K)
    a = rng.standard_normal((M, K)) * np.exp2(rng.integers(-18, 2, (M, K)))
    A = bf16(a.astype(np.float32))
    ref = orc.gemv_fp8_bf16(A, codes, scale)
    out, gA, gB, gS = gk.zeros((M, N), gk.bfloat16), gk.from_numpy(A), gk.from_numpy(codes), gk.from_numpy(scale)
    api.gemm_fp8_bf16_f8mma(gA.device_ptr, gB.device_ptr, gS.device_ptr, out.device_ptr, M, K, N, None, None, None)
    assert f32acc_mismatch(out.to_numpy(), ref, TOL) == 0
