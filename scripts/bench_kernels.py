"""Per-kernel achieved HBM GB/s for the decode GEMV family (run on the GPU box).

Each kernel cycles through a ring of distinct weight buffers larger than L2 (126 MB) so every
launch streams from HBM; timing = CUDA events around `iters` back-to-back launches."""
import ctypes as C
import json
import sys

import numpy as np

import pygpukit_b200 as gk
from pygpukit_b200._lib import api

PEAK = 6539.2
try:
    PEAK = json.load(open("MEASURED_PEAKS.json"))["hbm_gbs"]
except Exception:
    pass


def rand_u8(shape, rng):
    a = rng.integers(0, 0x7E, size=shape, dtype=np.uint8)
    return a


def ring(nbytes_one, make):
    n = max(2, int(300e6 // nbytes_one) + 1)
    return [make() for _ in range(min(n, 64))]


def timeit(fn, iters=50, warm=5):
    for i in range(warm):
        fn(i)
    api.device_sync()
    e0, e1 = gk.CudaEvent(), gk.CudaEvent()
    e0.record()
    for i in range(iters):
        fn(i)
    e1.record()
    e1.synchronize()
    return gk.event_elapsed_ms(e0, e1) / iters * 1e3  # us


def bench_fp8(K, N, M, rng, swiglu=False):
    A = gk.from_numpy(rng.integers(0x3C00, 0x4000, size=(M, K)).astype(np.uint16))
    mk = lambda: (gk.from_numpy(rand_u8((N, K), rng)), gk.from_numpy(np.full((N // 128 + 1, K // 128 + 1), 0x3C00, np.uint16)))
    nmat = 2 if swiglu else 1
    bufs = ring(N * K * nmat, lambda: [mk() for _ in range(nmat)])
    out = gk.empty((M, N), gk.bfloat16)

    def run(i):
        b = bufs[i % len(bufs)]
        if swiglu:
            api.gemv_fp8_swiglu_bf16(A.device_ptr, b[0][0].device_ptr, b[0][1].device_ptr, b[1][0].device_ptr,
                                     b[1][1].device_ptr, out.device_ptr, M, K, N, None)
        else:
            api.gemv_fp8_bf16(A.device_ptr, b[0][0].device_ptr, b[0][1].device_ptr, out.device_ptr, M, K, N, None)
    us = timeit(run)
    byts = nmat * (N * K + 2 * (N // 128) * (K // 128)) + 2 * M * (K + N)
    return us, byts


def bench_bf16(K, N, M, rng):
    A = gk.from_numpy(rng.integers(0x3C00, 0x4000, size=(M, K)).astype(np.uint16))
    bufs = ring(N * K * 2, lambda: gk.from_numpy(rng.integers(0x3800, 0x3C00, size=(N, K)).astype(np.uint16)))
    out = gk.empty((M, N), gk.bfloat16)
    us = timeit(lambda i: api.gemv_bf16(A.device_ptr, bufs[i % len(bufs)].device_ptr, out.device_ptr, M, K, N, None), iters=20)
    return us, N * K * 2 + 2 * M * (K + N)


def bench_nvf4(K, N, M, rng, swiglu=False):
    A = gk.from_numpy(rng.integers(0x3C00, 0x4000, size=(M, K)).astype(np.uint16))
    mk = lambda: (gk.from_numpy(rng.integers(0, 256, size=(K // 2, N), dtype=np.uint8)),
                  gk.from_numpy(np.full((K // 32, N), 0x38, np.uint8)))
    bufs = ring(N * K // 2, mk)
    out = gk.empty((M, N), gk.bfloat16)

    def run(i):
        d, s = bufs[i % len(bufs)]
        if swiglu:
            api.gemv_nvf4_swiglu_bf16(A.device_ptr, d.device_ptr, s.device_ptr, out.device_ptr, M, K, N, None)
        else:
            api.gemv_nvf4_bf16(A.device_ptr, d.device_ptr, s.device_ptr, out.device_ptr, M, K, N, C.c_float(1.0), None)
    us = timeit(run)
    return us, N * K // 2 + N * K // 32 + 2 * M * (K + N)


def bench_attn(H, Hkv, hd, ctx, rng):
    S = ctx
    n = max(2, int(300e6 // (2 * Hkv * S * hd * 2)) + 1)
    kv = [(gk.from_numpy(rng.integers(0x3000, 0x3C00, size=(Hkv, S, hd)).astype(np.uint16)),
           gk.from_numpy(rng.integers(0x3000, 0x3C00, size=(Hkv, S, hd)).astype(np.uint16))) for _ in range(min(n, 16))]
    Q = gk.from_numpy(rng.integers(0x3C00, 0x4000, size=(H, 1, hd)).astype(np.uint16))
    out = gk.empty((H, 1, hd), gk.bfloat16)

    def run(i):
        k, v = kv[i % len(kv)]
        api.sdpa_decode(Q.device_ptr, k.device_ptr, v.device_ptr, out.device_ptr, H, Hkv, 1, hd, S, ctx, None,
                        C.c_float(0.0), hd, hd, hd, hd, 2, None)
    us = timeit(run)
    return us, 2 * Hkv * ctx * hd * 2 + 4 * H * hd


def main():
    rng = np.random.default_rng(0)
    rows = []
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    if which in ("all", "fp8"):
        for (K, N) in [(4096, 6144), (4096, 4096), (12288, 4096), (14336, 4096), (4096, 1024), (8192, 8192)]:
            for M in (1, 2, 4, 8):
                us, b = bench_fp8(K, N, M, rng)
                rows.append(("fp8", K, N, M, us, b))
        for (K, N) in [(4096, 12288), (4096, 14336)]:
            for M in (1, 2):
                us, b = bench_fp8(K, N, M, rng, swiglu=True)
                rows.append(("fp8_swiglu", K, N, M, us, b))
    if which in ("all", "bf16"):
        for (K, N) in [(4096, 151936), (4096, 128256), (2048, 128)]:
            us, b = bench_bf16(K, N, 1, rng)
            rows.append(("bf16", K, N, 1, us, b))
    if which in ("all", "nvf4"):
        for (K, N) in [(8192, 10240), (8192, 8192), (28672, 8192), (8192, 28672), (2048, 768), (768, 2048)]:
            for M in (1, 2, 4):
                us, b = bench_nvf4(K, N, M, rng)
                rows.append(("nvf4", K, N, M, us, b))
        us, b = bench_nvf4(8192, 57344, 1, rng, swiglu=True)
        rows.append(("nvf4_swiglu", 8192, 57344, 1, us, b))
    if which in ("all", "attn"):
        for (H, Hkv, ctx) in [(32, 8, 4096), (32, 8, 32768), (64, 8, 8192), (32, 4, 4096)]:
            us, b = bench_attn(H, Hkv, 128, ctx, rng)
            rows.append(("attn", ctx, H, Hkv, us, b))
    print(f"{'kernel':12s} {'K':>6s} {'N':>7s} {'M':>3s} {'us':>9s} {'GB/s':>8s} {'frac':>6s}")
    for name, K, N, M, us, b in rows:
        gbs = b / us / 1e3
        print(f"{name:12s} {K:6d} {N:7d} {M:3d} {us:9.2f} {gbs:8.1f} {gbs / PEAK:6.3f}")


if __name__ == "__main__":
    main()
