"""Kernel against kernel on the same B200: the REFERENCE's own CUDA kernels (oracle/_ref/libref_cuda.so, compiled
in place from /root/reference by oracle/build_ref.sh for sm_100a with the reference's flags) next to ours, on the
shapes of BASELINE.json's configs (SURVEY.md section 8d "Reference-on-B200 baseline").

Timing as src/pygpukit/benchmark/base.py:31-60 prescribes (warm-up, then CUDA events around back-to-back launches on
the default stream), every launch on a different weight set so the stream comes from HBM (ring >> 126 MB L2).
    python scripts/bench_vs_reference.py [out.json]
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import pygpukit_b200 as gk  # noqa: E402
from oracle import refcuda  # noqa: E402
from pygpukit_b200._lib import api  # noqa: E402
from pygpukit_b200.core.stream import CudaEvent, event_elapsed_ms  # noqa: E402

if not refcuda.available():
    print(json.dumps({"unavailable": "oracle/_ref/libref_cuda.so missing (build it where /root/reference is mounted)"}))
    sys.exit(0)
api.set_pdl(0)                       # plain back-to-back launches on both sides
refcuda.set_sync(False)
RING_BYTES = 320e6
rows = []


def fill(a, seed, std=1.0):
    api.fill_normal_bf16(a.device_ptr, a.nbytes // 2, seed, std, None)
    return a


def ring(nbytes, make):
    n = max(2, min(48, int(RING_BYTES // max(nbytes, 1)) + 1))
    return [make(i) for i in range(n)]


def timed(fn, n_sets, reps=3):
    for i in range(n_sets):
        fn(i)
    api.device_sync()
    e0, e1 = CudaEvent(), CudaEvent()
    api.event_record(e0.handle, None)
    for _ in range(reps):
        for i in range(n_sets):
            fn(i)
    api.event_record(e1.handle, None)
    e1.synchronize()
    return event_elapsed_ms(e0, e1) * 1e3 / (reps * n_sets)


def report(op, shape, t_ref, t_ours, nbytes, note=""):
    r = {"op": op, "shape": shape, "ref_us": round(t_ref, 2), "ours_us": round(t_ours, 2),
         "speedup": round(t_ref / t_ours, 2), "ours_GBs": round(nbytes / t_ours / 1e3, 0),
         "ref_GBs": round(nbytes / t_ref / 1e3, 0), "note": note}
    rows.append(r)
    print(f"{op:26s} {shape:26s} ref {t_ref:9.2f} us  ours {t_ours:8.2f} us  x{t_ref / t_ours:6.2f}  "
          f"ours {nbytes / t_ours / 1e3:6.0f} GB/s  {note}", flush=True)


# ---------------------------------------------------------------- FP8 GEMV, batch 1 (Qwen3-8B / Llama-3-8B shapes)
def fp8_set(N, K, i):
    w = fill(gk.empty((N, K), gk.uint8), 100 + i)
    s = fill(gk.empty((N // 128, K // 128), gk.bfloat16), 900 + i, 0.01)
    return w, s


for label, K, N in (("q/o 4096x4096", 4096, 4096), ("k/v 4096x1024", 4096, 1024), ("gate/up 4096x12288", 4096, 12288),
                    ("down 12288x4096", 12288, 4096), ("gate/up 4096x14336", 4096, 14336), ("down 14336x4096", 14336, 4096)):
    sets = ring(N * K, lambda i: fp8_set(N, K, i))
    a = fill(gk.empty((1, K), gk.bfloat16), 7)
    c = gk.empty((1, N), gk.bfloat16)
    t_ref = timed(lambda i: refcuda.gemv_fp8(a, sets[i][0], sets[i][1], c, K, N), len(sets))
    t_our = timed(lambda i: api.gemv_fp8_bf16(a.device_ptr, sets[i][0].device_ptr, sets[i][1].device_ptr, c.device_ptr,
                                              1, K, N, None), len(sets))
    report("gemv_fp8 M=1", label, t_ref, t_our, N * K)
    if label.startswith("q/o"):
        for M in (4, 8, 16, 64):
            am = fill(gk.empty((M, K), gk.bfloat16), 8)
            cm = gk.empty((M, N), gk.bfloat16)
            t_ref = timed(lambda i: refcuda.gemv_fp8_batched(am, sets[i][0], sets[i][1], cm, K, N, M), len(sets))
            t_our = timed(lambda i: api.gemv_fp8_bf16(am.device_ptr, sets[i][0].device_ptr, sets[i][1].device_ptr,
                                                      cm.device_ptr, M, K, N, None), len(sets))
            report(f"gemv_fp8_batched M={M}", label, t_ref, t_our, N * K,
                   "ours: tcgen05 GEMM" if M >= 5 else "ours: CUDA-core slab")
            if M >= 16:
                # the reference's tensor-core path (w8a16_gemm_sm120, mma.sync) takes B as [K, N]: same bytes
                skn = fill(gk.empty((K // 128, N // 128), gk.bfloat16), 901, 0.01)
                t_ref2 = timed(lambda i: refcuda.w8a16_gemm(am, sets[i][0], skn, cm, M, N, K), len(sets))
                report(f"w8a16_gemm_sm120 M={M}", label, t_ref2, t_our, N * K, "reference mma.sync path vs our tcgen05")
    del sets

# ---------------------------------------------------------------- NVF4 GEMV (Llama-3-70B shapes)
for label, K, N in (("q/o 8192x8192", 8192, 8192), ("gate/up 8192x28672", 8192, 28672), ("down 28672x8192", 28672, 8192)):
    nb = K * N // 2 + (K // 32) * N
    sets = ring(nb, lambda i: (fill(gk.empty((K // 2, N), gk.uint8), 300 + i),
                               gk.from_numpy(__import__("numpy").full((K // 32, N), 0x38, "uint8"))))
    a = fill(gk.empty((1, K), gk.bfloat16), 9)
    c = gk.empty((1, N), gk.bfloat16)
    t_ref = timed(lambda i: refcuda.gemv_nvf4(a, sets[i][0], sets[i][1], c, K, N), len(sets))
    t_our = timed(lambda i: api.gemv_nvf4_bf16(a.device_ptr, sets[i][0].device_ptr, sets[i][1].device_ptr, c.device_ptr,
                                               1, K, N, 1.0, None), len(sets))
    report("gemv_nvf4 M=1", label, t_ref, t_our, nb)
    del sets

# ---------------------------------------------------------------- BF16 lm_head GEMV
K, N = 4096, 151936
sets = ring(N * K * 2, lambda i: fill(gk.empty((N, K), gk.bfloat16), 500 + i, 0.02))
a = fill(gk.empty((1, K), gk.bfloat16), 10)
c = gk.empty((1, N), gk.bfloat16)
t_ref = timed(lambda i: refcuda.gemv_bf16(a, sets[i], c, K, N), len(sets), reps=2)
t_our = timed(lambda i: api.gemv_bf16(a.device_ptr, sets[i].device_ptr, c.device_ptr, 1, K, N, None), len(sets), reps=2)
report("gemv_bf16 lm_head", "4096x151936", t_ref, t_our, N * K * 2)
del sets

# ---------------------------------------------------------------- decode attention (32 q heads / 8 kv heads, hd 128)
H, Hkv, hd = 32, 8, 128
for ctx in (4096, 32768):
    n_layers = max(2, int(RING_BYTES // (2 * H * ctx * hd * 2)) + 1)
    # reference layout: GQA-expanded cache [H, max_seq, hd]; ours: un-expanded [Hkv, max_seq, hd]
    ref_kv = [(fill(gk.empty((H, ctx, hd), gk.bfloat16), 600 + 2 * i, 0.1), fill(gk.empty((H, ctx, hd), gk.bfloat16), 601 + 2 * i, 0.1))
              for i in range(n_layers)]
    q = fill(gk.empty((H, 1, hd), gk.bfloat16), 11)
    o = gk.empty((H, 1, hd), gk.bfloat16)
    try:
        t_ref = timed(lambda i: refcuda.sdpa(q, ref_kv[i][0], ref_kv[i][1], o, H, 1, ctx, ctx, hd), n_layers)
    except RuntimeError as e:        # the reference kernel keeps kv_len floats in shared memory (sdpa_causal.inl:636-666)
        print(f"reference sdpa at ctx {ctx}: {e}")
        api.device_sync()
        t_ref = float("nan")
    del ref_kv
    n_layers = max(2, int(RING_BYTES // (2 * Hkv * ctx * hd * 2)) + 1)
    our_kv = [(fill(gk.empty((Hkv, ctx, hd), gk.bfloat16), 700 + 2 * i, 0.1), fill(gk.empty((Hkv, ctx, hd), gk.bfloat16), 701 + 2 * i, 0.1))
              for i in range(n_layers)]
    t_our = timed(lambda i: api.sdpa_decode(q.device_ptr, our_kv[i][0].device_ptr, our_kv[i][1].device_ptr, o.device_ptr,
                                            H, Hkv, 1, hd, ctx, ctx, None, 0.0, hd, H * hd, hd, H * hd, 2, None), n_layers)
    report("sdpa_causal_fixed_cache", f"H32/kv8 hd128 ctx{ctx}", t_ref, t_our, 2 * Hkv * ctx * hd * 2,
           "ref reads the 4x GQA-expanded cache; bytes = un-expanded")
    del our_kv

# ---------------------------------------------------------------- RMSNorm
x = fill(gk.empty((1, 4096), gk.bfloat16), 12)
g = fill(gk.empty((4096,), gk.bfloat16), 13)
y = gk.empty((1, 4096), gk.bfloat16)
t_ref = timed(lambda i: refcuda.rmsnorm_bf16(x, g, y, 1, 4096, 1e-6), 8, reps=10)
t_our = timed(lambda i: api.rmsnorm(x.device_ptr, g.device_ptr, y.device_ptr, 1, 4096, 1e-6, 2, None), 8, reps=10)
report("rmsnorm", "1x4096", t_ref, t_our, 3 * 4096 * 2, "launch-bound; fused into the GEMV prologue in the step")

refcuda.set_sync(True)
out = sys.argv[1] if len(sys.argv) > 1 else None
if out:
    with open(out, "w") as f:
        json.dump({"device": "B200", "method": "CUDA events, back-to-back launches, distinct weight set per launch (ring >> L2)",
                   "reference": "oracle/_ref/libref_cuda.so = the reference's kernels compiled for sm_100a (-O3 --use_fast_math)",
                   "rows": rows}, f, indent=1)
