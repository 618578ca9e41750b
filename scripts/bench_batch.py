"""Decode batch 1..64 through the W8A16 entry point (pgk_gemv_fp8_bf16_ex): CUDA-core GEMV for M <= 8 (slabs of 4
above), tcgen05 skinny GEMM for 16 <= M <= 64 (PGK_TC=0 keeps the slab GEMV there: run twice to compare).
L distinct weight sets per shape (>> L2), CUDA events, reports us / weight GB/s / TFLOP/s.
    python scripts/bench_batch.py [L]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402

import pygpukit_b200 as gk  # noqa: E402
from pygpukit_b200._lib import api  # noqa: E402
from pygpukit_b200.core.stream import CudaEvent, Stream, event_elapsed_ms, use_stream  # noqa: E402
from pygpukit_b200.llm.random_init import normal_bf16, quantize_fp8  # noqa: E402

L = int(sys.argv[1]) if len(sys.argv) > 1 else 12
api.set_pdl(1)
st = Stream()
shapes = [("qkv  K=4096 N=6144", 4096, 6144), ("o    K=4096 N=4096", 4096, 4096),
          ("gate K=4096 N=12288", 4096, 12288), ("down K=12288 N=4096", 12288, 4096)]
print(f"PGK_TC={os.environ.get('PGK_TC', '1')}  L={L} weight sets per shape")
for label, K, N in shapes:
    Ws, Ss = [], []
    for i in range(L):
        # real block-quantised weights (random bytes would hold the codes 0x7F / 0xFF, which the entry point serves with the
        # exact table kernel instead of the streaming ones)
        lin = quantize_fp8(normal_bf16((N, K), 100 + i, 0.02))
        Ws.append(lin.weight_fp8)
        Ss.append(lin.scale_inv)
    for M in (1, 2, 4, 5, 8, 12, 16, 32, 64):
        a = gk.empty((M, K), gk.bfloat16)
        api.fill_normal_bf16(a.device_ptr, a.size, 7, 1.0, None)
        c = gk.empty((M, N), gk.bfloat16)
        api.device_sync()
        with use_stream(st):
            def run():
                for w, s in zip(Ws, Ss):
                    api.gemv_fp8_bf16_ex(a.device_ptr, w.device_ptr, s.device_ptr, c.device_ptr, M, K, N, None, None, st.handle)
            run()
            st.synchronize()
            e0, e1 = CudaEvent(), CudaEvent()
            api.event_record(e0.handle, st.handle)
            for _ in range(3):
                run()
            api.event_record(e1.handle, st.handle)
            e1.synchronize()
        us = event_elapsed_ms(e0, e1) * 1e3 / (3 * L)
        print(f"  {label}  M={M:2d}  {us:8.2f} us   {N * K / us / 1e3:7.0f} GB/s   {2.0 * M * N * K / us / 1e6:7.1f} TFLOP/s")
_ = np
