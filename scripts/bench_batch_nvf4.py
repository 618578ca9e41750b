"""Decode batch 1..64 through the NVF4 entry point (pgk_gemv_nvf4_bf16_ex): CUDA-core GEMV for M <= 4 (slabs of 4 above
with PGK_TC=0), tcgen05 skinny GEMM (exact BF16 dequantisation) for 5 <= M <= 64.  L distinct weight sets per shape
(>> L2), CUDA events; reports us / packed-weight GB/s (0.53125 B per weight) / TFLOP/s.
    python scripts/bench_batch_nvf4.py [L]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import pygpukit_b200 as gk  # noqa: E402
from pygpukit_b200._lib import api  # noqa: E402
from pygpukit_b200.core.stream import CudaEvent, Stream, event_elapsed_ms, use_stream  # noqa: E402

L = int(sys.argv[1]) if len(sys.argv) > 1 else 8
api.set_pdl(1)
st = Stream()
shapes = [("70B qkv  K=8192 N=10240", 8192, 10240), ("70B o    K=8192 N=8192", 8192, 8192),
          ("70B gate K=8192 N=28672", 8192, 28672), ("70B down K=28672 N=8192", 28672, 8192)]
print(f"PGK_TC={os.environ.get('PGK_TC', '1')}  L={L} weight sets per shape")
for label, K, N in shapes:
    Ds, Ss = [], []
    for i in range(L):
        d = gk.empty((K // 2, N), gk.uint8)
        api.fill_normal_bf16(d.device_ptr, d.size // 2, 100 + i, 1.0, None)     # random bytes: every nibble is a valid weight
        s = gk.empty((K // 32, N), gk.uint8)
        api.memset(s.device_ptr, 0x38, s.size, None)                             # scale 1.0
        Ds.append(d)
        Ss.append(s)
    wbytes = K // 2 * N + K // 32 * N
    for M in (1, 4, 5, 8, 16, 32, 64):
        a = gk.empty((M, K), gk.bfloat16)
        api.fill_normal_bf16(a.device_ptr, a.size, 7, 1.0, None)
        c = gk.empty((M, N), gk.bfloat16)
        api.device_sync()
        with use_stream(st):
            def run():
                for d, s in zip(Ds, Ss):
                    api.gemv_nvf4_bf16_ex(a.device_ptr, d.device_ptr, s.device_ptr, c.device_ptr, M, K, N, 1.0, None, None, st.handle)
            run()
            st.synchronize()
            e0, e1 = CudaEvent(), CudaEvent()
            api.event_record(e0.handle, st.handle)
            for _ in range(3):
                run()
            api.event_record(e1.handle, st.handle)
            e1.synchronize()
        us = event_elapsed_ms(e0, e1) * 1e3 / (3 * L)
        print(f"  {label}  M={M:2d}  {us:8.2f} us   {wbytes / us / 1e3:7.0f} GB/s   {2.0 * M * N * K / us / 1e6:7.1f} TFLOP/s")
