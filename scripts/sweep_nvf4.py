"""Sweep the NVF4 GEMV launch configuration (PGK_NVF4_CFG=warps,k_splits) for one shape; one process per config.
    python scripts/sweep_nvf4.py K N        (driver)      python scripts/sweep_nvf4.py K N --one   (worker)"""
import os
import subprocess
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
K, N = int(sys.argv[1]), int(sys.argv[2])
if "--one" in sys.argv:
    import pygpukit_b200 as gk
    from pygpukit_b200._lib import api
    from pygpukit_b200.core.stream import CudaEvent, Stream, event_elapsed_ms, use_stream
    api.set_pdl(1)
    L = 10
    st = Stream()
    Ws = []
    for i in range(L):
        d = gk.empty((K // 2, N), gk.uint8); api.fill_normal_bf16(d.device_ptr, d.size // 2, 10 + i, 1.0, None)
        s = gk.empty((K // 32, N), gk.uint8); api.memset(s.device_ptr, 0x38, s.size, None)
        Ws.append((d, s))
    a = gk.empty((1, K), gk.bfloat16); api.fill_normal_bf16(a.device_ptr, a.size, 7, 1.0, None)
    c = gk.empty((1, N), gk.bfloat16)
    api.device_sync()
    with use_stream(st):
        def run():
            for d, s in Ws:
                api.gemv_nvf4_bf16_ex(a.device_ptr, d.device_ptr, s.device_ptr, c.device_ptr, 1, K, N, 1.0, None, None, st.handle)
        run(); st.synchronize()
        e0, e1 = CudaEvent(), CudaEvent()
        api.event_record(e0.handle, st.handle)
        for _ in range(5):
            run()
        api.event_record(e1.handle, st.handle); e1.synchronize()
    us = event_elapsed_ms(e0, e1) * 1e3 / (5 * L)
    print(f"{os.environ.get('PGK_NVF4_CFG', 'default'):>10s}  {us:8.2f} us  {(K * N * 0.53125) / us / 1e3:7.0f} GB/s")
else:
    cfgs = ["default"] + [f"{w},{ks}" for w in (8, 12, 16) for ks in (1, 2, 4, 6, 9, 12, 18, 27)]
    for cfg in cfgs:
        env = dict(os.environ)
        if cfg != "default":
            env["PGK_NVF4_CFG"] = cfg
        r = subprocess.run([sys.executable, __file__, str(K), str(N), "--one"], env=env, capture_output=True, text=True, timeout=120)
        print(r.stdout.strip() or r.stderr.strip()[-200:], flush=True)
