"""Launch time of pgk_sample_token_bf16 (CUDA events, logits resident) next to the host path it replaces.
    python scripts/bench_sampling.py [vocab]"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pygpukit_b200 as gk  # noqa: E402
from pygpukit_b200 import ops  # noqa: E402
from pygpukit_b200._lib import api  # noqa: E402
from pygpukit_b200.core.dtypes import f32_to_bf16_bits  # noqa: E402
from pygpukit_b200.core.stream import CudaEvent, Stream, event_elapsed_ms, use_stream  # noqa: E402
from pygpukit_b200.llm.sampling import sample_token  # noqa: E402

V = int(sys.argv[1]) if len(sys.argv) > 1 else 151936
rng = np.random.default_rng(0)
lg = (rng.standard_normal(V) * 2.5).astype(np.float32)
g = gk.from_numpy(f32_to_bf16_bits(lg))
out = gk.zeros((1,), gk.int32)
st = Stream()
print(f"vocab {V}")
for name, (t, k, p) in {"temperature only": (0.8, 0, 1.0), "top-k 50": (0.8, 50, 1.0), "top-p 0.9": (0.8, 0, 0.9),
                        "top-k 50 + top-p 0.9": (0.8, 50, 0.9), "greedy": (0.0, 0, 1.0)}.items():
    with use_stream(st):
        for _ in range(3):
            ops.sample_token_to_buf(g, out, t, k, p, random_val=0.4)
        st.synchronize()
        e0, e1 = CudaEvent(), CudaEvent()
        api.event_record(e0.handle, st.handle)
        for _ in range(20):
            ops.sample_token_to_buf(g, out, t, k, p, random_val=0.4)
        api.event_record(e1.handle, st.handle)
        e1.synchronize()
    us = event_elapsed_ms(e0, e1) * 1e3 / 20
    t0 = time.perf_counter()
    for _ in range(5):
        host = g.to_numpy()
        from pygpukit_b200.core.dtypes import bf16_bits_to_f32
        sample_token(bf16_bits_to_f32(host), t, k, p)
    host_us = (time.perf_counter() - t0) * 1e6 / 5
    print(f"  {name:24s} device {us:8.1f} us    host path (D2H + NumPy sample_token) {host_us:9.1f} us")
