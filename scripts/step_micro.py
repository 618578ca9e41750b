"""Micro-experiment on the step kernel's op boundary: a chain of identical FP8 GEMV ops (x -> W x -> W x ...), traced.
    python scripts/step_micro.py"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pygpukit_b200 as gk  # noqa: E402
from pygpukit_b200._lib import api  # noqa: E402
from pygpukit_b200.llm import step_program as sp  # noqa: E402
from pygpukit_b200.llm.layers.linear import LinearFP8  # noqa: E402
from pygpukit_b200.llm.random_init import normal_bf16, quantize_fp8  # noqa: E402


def chain(K, N_ops, norm, distinct):
    ws = [quantize_fp8(normal_bf16((K, K), 100 + i, 0.02)) for i in range(N_ops if distinct else 1)]
    bufs = [gk.zeros((1, K), gk.bfloat16) for _ in range(2)]
    api.fill_normal_bf16(bufs[0].device_ptr, K, 5, 1.0, None)
    gamma = gk.ones((K,), gk.bfloat16)
    ops = []
    for i in range(N_ops):
        ops.append(sp.gemv_op(ws[i % len(ws)], bufs[i & 1], bufs[(i + 1) & 1], gamma=gamma if norm else None, eps=1e-5))
    st = gk.from_numpy(np.array([0, 1, 0], np.int32))
    prog = sp.StepProgram(ops, st.narrow(2, 1), st.narrow(0, 1), st.narrow(1, 1), gk.zeros((1,), gk.int32), 1 << 30, False,
                          keep=[ws, bufs, gamma, st])
    for _ in range(3):
        prog.run(None)
    assert prog.error() == 0
    tr = prog.trace().astype(np.int64).astype(np.float64) / 1e3
    d = np.diff(tr[:, 0])
    o = tr[4:-1]
    print(f"K={K} ops={N_ops} norm={norm} distinct_weights={distinct}: per op {d[4:].mean():.2f} us | barrier->loads back "
          f"{((o[:, 5] if norm else o[:, 7]) - o[:, 0]).mean():.2f}  staged {(o[:, 1] - o[:, 0]).mean():.2f}  warp0 items "
          f"{(o[:, 2] - o[:, 1]).mean():.2f}  CTA0 busy {(o[:, 4] - o[:, 0]).mean():.2f}  barrier wait {(tr[5:, 0] - o[:, 4]).mean():.2f}")


api.set_pdl(1)
print("bar mode", os.environ.get("PGK_STEP_BAR", "0"))
chain(1024, 24, False, True)
chain(4096, 24, True, True)
