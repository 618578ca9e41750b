"""One small invocation of every kernel family of the library, for compute-sanitizer (one tool per run):

    compute-sanitizer --tool memcheck  python scripts/sanitize_smoke.py
    compute-sanitizer --tool racecheck python scripts/sanitize_smoke.py

Covers the hand-rolled synchronisation protocols the tools are meant for: the split-KV ticket + last-CTA fold of the
attention kernels, the split-K tickets of the tensor-core GEMMs (mbarrier / TMEM hand-offs included), the NVF4 K-split
fold, and the persistent step kernel (TMA ring, per-slot mbarriers, device-side grid barrier)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pygpukit_b200 as gk  # noqa: E402
from pygpukit_b200 import llm  # noqa: E402
from pygpukit_b200._lib import api  # noqa: E402
from pygpukit_b200.llm.random_init import normal_bf16, quantize_fp8, quantize_nvf4  # noqa: E402

api.set_pdl(int(os.environ.get("PGK_PDL", "1")))
K, N = 512, 512
x1, x8, x17 = (normal_bf16((m, K), 5 + m, 1.0) for m in (1, 8, 17))
fp8 = quantize_fp8(normal_bf16((N, K), 1, 0.05))
nv = quantize_nvf4(normal_bf16((K, N), 2, 0.05))
wb = normal_bf16((N, K), 3, 0.05)
res = normal_bf16((17, N), 4, 1.0)
out = gk.zeros((17, N), gk.bfloat16)
tok = gk.zeros((1,), gk.int32)
for x, m in ((x1, 1), (x8, 8), (x17, 17)):
    api.gemv_fp8_bf16_ex(x.device_ptr, fp8.weight_fp8.device_ptr, fp8.scale_inv.device_ptr, out.device_ptr, m, K, N, None, res.device_ptr, None)
    api.gemm_fp8_bf16_f8mma(x.device_ptr, fp8.weight_fp8.device_ptr, fp8.scale_inv.device_ptr, out.device_ptr, m, K, N, None, res.device_ptr, None)
    api.gemv_nvf4_bf16_ex(x.device_ptr, nv.data.device_ptr, nv.scale.device_ptr, out.device_ptr, m, K, N, 1.0, None, res.device_ptr, None)
    api.gemv_bf16_ex(x.device_ptr, wb.device_ptr, out.device_ptr, m, K, N, None, None, None)
api.gemv_bf16_argmax(x1.device_ptr, wb.device_ptr, out.device_ptr, tok.device_ptr, K, N, None)
api.device_sync()
print("kernels ok", int(tok.to_numpy()[0]))
for fmt in ("fp8", "nvf4"):
    m = llm.build_random_model("small-qwen3", fmt, max_seq=96, seed=1, std=0.05)
    m.init_decode(96, max_batch_size=8)
    a = m.generate([3, 141, 59, 26, 535], 6, temperature=0.0)
    m.use_step_program = True                       # the persistent step kernel
    m.init_decode(96, max_batch_size=8)
    b = m.generate([3, 141, 59, 26, 535], 6, temperature=0.0)
    assert m.step_program() is not None and m.step_program().error() == 0
    m.use_step_program = False
    m.init_decode(96, max_batch_size=8, num_sequences=3)    # multi-sequence decode
    m.init_decode_graph()
    m.set_decode_state([5, 6, 7], [0, 0, 0])
    for _ in range(3):
        m.decode_graph_replay()
    api.device_sync()
    print(fmt, "per-op", a[5:], "step", b[5:], "multi", m.buffers.next_token.to_numpy()[:3].tolist())
moe = llm.build_random_model("tiny-moe", "fp8", max_seq=64, seed=1, std=0.05)
moe.init_decode(64, max_batch_size=8)
print("moe", moe.generate([3, 14, 15, 92], 4, temperature=0.0)[4:])
_ = np
