"""Generate golden vectors from the REFERENCE's own Python CPU code (run in the build container,
where /root/reference is mounted; the GPU box only sees the committed .npz).

    python tests/golden/make_golden.py

Sources (paths under /root/reference/src/pygpukit): llm/quant.py:292-367
(dequantize_fp8_e4m3_block), ops/nn/norm.py:173-195 (_rmsnorm_cpu), ops/nn/rope.py:49-89
(_rope_inplace_cpu), ops/nn/attention.py:89-131 (_sdpa_causal_cpu), llm/layers/rope.py:13-24
(precompute_freqs_cis), llm/sampling.py:11-63 (sample_token, temperature 0).
"""
import os
import sys

import numpy as np

sys.path.insert(0, "/root/reference/src")
import pygpukit as ref  # noqa: E402
from pygpukit.core.backend import CPUSimulationBackend, set_backend  # noqa: E402
from pygpukit.llm.layers.rope import precompute_freqs_cis  # noqa: E402
from pygpukit.llm.quant import _get_fp8_e4m3_table, dequantize_fp8_e4m3_block  # noqa: E402
from pygpukit.llm.sampling import sample_token  # noqa: E402
from pygpukit.ops.nn.attention import _sdpa_causal_cpu  # noqa: E402
from pygpukit.ops.nn.norm import _rmsnorm_cpu  # noqa: E402
from pygpukit.ops.nn.rope import _rope_inplace_cpu  # noqa: E402

set_backend(CPUSimulationBackend())
rng = np.random.default_rng(1234)
out = {}

# FP8 E4M3 table + block dequant
out["fp8_table_py"] = _get_fp8_e4m3_table().copy()
codes = rng.integers(0, 256, size=(256, 384), dtype=np.uint8)
codes[(codes & 0x7F) == 0x7F] = 0x3A          # the Python table has NaN there; CUDA LUT says +-480
scale_f32 = rng.uniform(1e-3, 2e-2, size=(2, 3)).astype(np.float32)
scale_bits = ((scale_f32.view(np.uint32) + 0x7FFF + ((scale_f32.view(np.uint32) >> 16) & 1)) >> 16).astype(np.uint16)
out["fp8_codes"] = codes
out["fp8_scale_bf16"] = scale_bits
out["fp8_dequant"] = dequantize_fp8_e4m3_block(codes, scale_bits).astype(np.float32)

# rmsnorm f32
x = rng.standard_normal((4, 96)).astype(np.float32)
g = rng.uniform(0.5, 1.5, 96).astype(np.float32)
out["rms_x"], out["rms_g"] = x, g
out["rms_y"] = _rmsnorm_cpu(ref.from_numpy(x), ref.from_numpy(g), 1e-6).to_numpy()

# rope tables + rope f32
cos, sin = precompute_freqs_cis(16, 8, 10000.0)
out["rope_cos16"], out["rope_sin16"] = cos.astype(np.float32), sin.astype(np.float32)
q = rng.standard_normal((3, 4, 16)).astype(np.float32)
k = rng.standard_normal((3, 2, 16)).astype(np.float32)
out["rope_q"], out["rope_k"] = q.copy(), k.copy()
qa, ka = ref.from_numpy(q), ref.from_numpy(k)
_rope_inplace_cpu(qa, ka, ref.from_numpy(cos[:3].astype(np.float32)), ref.from_numpy(sin[:3].astype(np.float32)))
out["rope_q_out"], out["rope_k_out"] = qa.to_numpy(), ka.to_numpy()
for hd, theta, tag in ((128, 1e6, "qwen3"), (128, 5e5, "llama3"), (64, 1e4, "gpt")):
    c, s = precompute_freqs_cis(hd, 4100, theta)
    rows = np.array([0, 1, 2, 17, 1023, 4095])
    out[f"freqs_{tag}_rows"] = rows
    out[f"freqs_{tag}_cos"] = c[rows].astype(np.float32)
    out[f"freqs_{tag}_sin"] = s[rows].astype(np.float32)
    out[f"freqs_{tag}_dtype"] = np.array(str(c.dtype))

# sdpa f32 (causal, q_len 2 over kv_len 6)
Q = rng.standard_normal((4, 2, 16)).astype(np.float32)
K = rng.standard_normal((4, 6, 16)).astype(np.float32)
V = rng.standard_normal((4, 6, 16)).astype(np.float32)
out["sdpa_q"], out["sdpa_k"], out["sdpa_v"] = Q, K, V
out["sdpa_out"] = _sdpa_causal_cpu(ref.from_numpy(Q), ref.from_numpy(K), ref.from_numpy(V), 0.0).to_numpy()

# greedy sampling
logits = rng.standard_normal(1000).astype(np.float32)
logits[[17, 400]] = logits.max() + 1.0      # tie: first index wins
out["sample_logits"] = logits
out["sample_token"] = np.array(sample_token(logits, temperature=0.0))

path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden_ops.npz")
np.savez_compressed(path, **out)
print("wrote", path, {k: getattr(v, "shape", None) for k, v in out.items()})
