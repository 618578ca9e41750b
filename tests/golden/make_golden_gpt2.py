"""Golden run of BASELINE.json configs[0]'s shape of workload (GPT-2, random init, greedy decode on the
reference's NumPy/CPU backend) at test size: the REFERENCE's own CausalTransformerModel.generate on
CPUSimulationBackend (FP32), in the build container where /root/reference is mounted.

    python tests/golden/make_golden_gpt2.py        -> tests/golden/golden_gpt2.npz

Records the greedy tokens, the top-1/top-2 logit margin at every generated step and the logits of the
last prompt position, so the GPU test can tell a BF16 near-tie from a real mismatch."""
import os
import sys

import numpy as np

sys.path.insert(0, "/root/reference/src")
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import pygpukit as ref  # noqa: E402
from pygpukit.core.backend import CPUSimulationBackend, set_backend  # noqa: E402

set_backend(CPUSimulationBackend())
from pygpukit.llm import (GPT2_SPEC, Attention, CausalTransformerModel, GPT2Config, MLP, Norm,  # noqa: E402
                          TransformerBlock)

import gpt2_tiny as T  # noqa: E402
from pygpukit.core.array import GPUArray as _RefArray  # noqa: E402
from pygpukit.core.backend import get_backend  # noqa: E402

# The reference's CPU in-place branches assign `output._data = from_numpy(x)._data`, an attribute its
# GPUArray lacks (ops/nn/linear.py:51-59, ops/nn/norm.py:190-194), so biased models cannot run on the
# CPU backend unmodified.  In-memory shim (nothing under /root/reference is touched): `_data` reads as
# the array itself and assigning it copies the bytes into this array's buffer -- the in-place
# semantics the native branch has.
_RefArray._data = property(lambda self: self,
                           lambda self, other: get_backend().copy_host_to_device(other.to_numpy(), self._device_ptr))

w = T.weights()
cfg = GPT2Config(**T.CFG).to_transformer_config()
g = ref.from_numpy


def norm(p):
    return Norm(g(w[p + ".g"]), g(w[p + ".b"]), "layernorm", cfg.norm_eps)


blocks = []
for i in range(cfg.num_layers):
    p = f"h{i}."
    attn = Attention(g(w[p + "q.w"]), g(w[p + "k.w"]), g(w[p + "v.w"]), g(w[p + "o.w"]), cfg,
                     q_bias=g(w[p + "q.b"]), k_bias=g(w[p + "k.b"]), v_bias=g(w[p + "v.b"]), o_bias=g(w[p + "o.b"]))
    mlp = MLP(cfg, fc1_weight=g(w[p + "fc1.w"]), fc1_bias=g(w[p + "fc1.b"]), fc2_weight=g(w[p + "fc2.w"]),
              fc2_bias=g(w[p + "fc2.b"]))
    blocks.append(TransformerBlock(norm(p + "ln_1"), attn, norm(p + "ln_2"), mlp))
model = CausalTransformerModel(cfg, g(w["wte"]), blocks, norm("ln_f"), lm_head=None, position_embed=g(w["wpe"]),
                               spec=GPT2_SPEC)

tokens = model.generate(list(T.PROMPT), max_new_tokens=T.N_NEW, temperature=0.0, use_cache=True)
# margins + logits by re-running the full sequence without cache (teacher forcing on the reference's tokens)
hidden, _ = model(tokens[:-1])
logits = model.get_logits(hidden).to_numpy().astype(np.float32)          # [S-1, V]
gen = logits[len(T.PROMPT) - 1:]
srt = np.sort(gen, axis=-1)
out = {"tokens": np.asarray(tokens, np.int32), "margins": (srt[:, -1] - srt[:, -2]).astype(np.float32),
       "argmax_teacher_forced": gen.argmax(-1).astype(np.int32), "logits_teacher_forced": gen,
       "logit_scale": np.float32(np.abs(gen).max())}
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden_gpt2.npz")
np.savez_compressed(path, **out)
print("wrote", path, tokens, "min margin", out["margins"].min(), "scale", out["logit_scale"], "distinct", len(set(tokens[len(T.PROMPT):])))
assert np.array_equal(out["argmax_teacher_forced"], out["tokens"][len(T.PROMPT):]), "cache vs no-cache disagree"
