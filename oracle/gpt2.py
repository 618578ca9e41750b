"""TEST INFRASTRUCTURE -- NumPy f32 restatement of the reference's GPT-2 forward on its CPU backend.

Follows CausalTransformerModel.__call__ (src/pygpukit/llm/models/causal.py:108-160: wte[ids] + wpe[pos]),
TransformerBlock (llm/layers/block.py:18-62: pre-norm, residual adds), Attention.__call__ with biases and no
RoPE (llm/layers/attention.py:173-277), _sdpa_causal_cpu (ops/nn/attention.py:89-131: scale 1/sqrt(hd),
causal mask, f32 softmax), _layernorm_cpu (ops/nn/norm.py:78-98: biased variance), _gelu_cpu
(ops/nn/activation.py:37-44: tanh form, c1 = 0.7978845608), get_logits with the tied head
(models/causal.py:167-177).

Parity pin: tests/test_oracle_golden.py checks it against tests/golden/golden_gpt2.npz, which is the
REFERENCE's own CausalTransformerModel.generate run on CPUSimulationBackend (tests/golden/make_golden_gpt2.py).
Only tests/ imports this."""
from __future__ import annotations

import numpy as np


def layernorm(x, g, b, eps):
    mean = x.mean(axis=1, keepdims=True)
    var = x.var(axis=1, keepdims=True)
    return (x - mean) / np.sqrt(var + eps) * g + b


def gelu(x):
    return x * 0.5 * (1 + np.tanh(np.float32(0.7978845608) * (x + np.float32(0.044715) * x ** 3)))


def forward_logits(w: dict, cfg: dict, ids: list[int], eps: float = 1e-5) -> np.ndarray:
    """Logits [S, V] (f32) of the whole sequence `ids` (no cache: causal attention over the prefix)."""
    H, L, nh = cfg["n_embd"], cfg["n_layer"], cfg["n_head"]
    hd = H // nh
    S = len(ids)
    x = (w["wte"][ids] + w["wpe"][:S]).astype(np.float32)
    mask = np.triu(np.full((S, S), -np.inf, np.float32), 1)
    for i in range(L):
        p = f"h{i}."
        h = layernorm(x, w[p + "ln_1.g"], w[p + "ln_1.b"], eps)
        q = (h @ w[p + "q.w"].T + w[p + "q.b"]).reshape(S, nh, hd).transpose(1, 0, 2)
        k = (h @ w[p + "k.w"].T + w[p + "k.b"]).reshape(S, nh, hd).transpose(1, 0, 2)
        v = (h @ w[p + "v.w"].T + w[p + "v.b"]).reshape(S, nh, hd).transpose(1, 0, 2)
        sc = q @ k.transpose(0, 2, 1) / np.float32(np.sqrt(hd)) + mask
        sc = np.exp(sc - sc.max(-1, keepdims=True))
        a = (sc / sc.sum(-1, keepdims=True)) @ v
        a = a.transpose(1, 0, 2).reshape(S, H)
        x = x + (a @ w[p + "o.w"].T + w[p + "o.b"])
        h = layernorm(x, w[p + "ln_2.g"], w[p + "ln_2.b"], eps)
        x = x + (gelu(h @ w[p + "fc1.w"].T + w[p + "fc1.b"]) @ w[p + "fc2.w"].T + w[p + "fc2.b"])
    x = layernorm(x, w["ln_f.g"], w["ln_f.b"], eps)
    return (x @ w["wte"].T).astype(np.float32)


def generate(w: dict, cfg: dict, prompt: list[int], n_new: int) -> list[int]:
    """Greedy decode (sample_token at temperature 0 = first argmax, llm/sampling.py:59-61)."""
    ids = list(prompt)
    for _ in range(n_new):
        ids.append(int(np.argmax(forward_logits(w, cfg, ids)[-1])))
    return ids
