"""Host-side token sampling (behaviour of pygpukit/llm/sampling.py:11-63: temperature, top-k, top-p
over a NumPy logits vector; temperature == 0 means greedy argmax, first maximum wins).  The
decode fast path never calls this: greedy sampling is fused into the lm_head GEMV
(pgk_gemv_bf16_argmax) and the token stays on the device."""
from __future__ import annotations

import numpy as np


def _softmax(v: np.ndarray) -> np.ndarray:
    e = np.exp(v - v.max())
    return e / e.sum()


def _keep(probs: np.ndarray, idx: np.ndarray) -> np.ndarray:
    kept = np.zeros_like(probs)
    kept[idx] = probs[idx]
    return kept / kept.sum()


def sample_token(logits: np.ndarray, temperature: float = 1.0, top_k: int = 0, top_p: float = 1.0) -> int:
    scaled = logits / temperature if (temperature > 0 and temperature != 1.0) else logits
    probs = _softmax(scaled)
    n = len(probs)
    if 0 < top_k < n:
        probs = _keep(probs, np.argsort(probs)[n - top_k:])
    if top_p < 1.0:
        order = np.argsort(probs)[::-1]
        cut = min(int(np.searchsorted(np.cumsum(probs[order]), top_p)) + 1, n)
        probs = _keep(probs, order[:cut])
    if temperature == 0:
        return int(np.argmax(probs))
    return int(np.random.choice(n, p=probs))


__all__ = ["sample_token"]
