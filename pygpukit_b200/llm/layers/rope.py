"""RoPE tables (host, NumPy float32) -- must be bit-identical to the reference's
llm/layers/rope.py:13-24 because token parity at long context depends on the table rounding
(SURVEY.md Appendix A.3): float32 inverse frequencies, float32 positions, outer product,
cos/sin, each duplicated across the two halves of head_dim."""
from __future__ import annotations

import numpy as np


def precompute_freqs_cis(head_dim: int, max_seq_len: int, theta: float = 10000.0) -> tuple[np.ndarray, np.ndarray]:
    exponents = np.arange(0, head_dim, 2, dtype=np.float32) / head_dim
    inv_freq = 1.0 / (theta ** exponents)
    angles = np.outer(np.arange(max_seq_len, dtype=np.float32), inv_freq)
    cos_half, sin_half = np.cos(angles), np.sin(angles)
    return np.concatenate([cos_half, cos_half], axis=-1), np.concatenate([sin_half, sin_half], axis=-1)


def apply_rotary_pos_emb_numpy(q: np.ndarray, k: np.ndarray, cos: np.ndarray, sin: np.ndarray):
    """NumPy reference rotation (llm/layers/rope.py:27-42), used by tests."""
    def rot(x):
        h = x.shape[-1] // 2
        return np.concatenate([-x[..., h:], x[..., :h]], axis=-1)
    c, s = cos[:, None, :], sin[:, None, :]
    return q * c + rot(q) * s, k * c + rot(k) * s
