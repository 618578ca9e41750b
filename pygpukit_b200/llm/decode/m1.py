"""Single-token decode without CUDA graph (mirror of pygpukit/llm/decode/m1.py:28-121): returns the
logits buffer [1, vocab].  ~8 launches per layer instead of the reference's ~25 launches + syncs."""
from __future__ import annotations

from .base import DecodeStrategy


class DecodeM1(DecodeStrategy):
    def step(self, token_id: int, position: int, context_len: int, buffers=None):
        model = self.model
        hidden = model._decode_step_fixed_cache(token_id, position, context_len)
        b = model.buffers
        logits = b.logits.slice_rows(1) if b.logits.shape[0] != 1 else b.logits
        return model.get_logits(hidden, out=logits)


__all__ = ["DecodeM1"]
