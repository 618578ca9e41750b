"""Batch decode (mirror of pygpukit/llm/decode/batch.py:30-363): M consecutive tokens of ONE sequence
sharing one KV cache (speculative verification), causal inside the batch; returns hidden [M, H]."""
from __future__ import annotations

from .base import DecodeStrategy


class DecodeBatch(DecodeStrategy):
    def __init__(self, batch_size: int = 8) -> None:
        super().__init__()
        self._batch_size = batch_size

    def step(self, token_id: int, position: int, context_len: int, buffers=None):
        return self.step_batch([token_id], position, context_len, buffers)

    def step_batch(self, token_ids: list[int], start_position: int, context_len: int, buffers=None):
        return self.model._decode_step_fixed_cache_batch(token_ids, start_position, context_len)


__all__ = ["DecodeBatch"]
