"""Self-speculative decode strategy (mirror of pygpukit/llm/decode/speculative.py:19-200): the model's first layers draft
tokens, the full model verifies them in one batched pass.  Both entry points use the fixed cache's lookahead pointers
(`CausalTransformerModel.decode_step_self_speculative_lookahead`); the reference's `step_speculative` snapshots the whole
KV cache to the HOST and restores it twice per step (`snapshot_kv_cache` / `restore_kv_cache`, causal.py:1003-1034),
which the fixed cache does not need: speculative rows beyond the committed position are simply never read."""
from __future__ import annotations

from .base import DecodeStrategy


class DecodeSpeculative(DecodeStrategy):
    def __init__(self, max_draft_tokens: int = 4, draft_layers: int = 8) -> None:
        super().__init__()
        self._max_draft_tokens = max_draft_tokens
        self._draft_layers = draft_layers

    def step(self, token_id: int, position: int, context_len: int, buffers):
        """Plain decode step (speculative.py:45-68): hidden [1, H] of one token."""
        return self.model._decode_step_fixed_cache(token_id, position, context_len)

    def step_speculative(self, token_id: int, position: int, context_len: int) -> tuple[list[int], int, dict]:
        """(accepted tokens, new position, stats) for one speculative step starting at `position`
        (speculative.py:70-167); `context_len` must be position + 1, as upstream."""
        if context_len != position + 1:
            raise ValueError("step_speculative expects context_len == position + 1")
        self.model.set_lookahead_confirmed_pos(position)
        accepted, stats = self.step_lookahead(token_id)
        return accepted, position + len(accepted), stats

    def step_lookahead(self, token_id: int) -> tuple[list[int], dict]:
        """speculative.py:169-190: the model tracks the committed position itself."""
        return self.model.decode_step_self_speculative_lookahead(
            token_id, max_draft_tokens=self._max_draft_tokens, draft_layers=self._draft_layers)

    @property
    def max_draft_tokens(self) -> int:
        return self._max_draft_tokens

    @property
    def draft_layers(self) -> int:
        return self._draft_layers


__all__ = ["DecodeSpeculative"]
