"""Jacobi decode strategy (mirror of pygpukit/llm/decode/jacobi.py:20-218): n future tokens are guessed and refined by
batched forward passes until they stop changing.  Both entry points run on the fixed cache's lookahead pointers
(`CausalTransformerModel.decode_step_jacobi_lookahead`); the reference's `step_jacobi` snapshots the KV cache to the
host around every iteration, which the fixed cache does not need."""
from __future__ import annotations

from .base import DecodeStrategy


class DecodeJacobi(DecodeStrategy):
    def __init__(self, n_tokens: int = 8, max_iter: int = 3, init_strategy: str = "repeat") -> None:
        super().__init__()
        if init_strategy not in ("repeat", "ngram", "greedy"):
            raise ValueError(f"Unknown init strategy: {init_strategy}")
        self._n_tokens, self._max_iter, self._init_strategy = n_tokens, max_iter, init_strategy

    def step(self, token_id: int, position: int, context_len: int, buffers):
        """Plain decode step (jacobi.py:55-78): hidden [1, H] of one token."""
        return self.model._decode_step_fixed_cache(token_id, position, context_len)

    def step_jacobi(self, token_id: int, position: int, context_len: int) -> tuple[list[int], int, dict]:
        """(accepted tokens, new position, stats) for one Jacobi step starting at `position` (jacobi.py:80-177);
        `context_len` must be position + 1, as upstream."""
        if context_len != position + 1:
            raise ValueError("step_jacobi expects context_len == position + 1")
        self.model.set_lookahead_confirmed_pos(position)
        accepted, stats = self.step_lookahead(token_id)
        return accepted, position + len(accepted), stats

    def step_lookahead(self, token_id: int) -> tuple[list[int], dict]:
        return self.model.decode_step_jacobi_lookahead(token_id, n_tokens=self._n_tokens, max_iter=self._max_iter,
                                                       init_strategy=self._init_strategy)

    @property
    def n_tokens(self) -> int:
        return self._n_tokens

    @property
    def max_iter(self) -> int:
        return self._max_iter

    @property
    def init_strategy(self) -> str:
        return self._init_strategy


__all__ = ["DecodeJacobi"]
