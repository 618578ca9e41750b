"""Single-token decode with ONE CUDA graph per step (API of pygpukit/llm/decode/m1_graph.py:63-594:
init_graph / has_graph / step_graph / buffers).  The reference captures per-layer pre/post-SDPA
graphs and synchronises the device between replays (m1_graph.py:509-577); here embedding, all
layers (attention included), final norm, lm_head and greedy argmax are one graph whose position /
context / token inputs are int32 device buffers."""
from __future__ import annotations

from .base import DecodeStrategy


class DecodeM1Graph(DecodeStrategy):
    def __init__(self) -> None:
        super().__init__()
        self._graph_ready = False

    def init_graph(self, max_seq_len: int = 512) -> None:
        model = self.model
        if model._buffers is None or model._max_seq < max_seq_len:
            model.init_decode(max_seq_len)
        model.init_decode_graph()
        self._graph_ready = True

    def has_graph(self) -> bool:
        return self._graph_ready

    @property
    def buffers(self):
        return self.model._buffers

    def step(self, token_id: int, position: int, context_len: int, buffers=None):
        return self.step_graph(token_id, position, context_len)

    def step_graph(self, token_id: int, position: int, context_len: int):
        """Replay the step graph for (token_id, position); returns the logits buffer [1, vocab].
        The greedy token of this step is left in buffers.next_token."""
        assert self._graph_ready, "Call init_graph() first"
        from ...core.stream import use_stream
        model = self.model
        with use_stream(model._graph_stream):
            model.set_decode_state(token_id, position)
            model.decode_graph_replay()
            model._graph_stream.synchronize()
        return model.buffers.logits


__all__ = ["DecodeM1Graph"]
