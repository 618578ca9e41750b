"""Greedy sampling on the device (mirror of pygpukit/ops/sampling.py:75-88)."""
from __future__ import annotations

from .._lib import api
from ..core.array import GPUArray
from ..core.dtypes import float_code, int32
from ._common import _st, _validate_float_dtype


def sample_greedy(logits: GPUArray) -> int:
    """argmax over [vocab] or [1, vocab] logits; the first maximum wins (np.argmax)."""
    _validate_float_dtype(logits, "sample_greedy")
    if logits.ndim == 2 and logits.shape[0] != 1:
        raise ValueError("sample_greedy expects [vocab_size] or [1, vocab_size] logits")
    out = GPUArray((1,), int32)
    api.argmax(logits.device_ptr, out.device_ptr, logits.size, float_code(logits.dtype), _st())
    return int(out.to_numpy()[0])


def sample_greedy_to_buf(logits: GPUArray, out_buf: GPUArray) -> None:
    """argmax written to an int32 device buffer (graph-friendly variant)."""
    api.argmax(logits.device_ptr, out_buf.device_ptr, logits.size, float_code(logits.dtype), _st())


__all__ = ["sample_greedy", "sample_greedy_to_buf"]
