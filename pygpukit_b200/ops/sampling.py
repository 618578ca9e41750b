"""Sampling on the device (mirror of pygpukit/ops/sampling.py:11-141): greedy argmax, and temperature / top-k / top-p
sampling with the semantics of pygpukit.llm.sampling.sample_token (llm/sampling.py:11-63)."""
from __future__ import annotations

import numpy as np

from .._lib import api
from ..core.array import GPUArray
from ..core.dtypes import bfloat16, float_code, int32
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


_rng = np.random.default_rng()


def set_sampling_seed(seed: int) -> None:
    """Seed of the host generator that supplies the uniform numbers (native.set_sampling_seed upstream)."""
    global _rng
    _rng = np.random.default_rng(seed)


def _check_sampling(logits: GPUArray, temperature: float, top_p: float, fn: str) -> None:
    if logits.dtype != bfloat16:
        raise ValueError(f"{fn}: logits must be bfloat16 on the B200 path, got {logits.dtype}")
    if logits.ndim not in (1, 2) or (logits.ndim == 2 and logits.shape[0] != 1):
        raise ValueError(f"{fn} expects [vocab_size] or [1, vocab_size] logits")
    if temperature < 0:
        raise ValueError(f"{fn}: temperature must be >= 0")


def sample_token_to_buf(logits: GPUArray, out_buf: GPUArray, temperature: float = 1.0, top_k: int = 0, top_p: float = 1.0,
                        random_val: float | None = None, random_buf: GPUArray | None = None) -> None:
    """Sampled token written to an int32 device buffer (graph-friendly: pass `random_buf`, a float32 [1] device array,
    update it before each replay).  `random_val` in [0, 1) makes the call a pure function of its inputs."""
    _check_sampling(logits, temperature, top_p, "sample_token_to_buf")
    u = float(_rng.random()) if (random_val is None and random_buf is None) else float(random_val or 0.0)
    api.sample_token_bf16(logits.device_ptr, logits.size, float(temperature), int(top_k), float(top_p), u,
                          None if random_buf is None else random_buf.device_ptr, out_buf.device_ptr, _st())


def sample_token_gpu(logits: GPUArray, temperature: float = 1.0, top_k: int = 0, top_p: float = 1.0,
                     random_val: float | None = None) -> int:
    """ops/sampling.py:11-41 signature; only the 4-byte token id comes back to the host."""
    out = GPUArray((1,), int32)
    sample_token_to_buf(logits, out, temperature, top_k, top_p, random_val)
    return int(out.to_numpy()[0])


def sample_multinomial(logits: GPUArray, temperature: float) -> int:
    """ops/sampling.py:91-105: draw from softmax(logits / temperature)."""
    if temperature <= 0:
        raise ValueError("sample_multinomial: temperature must be > 0")
    return sample_token_gpu(logits, temperature, 0, 1.0)


def sample_topk(logits: GPUArray, top_k: int, temperature: float) -> int:
    """ops/sampling.py:108-123: draw among the top_k largest logits (ties at the cut: highest indices)."""
    if temperature <= 0 or top_k <= 0:
        raise ValueError("sample_topk: temperature and top_k must be > 0")
    return sample_token_gpu(logits, temperature, top_k, 1.0)


def sample_topp(logits: GPUArray, top_p: float, temperature: float) -> int:
    """ops/sampling.py:126-141 with llm.sampling.sample_token's nucleus rule: the smallest set of most probable tokens
    whose mass reaches top_p (the reference's kernel scales its threshold by top_p instead of truncating)."""
    if temperature <= 0 or not 0 < top_p <= 1:
        raise ValueError("sample_topp: temperature must be > 0 and top_p in (0, 1]")
    return sample_token_gpu(logits, temperature, 0, top_p)


def sample_topk_to_buf_ptr(logits: GPUArray, result_buf: GPUArray, random_val_buf: GPUArray, top_k: int,
                           temperature: float) -> None:
    """ops/sampling.py:44-73: top-k sampling for graph replay -- the uniform number is read from `random_val_buf`
    (float32 [1] on the device), the token goes to `result_buf` (int32 [1]); nothing returns to the host."""
    if temperature <= 0 or top_k <= 0:
        raise ValueError("sample_topk_to_buf_ptr: temperature and top_k must be > 0")
    sample_token_to_buf(logits, result_buf, temperature, top_k, 1.0, random_buf=random_val_buf)


__all__ = ["sample_greedy", "sample_greedy_to_buf", "sample_token_gpu", "sample_token_to_buf", "set_sampling_seed",
           "sample_multinomial", "sample_topk", "sample_topp", "sample_topk_to_buf_ptr"]
