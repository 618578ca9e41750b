"""Norm layer (mirror of pygpukit/llm/layers/norm.py:19-44)."""
from __future__ import annotations

from typing import Literal

from ...core.array import GPUArray
from ...ops.nn import layernorm, rmsnorm


class Norm:
    """Unified RMSNorm / LayerNorm."""

    def __init__(self, weight: GPUArray, bias: GPUArray | None = None,
                 norm_type: Literal["rmsnorm", "layernorm"] = "rmsnorm", eps: float = 1e-5):
        self.weight = weight
        self.bias = bias
        self.norm_type = norm_type
        self.eps = eps

    def __call__(self, x: GPUArray, *, out: GPUArray | None = None) -> GPUArray:
        if self.norm_type == "rmsnorm":
            return rmsnorm(x, self.weight, self.eps, out=out)
        if self.bias is None:
            raise ValueError("LayerNorm requires bias")
        return layernorm(x, self.weight, self.bias, self.eps, out=out)


__all__ = ["Norm"]
