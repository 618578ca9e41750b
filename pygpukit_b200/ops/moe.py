"""MoE routing ops (mirror of the native.moe_* bindings, native/bindings/moe.cpp:43-223, as used by
llm/layers/moe.py:153-318)."""
from __future__ import annotations

from .._lib import api
from ..core.array import GPUArray
from ..core.dtypes import bfloat16, float_code, int32
from ._common import _st


def moe_topk_with_indices(logits: GPUArray, values: GPUArray, indices: GPUArray, k: int) -> None:
    if logits.ndim != 2:
        raise ValueError("moe_topk_with_indices expects logits [num_tokens, num_experts]")
    T, E = logits.shape
    if values.size < T * k or indices.size < T * k or indices.dtype != int32 or values.dtype != logits.dtype:
        raise ValueError("moe_topk_with_indices: values/indices buffers do not match [num_tokens, k]")
    api.moe_topk(logits.device_ptr, values.device_ptr, indices.device_ptr, T, E, int(k), float_code(logits.dtype), _st())


def moe_softmax_topk(values: GPUArray, k: int) -> None:
    T = values.size // k
    api.moe_softmax_topk(values.device_ptr, T, int(k), float_code(values.dtype), _st())


def moe_compute_permutation(expert_indices: GPUArray, expert_counts: GPUArray, expert_offsets: GPUArray,
                            permute_indices: GPUArray, reverse_perm: GPUArray, num_experts: int, k: int) -> None:
    T = expert_indices.size // k
    api.moe_compute_permutation(expert_indices.device_ptr, expert_counts.device_ptr, expert_offsets.device_ptr,
                                permute_indices.device_ptr, reverse_perm.device_ptr, T, int(num_experts), int(k), _st())


def moe_expand_expert_offsets(expert_offsets: GPUArray, row_expert_ids: GPUArray, num_experts: int) -> None:
    api.moe_expand_expert_offsets(expert_offsets.device_ptr, row_expert_ids.device_ptr, int(num_experts),
                                  row_expert_ids.size, _st())


def moe_gather(hidden: GPUArray, permute_indices: GPUArray, gathered: GPUArray, k: int) -> None:
    if hidden.dtype != bfloat16:
        raise ValueError("moe_gather requires bfloat16 hidden states")
    T, H = hidden.shape
    api.moe_gather(hidden.device_ptr, permute_indices.device_ptr, gathered.device_ptr, T, H, int(k), _st())


def moe_scatter(expert_out: GPUArray, router_weights: GPUArray, reverse_perm: GPUArray, output: GPUArray, k: int) -> None:
    if expert_out.dtype != bfloat16:
        raise ValueError("moe_scatter requires bfloat16 expert outputs")
    T, H = output.shape
    api.moe_scatter(expert_out.device_ptr, router_weights.device_ptr, reverse_perm.device_ptr, output.device_ptr,
                    T, H, int(k), _st())


__all__ = ["moe_topk_with_indices", "moe_softmax_topk", "moe_compute_permutation", "moe_expand_expert_offsets",
           "moe_gather", "moe_scatter"]
