"""Embedding gather and KV-cache writes (mirror of pygpukit/ops/embedding.py:15-190)."""
from __future__ import annotations

from .._lib import api
from ..core.array import GPUArray
from ..core.dtypes import int32
from ._common import _st


def embedding_lookup(embed_matrix: GPUArray, out: GPUArray, token_id: int) -> None:
    """out[hidden] = embed_matrix[token_id] (ops/embedding.py:15-30)."""
    if embed_matrix.ndim != 2:
        raise ValueError("embedding_lookup expects 2D embed_matrix [vocab, hidden]")
    if not 0 <= int(token_id) < embed_matrix.shape[0]:
        raise ValueError(f"embedding_lookup: token_id {token_id} out of range [0, {embed_matrix.shape[0]})")
    if out.size != embed_matrix.shape[1] or out.dtype != embed_matrix.dtype:
        raise ValueError("embedding_lookup: out must hold one row of embed_matrix")
    api.embedding_lookup(embed_matrix.device_ptr, out.device_ptr, embed_matrix.shape[1], int(token_id),
                         embed_matrix.itemsize, _st())


def embedding_lookup_ptr(embed_matrix: GPUArray, out: GPUArray, token_id_buf: GPUArray) -> None:
    """Token id read from an int32 device buffer (ops/embedding.py:33-50)."""
    if token_id_buf.dtype != int32:
        raise ValueError("embedding_lookup_ptr: token_id_buf must be int32")
    api.embedding_lookup_ptr(embed_matrix.device_ptr, out.device_ptr, embed_matrix.shape[1],
                             token_id_buf.device_ptr, 1, embed_matrix.itemsize, _st())


def embedding_lookup_batch(embed_matrix: GPUArray, out: GPUArray, token_ids_buf: GPUArray, batch_size: int) -> None:
    """out[batch, hidden] = embed_matrix[token_ids[batch]] (ops/embedding.py:53-81)."""
    if token_ids_buf.dtype != int32:
        raise ValueError("embedding_lookup_batch: token_ids_buf must be int32")
    if batch_size > token_ids_buf.size or batch_size * embed_matrix.shape[1] > out.size:
        raise ValueError("embedding_lookup_batch: batch_size exceeds buffers")
    api.embedding_lookup_ptr(embed_matrix.device_ptr, out.device_ptr, embed_matrix.shape[1],
                             token_ids_buf.device_ptr, int(batch_size), embed_matrix.itemsize, _st())


def _kv(new_kv: GPUArray, cache: GPUArray, num_heads: int, pos: int, pos_buf, name: str) -> None:
    if new_kv.ndim != 3 or cache.ndim != 3:
        raise ValueError(f"{name} expects new_kv [seq, num_kv_heads, head_dim] and cache [num_heads, max_seq_len, head_dim]")
    if new_kv.dtype != cache.dtype:
        raise ValueError(f"{name}: new_kv and cache must have same dtype")
    seq, n_kv, hd = new_kv.shape
    if cache.shape[0] != num_heads or cache.shape[2] != hd:
        raise ValueError(f"{name}: cache shape {cache.shape} does not match num_heads={num_heads}, head_dim={hd}")
    if num_heads % n_kv != 0:
        raise ValueError(f"{name}: num_heads {num_heads} must be a multiple of num_kv_heads {n_kv}")
    if pos_buf is None and (pos < 0 or pos + seq > cache.shape[1]):
        raise ValueError(f"{name}: position {pos} + {seq} exceeds max_seq_len {cache.shape[1]}")
    api.kv_cache_update(new_kv.device_ptr, cache.device_ptr, seq, n_kv, num_heads, hd, cache.shape[1], int(pos),
                        pos_buf.device_ptr if pos_buf is not None else None, new_kv.itemsize, _st())


def kv_cache_update(new_kv: GPUArray, cache: GPUArray, position: int) -> None:
    """cache [num_kv_heads... ] un-expanded single-position write (ops/embedding.py:84-103)."""
    _kv(new_kv, cache, cache.shape[0], position, None, "kv_cache_update")


def kv_cache_prefill(new_kv: GPUArray, cache: GPUArray, start_pos: int = 0) -> None:
    _kv(new_kv, cache, cache.shape[0], start_pos, None, "kv_cache_prefill")


def kv_cache_update_gqa(new_kv: GPUArray, cache: GPUArray, num_heads: int, position: int) -> None:
    """new_kv [1, num_kv_heads, hd] -> GQA-expanded cache [num_heads, max_seq_len, hd] (ops/embedding.py:128-145)."""
    _kv(new_kv, cache, num_heads, position, None, "kv_cache_update_gqa")


def kv_cache_prefill_gqa(new_kv: GPUArray, cache: GPUArray, num_heads: int, start_pos: int = 0) -> None:
    _kv(new_kv, cache, num_heads, start_pos, None, "kv_cache_prefill_gqa")


def kv_cache_update_gqa_ptr(new_kv: GPUArray, cache: GPUArray, num_heads: int, position_buf: GPUArray) -> None:
    if position_buf.dtype != int32:
        raise ValueError("kv_cache_update_gqa_ptr: position_buf must be int32")
    _kv(new_kv, cache, num_heads, 0, position_buf, "kv_cache_update_gqa_ptr")


__all__ = ["embedding_lookup", "embedding_lookup_ptr", "embedding_lookup_batch", "kv_cache_update",
           "kv_cache_prefill", "kv_cache_update_gqa", "kv_cache_prefill_gqa", "kv_cache_update_gqa_ptr"]
