"""Host-side weight quantisation helpers for synthetic / random-init models.

FP8 block format = the reference's LinearFP8 (llm/layers/linear.py:101-179, llm/quant.py:30-57):
E4M3 codes [out, in] + bf16 `scale_inv` [out/128, in/128], weight = decode(code) * scale_inv.
The reference only *reads* FP8 checkpoints (it has no quantiser); this one is used to build
random-init weights of the benchmark shapes and never emits the 0x7F/0xFF codes.
"""
from __future__ import annotations

import numpy as np

from ..core.dtypes import bf16_bits_to_f32, f32_to_bf16_bits


def fp8_e4m3_table() -> np.ndarray:
    """256-entry decode table, CUDA-LUT semantics (fp8.cuh:57-122: 0x7F/0xFF -> +-480)."""
    i = np.arange(256)
    e, m = (i >> 3) & 0xF, i & 7
    v = np.where(e == 0, m * 2.0 ** -9, (1.0 + m / 8.0) * 2.0 ** (e.astype(np.float64) - 7))
    return np.where(i & 0x80, -v, v).astype(np.float32)


_POS = fp8_e4m3_table()[:0x7F].astype(np.float64)          # codes 0x00..0x7E ascending, max 448
_MID = (_POS[1:] + _POS[:-1]) / 2.0


def encode_fp8_e4m3(x: np.ndarray) -> np.ndarray:
    """float -> nearest E4M3 code (saturating at +-448, ties to even mantissa)."""
    a = np.abs(x).astype(np.float64)
    idx = np.searchsorted(_MID, a, side="left")             # a == mid -> lower code
    tie = (idx < len(_MID)) & (a == _MID[np.minimum(idx, len(_MID) - 1)])
    idx = np.where(tie & (idx % 2 == 1), idx + 1, idx)      # tie: pick the even code
    idx = np.minimum(idx, 0x7E)
    return (idx | np.where(np.signbit(x), 0x80, 0)).astype(np.uint8)


def quantize_fp8_e4m3_block(w: np.ndarray, block: int = 128) -> tuple[np.ndarray, np.ndarray]:
    """w [N,K] float32 -> (codes u8 [N,K], scale_inv bf16 bits [ceil(N/128), ceil(K/128)])."""
    N, K = w.shape
    bn, bk = (N + block - 1) // block, (K + block - 1) // block
    pad = np.zeros((bn * block, bk * block), np.float32)
    pad[:N, :K] = w
    blocks = pad.reshape(bn, block, bk, block)
    absmax = np.abs(blocks).max(axis=(1, 3))
    scale = np.where(absmax > 0, absmax / 448.0, 1.0).astype(np.float32)
    scale_bits = f32_to_bf16_bits(scale)
    s = bf16_bits_to_f32(scale_bits).reshape(bn, 1, bk, 1)
    codes = encode_fp8_e4m3(blocks / s).reshape(bn * block, bk * block)[:N, :K]
    return np.ascontiguousarray(codes), scale_bits.reshape(bn, bk)


def dequantize_fp8_e4m3_block(codes: np.ndarray, scale_bits: np.ndarray, block: int = 128) -> np.ndarray:
    """Inverse of the above (f32); same result as the reference's llm/quant.py:323-367 for
    codes other than 0x7F/0xFF."""
    N, K = codes.shape
    w = fp8_e4m3_table()[codes]
    s = bf16_bits_to_f32(scale_bits)
    s_full = np.repeat(np.repeat(s, block, axis=0), block, axis=1)[:N, :K]
    return (w * s_full).astype(np.float32)
