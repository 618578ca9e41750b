"""Pre-allocated decode buffers (mirror of pygpukit/llm/buffers.py:26-466, reduced to what the
fused B200 step touches).  `allocate(config, dtype, use_qk_norm, vocab_size, max_batch_size)` keeps
the reference's signature; rows = max(1, max_batch_size).  Extra (net-new) members carry the
device-side step state for graph replay: token_buf / position_buf / context_buf (int32)."""
from __future__ import annotations

from dataclasses import dataclass

from ..core.array import GPUArray
from ..core.dtypes import bfloat16, int32
from ..core.factory import zeros
from .config import TransformerConfig


@dataclass
class DecodeBuffers:
    hidden: GPUArray          # [M, H]   residual stream, updated in place
    norm_out: GPUArray        # [M, H]
    qkv_proj_out: GPUArray    # [M, (Hq + 2 Hkv) * hd]  (local heads under TP)
    attn_flat: GPUArray       # [M, Hq * hd]
    mlp_act: GPUArray         # [M, I_local]
    partial: GPUArray         # [M, H]   row-parallel partial sums (TP only)
    logits: GPUArray | None   # [M, V_local]
    token_buf: GPUArray       # int32 [M]  current token ids
    position_buf: GPUArray    # int32 [1]
    context_buf: GPUArray     # int32 [1]
    next_token: GPUArray      # int32 [M]  greedy result
    max_batch_size: int = 1
    num_sequences: int = 1
    state_buf: GPUArray | None = None   # int32 [2 S + M] = [positions, contexts, tokens...]: ONE H2D copy per step

    # aliases with the reference's member names (buffers.py:26-188)
    @property
    def residual(self) -> GPUArray:
        return self.partial

    @property
    def attn_out(self) -> GPUArray:
        return self.attn_flat

    @property
    def mlp_gate(self) -> GPUArray:
        return self.mlp_act

    @classmethod
    def allocate(cls, config: TransformerConfig, dtype: str = "bfloat16", use_qk_norm: bool = False,
                 vocab_size: int | None = None, max_batch_size: int = 0, moe_config: dict | None = None,
                 num_sequences: int = 1) -> "DecodeBuffers":
        if dtype != "bfloat16":
            raise ValueError("the B200 decode path runs in bfloat16")
        M = max(1, max_batch_size)
        tp = max(1, config.tp_size)
        H, hd = config.hidden_size, config.head_dim
        # tp > num_kv_heads: ranks replicate a whole KV head (same rule as Attention.__init__)
        hq, hkv = config.num_heads // tp, max(1, config.num_kv_heads // tp)
        inter = (config.moe_intermediate_size if config.is_moe else config.intermediate_size) // (1 if config.is_moe else tp)
        V = vocab_size if vocab_size is not None else config.vocab_size
        S = max(1, num_sequences)
        # decode state, ONE host-to-device copy per step: [position x S | context x S | token x M]
        state = zeros((2 * S + M,), int32)
        return cls(
            state_buf=state,
            hidden=zeros((M, H), bfloat16), norm_out=zeros((M, H), bfloat16),
            qkv_proj_out=zeros((M, (hq + 2 * hkv) * hd), bfloat16), attn_flat=zeros((M, hq * hd), bfloat16),
            mlp_act=zeros((M, inter), bfloat16), partial=zeros((M, H), bfloat16),
            logits=zeros((M, V), bfloat16) if V else None,
            token_buf=state.narrow(2 * S, M), position_buf=state.narrow(0, S), context_buf=state.narrow(S, S),
            next_token=zeros((M,), int32), max_batch_size=M, num_sequences=S)


@dataclass
class PrefillBuffers:
    """Kept for API parity (buffers.py:467-621); prefill reuses DecodeBuffers in slabs of rows."""
    max_seq_len: int = 0

    @classmethod
    def allocate(cls, config: TransformerConfig, max_seq_len: int, dtype: str = "bfloat16", use_qk_norm: bool = False):
        return cls(max_seq_len=max_seq_len)


__all__ = ["DecodeBuffers", "PrefillBuffers"]
