"""Transformer block (mirror of pygpukit/llm/layers/block.py:18-62): Norm -> Attention -> residual,
Norm -> MLP/MoE -> residual.  Residual adds ride in the o_proj / down_proj GEMV epilogues with the
fixed-cache path's rounding (bf16 + bf16 -> one rounding, elementwise_kernels.cuh:381-390).  Under
tensor parallelism the row-parallel partial sums are all-reduced BEFORE the residual is added."""
from __future__ import annotations

from ...core.array import GPUArray
from .attention import Attention
from .linear import can_fuse_norm
from .mlp import MLP
from .moe import MoELayer
from .norm import Norm


class TransformerBlock:
    def __init__(self, attn_norm: Norm, attn: Attention, mlp_norm: Norm, mlp: "MLP | MoELayer", comm=None):
        self.attn_norm = attn_norm
        self.attn = attn
        self.mlp_norm = mlp_norm
        self.mlp = mlp
        self.comm = comm   # TensorParallel communicator or None

    def decode(self, hidden: GPUArray, position: int, context_len: int, bufs=None, pos_buf=None, ctx_buf=None,
               multi: bool = False) -> GPUArray:
        """hidden [M, H] -> hidden [M, H], updated IN PLACE (M consecutive tokens of one sequence; multi = True: the
        new tokens of M independent sequences, pos_buf / ctx_buf are then arrays of M)."""
        M = hidden.shape[0]

        def buf(name):
            b = getattr(bufs, name, None) if bufs is not None else None
            return b if (b is None or b.shape[0] == M) else b.slice_rows(M)

        tp = self.comm is not None and self.comm.world > 1
        # pre-norms ride inside the qkv / gate-up GEMV launches when the formats allow it
        if multi:
            fuse = can_fuse_norm(self.attn_norm, hidden.shape[1])
            x = hidden if fuse else self.attn_norm(hidden, out=buf("norm_out"))
            a = self.attn.attend_multi(x, pos_buf, ctx_buf, qkv_buf=buf("qkv_proj_out"), attn_buf=buf("attn_flat"),
                                       norm=self.attn_norm if fuse else None)
        elif can_fuse_norm(self.attn_norm, hidden.shape[1]):
            a = self.attn.attend(hidden, position, context_len, qkv_buf=buf("qkv_proj_out"), attn_buf=buf("attn_flat"),
                                 pos_buf=pos_buf, ctx_buf=ctx_buf, norm=self.attn_norm)
        else:
            x = self.attn_norm(hidden, out=buf("norm_out"))
            a = self.attn.attend(x, position, context_len, qkv_buf=buf("qkv_proj_out"), attn_buf=buf("attn_flat"),
                                 pos_buf=pos_buf, ctx_buf=ctx_buf)
        if tp and self.comm.fuses_allreduce(self.attn.o_proj, M):
            self.attn.o_proj.forward_allreduce(a, hidden)          # GEMV + all-reduce + residual: one launch
        elif tp:
            part = self.attn.o_proj.forward(a, out=buf("partial"))
            self.comm.allreduce_add(part, hidden)
        else:
            self.attn.o_proj.forward(a, out=hidden, residual=hidden)
        fuse_mlp = isinstance(self.mlp, MLP) and self.mlp.fuses_norm(self.mlp_norm)
        x, nrm = (hidden, self.mlp_norm) if fuse_mlp else (self.mlp_norm(hidden, out=buf("norm_out")), None)
        kw = {"norm": nrm} if fuse_mlp else {}
        if tp and isinstance(self.mlp, MoELayer):
            # expert parallel: the exchange happens inside the layer, on the un-combined rows (bit-identical to 1 GPU)
            self.mlp.forward(x, out=hidden, residual=hidden, act_buf=buf("mlp_act"), comm=self.comm)
        elif tp and isinstance(self.mlp, MLP) and self.mlp.activation != "gelu" and self.comm.fuses_allreduce(self.mlp.down_proj, M):
            self.mlp.forward_allreduce(x, hidden, act_buf=buf("mlp_act"), **kw)
        elif tp:
            part = self.mlp.forward(x, out=buf("partial"), act_buf=buf("mlp_act"), **kw)
            self.comm.allreduce_add(part, hidden)
        else:
            self.mlp.forward(x, out=hidden, residual=hidden, act_buf=buf("mlp_act"), **kw)
        return hidden

    def __call__(self, x: GPUArray, position_ids: list[int] | None = None, past_kv=None, use_cache: bool = False):
        start = position_ids[0] if position_ids else (past_kv or 0)
        h = x.clone()
        self.decode(h, start, start + x.shape[0])
        return h, (start + x.shape[0] if use_cache else None)

    @property
    def weight_bytes(self) -> int:
        return (self.attn.weight_bytes + self.mlp.weight_bytes + self.attn_norm.weight.nbytes
                + self.mlp_norm.weight.nbytes)


__all__ = ["TransformerBlock"]
