"""Mixture-of-experts layer (mirror of pygpukit/llm/layers/moe.py:26-458).

Same pipeline as the reference's `MoELayer.__call__` (moe.py:153-318): router GEMV -> top-k with
indices (strict >, lowest index wins ties) -> softmax over the selected logits -> permutation by
expert -> gather -> expert SwiGLU as three expert-grouped GEMVs (row m uses expert
row_expert_ids[m]) -> weighted scatter-combine in slot order.  Every step is a libpgk_b200 kernel
on the current stream; no host synchronisation, no per-call allocation when `bufs` are given.

Experts are held STACKED ([E, N, K] FP8 + [E, N/128, K/128] scales as the reference's
`_stack_fp8_weights`, moe.py:86-151; or [E, K/2, N] + [E, K/32, N] for the net-new NVF4 experts).
Under tensor parallelism every rank holds E/tp experts and the router is replicated.  Rows routed to
remote experts come out as zeros; the UN-COMBINED down-projection rows [T*k, H] are then all-reduced
(every row has exactly one non-zero contributor, so the sum is exact in any order) and the
slot-ordered weighted combine + the single BF16 rounding + the residual add run identically on every
rank -- the same arithmetic, in the same order, as on one GPU (scatter_with_reverse_perm_kernel,
permute_kernels.cuh:218-242), so greedy tokens cannot depend on the number of ranks.
"""
from __future__ import annotations

import os
from typing import TYPE_CHECKING

import numpy as np

from ..._lib import api
from ...core.array import GPUArray
from ...core.dtypes import bfloat16, int32, uint8
from ...core.factory import from_numpy
from ...core.stream import current_stream_handle as _st
from .linear import LinearBF16, LinearFP8, LinearNVF4
from .mlp import MLP

if TYPE_CHECKING:
    from ..config import TransformerConfig


_DECODE_FORM = os.environ.get("PGK_MOE_T1", "1") != "0"      # 0: always the generic pipeline (A/B, parity tests)


def _stack(arrs: list[GPUArray]) -> GPUArray:
    first = arrs[0]
    out = GPUArray((len(arrs),) + tuple(first.shape), first.dtype)
    for i, a in enumerate(arrs):
        api.memcpy_d2d(out.device_ptr + i * first.nbytes, a.device_ptr, first.nbytes, _st())
    return out


class MoELayer:
    def __init__(self, config: "TransformerConfig", gate_weight: GPUArray, expert_weights: list | None = None,
                 stacked: dict | None = None, expert_offset: int = 0):
        """expert_weights: [(gate, up, down), ...] of LinearFP8 / LinearNVF4 (reference signature), or
        stacked = {"format": "fp8"|"nvf4", "gate": (w, s), "up": (w, s), "down": (w, s)} device tensors.
        expert_offset: global id of this rank's first expert (tensor/expert parallel)."""
        self.config = config
        self.num_experts = config.num_experts or len(expert_weights or [])
        self.num_experts_per_tok = config.num_experts_per_tok
        self.hidden_size = config.hidden_size
        self.intermediate_size = config.moe_intermediate_size or config.intermediate_size
        self.gate = LinearBF16(gate_weight)
        self.expert_offset = expert_offset
        self.experts: list[MLP] = []
        if stacked is None:
            if not expert_weights:
                raise ValueError("MoELayer needs expert_weights or stacked tensors")
            self.experts = [MLP(config, gate_proj=g, up_proj=u, down_proj=d) for g, u, d in expert_weights]
            g0 = self.experts[0].gate_proj
            if isinstance(g0, LinearFP8):
                stacked = {"format": "fp8"}
                for name in ("gate", "up", "down"):
                    ls = [getattr(e, name + "_proj") for e in self.experts]
                    stacked[name] = (_stack([l.weight_fp8 for l in ls]), _stack([l.scale_inv for l in ls]))
            elif isinstance(g0, LinearNVF4):
                stacked = {"format": "nvf4"}
                for name in ("gate", "up", "down"):
                    ls = [getattr(e, name + "_proj") for e in self.experts]
                    stacked[name] = (_stack([l.data for l in ls]), _stack([l.scale for l in ls]))
            else:
                raise ValueError("MoELayer experts must be LinearFP8 or LinearNVF4 on the B200 path")
        self.format = stacked["format"]
        self._w = stacked
        self.num_local_experts = stacked["gate"][0].shape[0]
        self._use_grouped_gemm = True
        self._bufs: dict[int, dict] = {}

    def _buffers(self, T: int) -> dict:
        b = self._bufs.get(T)
        if b is None:
            k, H, I, E = self.num_experts_per_tok, self.hidden_size, self.intermediate_size, self.num_experts
            R = T * k
            b = {
                "logits": GPUArray((T, E), bfloat16), "weights": GPUArray((T, k), bfloat16),
                "indices": GPUArray((T, k), int32), "counts": GPUArray((E,), int32),
                "offsets": GPUArray((E + 1,), int32), "permute": GPUArray((R,), int32),
                "reverse": GPUArray((R,), int32), "row_ids": GPUArray((R,), int32),
                "gathered": GPUArray((R, H), bfloat16), "act": GPUArray((R, I), bfloat16),
                "up": GPUArray((R, I), bfloat16), "down": GPUArray((R, H), bfloat16),
                "combine": GPUArray((T, H), bfloat16),
            }
            self._bufs[T] = b
        return b

    def _grouped(self, name: str, a: GPUArray, out: GPUArray, row_ids: GPUArray, K: int, N: int, mode: int = 0,
                 second: str | None = None) -> None:
        w, s = self._w[name]
        R = a.shape[0]
        if self.format == "fp8":
            api.moe_grouped_gemv_fp8_ep(a.device_ptr, w.device_ptr, s.device_ptr, out.device_ptr, row_ids.device_ptr,
                                        R, K, N, self.expert_offset, self.num_local_experts, _st())
        else:
            api.moe_grouped_gemv_nvf4(a.device_ptr, w.device_ptr, s.device_ptr, out.device_ptr, row_ids.device_ptr,
                                      R, K, N, self.expert_offset, self.num_local_experts, _st())

    def forward(self, x: GPUArray, *, out: GPUArray | None = None, residual: GPUArray | None = None,
                act_buf: GPUArray | None = None, comm=None) -> GPUArray:
        """comm: the TensorParallel communicator when the experts are sharded over ranks -- the result is then the
        COMPLETE layer output on every rank (the exchange happens inside, on the un-combined rows)."""
        if x.ndim != 2:
            raise ValueError("MoELayer expects [tokens, hidden]")
        T, H = x.shape
        k, I, E = self.num_experts_per_tok, self.intermediate_size, self.num_experts
        b = self._buffers(T)
        st = _st()
        tp = comm is not None and comm.world > 1
        self.gate.forward(x, out=b["logits"])
        if T == 1 and _DECODE_FORM and k <= 32:
            return self._forward_t1(x, b, out, residual, st, comm if tp else None)
        api.moe_topk(b["logits"].device_ptr, b["weights"].device_ptr, b["indices"].device_ptr, T, E, k, 2, st)
        api.moe_softmax_topk(b["weights"].device_ptr, T, k, 2, st)
        api.moe_compute_permutation(b["indices"].device_ptr, b["counts"].device_ptr, b["offsets"].device_ptr,
                                    b["permute"].device_ptr, b["reverse"].device_ptr, T, E, k, st)
        api.moe_expand_expert_offsets(b["offsets"].device_ptr, b["row_ids"].device_ptr, E, T * k, st)
        api.moe_gather(x.device_ptr, b["permute"].device_ptr, b["gathered"].device_ptr, T, H, k, st)
        # expert SwiGLU: gate -> silu, up, mul (reference rounding: silu rounded, then bf16 multiply)
        self._grouped("gate", b["gathered"], b["act"], b["row_ids"], H, I)
        api.silu(b["act"].device_ptr, b["act"].device_ptr, b["act"].size, 2, st)
        self._grouped("up", b["gathered"], b["up"], b["row_ids"], H, I)
        api.mul_inplace(b["act"].device_ptr, b["up"].device_ptr, b["act"].size, 2, st)
        self._grouped("down", b["act"], b["down"], b["row_ids"], I, H)
        if tp:
            comm.allreduce_rows(b["down"])       # rows of remote experts are zeros here: exact sum, any order
        y = out if out is not None else GPUArray((T, H), bfloat16)
        target = y if residual is None else b["combine"]
        api.moe_scatter(b["down"].device_ptr, b["weights"].device_ptr, b["reverse"].device_ptr, target.device_ptr,
                        T, H, k, st)
        if residual is not None:
            api.add(target.device_ptr, residual.device_ptr, y.device_ptr, T * H, 2, st)
        return y

    def _forward_t1(self, x: GPUArray, b: dict, out: GPUArray | None, residual: GPUArray | None, st,
                    comm=None) -> GPUArray:
        """One token: rows stay in slot order (the combine sums in slot order anyway), so routing is one
        launch and permutation / expand / gather disappear; the experts run as one gate+up+silu+mul kernel and
        (FP8) one down+combine(+residual) kernel, (NVF4) down then combine(+residual).  Bit-identical to the generic pipeline (tests/test_gpu_moe.py)."""
        H, I, E, k = self.hidden_size, self.intermediate_size, self.num_experts, self.num_experts_per_tok
        off, nl = self.expert_offset, self.num_local_experts
        api.moe_route_t1(b["logits"].device_ptr, b["weights"].device_ptr, b["indices"].device_ptr,
                         b["row_ids"].device_ptr, E, k, st)
        y = out if out is not None else GPUArray((1, H), bfloat16)
        res = residual.device_ptr if residual is not None else None
        if self.format == "fp8":
            (wg, sg), (wu, su), (wd, sd) = self._w["gate"], self._w["up"], self._w["down"]
            api.moe_gateup_fp8_t1(x.device_ptr, wg.device_ptr, sg.device_ptr, wu.device_ptr, su.device_ptr,
                                  b["act"].device_ptr, b["row_ids"].device_ptr, k, H, I, off, nl, st)
            if comm is None:
                api.moe_down_combine_fp8_t1(b["act"].device_ptr, wd.device_ptr, sd.device_ptr, b["weights"].device_ptr,
                                            b["row_ids"].device_ptr, res, y.device_ptr, k, I, H, off, nl, st)
                return y
            # expert parallel: down rows (zeros for remote experts) -> exact all-reduce -> the same combine everywhere
            api.moe_grouped_gemv_fp8_ex(b["act"].device_ptr, I, wd.device_ptr, sd.device_ptr, b["down"].device_ptr,
                                        b["row_ids"].device_ptr, k, I, H, off, nl, st)
            comm.allreduce_rows(b["down"])
            api.moe_combine_t1(b["down"].device_ptr, b["weights"].device_ptr, res, y.device_ptr, H, k, st)
            return y
        (wg, sg), (wu, su), (wd, sd) = self._w["gate"], self._w["up"], self._w["down"]
        rid = b["row_ids"].device_ptr
        api.moe_gateup_nvf4_t1(x.device_ptr, wg.device_ptr, sg.device_ptr, wu.device_ptr, su.device_ptr,
                               b["act"].device_ptr, rid, k, H, I, off, nl, st)
        api.moe_grouped_gemv_nvf4_ex(b["act"].device_ptr, I, wd.device_ptr, sd.device_ptr, b["down"].device_ptr, rid, k, I, H, off, nl, st)
        if comm is not None:
            comm.allreduce_rows(b["down"])
        api.moe_combine_t1(b["down"].device_ptr, b["weights"].device_ptr, res, y.device_ptr, H, k, st)
        return y

    def __call__(self, x: GPUArray) -> GPUArray:
        return self.forward(x)

    @property
    def weight_bytes(self) -> int:
        """Bytes one token touches: router + top-k experts (not the whole expert bank)."""
        per = sum(self._w[n][0].nbytes + self._w[n][1].nbytes for n in ("gate", "up", "down")) // self.num_local_experts
        return self.gate.weight_bytes + per * self.num_experts_per_tok


__all__ = ["MoELayer"]
_ = (np, uint8, from_numpy)
