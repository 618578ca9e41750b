"""MLP layer (mirror of pygpukit/llm/layers/mlp.py:25-103).

SwiGLU: down( bf16(bf16(silu(gate(x))) * up(x)) ) -- the reference runs gate GEMV, silu, up GEMV,
mul_inplace, down GEMV as five launches (models/causal.py:489-514); here gate+up+silu+mul is ONE
weight-streaming kernel with the same rounding points, and the residual add rides in the down
GEMV's epilogue.  GELU (GPT-2): fc1 -> gelu -> fc2.
"""
from __future__ import annotations

from typing import TYPE_CHECKING

import numpy as np

from ..._lib import api
from ...core.array import GPUArray
from ...core.dtypes import bfloat16, uint8
from ...core.factory import from_numpy
from ...core.stream import current_stream_handle as _st
from ...ops.elementwise import mul
from ...ops.nn import gelu, silu
from .linear import LinearBF16, LinearFP8, LinearNVF4, can_fuse_norm

if TYPE_CHECKING:
    from ..config import TransformerConfig

_LINEARS = (LinearBF16, LinearFP8, LinearNVF4)


def fuse_gate_up_nvf4(gate: LinearNVF4, up: LinearNVF4) -> LinearNVF4:
    """Interleave the columns of two NVF4 matrices in groups of 8 ([g0..g7, u0..u7, g8..]) -- the
    layout pgk_gemv_nvf4_swiglu_bf16 streams (include/pgk_b200.h).  Init-time, through the host."""
    K, I = gate.in_features, gate.out_features
    if (up.in_features, up.out_features) != (K, I) or I % 8:
        raise ValueError("fuse_gate_up_nvf4: gate/up shapes must match and I must be a multiple of 8")

    def il(a: np.ndarray, b: np.ndarray) -> np.ndarray:
        r = a.shape[0]
        return np.stack([a.reshape(r, I // 8, 8), b.reshape(r, I // 8, 8)], axis=2).reshape(r, 2 * I)
    data = il(gate.data.to_numpy().reshape(K // 2, I), up.data.to_numpy().reshape(K // 2, I))
    scale = il(gate.scale.to_numpy().reshape(-1, I), up.scale.to_numpy().reshape(-1, I))
    return LinearNVF4(from_numpy(np.ascontiguousarray(data)), from_numpy(np.ascontiguousarray(scale)), K, 2 * I)


class MLP:
    def __init__(self, config: "TransformerConfig", fc1_weight=None, fc1_bias: GPUArray | None = None,
                 fc2_weight=None, fc2_bias: GPUArray | None = None, gate_proj=None, up_proj=None, down_proj=None,
                 gate_up_proj: LinearNVF4 | None = None):
        self.config = config
        self.activation = config.activation
        self._scratch: dict[int, tuple] = {}

        def wrap(p, bias=None):
            if p is None or isinstance(p, _LINEARS):
                return p
            return LinearBF16(p, bias)

        if config.activation == "gelu":
            if fc1_weight is None or fc2_weight is None:
                raise ValueError("GELU MLP requires fc1_weight and fc2_weight")
            self.fc1, self.fc2 = wrap(fc1_weight, fc1_bias), wrap(fc2_weight, fc2_bias)
            self.intermediate_size = self.fc1.out_features
        else:
            if down_proj is None or (gate_up_proj is None and (gate_proj is None or up_proj is None)):
                raise ValueError("SwiGLU MLP requires gate_proj, up_proj, down_proj")
            self.gate_proj, self.up_proj, self.down_proj = wrap(gate_proj), wrap(up_proj), wrap(down_proj)
            # NVF4: one fused, column-interleaved gate/up tensor (net-new layout)
            self.gate_up_proj = gate_up_proj
            if gate_up_proj is None and isinstance(self.gate_proj, LinearNVF4):
                self.gate_up_proj = fuse_gate_up_nvf4(self.gate_proj, self.up_proj)
            self.intermediate_size = (self.gate_up_proj.out_features // 2 if self.gate_up_proj is not None
                                      else self.gate_proj.out_features)

    # ---- fused activation( gate/up ) -> act_out [M, I]
    def fuses_norm(self, norm) -> bool:
        """True when `norm` can ride inside the gate/up launch (FP8 / NVF4 fused SwiGLU kernels)."""
        if self.activation != "silu" or not can_fuse_norm(norm, self.config.hidden_size):
            return False
        return self.gate_up_proj is not None or (
            isinstance(self.gate_proj, LinearFP8) and isinstance(self.up_proj, LinearFP8)
            and self.gate_proj.bias is None and self.up_proj.bias is None)

    def gate_up_act(self, x: GPUArray, *, out: GPUArray | None = None, norm=None) -> GPUArray:
        M, K, I = x.shape[0], x.shape[1], self.intermediate_size
        if out is None:
            out = GPUArray((M, I), bfloat16)
        if (self.gate_up_proj is None and isinstance(self.gate_proj, LinearFP8) and isinstance(self.up_proj, LinearFP8)
                and self.gate_proj.uses_tensor_cores(M) and self.up_proj.uses_tensor_cores(M)):
            # batch >= TC_MIN_M: gate and up on the tcgen05 GEMM, then silu -> mul with the layer path's two roundings
            if norm is not None:
                xs = self._scratch.get(("x", M))
                if xs is None:
                    xs = self._scratch[("x", M)] = GPUArray((M, K), bfloat16)
                x = norm(x, out=xs)
            sc = self._scratch.get(M)
            if sc is None:
                sc = self._scratch[M] = (GPUArray((M, I), bfloat16), GPUArray((M, I), bfloat16))
            gate = silu(self.gate_proj.forward(x, out=sc[0]), out=sc[0])
            return mul(gate, self.up_proj.forward(x, out=sc[1]), out=out)
        if self.gate_up_proj is not None and self.gate_up_proj.uses_tensor_cores(M):
            # batch >= TC_MIN_M: the fused gate/up tensor on the tcgen05 GEMM (weights read once for all rows), then
            # silu -> mul over the interleaved columns with the layer path's two roundings
            sc = self._scratch.get(("gu", M))
            if sc is None:
                sc = self._scratch[("gu", M)] = GPUArray((M, 2 * I), bfloat16)
            self.gate_up_proj.forward(x, out=sc, norm=norm)
            api.swiglu_interleaved8_bf16(sc.device_ptr, out.device_ptr, M, I, _st())
            return out
        if norm is not None:
            g, e = norm.weight.device_ptr, float(norm.eps)
            if self.gate_up_proj is not None:
                gu = self.gate_up_proj
                api.rmsnorm_gemv_nvf4_swiglu_bf16(x.device_ptr, g, e, gu.data.device_ptr, gu.scale.device_ptr,
                                                  out.device_ptr, M, K, 2 * I, _st())
            else:
                gp, up = self.gate_proj, self.up_proj
                api.rmsnorm_gemv_fp8_swiglu_bf16(x.device_ptr, g, e, gp.weight_fp8.device_ptr, gp.scale_inv.device_ptr,
                                                 up.weight_fp8.device_ptr, up.scale_inv.device_ptr, out.device_ptr,
                                                 M, K, I, _st())
            return out
        if self.gate_up_proj is not None:
            gu = self.gate_up_proj
            api.gemv_nvf4_swiglu_bf16(x.device_ptr, gu.data.device_ptr, gu.scale.device_ptr, out.device_ptr, M, K,
                                      2 * I, _st())
        elif isinstance(self.gate_proj, LinearFP8) and isinstance(self.up_proj, LinearFP8) \
                and self.gate_proj.bias is None and self.up_proj.bias is None:
            g, u = self.gate_proj, self.up_proj
            api.gemv_fp8_swiglu_bf16(x.device_ptr, g.weight_fp8.device_ptr, g.scale_inv.device_ptr,
                                     u.weight_fp8.device_ptr, u.scale_inv.device_ptr, out.device_ptr, M, K, I, _st())
        else:
            # unfused formats (BF16): gate -> silu -> up -> mul with cached scratch (graph-capture safe)
            sc = self._scratch.get(M)
            if sc is None:
                sc = self._scratch[M] = (GPUArray((M, I), bfloat16), GPUArray((M, I), bfloat16))
            gate = silu(self.gate_proj(x, out=sc[0]), out=sc[0])
            mul(gate, self.up_proj(x, out=sc[1]), out=out)
        return out

    def forward(self, x: GPUArray, *, out: GPUArray | None = None, residual: GPUArray | None = None,
                act_buf: GPUArray | None = None, norm=None) -> GPUArray:
        if self.activation == "gelu":
            M = x.shape[0]
            h = act_buf
            if h is None or tuple(h.shape) != (M, self.intermediate_size):   # cached scratch: graph-capture safe
                h = self._scratch.get(M)
                if h is None:
                    h = self._scratch[M] = GPUArray((M, self.intermediate_size), bfloat16)
            self.fc1.forward(x, out=h)
            gelu(h, out=h)
            return self.fc2.forward(h, out=out, residual=residual)
        h = self.gate_up_act(x, out=act_buf, norm=norm)
        return self.down_proj.forward(h, out=out, residual=residual)

    def forward_allreduce(self, x: GPUArray, hidden: GPUArray, *, act_buf: GPUArray | None = None, norm=None) -> GPUArray:
        """Tensor parallel, one token: hidden <- hidden + sum over ranks of the shard's MLP output, the down projection and the
        all-reduce being one launch (TensorParallel.fuses_allreduce(self.down_proj, 1) must hold)."""
        h = self.gate_up_act(x, out=act_buf, norm=norm)
        return self.down_proj.forward_allreduce(h, hidden)

    def __call__(self, x: GPUArray) -> GPUArray:
        return self.forward(x)

    @property
    def weight_bytes(self) -> int:
        if self.activation == "gelu":
            return self.fc1.weight_bytes + self.fc2.weight_bytes
        gu = (self.gate_up_proj.weight_bytes if self.gate_up_proj is not None
              else self.gate_proj.weight_bytes + self.up_proj.weight_bytes)
        return gu + self.down_proj.weight_bytes


__all__ = ["MLP", "fuse_gate_up_nvf4"]
_ = uint8
