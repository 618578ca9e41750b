"""Linear layers of the decode path (mirror of pygpukit/llm/layers/linear.py:25-267).

LinearBF16 / LinearFP8 keep the reference's constructor arguments, attributes (`weight`,
`weight_fp8`, `scale_inv`, `bias`, `in_features`, `out_features`, `_use_gemv`) and call
signature `layer(x, *, out=None)`; LinearNVF4 is net-new (the reference has NVF4 kernels but no
NVF4 layer).  Differences, all on the dispatch side:
  * every batch size 1..64 goes to the weight-streaming GEMV family of libpgk_b200.so, which
    reads the weights once for all rows -- there is no [K,N] transposed copy, no cuBLASLt matmul,
    no CPU dequantisation fallback;
  * bias and residual adds are fused into the GEMV epilogue with the reference's rounding points
    (`forward(x, out=, residual=)`).
"""
from __future__ import annotations

import os

import numpy as np

from ..._lib import api
from ...core.array import GPUArray
from ...core.dtypes import bfloat16, uint8
from ...core.stream import current_stream_handle as _st


def _check_x(x: GPUArray, in_features: int) -> None:
    if x.ndim != 2:
        raise ValueError(f"input must be 2D [batch, in_features], got {x.ndim}D")
    if x.shape[1] != in_features:
        raise ValueError(f"input features {x.shape[1]} != weight {in_features}")
    if x.dtype != bfloat16:
        raise ValueError(f"the B200 decode path requires bfloat16 activations, got {x.dtype}")


def _out2d(out: GPUArray | None, M: int, N: int) -> GPUArray:
    if out is None:
        return GPUArray((M, N), bfloat16)
    if out.size != M * N or out.dtype != bfloat16:
        raise ValueError(f"out must be bfloat16 with {M * N} elements, got {out.shape} {out.dtype}")
    return out


def _ptr(a: GPUArray | None):
    return None if a is None else a.device_ptr


# smallest batch served by the tcgen05 skinny GEMM (same default and override as gemm_tc.cu: measured crossover)
TC_MIN_M = max(1, int(os.environ.get("PGK_TC_MIN_M", "5")))


def can_fuse_norm(norm, in_features: int) -> bool:
    """RMSNorm (no bias) can be folded into the GEMV's activation staging (pgk_rmsnorm_gemv_*)."""
    return (norm is not None and getattr(norm, "norm_type", None) == "rmsnorm" and norm.bias is None
            and in_features >= 512 and in_features % 32 == 0)


class LinearBF16:
    """y = x W^T + b, W [out_features, in_features] bf16 (layers/linear.py:25-99)."""

    _use_gemv: bool = True

    def __init__(self, weight: GPUArray, bias: GPUArray | None = None):
        if weight.ndim != 2:
            raise ValueError(f"weight must be 2D, got {weight.ndim}D")
        self.weight = weight
        self.bias = bias
        self.out_features = weight.shape[0]
        self.in_features = weight.shape[1]

    def forward(self, x: GPUArray, *, out: GPUArray | None = None, residual: GPUArray | None = None,
                norm=None) -> GPUArray:
        """norm: an RMSNorm layer applied to x inside the same launch (see can_fuse_norm)."""
        _check_x(x, self.in_features)
        M = x.shape[0]
        y = _out2d(out, M, self.out_features)
        if norm is not None:
            api.rmsnorm_gemv_bf16(x.device_ptr, norm.weight.device_ptr, float(norm.eps), self.weight.device_ptr,
                                  y.device_ptr, M, self.in_features, self.out_features, _ptr(self.bias),
                                  _ptr(residual), _st())
            return y
        api.gemv_bf16_ex(x.device_ptr, self.weight.device_ptr, y.device_ptr, M, self.in_features, self.out_features,
                         _ptr(self.bias), _ptr(residual), _st())
        return y

    def __call__(self, x: GPUArray, *, out: GPUArray | None = None) -> GPUArray:
        return self.forward(x, out=out)

    @property
    def weight_bytes(self) -> int:
        return self.weight.nbytes


Linear = LinearBF16


class LinearFP8:
    """y = x dequant(W)^T + b; W = E4M3 codes [out, in] + bf16 scale_inv [out/128, in/128]
    (layers/linear.py:101-260)."""

    _use_gemv: bool = True
    _FP8_TABLE: np.ndarray | None = None

    @classmethod
    def _get_fp8_table(cls) -> np.ndarray:
        """E4M3 -> f32 table with the CUDA LUT's values (fp8.cuh:57-122; the reference's Python
        table marks 0x7F/0xFF as NaN, layers/linear.py:136-137 -- the hot path uses +-480)."""
        if cls._FP8_TABLE is None:
            from ..quant import fp8_e4m3_table
            cls._FP8_TABLE = fp8_e4m3_table()
        return cls._FP8_TABLE

    def __init__(self, weight_fp8: GPUArray, scale_inv: GPUArray, bias: GPUArray | None = None,
                 block_size: tuple[int, int] = (128, 128)):
        if weight_fp8.ndim != 2:
            raise ValueError(f"weight must be 2D, got {weight_fp8.ndim}D")
        if weight_fp8.dtype != uint8:
            raise ValueError(f"weight_fp8 must be uint8 (E4M3 codes), got {weight_fp8.dtype}")
        if tuple(block_size) != (128, 128):
            raise ValueError("LinearFP8 supports 128x128 block scales only")
        self.weight_fp8 = weight_fp8
        self.scale_inv = scale_inv
        self.bias = bias
        self.block_size = block_size
        self.out_features = weight_fp8.shape[0]
        self.in_features = weight_fp8.shape[1]
        self._norm_scratch: dict[int, GPUArray] = {}

    def uses_tensor_cores(self, M: int) -> bool:
        """True when a batch of M rows is served by the tcgen05 skinny GEMM (pgk_gemm_fp8_bf16_tc) instead of
        the CUDA-core GEMV: TC_MIN_M <= M <= 64 and 128-aligned shapes (every model shape of BASELINE.json)."""
        return (TC_MIN_M <= M <= 64 and self.in_features % 128 == 0 and self.out_features % 128 == 0
                and os.environ.get("PGK_TC", "1") != "0")

    def _dequantize_cpu(self) -> np.ndarray:
        """f32 dequantised weight on the host (testing aid; layers/linear.py:181-211)."""
        from ..quant import dequantize_fp8_e4m3_block
        return dequantize_fp8_e4m3_block(self.weight_fp8.to_numpy(), self.scale_inv.to_numpy())

    def forward(self, x: GPUArray, *, out: GPUArray | None = None, residual: GPUArray | None = None,
                norm=None) -> GPUArray:
        _check_x(x, self.in_features)
        M = x.shape[0]
        y = _out2d(out, M, self.out_features)
        if norm is not None and self.uses_tensor_cores(M):
            # batch >= TC_MIN_M goes to the tcgen05 GEMM, which takes plain activations: normalise first (cached scratch)
            sc = self._norm_scratch.get(M)
            if sc is None:
                sc = self._norm_scratch[M] = GPUArray((M, self.in_features), bfloat16)
            x, norm = norm(x, out=sc), None
        if norm is not None:
            api.rmsnorm_gemv_fp8_bf16(x.device_ptr, norm.weight.device_ptr, float(norm.eps),
                                      self.weight_fp8.device_ptr, self.scale_inv.device_ptr, y.device_ptr, M,
                                      self.in_features, self.out_features, _ptr(self.bias), _ptr(residual), _st())
            return y
        api.gemv_fp8_bf16_ex(x.device_ptr, self.weight_fp8.device_ptr, self.scale_inv.device_ptr, y.device_ptr, M,
                             self.in_features, self.out_features, _ptr(self.bias), _ptr(residual), _st())
        return y

    def forward_allreduce(self, x: GPUArray, hidden: GPUArray) -> GPUArray:
        """Row-parallel layer of one token under tensor parallelism: hidden <- hidden + sum over ranks of (x . W^T), GEMV and
        all-reduce in ONE launch (`pgk_gemv_fp8_allreduce_bf16`; see TensorParallel.fuses_allreduce)."""
        _check_x(x, self.in_features)
        api.gemv_fp8_allreduce_bf16(x.device_ptr, self.weight_fp8.device_ptr, self.scale_inv.device_ptr, hidden.device_ptr,
                                    hidden.device_ptr, self.in_features, self.out_features, _st())
        return hidden

    def __call__(self, x: GPUArray, *, out: GPUArray | None = None) -> GPUArray:
        return self.forward(x, out=out)

    @property
    def weight_bytes(self) -> int:
        return self.weight_fp8.nbytes + self.scale_inv.nbytes


class LinearNVF4:
    """y = alpha * x dequant(W) + b with the reference's NVF4 storage (ops/matmul/nvf4.py:14-194):
    data u8 [in/2, out] (low nibble = even k), scale UE4M3 u8 [in/32, out]."""

    def __init__(self, data: GPUArray, scale: GPUArray, in_features: int, out_features: int,
                 bias: GPUArray | None = None, alpha: float = 1.0):
        if data.dtype != uint8 or scale.dtype != uint8:
            raise ValueError("LinearNVF4 requires uint8 data and scale")
        if data.size != (in_features // 2) * out_features:
            raise ValueError(f"data size {data.size} != (in/2)*out = {(in_features // 2) * out_features}")
        if scale.size != ((in_features + 31) // 32) * out_features:
            raise ValueError(f"scale size {scale.size} != ceil(in/32)*out")
        self.data, self.scale = data, scale
        self.in_features, self.out_features = in_features, out_features
        self.bias, self.alpha = bias, float(alpha)
        self._norm_scratch: dict[int, GPUArray] = {}

    @classmethod
    def from_bf16(cls, weight_kn: GPUArray, bias: GPUArray | None = None) -> "LinearNVF4":
        """Quantise a bf16 [in, out] matrix with the bit-exact device quantiser."""
        from ...ops.matmul import nvf4_get_sizes, quantize_bf16_to_nvf4
        K, N = weight_kn.shape
        dsz, ssz = nvf4_get_sizes(K, N)
        data, scale = GPUArray((K // 2, N), uint8), GPUArray(((K + 31) // 32, N), uint8)
        quantize_bf16_to_nvf4(weight_kn, data, scale)
        return cls(data, scale, K, N, bias)

    def uses_tensor_cores(self, M: int) -> bool:
        """True when a batch of M rows is served by the tcgen05 skinny GEMM (pgk_gemm_nvf4_bf16_tc): one row later
        than FP8 (measured crossover against 4 + 1 rows on the CUDA cores)."""
        return (TC_MIN_M + 1 <= M <= 64 and self.in_features % 128 == 0 and self.out_features % 128 == 0
                and os.environ.get("PGK_TC", "1") != "0")

    def forward(self, x: GPUArray, *, out: GPUArray | None = None, residual: GPUArray | None = None,
                norm=None) -> GPUArray:
        _check_x(x, self.in_features)
        M = x.shape[0]
        y = _out2d(out, M, self.out_features)
        if norm is not None and self.uses_tensor_cores(M):
            # the tensor-core GEMM takes plain activations: normalise first (cached scratch, graph-capture safe)
            sc = self._norm_scratch.get(M)
            if sc is None:
                sc = self._norm_scratch[M] = GPUArray((M, self.in_features), bfloat16)
            x, norm = norm(x, out=sc), None
        if norm is not None:
            api.rmsnorm_gemv_nvf4_bf16(x.device_ptr, norm.weight.device_ptr, float(norm.eps), self.data.device_ptr,
                                       self.scale.device_ptr, y.device_ptr, M, self.in_features, self.out_features,
                                       self.alpha, _ptr(self.bias), _ptr(residual), _st())
            return y
        api.gemv_nvf4_bf16_ex(x.device_ptr, self.data.device_ptr, self.scale.device_ptr, y.device_ptr, M,
                              self.in_features, self.out_features, self.alpha, _ptr(self.bias), _ptr(residual), _st())
        return y

    def forward_allreduce(self, x: GPUArray, hidden: GPUArray) -> GPUArray:
        """NVF4 form of LinearFP8.forward_allreduce (`pgk_gemv_nvf4_allreduce_bf16`)."""
        _check_x(x, self.in_features)
        api.gemv_nvf4_allreduce_bf16(x.device_ptr, self.data.device_ptr, self.scale.device_ptr, hidden.device_ptr,
                                     hidden.device_ptr, self.in_features, self.out_features, self.alpha, _st())
        return hidden

    def __call__(self, x: GPUArray, *, out: GPUArray | None = None) -> GPUArray:
        return self.forward(x, out=out)

    @property
    def weight_bytes(self) -> int:
        return self.data.nbytes + self.scale.nbytes


__all__ = ["LinearBF16", "LinearFP8", "LinearNVF4", "Linear", "can_fuse_norm"]
