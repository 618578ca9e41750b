"""One-launch decode step (net-new orchestration, same math): the whole token -- every layer of
`DecodeM1.step` (pygpukit/llm/decode/m1.py:40-121) plus the final norm, lm_head and greedy argmax -- as ONE
persistent kernel (`pgk_step_run`, csrc/step.cu) instead of 5 launches per layer.

`StepProgram.from_model(model)` describes the model's op sequence to the library once (pointers to the
model's own weights, caches and decode buffers; nothing is copied except NVF4 weights, which are re-packed
row-major once); `run()` launches it.  A model whose layers do not fit the kernel (GPT-2 / BF16 / MoE /
tensor-parallel shards, biases, head_dim != 128) raises `Unsupported`, and the caller keeps the per-op path.
"""
from __future__ import annotations

import ctypes as C

from .._lib import api, load
from ..core.array import GPUArray
from ..core.dtypes import uint8
from ..core.stream import current_stream_handle as _st
from .layers.linear import LinearBF16, LinearFP8, LinearNVF4
from .layers.mlp import MLP

OP_FP8, OP_BF16, OP_NVF4R, OP_ATTN = 1, 2, 3, 4
F_SWIGLU, F_ACT_EMBED, F_RES_EMBED, F_ARGMAX, F_XCHG = 1, 2, 4, 8, 16


class StepOp(C.Structure):
    """pgk_step_op_t (include/pgk_b200.h)."""
    _fields_ = [("kind", C.c_int32), ("flags", C.c_int32), ("K", C.c_int32), ("N", C.c_int32),
                ("act", C.c_void_p), ("gamma", C.c_void_p), ("eps", C.c_float), ("alpha", C.c_float),
                ("W0", C.c_void_p), ("S0", C.c_void_p), ("W1", C.c_void_p), ("S1", C.c_void_p),
                ("residual", C.c_void_p), ("out", C.c_void_p),
                ("q_gamma", C.c_void_p), ("k_gamma", C.c_void_p), ("cos_table", C.c_void_p), ("sin_table", C.c_void_p),
                ("k_cache", C.c_void_p), ("v_cache", C.c_void_p),
                ("hq", C.c_int32), ("n_kv", C.c_int32), ("max_seq", C.c_int32), ("reserved", C.c_int32),
                ("scale", C.c_float), ("attn_eps", C.c_float)]


class Unsupported(ValueError):
    """The model (or an op) does not fit the one-launch step kernel; use the per-op path."""


def _p(a):
    return None if a is None else a.device_ptr


def pack_nvf4_rows(lin: LinearNVF4, n_out: int | None = None, col0: int = 0, group: int | None = None,
                   stride: int | None = None) -> GPUArray:
    """Row-packed copy [n_out, K/2 + K/32] of (a column subset of) an NVF4 matrix in the reference layout."""
    K, ld = lin.in_features, lin.out_features
    n_out = ld if n_out is None else n_out
    group = n_out if group is None else group
    stride = n_out if stride is None else stride
    out = GPUArray((n_out, K // 2 + K // 32), uint8)
    api.nvf4_pack_rows(lin.data.device_ptr, lin.scale.device_ptr, out.device_ptr, K, n_out, ld, col0, group, stride, _st())
    return out


def gemv_op(lin, act, out, *, residual=None, gamma=None, eps: float = 0.0, flags: int = 0, lin_up=None,
            packed=None, packed_up=None) -> StepOp:
    """One GEMV-type op over a Linear layer (`packed*`: row-packed NVF4 weights from pack_nvf4_rows).
    act / out / residual / gamma: GPUArray or raw device pointer (int) or None."""
    def ptr(x):
        return x if (x is None or isinstance(x, int)) else x.device_ptr
    op = StepOp()
    op.K, op.N = lin.in_features, lin.out_features
    op.act, op.out, op.residual, op.gamma, op.eps, op.alpha = ptr(act), ptr(out), ptr(residual), ptr(gamma), eps, 1.0
    op.flags = flags
    if isinstance(lin, LinearFP8):
        op.kind, op.W0, op.S0 = OP_FP8, lin.weight_fp8.device_ptr, lin.scale_inv.device_ptr
        if lin_up is not None:
            op.W1, op.S1 = lin_up.weight_fp8.device_ptr, lin_up.scale_inv.device_ptr
    elif isinstance(lin, LinearBF16):
        op.kind, op.W0 = OP_BF16, lin.weight.device_ptr
    elif isinstance(lin, LinearNVF4):
        if packed is None:
            raise Unsupported("NVF4 ops need row-packed weights")
        op.kind, op.W0, op.alpha = OP_NVF4R, packed.device_ptr, float(lin.alpha)
        op.N = packed.shape[0]
        if packed_up is not None:
            op.W1 = packed_up.device_ptr
    else:
        raise Unsupported(f"unsupported linear layer {type(lin).__name__}")
    if getattr(lin, "bias", None) is not None:
        raise Unsupported("biased projections are not part of the step kernel")
    return op


def attn_op(attn, qkv, out) -> StepOp:
    if attn.head_dim != 128 or not attn.config.use_rope or attn.num_kv_groups > 8:
        raise Unsupported("attention op needs head_dim 128, RoPE and at most 8 query heads per KV head")
    op = StepOp()
    op.kind = OP_ATTN
    op.act, op.out = qkv.device_ptr, out.device_ptr
    op.q_gamma, op.k_gamma = _p(attn.q_norm.weight if attn.q_norm is not None else None), \
        _p(attn.k_norm.weight if attn.k_norm is not None else None)
    op.attn_eps = float(attn.q_norm.eps) if attn.q_norm is not None else 0.0
    op.cos_table, op.sin_table = attn._cos_dev.device_ptr, attn._sin_dev.device_ptr
    op.k_cache, op.v_cache = attn._k_cache.device_ptr, attn._v_cache.device_ptr
    op.hq, op.n_kv, op.max_seq = attn.num_heads, attn.num_kv_heads, attn._max_cache_len
    op.scale = 0.0
    return op


class StepProgram:
    """A compiled op sequence + its persistent-kernel launch."""

    def __init__(self, ops: list[StepOp], token_buf: GPUArray, pos_buf: GPUArray, ctx_buf: GPUArray,
                 next_token: GPUArray, vocab: int, advance: bool, keep=()):
        load()
        arr = (StepOp * len(ops))(*ops)
        h = C.c_void_p()
        api.step_create(C.cast(arr, C.c_void_p), len(ops), token_buf.device_ptr, pos_buf.device_ptr, ctx_buf.device_ptr,
                        next_token.device_ptr, int(vocab), 1 if advance else 0, C.byref(h))
        self._h = h
        self._keep = (list(keep), token_buf, pos_buf, ctx_buf, next_token)   # arrays the program points into
        self.n_ops = len(ops)

    def set_vocab_offset(self, offset: int) -> None:
        api.step_set_vocab_offset(self._h, int(offset))

    def run(self, stream=None) -> None:
        api.step_run(self._h, stream if stream is not None else _st())

    def error(self) -> int:
        code = C.c_int32(0)
        api.step_error(self._h, C.byref(code))
        return int(code.value)

    def trace(self):
        """[n_ops + 1, 8] GPU globaltimer stamps (ns) of CTA 0 for the last run: per op [0] start, [1] staged,
        [2] warp 0 done, [3] attention fold done, [4] arrived at the closing barrier, [5]/[6] attention inner barrier /
        combine start; the last row's [0] is the end of the step."""
        import numpy as np
        buf = np.zeros((self.n_ops + 1) * 8, np.uint64)
        api.step_trace(self._h, buf.ctypes.data_as(C.c_void_p), buf.size)
        return buf.reshape(self.n_ops + 1, 8)

    def info(self) -> dict:
        v = [C.c_int32(0) for _ in range(4)]
        api.step_info(self._h, *[C.byref(x) for x in v])
        return dict(zip(("ring_slots", "act_bytes", "smem_bytes", "ctas"), (int(x.value) for x in v)))

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                api.step_destroy(h)
            except Exception:  # noqa: BLE001 -- interpreter shutdown
                pass

    # ------------------------------------------------------------------ whole model
    @classmethod
    def from_model(cls, model) -> "StepProgram":
        """The greedy decode step of `model` (init_decode must have run): embedding -> layers -> final norm ->
        lm_head -> argmax -> advance position.  Raises Unsupported when a layer does not fit."""
        cfg, b = model.config, model.buffers
        tp = model.tp
        if tp > 1 and not getattr(model.comm, "_p2p", False):
            raise Unsupported("tensor-parallel shards need the NVLink peer-memory buffers (TensorParallel.init(allgather=...))")
        xf = F_XCHG if tp > 1 else 0          # row-parallel o_proj / down_proj exchange their partial sums in the kernel
        if model.position_embed is not None or not cfg.use_rope:
            raise Unsupported("learned position embeddings (GPT-2 family)")
        H = cfg.hidden_size
        fn = model.final_norm

        def rms(n):
            return n is not None and getattr(n, "norm_type", None) == "rmsnorm" and n.bias is None
        if not rms(fn):
            raise Unsupported("final norm must be a bias-free RMSNorm")
        hidden = b.hidden.slice_rows(1) if b.hidden.shape[0] != 1 else b.hidden
        qkv_buf = b.qkv_proj_out.slice_rows(1) if b.qkv_proj_out.shape[0] != 1 else b.qkv_proj_out
        attn_buf = b.attn_flat.slice_rows(1) if b.attn_flat.shape[0] != 1 else b.attn_flat
        act_buf = b.mlp_act.slice_rows(1) if b.mlp_act.shape[0] != 1 else b.mlp_act
        embed = model.embed_tokens
        ops, keep = [], []
        for li, blk in enumerate(model.blocks):
            at, mlp = blk.attn, blk.mlp
            if not isinstance(mlp, MLP) or mlp.activation != "silu":
                raise Unsupported("only dense SwiGLU layers")
            if not (rms(blk.attn_norm) and rms(blk.mlp_norm)):
                raise Unsupported("layer norms must be bias-free RMSNorm")
            first = li == 0
            kw_qkv, kw_o, kw_down = {}, {}, {}
            gate, up = mlp.gate_proj, mlp.up_proj
            if isinstance(at.qkv_proj, LinearNVF4):
                gu = mlp.gate_up_proj
                I = mlp.intermediate_size
                pk = getattr(blk, "_step_packed", None)
                if pk is None:
                    pk = blk._step_packed = (
                        pack_nvf4_rows(at.qkv_proj), pack_nvf4_rows(at.o_proj),
                        pack_nvf4_rows(gu, I, 0, 8, 16), pack_nvf4_rows(gu, I, 8, 8, 16), pack_nvf4_rows(mlp.down_proj))
                kw_qkv, kw_o, kw_down = {"packed": pk[0]}, {"packed": pk[1]}, {"packed": pk[4]}
                kw_gu = {"packed": pk[2], "packed_up": pk[3]}
                gate = gu
            else:
                if not (isinstance(gate, LinearFP8) and isinstance(up, LinearFP8)):
                    raise Unsupported("gate/up must both be FP8 (or NVF4)")
                kw_gu = {"lin_up": up}
            ops.append(gemv_op(at.qkv_proj, embed if first else hidden, qkv_buf, gamma=blk.attn_norm.weight,
                               eps=float(blk.attn_norm.eps), flags=F_ACT_EMBED if first else 0, **kw_qkv))
            ops.append(attn_op(at, qkv_buf, attn_buf))
            ops.append(gemv_op(at.o_proj, attn_buf, hidden, residual=embed if first else hidden,
                               flags=(F_RES_EMBED if first else 0) | xf, **kw_o))
            ops.append(gemv_op(gate, hidden, act_buf, gamma=blk.mlp_norm.weight, eps=float(blk.mlp_norm.eps),
                               flags=F_SWIGLU, **kw_gu))
            ops.append(gemv_op(mlp.down_proj, act_buf, hidden, residual=hidden, flags=xf, **kw_down))
        head = LinearBF16(model._head)
        ops.append(gemv_op(head, hidden, b.logits, gamma=fn.weight, eps=float(fn.eps), flags=F_ARGMAX))
        keep += [hidden, qkv_buf, attn_buf, act_buf, b.logits]
        v_global = cfg.vocab_size if tp > 1 else model._head.shape[0]
        prog = cls(ops, b.token_buf, b.position_buf, b.context_buf, b.next_token, v_global, True, keep)
        if tp > 1:
            prog.set_vocab_offset(model.comm.rank * (cfg.vocab_size // tp))   # consecutive vocabulary shards (random_init)
        return prog


__all__ = ["StepProgram", "StepOp", "Unsupported", "gemv_op", "attn_op", "pack_nvf4_rows",
           "OP_FP8", "OP_BF16", "OP_NVF4R", "OP_ATTN", "F_SWIGLU", "F_ACT_EMBED", "F_RES_EMBED", "F_ARGMAX", "F_XCHG"]
