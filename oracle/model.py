"""TEST INFRASTRUCTURE -- CPU restatement of the reference's fixed-cache decode step, op by op.

Follows `DecodeM1.step` (src/pygpukit/llm/decode/m1.py:40-121) -> `Attention.forward_fixed_cache`
(llm/layers/attention.py:279-370) -> `_mlp_forward_zero_alloc` (llm/models/causal.py:489-514) ->
`MoELayer.__call__` (llm/layers/moe.py:153-318), calling the per-kernel restatements of oracle.c in
the reference's order so that every rounding point of SURVEY.md Appendix A.1 is reproduced:
rmsnorm -> q/k/v GEMV -> [QK-norm] -> RoPE (f32 tables) -> KV write -> SDPA -> o GEMV -> add_inplace
-> rmsnorm -> gate GEMV -> silu -> up GEMV -> mul_inplace -> down GEMV -> add_inplace -> ... ->
final rmsnorm -> lm_head -> argmax (first maximum).  lm_head uses the BF16 GEMV restatement
(bf16_opt.cuh) because the reference's cuBLASLt call has no defined accumulation order.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg import this.
Parity pin: the per-kernel functions are pinned against the reference's own CUDA kernels
(oracle/_ref/libref_cuda.so, tests/test_gpu_*.py::test_oracle_vs_reference_*) and the golden vectors
generated from the reference's Python (tests/golden/); the step ORDER is pinned by the citations above.
"""
from __future__ import annotations

import numpy as np

from . import oracle as orc


def _f32(bits):
    return (np.asarray(bits).astype(np.uint32) << 16).view(np.float32)


class OracleLinear:
    def __init__(self, fmt: str, **kw):
        self.fmt = fmt
        self.__dict__.update(kw)

    def __call__(self, x_bits: np.ndarray) -> np.ndarray:
        """x [M, K] bf16 bits -> [M, N] bf16 bits (one reference GEMV per row)."""
        x = np.ascontiguousarray(x_bits).reshape(-1, x_bits.shape[-1])
        if self.fmt == "fp8":
            return orc.gemv_fp8_bf16(x, self.codes, self.scale)
        if self.fmt == "nvf4":
            return orc.gemv_nvf4_bf16(x, self.data, self.nscale)
        return orc.gemv_bf16(x, self.weight)


def _lin_from(layer) -> OracleLinear:
    """pygpukit_b200 Linear* -> host copy."""
    n = type(layer).__name__
    if n == "LinearFP8":
        return OracleLinear("fp8", codes=layer.weight_fp8.to_numpy(), scale=layer.scale_inv.to_numpy())
    if n == "LinearNVF4":
        K, N = layer.in_features, layer.out_features
        return OracleLinear("nvf4", data=layer.data.to_numpy().reshape(K // 2, N),
                            nscale=layer.scale.to_numpy().reshape(-1, N))
    return OracleLinear("bf16", weight=layer.weight.to_numpy())


def _deinterleave(a: np.ndarray):
    r, n2 = a.shape
    g = a.reshape(r, n2 // 16, 2, 8)
    return np.ascontiguousarray(g[:, :, 0].reshape(r, n2 // 2)), np.ascontiguousarray(g[:, :, 1].reshape(r, n2 // 2))


class OracleModel:
    """Host mirror of a pygpukit_b200 CausalTransformerModel (single rank)."""

    def __init__(self, model, max_seq: int):
        cfg = model.config
        self.cfg = cfg
        self.hd = cfg.head_dim
        self.Hq, self.Hkv = cfg.num_heads, cfg.num_kv_heads
        self.eps = cfg.norm_eps
        self.embed = model.embed_tokens.to_numpy()
        self.head = model._head.to_numpy()
        self.final_gamma = model.final_norm.weight.to_numpy()
        self.layers = []
        a0 = model.blocks[0].attn
        self.cos, self.sin = a0._cos.astype(np.float32), a0._sin.astype(np.float32)
        for blk in model.blocks:
            L = {"attn_gamma": blk.attn_norm.weight.to_numpy(), "mlp_gamma": blk.mlp_norm.weight.to_numpy(),
                 "qkv": _lin_from(blk.attn.qkv_proj), "o": _lin_from(blk.attn.o_proj),
                 "q_gamma": None if blk.attn.q_norm is None else blk.attn.q_norm.weight.to_numpy(),
                 "k_gamma": None if blk.attn.k_norm is None else blk.attn.k_norm.weight.to_numpy()}
            mlp = blk.mlp
            if type(mlp).__name__ == "MoELayer":
                L["moe"] = self._moe_from(mlp)
            else:
                if mlp.gate_up_proj is not None:
                    gu = mlp.gate_up_proj
                    K = gu.in_features
                    dg, du = _deinterleave(gu.data.to_numpy().reshape(K // 2, -1))
                    sg, su = _deinterleave(gu.scale.to_numpy().reshape(K // 32, -1))
                    L["gate"] = OracleLinear("nvf4", data=dg, nscale=sg)
                    L["up"] = OracleLinear("nvf4", data=du, nscale=su)
                else:
                    L["gate"], L["up"] = _lin_from(mlp.gate_proj), _lin_from(mlp.up_proj)
                L["down"] = _lin_from(mlp.down_proj)
            # the reference's cache is GQA-expanded [num_heads, max_seq, hd]; values identical
            L["k_cache"] = np.zeros((self.Hq, max_seq, self.hd), np.uint16)
            L["v_cache"] = np.zeros((self.Hq, max_seq, self.hd), np.uint16)
            self.layers.append(L)

    @staticmethod
    def _moe_from(moe):
        w = moe._w
        d = {"router": moe.gate.weight.to_numpy(), "k": moe.num_experts_per_tok, "E": moe.num_experts,
             "format": w["format"]}
        for n in ("gate", "up", "down"):
            d[n] = (w[n][0].to_numpy(), w[n][1].to_numpy())
        return d

    # ---- one expert-grouped GEMM (grouped_gemm.cu:77-142 for FP8; per-row NVF4 GEMV for NVF4)
    def _grouped(self, moe, name, rows_bits, row_ids):
        w, s = moe[name]
        if moe["format"] == "fp8":
            return orc.grouped_gemm_fp8_bf16(rows_bits, w, s, row_ids)
        out = []
        for r, e in enumerate(row_ids):
            out.append(orc.gemv_nvf4_bf16(rows_bits[r:r + 1], w[e], s[e])[0])
        return np.stack(out)

    def _moe(self, moe, x_bits):
        T = x_bits.shape[0]
        k = moe["k"]
        logits = orc.gemv_bf16(x_bits, moe["router"])
        vals, idx = orc.moe_topk_bf16(logits, k)
        wts = orc.moe_softmax_topk_bf16(vals)
        counts, offsets, perm, rev = orc.moe_compute_permutation(idx, moe["E"])
        row_ids = orc.moe_expand_expert_offsets(offsets, T * k)
        gathered = orc.moe_gather_bf16(x_bits, perm, k)
        g = orc.silu_bf16(self._grouped(moe, "gate", gathered, row_ids))
        u = self._grouped(moe, "up", gathered, row_ids)
        h = orc.mul_inplace_bf16(g, u)
        dn = self._grouped(moe, "down", h, row_ids)
        return orc.moe_scatter_bf16(dn, wts, rev, T, k)

    def step(self, token: int, position: int, context_len: int):
        """One decode token; returns (logits bf16 bits [V], greedy token)."""
        H, hd, Hq, Hkv = self.cfg.hidden_size, self.hd, self.Hq, self.Hkv
        hidden = self.embed[token:token + 1].copy()                                  # [1, H]
        for L in self.layers:
            x = orc.rmsnorm_bf16(hidden, L["attn_gamma"], self.eps)
            qkv = L["qkv"](x)[0]
            q = qkv[:Hq * hd].reshape(1, Hq, hd)
            k = qkv[Hq * hd:(Hq + Hkv) * hd].reshape(1, Hkv, hd)
            v = qkv[(Hq + Hkv) * hd:].reshape(1, Hkv, hd)
            if L["q_gamma"] is not None:
                q = orc.rmsnorm_bf16(q.reshape(Hq, hd), L["q_gamma"], self.eps).reshape(1, Hq, hd)
                k = orc.rmsnorm_bf16(k.reshape(Hkv, hd), L["k_gamma"], self.eps).reshape(1, Hkv, hd)
            q, k = orc.rope_bf16_f32table(q, k, self.cos[position:position + 1], self.sin[position:position + 1])
            orc.kv_cache_update_gqa(k, L["k_cache"], Hq, position)
            orc.kv_cache_update_gqa(np.ascontiguousarray(v), L["v_cache"], Hq, position)
            att = orc.sdpa_causal_fixed_cache(np.ascontiguousarray(q.reshape(Hq, 1, hd)), L["k_cache"], L["v_cache"],
                                              context_len)
            o = L["o"](att.reshape(1, Hq * hd))
            hidden = orc.add_inplace_bf16(o, hidden)
            x = orc.rmsnorm_bf16(hidden, L["mlp_gamma"], self.eps)
            if "moe" in L:
                m = self._moe(L["moe"], x)
            else:
                g = orc.silu_bf16(L["gate"](x))
                m = L["down"](orc.mul_inplace_bf16(g, L["up"](x)))
            hidden = orc.add_inplace_bf16(m, hidden)
        x = orc.rmsnorm_bf16(hidden, self.final_gamma, self.eps)
        logits = orc.gemv_bf16(x, self.head)[0]
        return logits, orc.argmax_bf16(logits)

    def generate(self, prompt: list[int], n_new: int):
        """Greedy generation token by token (the prompt is also fed one token at a time, which the
        fixed-cache path defines as the same computation as a batched prefill).  Returns
        (tokens, per-step logits margins top1-top2 in f32); `self.top_logits` keeps the top-1 logit of every step
        (a margin of one BF16 ulp of that value means the two candidates are ADJACENT BF16 numbers)."""
        toks = list(prompt)
        margins = []
        self.top_logits = []
        logits = None
        for pos, t in enumerate(prompt):
            logits, nxt = self.step(t, pos, pos + 1)
        for _ in range(n_new):
            f = _f32(logits)
            top2 = np.partition(f, -2)[-2:]
            margins.append(float(top2[1] - top2[0]))
            self.top_logits.append(float(top2[1]))
            nxt = orc.argmax_bf16(logits)
            toks.append(nxt)
            logits, _ = self.step(nxt, len(toks) - 1, len(toks))
        return toks, margins
