"""Random-init models of the benchmark shapes (net-new; the reference only loads checkpoints,
llm/loader.py:132-612, which is out of scope -- SURVEY.md section 8f-3).

BF16 masters N(0, std^2) are drawn on the device (pgk_fill_normal_bf16) and quantised there
(pgk_quantize_bf16_to_fp8_block / the bit-exact pgk_quantize_bf16_to_nvf4), so an 8B..70B model is
ready in seconds.  Every sharded tensor is generated as `shards` independent pieces whose seeds
depend on (layer, tensor, piece): a tensor-parallel rank keeps piece `rank`; a single process asked
for `shards > 1` concatenates the pieces, which yields exactly the global model the TP ranks
jointly hold (used by the TP-vs-single-GPU parity tests).
"""
from __future__ import annotations

import numpy as np

from .._lib import api
from ..core.array import GPUArray
from ..core.dtypes import bfloat16, uint8
from ..core.factory import from_numpy, ones
from ..core.stream import current_stream_handle as _st
from .config import ModelSpec, TransformerConfig, named_config
from .layers.attention import Attention
from .layers.block import TransformerBlock
from .layers.linear import LinearBF16, LinearFP8, LinearNVF4
from .layers.mlp import MLP
from .layers.moe import MoELayer
from .layers.norm import Norm
from .models.causal import CausalTransformerModel

_TENSOR_IDS = {"embed": 1, "lm_head": 2, "q": 3, "k": 4, "v": 5, "o": 6, "gate": 7, "up": 8, "down": 9,
               "router": 10, "e_gate": 11, "e_up": 12, "e_down": 13,
               # GPT-2 family
               "wpe": 14, "qb": 15, "kb": 16, "vb": 17, "ob": 18, "fc1": 19, "fc1b": 20, "fc2": 21, "fc2b": 22,
               "ln1_b": 23, "ln2_b": 24, "lnf_b": 25}


def _seed(base: int, layer: int, tensor: str, piece: int) -> int:
    return (base * 1000003 + layer * 8191 + _TENSOR_IDS[tensor] * 131 + piece * 17) & 0x7FFFFFFF


def normal_bf16(shape, seed: int, std: float) -> GPUArray:
    a = GPUArray(shape, bfloat16)
    api.fill_normal_bf16(a.device_ptr, a.size, seed, float(std), _st())
    return a


def _pieces(shape_piece, axis: int, n: int, seeds: list[int], std: float) -> GPUArray:
    """n pieces of `shape_piece` concatenated along `axis` (0 or 1); n == 1 returns the piece itself."""
    if n == 1:
        return normal_bf16(shape_piece, seeds[0], std)
    parts = [normal_bf16(shape_piece, s, std).to_numpy() for s in seeds]     # test-sized models only
    return from_numpy(np.ascontiguousarray(np.concatenate(parts, axis=axis)))


def quantize_fp8(master_nk: GPUArray) -> LinearFP8:
    N, K = master_nk.shape
    codes = GPUArray((N, K), uint8)
    scale = GPUArray(((N + 127) // 128, (K + 127) // 128), bfloat16)
    api.quantize_bf16_to_fp8_block(master_nk.device_ptr, codes.device_ptr, scale.device_ptr, N, K, _st())
    return LinearFP8(codes, scale)


def quantize_nvf4(master_kn: GPUArray) -> LinearNVF4:
    return LinearNVF4.from_bf16(master_kn)


class _Maker:
    """Creates one (possibly sharded) linear layer in the requested weight format."""

    def __init__(self, fmt: str, base_seed: int, std: float, world: int, rank: int, shards: int):
        self.fmt, self.base, self.std = fmt, base_seed, std
        self.world, self.rank, self.shards = world, rank, shards

    def _which(self) -> tuple[int, list[int]]:
        """(number of pieces to materialise, piece ids)."""
        if self.world > 1:
            return 1, [self.rank]
        return self.shards, list(range(self.shards))

    def linear(self, layer: int, name: str, out_f: int, in_f: int, split: str | None, fmt: str | None = None):
        """split: 'out' (column-parallel), 'in' (row-parallel) or None (replicated); out_f / in_f are the
        GLOBAL sizes.  Returns a layer holding this process's share."""
        fmt = fmt or self.fmt
        div = max(self.world, self.shards)
        n, ids = self._which() if split else (1, [0])
        po, pi = (out_f // div if split == "out" else out_f), (in_f // div if split == "in" else in_f)
        seeds = [_seed(self.base, layer, name, p) for p in ids]
        if fmt == "nvf4":       # master in the reference's [K, N] = [in, out] orientation
            m = _pieces((pi, po), 1 if split == "out" else 0, n, seeds, self.std)
            return quantize_nvf4(m)
        m = _pieces((po, pi), 0 if split == "out" else 1, n, seeds, self.std)
        return quantize_fp8(m) if fmt == "fp8" else LinearBF16(m)


def build_random_model(name: str = "qwen3-8b", weight_format: str = "fp8", max_seq: int = 4096, comm=None,
                       seed: int = 1234, shards: int = 1, std: float = 0.02, config: TransformerConfig | None = None,
                       spec: ModelSpec | None = None, num_layers: int | None = None) -> CausalTransformerModel:
    """weight_format: 'fp8' (w8a16 block-scaled), 'nvf4' or 'bf16'.  Embedding, lm_head, norms and the
    MoE router stay BF16 (as in the FP8 checkpoints the reference loads, llm/loader.py:197-252)."""
    world = comm.world if comm is not None else 1
    rank = comm.rank if comm is not None else 0
    if config is None:
        config, spec = named_config(name, max_seq=max_seq, tp_size=world, tp_rank=rank)
    if num_layers is not None:
        config.num_layers = num_layers
    if spec is not None and spec.use_position_embed:
        return _gpt2(name, config, spec, weight_format, seed, std, world, shards)
    mk = _Maker(weight_format, seed, std, world, rank, shards)
    div = max(world, shards)
    H, hd = config.hidden_size, config.head_dim
    eps = config.norm_eps

    def gamma(n):
        return ones((n,), bfloat16)

    embed = normal_bf16((config.vocab_size, H), _seed(seed, 0, "embed", 0), std)
    blocks = []
    for L in range(config.num_layers):
        # q/k/v as ONE fused column-parallel projection: piece p = [q_p | k_p | v_p] rows
        q_rows, kv_rows = config.num_heads * hd, config.num_kv_heads * hd
        qkv = _fused_qkv(mk, L, q_rows, kv_rows, H, config.num_kv_heads)
        o = mk.linear(L, "o", H, q_rows, "in")
        qn = Norm(gamma(hd), None, "rmsnorm", eps) if spec.use_qk_norm else None
        kn = Norm(gamma(hd), None, "rmsnorm", eps) if spec.use_qk_norm else None
        attn = Attention(None, None, None, o, config, q_norm=qn, k_norm=kn, qkv_proj=qkv)
        if config.is_moe:
            mlp = _moe(mk, config, L, H, world, rank, div)
        else:
            I = config.intermediate_size
            down = mk.linear(L, "down", H, I, "in")
            if weight_format == "nvf4":
                gu = _fused_gate_up_nvf4(mk, L, I, H)
                mlp = MLP(config, down_proj=down, gate_up_proj=gu)
            else:
                mlp = MLP(config, gate_proj=mk.linear(L, "gate", I, H, "out"), up_proj=mk.linear(L, "up", I, H, "out"),
                          down_proj=down)
        blocks.append(TransformerBlock(Norm(gamma(H), None, "rmsnorm", eps), attn, Norm(gamma(H), None, "rmsnorm", eps),
                                       mlp, comm=comm))
    # lm_head: column-parallel over the vocabulary (rank keeps V/world rows)
    V = config.vocab_size
    if V % max(world, shards, 1):
        raise ValueError(f"vocab_size {V} is not a multiple of the tensor-parallel world size {max(world, shards)} "
                         "(equal lm_head shards only: pad the vocabulary)")
    if world > 1:
        lo = rank * (V // world)
        hi = lo + V // world
        head = normal_bf16((hi - lo, H), _seed(seed, 0, "lm_head", rank), std)
    elif shards > 1:
        per = V // shards
        parts = [normal_bf16(((V - p * per) if p == shards - 1 else per, H), _seed(seed, 0, "lm_head", p), std).to_numpy()
                 for p in range(shards)]
        head = from_numpy(np.concatenate(parts, axis=0))
    else:
        head = normal_bf16((V, H), _seed(seed, 0, "lm_head", 0), std)
    model = CausalTransformerModel(config, embed, blocks, Norm(gamma(H), None, "rmsnorm", eps), lm_head=head,
                                   spec=spec, comm=comm)
    model.weight_format = weight_format
    model.name = name
    return model


def _gpt2(name: str, config: TransformerConfig, spec: ModelSpec, weight_format: str, seed: int, std: float,
          world: int, shards: int) -> CausalTransformerModel:
    """GPT-2 family (BASELINE.json configs[0]): LayerNorm with bias, biased q/k/v/o and fc1/fc2, GELU, learned
    positions, tied lm_head -- BF16 weights on one GPU (the reference runs it in FP32/FP16; there is no quantised
    GPT-2 checkpoint format upstream)."""
    if weight_format != "bf16" or world > 1 or shards > 1:
        raise ValueError("GPT-2 runs with weight_format='bf16' on a single GPU")
    H, L = config.hidden_size, config.num_layers
    I, eps = config.intermediate_size, config.norm_eps

    def w(tag, layer, *shape, s=std):
        return normal_bf16(shape, _seed(seed, layer, tag, 0), s)

    def ln(tag, layer):
        return Norm(ones((H,), bfloat16), w(tag + "_b", layer, H, s=0.01), "layernorm", eps)

    blocks = []
    for i in range(L):
        attn = Attention(w("q", i, H, H), w("k", i, H, H), w("v", i, H, H), w("o", i, H, H), config,
                         q_bias=w("qb", i, H), k_bias=w("kb", i, H), v_bias=w("vb", i, H), o_bias=w("ob", i, H))
        mlp = MLP(config, fc1_weight=w("fc1", i, I, H), fc1_bias=w("fc1b", i, I), fc2_weight=w("fc2", i, H, I),
                  fc2_bias=w("fc2b", i, H))
        blocks.append(TransformerBlock(ln("ln1", i), attn, ln("ln2", i), mlp))
    model = CausalTransformerModel(config, w("embed", 0, config.vocab_size, H), blocks, ln("lnf", 0), lm_head=None,
                                   position_embed=w("wpe", 0, config.max_position_embeddings, H), spec=spec)
    model.weight_format = weight_format
    model.name = name
    return model


def _fused_qkv(mk: _Maker, L: int, q_rows: int, kv_rows: int, H: int, n_kv: int = 0):
    """[q | k | v] projection of this process: for each piece p, q/k/v pieces are generated separately
    and stacked so that piece p holds heads [p*Hq/div, ...) and KV heads [p*Hkv/div, ...)."""
    div = max(mk.world, mk.shards)
    n, ids = mk._which()
    qp, kp = q_rows // div, kv_rows // div
    if n_kv and div > n_kv:
        # more ranks than KV heads: div / n_kv ranks replicate one whole KV head (same seed -> same weights)
        if div % n_kv:
            raise ValueError(f"cannot place {n_kv} KV heads on {div} shards")
        if n != 1:
            # the one-process model the ranks jointly hold: every rank's query piece, each KV head ONCE (piece j of the
            # k / v seeds is KV head j, the one ranks j * rep .. j * rep + rep - 1 replicate)
            kph = kv_rows // n_kv
            if mk.fmt == "nvf4":
                cols = [normal_bf16((H, qp), _seed(mk.base, L, "q", p), mk.std).to_numpy() for p in ids]
                cols += [normal_bf16((H, kph), _seed(mk.base, L, t, j), mk.std).to_numpy() for t in ("k", "v") for j in range(n_kv)]
                return quantize_nvf4(from_numpy(np.ascontiguousarray(np.concatenate(cols, axis=1))))
            rows = [normal_bf16((qp, H), _seed(mk.base, L, "q", p), mk.std).to_numpy() for p in ids]
            rows += [normal_bf16((kph, H), _seed(mk.base, L, t, j), mk.std).to_numpy() for t in ("k", "v") for j in range(n_kv)]
            m = from_numpy(np.ascontiguousarray(np.concatenate(rows, axis=0)))
            return quantize_fp8(m) if mk.fmt == "fp8" else LinearBF16(m)
        rep, kp, p = div // n_kv, kv_rows // n_kv, ids[0]
        width = qp + 2 * kp
        if mk.fmt == "nvf4":
            m = np.concatenate([normal_bf16((H, r), _seed(mk.base, L, t, pp), mk.std).to_numpy()
                                for t, r, pp in (("q", qp, p), ("k", kp, p // rep), ("v", kp, p // rep))], axis=1)
            return quantize_nvf4(from_numpy(np.ascontiguousarray(m)))
        m = GPUArray((width, H), bfloat16)
        off = 0
        for t, r, pp in (("q", qp, p), ("k", kp, p // rep), ("v", kp, p // rep)):
            api.fill_normal_bf16(m.device_ptr + off, r * H, _seed(mk.base, L, t, pp), float(mk.std), _st())
            off += r * H * 2
        return quantize_fp8(m) if mk.fmt == "fp8" else LinearBF16(m)
    if mk.fmt == "nvf4":
        # every piece is ONE draw of its fused width with the piece's "q" seed -- exactly what a TP rank draws
        mats = [normal_bf16((H, qp + 2 * kp), _seed(mk.base, L, "q", p), mk.std).to_numpy() if n > 1 else None
                for p in ids]
        if n == 1:
            p = ids[0]
            m = GPUArray((H, qp + 2 * kp), bfloat16)
            # single piece: one draw of the fused width (column order q|k|v is a naming of iid columns)
            api.fill_normal_bf16(m.device_ptr, m.size, _seed(mk.base, L, "q", p), float(mk.std), _st())
            return quantize_nvf4(m)
        # global order must be [q_0..q_n | k_0.. | v_0..]: regroup the per-piece columns
        qs = np.concatenate([m[:, :qp] for m in mats], axis=1)
        ks = np.concatenate([m[:, qp:qp + kp] for m in mats], axis=1)
        vs = np.concatenate([m[:, qp + kp:] for m in mats], axis=1)
        return quantize_nvf4(from_numpy(np.ascontiguousarray(np.concatenate([qs, ks, vs], axis=1))))
    if n == 1:
        p = ids[0]
        m = GPUArray((qp + 2 * kp, H), bfloat16)
        off = 0
        for t, r in (("q", qp), ("k", kp), ("v", kp)):
            api.fill_normal_bf16(m.device_ptr + off, r * H, _seed(mk.base, L, t, p), float(mk.std), _st())
            off += r * H * 2
        return quantize_fp8(m) if mk.fmt == "fp8" else LinearBF16(m)
    parts = {t: [normal_bf16((r, H), _seed(mk.base, L, t, p), mk.std).to_numpy() for p in ids]
             for t, r in (("q", qp), ("k", kp), ("v", kp))}
    m = from_numpy(np.ascontiguousarray(np.concatenate(parts["q"] + parts["k"] + parts["v"], axis=0)))
    return quantize_fp8(m) if mk.fmt == "fp8" else LinearBF16(m)


def _fused_gate_up_nvf4(mk: _Maker, L: int, I: int, H: int) -> LinearNVF4:
    """[K=H, 2*I_local] master whose columns are gate/up interleaved in groups of 8 (iid columns, so the
    interleaved tensor is drawn directly; the test-mode concatenation keeps piece order)."""
    div = max(mk.world, mk.shards)
    n, ids = mk._which()
    seeds = [_seed(mk.base, L, "gate", p) for p in ids]
    return quantize_nvf4(_pieces((H, 2 * (I // div)), 1, n, seeds, mk.std))


def _moe(mk: _Maker, config: TransformerConfig, L: int, H: int, world: int, rank: int, div: int) -> MoELayer:
    E, I = config.num_experts, config.moe_intermediate_size
    router = normal_bf16((E, H), _seed(mk.base, L, "router", 0), mk.std)
    e_local = E // world
    n, ids = mk._which()
    per = E // div
    stacked = {"format": "nvf4" if mk.fmt == "nvf4" else "fp8"}
    if mk.fmt not in ("nvf4", "fp8"):
        raise ValueError("MoE experts must be 'fp8' or 'nvf4'")
    for name, (n_out, k_in) in (("gate", (I, H)), ("up", (I, H)), ("down", (H, I))):
        seeds = [_seed(mk.base, L, "e_" + name, p) for p in ids]
        cnt = per * n      # experts materialised here
        if mk.fmt == "nvf4":
            m = _pieces((per * k_in, n_out), 0, n, seeds, mk.std)                 # [E*K, N]
            lin = quantize_nvf4(m)
            stacked[name] = (lin.data.view((cnt, k_in // 2, n_out)), lin.scale.view((cnt, k_in // 32, n_out)))
        else:
            m = _pieces((per * n_out, k_in), 0, n, seeds, mk.std)                 # [E*N, K]
            lin = quantize_fp8(m)
            stacked[name] = (lin.weight_fp8.view((cnt, n_out, k_in)),
                             lin.scale_inv.view((cnt, n_out // 128, (k_in + 127) // 128)))
    return MoELayer(config, router, stacked=stacked, expert_offset=rank * e_local if world > 1 else 0)


__all__ = ["build_random_model", "normal_bf16", "quantize_fp8", "quantize_nvf4"]
