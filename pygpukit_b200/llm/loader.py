"""Checkpoint loading for the decode path (mirror of pygpukit/llm/loader.py:132-612).

`load_model_from_safetensors(path, dtype="bfloat16", spec=None)` keeps the reference's entry point: it opens a
`model.safetensors` file or a `model.safetensors.index.json` shard index, detects the family from the tensor names
(`detect_model_spec`, config.py:391-428), reads `config.json` next to it, detects block-wise FP8 checkpoints
(`quantization_config.quant_method == "fp8"` and `<weight>_scale_inv` tensors, loader.py:197-252) and builds a
`CausalTransformerModel` on the B200 decode path:

  * FP8 weights stay FP8 (`LinearFP8`: E4M3 codes + BF16 128x128 inverse scales, no dequantisation);
  * everything else is BF16 on the device (F32 / F16 tensors are rounded to nearest even on the host);
  * `weight_format="nvf4"` (net-new) quantises BF16 linear weights to the reference's NVF4 layout at load time with the
    bit-exact device quantiser (`pgk_quantize_bf16_to_nvf4`), `"fp8"` with the block quantiser;
  * q/k/v are fused row-wise and gate/up stay a pair, as the decode kernels expect.

Families: Llama / Qwen2 (biases) / Qwen3 (QK-norm) / Qwen3-MoE / Mixtral names, GPT-2 (`wte` / `h.N.attn.c_attn`,
Conv1D weights transposed).  `save_model_to_safetensors` writes a model back under the Hugging Face names (used by the
round-trip tests; there are no real checkpoints on the build machine).
"""
from __future__ import annotations

import json
import os

import numpy as np

from ..core.array import GPUArray
from ..core.dtypes import bfloat16, f32_to_bf16_bits, uint8
from ..core.factory import from_numpy
from .config import (GPT2_SPEC, MODEL_SPECS, ModelSpec, TransformerConfig, detect_model_spec)
from .layers.attention import Attention
from .layers.block import TransformerBlock
from .layers.linear import LinearBF16, LinearFP8, LinearNVF4
from .layers.mlp import MLP
from .layers.moe import MoELayer
from .layers.norm import Norm
from .models.causal import CausalTransformerModel
from .safetensors import SafeTensorsFile, load_safetensors, save_safetensors


def _to_bf16_bits(a: np.ndarray, dtype: str) -> np.ndarray:
    """Storage array of a checkpoint tensor -> BF16 bit pattern (uint16)."""
    if dtype == "BF16":
        return np.ascontiguousarray(a)
    if dtype in ("F32", "F16", "F64"):
        return f32_to_bf16_bits(np.asarray(a, np.float32))
    raise ValueError(f"cannot convert a {dtype} tensor to bfloat16")


def config_from_hf(hf: dict, spec: ModelSpec, max_position_embeddings: int | None = None) -> TransformerConfig:
    """Hugging Face config.json -> TransformerConfig (the fields loader.py:440-612 reads)."""
    if spec.name == "gpt2":
        return TransformerConfig(vocab_size=hf.get("vocab_size", 50257), hidden_size=hf.get("n_embd", 768),
                                 num_layers=hf.get("n_layer", 12), num_heads=hf.get("n_head", 12),
                                 intermediate_size=hf.get("n_inner") or 4 * hf.get("n_embd", 768), norm_type="layernorm",
                                 activation="gelu", use_rope=False,
                                 max_position_embeddings=max_position_embeddings or hf.get("n_positions", 1024),
                                 norm_eps=hf.get("layer_norm_epsilon", 1e-5), tie_word_embeddings=True)
    H, heads = hf["hidden_size"], hf["num_attention_heads"]
    return TransformerConfig(
        vocab_size=hf["vocab_size"], hidden_size=H, num_layers=hf["num_hidden_layers"], num_heads=heads,
        num_kv_heads=hf.get("num_key_value_heads", heads), intermediate_size=hf.get("intermediate_size", 4 * H),
        _head_dim=hf.get("head_dim"), num_experts=hf.get("num_experts", hf.get("num_local_experts")),
        num_experts_per_tok=hf.get("num_experts_per_tok", 2), moe_intermediate_size=hf.get("moe_intermediate_size"),
        norm_type=spec.norm_type, activation=spec.activation, use_rope=spec.use_rope,
        max_position_embeddings=max_position_embeddings or hf.get("max_position_embeddings", 4096),
        norm_eps=hf.get("rms_norm_eps", spec.default_norm_eps), rope_theta=hf.get("rope_theta", spec.default_rope_theta),
        tie_word_embeddings=bool(hf.get("tie_word_embeddings", False)))


def _fp8_block_config(hf: dict) -> tuple[int, int] | None:
    """(128, 128) when config.json declares block-wise E4M3 weights (loader.py:197-252), else None."""
    q = hf.get("quantization_config") or {}
    if q.get("quant_method") != "fp8":
        return None
    if str(q.get("fmt", "e4m3")).lower() not in ("e4m3", "float8_e4m3fn"):
        raise ValueError(f"unsupported FP8 format {q.get('fmt')!r}")
    bs = tuple(q.get("weight_block_size") or (128, 128))
    if bs != (128, 128):
        raise ValueError(f"FP8 checkpoints need 128x128 weight blocks, got {bs}")
    return bs


class _Loader:
    def __init__(self, st: SafeTensorsFile, fp8_blocks, weight_format: str | None):
        self.st, self.fp8_blocks, self.fmt = st, fp8_blocks, weight_format

    def has(self, name: str) -> bool:
        return name in self.st

    def bf16(self, name: str, transpose: bool = False) -> GPUArray:
        a = _to_bf16_bits(self.st.tensor(name), self.st.tensor_info(name).dtype)
        if transpose:
            a = np.ascontiguousarray(a.T)
        g = from_numpy(np.ascontiguousarray(a))
        return g

    def linear(self, wname: str, bname: str | None = None, transpose: bool = False):
        """[out, in] linear layer `wname` (+ bias): LinearFP8 for block-wise FP8 checkpoints, else BF16, optionally
        re-quantised to the requested weight format."""
        bias = self.bf16(bname) if (bname and self.has(bname)) else None
        if self.fp8_blocks is not None and self.has(wname + "_scale_inv"):
            codes = np.ascontiguousarray(self.st.tensor(wname)).view(np.uint8)
            sinfo = self.st.tensor_info(wname + "_scale_inv")
            scale = _to_bf16_bits(self.st.tensor(wname + "_scale_inv"), sinfo.dtype)
            return LinearFP8(from_numpy(codes), from_numpy(scale), bias, self.fp8_blocks)
        w = self.bf16(wname, transpose)
        if self.fmt == "fp8":
            from .random_init import quantize_fp8
            lin = quantize_fp8(w)
            lin.bias = bias
            return lin
        if self.fmt == "nvf4" and w.shape[1] % 32 == 0 and w.shape[0] % 16 == 0:
            wt = from_numpy(np.ascontiguousarray(w.to_numpy().T))       # the reference's [K, N] orientation
            return LinearNVF4.from_bf16(wt, bias)
        return LinearBF16(w, bias)


def load_model_from_safetensors(model_path: str, dtype: str = "bfloat16", spec: ModelSpec | None = None,
                                repack_weights: bool = True, weight_format: str | None = None,
                                max_position_embeddings: int | None = None, config: dict | None = None) -> CausalTransformerModel:
    """See the module docstring.  dtype must be "bfloat16" (the B200 decode path); `config` overrides config.json."""
    if dtype != "bfloat16":
        raise ValueError("the B200 decode path loads checkpoints as bfloat16 (FP8 weights stay FP8)")
    st = load_safetensors(model_path)
    names = st.tensor_names
    if spec is None:
        spec = detect_model_spec(names)
    hf = dict(config or {})
    cfg_path = os.path.join(os.path.dirname(os.path.abspath(model_path)), "config.json")
    if not hf and os.path.exists(cfg_path):
        with open(cfg_path, encoding="utf-8") as f:
            hf = json.load(f)
    if not hf:
        raise ValueError(f"no config.json next to {model_path} (pass config=...)")
    cfg = config_from_hf(hf, spec, max_position_embeddings)
    ld = _Loader(st, _fp8_block_config(hf), weight_format)
    if spec.name == "gpt2":
        return _load_gpt2(ld, cfg, spec)
    pre = "model."
    eps = cfg.norm_eps

    def norm(name):
        return Norm(ld.bf16(name), None, "rmsnorm", eps)

    blocks = []
    for i in range(cfg.num_layers):
        L = f"{pre}layers.{i}."
        A = L + "self_attn."
        b = (lambda n: A + n + ".bias") if spec.qkv_bias else (lambda n: None)
        q, k, v = (ld.linear(A + n + ".weight", b(n)) for n in ("q_proj", "k_proj", "v_proj"))
        o = ld.linear(A + "o_proj.weight")
        qn = norm(A + "q_norm.weight") if spec.use_qk_norm else None
        kn = norm(A + "k_norm.weight") if spec.use_qk_norm else None
        attn = Attention(q, k, v, o, cfg, q_norm=qn, k_norm=kn)
        if spec.is_moe:
            mlp = _load_moe(ld, cfg, spec, L)
        else:
            M = L + "mlp."
            mlp = MLP(cfg, gate_proj=ld.linear(M + "gate_proj.weight"), up_proj=ld.linear(M + "up_proj.weight"),
                      down_proj=ld.linear(M + "down_proj.weight"))
        blocks.append(TransformerBlock(norm(L + "input_layernorm.weight"), attn, norm(L + "post_attention_layernorm.weight"), mlp))
    embed = ld.bf16(pre + "embed_tokens.weight")
    head = ld.bf16("lm_head.weight") if (ld.has("lm_head.weight") and not cfg.tie_word_embeddings) else None
    model = CausalTransformerModel(cfg, embed, blocks, norm(pre + "norm.weight"), lm_head=head, spec=spec)
    model.weight_format = "fp8" if ld.fp8_blocks is not None else (weight_format or "bf16")
    return model


def _load_moe(ld: _Loader, cfg: TransformerConfig, spec: ModelSpec, L: str) -> MoELayer:
    mix = spec.name == "mixtral"
    M = L + ("block_sparse_moe." if mix else "mlp.")
    router = ld.bf16(M + "gate.weight")
    names = ("w1", "w3", "w2") if mix else ("gate_proj", "up_proj", "down_proj")
    experts = []
    for e in range(cfg.num_experts):
        E = f"{M}experts.{e}."
        experts.append(tuple(ld.linear(E + n + ".weight") for n in names))
    if not isinstance(experts[0][0], (LinearFP8, LinearNVF4)):
        raise ValueError("MoE experts run as FP8 or NVF4 on the B200 path: load an FP8 checkpoint or pass weight_format=")
    return MoELayer(cfg, router, expert_weights=experts)


def _load_gpt2(ld: _Loader, cfg: TransformerConfig, spec: ModelSpec) -> CausalTransformerModel:
    pre = "transformer." if ld.has("transformer.wte.weight") else ""
    H, eps = cfg.hidden_size, cfg.norm_eps

    def ln(n):
        return Norm(ld.bf16(n + ".weight"), ld.bf16(n + ".bias"), "layernorm", eps)

    def rows(w: GPUArray, lo: int, hi: int) -> GPUArray:
        return from_numpy(np.ascontiguousarray(w.to_numpy()[lo:hi]))

    blocks = []
    for i in range(cfg.num_layers):
        B = f"{pre}h.{i}."
        qkv_w = ld.bf16(B + "attn.c_attn.weight", transpose=True)          # Conv1D [in, 3H] -> [3H, in]
        qkv_b = ld.bf16(B + "attn.c_attn.bias").to_numpy()
        q, k, v = (rows(qkv_w, j * H, (j + 1) * H) for j in range(3))
        qb, kb, vb = (from_numpy(np.ascontiguousarray(qkv_b[j * H:(j + 1) * H])) for j in range(3))
        attn = Attention(q, k, v, ld.bf16(B + "attn.c_proj.weight", transpose=True), cfg, q_bias=qb, k_bias=kb, v_bias=vb,
                         o_bias=ld.bf16(B + "attn.c_proj.bias"))
        mlp = MLP(cfg, fc1_weight=ld.bf16(B + "mlp.c_fc.weight", transpose=True), fc1_bias=ld.bf16(B + "mlp.c_fc.bias"),
                  fc2_weight=ld.bf16(B + "mlp.c_proj.weight", transpose=True), fc2_bias=ld.bf16(B + "mlp.c_proj.bias"))
        blocks.append(TransformerBlock(ln(B + "ln_1"), attn, ln(B + "ln_2"), mlp))
    model = CausalTransformerModel(cfg, ld.bf16(pre + "wte.weight"), blocks, ln(pre + "ln_f"), lm_head=None,
                                   position_embed=ld.bf16(pre + "wpe.weight"), spec=spec)
    model.weight_format = "bf16"
    return model


def load_gpt2_from_safetensors(model_path: str, dtype: str = "bfloat16") -> CausalTransformerModel:
    return load_model_from_safetensors(model_path, dtype, GPT2_SPEC)


def load_mixtral_from_safetensors(model_path: str, dtype: str = "bfloat16") -> CausalTransformerModel:
    """Mixtral MoE checkpoint (loader.py:111-130): the generic loader with the Mixtral tensor names."""
    from .config import MIXTRAL_SPEC
    return load_model_from_safetensors(model_path, dtype, MIXTRAL_SPEC)


def load_llama_from_safetensors(model_path: str, dtype: str = "bfloat16") -> CausalTransformerModel:
    return load_model_from_safetensors(model_path, dtype, MODEL_SPECS["llama"])


def load_qwen3_from_safetensors(model_path: str, dtype: str = "bfloat16") -> CausalTransformerModel:
    return load_model_from_safetensors(model_path, dtype, MODEL_SPECS["qwen3"])


# ------------------------------------------------------------------ export (round-trip tests)
def save_model_to_safetensors(model: CausalTransformerModel, directory: str) -> str:
    """Write `model` (dense Llama / Qwen3 family, BF16 or FP8 linears) as `model.safetensors` + `config.json` under the
    Hugging Face names; returns the file path.  The fused q|k|v projection is split back into its three row blocks."""
    cfg, spec = model.config, model.spec
    if spec is None or spec.name not in ("llama", "qwen3", "qwen2"):
        raise ValueError("save_model_to_safetensors handles the dense llama / qwen families")
    os.makedirs(directory, exist_ok=True)
    T: dict[str, tuple[np.ndarray, str]] = {}
    fp8 = False

    def put_linear(name, lin, r0=None, r1=None):
        nonlocal fp8
        if isinstance(lin, LinearFP8):
            fp8 = True
            w, s = lin.weight_fp8.to_numpy(), lin.scale_inv.to_numpy()
            if r0 is not None:
                w, s = w[r0:r1], s[r0 // 128:(r1 + 127) // 128]
            T[name] = (np.ascontiguousarray(w), "F8_E4M3")
            T[name + "_scale_inv"] = (np.ascontiguousarray(s), "BF16")
        elif isinstance(lin, LinearBF16):
            w = lin.weight.to_numpy()
            T[name] = (np.ascontiguousarray(w if r0 is None else w[r0:r1]), "BF16")
        else:
            raise ValueError(f"cannot export {type(lin).__name__}")

    for i, blk in enumerate(model.blocks):
        L = f"model.layers.{i}."
        at = blk.attn
        qd, kd = at.q_dim, at.k_dim
        put_linear(L + "self_attn.q_proj.weight", at.qkv_proj, 0, qd)
        put_linear(L + "self_attn.k_proj.weight", at.qkv_proj, qd, qd + kd)
        put_linear(L + "self_attn.v_proj.weight", at.qkv_proj, qd + kd, qd + 2 * kd)
        put_linear(L + "self_attn.o_proj.weight", at.o_proj)
        if at.q_norm is not None:
            T[L + "self_attn.q_norm.weight"] = (at.q_norm.weight.to_numpy(), "BF16")
            T[L + "self_attn.k_norm.weight"] = (at.k_norm.weight.to_numpy(), "BF16")
        put_linear(L + "mlp.gate_proj.weight", blk.mlp.gate_proj)
        put_linear(L + "mlp.up_proj.weight", blk.mlp.up_proj)
        put_linear(L + "mlp.down_proj.weight", blk.mlp.down_proj)
        T[L + "input_layernorm.weight"] = (blk.attn_norm.weight.to_numpy(), "BF16")
        T[L + "post_attention_layernorm.weight"] = (blk.mlp_norm.weight.to_numpy(), "BF16")
    T["model.embed_tokens.weight"] = (model.embed_tokens.to_numpy(), "BF16")
    T["model.norm.weight"] = (model.final_norm.weight.to_numpy(), "BF16")
    if model.lm_head is not None:
        T["lm_head.weight"] = (model.lm_head.to_numpy(), "BF16")
    hf = {"model_type": spec.hf_model_type, "vocab_size": cfg.vocab_size, "hidden_size": cfg.hidden_size,
          "num_hidden_layers": cfg.num_layers, "num_attention_heads": cfg.num_heads, "num_key_value_heads": cfg.num_kv_heads,
          "intermediate_size": cfg.intermediate_size, "head_dim": cfg.head_dim,
          "max_position_embeddings": cfg.max_position_embeddings, "rms_norm_eps": cfg.norm_eps, "rope_theta": cfg.rope_theta,
          "tie_word_embeddings": model.lm_head is None}
    if fp8:
        hf["quantization_config"] = {"quant_method": "fp8", "fmt": "e4m3", "weight_block_size": [128, 128]}
    path = os.path.join(directory, "model.safetensors")
    save_safetensors(path, T, {"format": "pt"})
    with open(os.path.join(directory, "config.json"), "w", encoding="utf-8") as f:
        json.dump(hf, f)
    return path


__all__ = ["load_model_from_safetensors", "load_gpt2_from_safetensors", "load_llama_from_safetensors", "load_mixtral_from_safetensors",
           "load_qwen3_from_safetensors", "save_model_to_safetensors", "config_from_hf"]
_ = (bfloat16, uint8)
