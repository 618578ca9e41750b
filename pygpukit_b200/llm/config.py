"""Model configuration (mirror of pygpukit/llm/config.py:20-615).

`ModelSpec` here keeps only the architecture flags the decode path reads (the reference's HF
tensor-name patterns belong to checkpoint loading, which is out of scope: SURVEY.md section 8f-3);
`TransformerConfig` has the reference's fields, defaults and derived properties
(config.py:440-510) plus `tp_size` / `tp_rank` for the net-new tensor-parallel path.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Literal


@dataclass(frozen=True)
class ModelSpec:
    name: str
    norm_type: Literal["rmsnorm", "layernorm"] = "rmsnorm"
    activation: Literal["gelu", "silu"] = "silu"
    use_rope: bool = True
    use_qk_norm: bool = False
    use_position_embed: bool = False
    qkv_combined: bool = False
    weight_transpose: bool = False
    is_moe: bool = False
    qkv_bias: bool = False
    default_norm_eps: float = 1e-5
    default_rope_theta: float = 10000.0
    hf_model_type: str = ""


GPT2_SPEC = ModelSpec(name="gpt2", norm_type="layernorm", activation="gelu", use_rope=False,
                      use_position_embed=True, qkv_combined=True, weight_transpose=True, qkv_bias=True,
                      hf_model_type="gpt2")
LLAMA_SPEC = ModelSpec(name="llama", hf_model_type="llama")
QWEN2_SPEC = ModelSpec(name="qwen2", qkv_bias=True, default_norm_eps=1e-6, default_rope_theta=1000000.0,
                       hf_model_type="qwen2")
QWEN3_SPEC = ModelSpec(name="qwen3", use_qk_norm=True, default_norm_eps=1e-6, default_rope_theta=1000000.0,
                       hf_model_type="qwen3")
QWEN3_MOE_SPEC = ModelSpec(name="qwen3_moe", use_qk_norm=True, is_moe=True, default_norm_eps=1e-6,
                           default_rope_theta=1000000.0, hf_model_type="qwen3_moe")
MIXTRAL_SPEC = ModelSpec(name="mixtral", is_moe=True, default_rope_theta=1000000.0, hf_model_type="mixtral")

MODEL_SPECS: dict[str, ModelSpec] = {
    "gpt2": GPT2_SPEC, "llama": LLAMA_SPEC, "qwen3": QWEN3_SPEC, "qwen3_moe": QWEN3_MOE_SPEC,
    "qwen2": QWEN2_SPEC, "mixtral": MIXTRAL_SPEC,
}


def detect_model_spec(tensor_names: list[str]) -> ModelSpec:
    """Same decision order as config.py:391-428."""
    if any("block_sparse_moe" in n for n in tensor_names):
        return MIXTRAL_SPEC
    has_moe = any("mlp.experts" in n for n in tensor_names)
    has_qk = any("q_norm" in n for n in tensor_names)
    if has_moe and has_qk:
        return QWEN3_MOE_SPEC
    if has_qk:
        return QWEN3_SPEC
    if "model.embed_tokens.weight" in tensor_names and "model.layers.0.self_attn.q_proj.bias" in tensor_names:
        return QWEN2_SPEC
    if "model.embed_tokens.weight" in tensor_names:
        return LLAMA_SPEC
    if "wte.weight" in tensor_names:
        return GPT2_SPEC
    raise ValueError(f"Cannot detect model type from tensor names. First 10 names: {tensor_names[:10]}")


@dataclass
class TransformerConfig:
    vocab_size: int = 32000
    hidden_size: int = 2048
    num_layers: int = 22
    num_heads: int = 32
    num_kv_heads: int | None = None
    intermediate_size: int | None = None
    _head_dim: int | None = None

    num_experts: int | None = None
    num_experts_per_tok: int = 2
    moe_intermediate_size: int | None = None

    norm_type: Literal["rmsnorm", "layernorm"] = "rmsnorm"
    activation: Literal["gelu", "silu"] = "silu"
    use_rope: bool = True
    causal: bool = True

    max_position_embeddings: int = 2048
    norm_eps: float = 1e-5
    rope_theta: float = 10000.0

    tie_word_embeddings: bool = True

    # net-new: tensor parallelism (the reference is single-GPU, docs/scheduler.md:358)
    tp_size: int = 1
    tp_rank: int = 0

    def __post_init__(self):
        if self.num_kv_heads is None:
            self.num_kv_heads = self.num_heads
        if self.intermediate_size is None:
            self.intermediate_size = 4 * self.hidden_size
        if self.moe_intermediate_size is None:
            self.moe_intermediate_size = self.intermediate_size

    @property
    def is_moe(self) -> bool:
        return self.num_experts is not None and self.num_experts > 1

    @property
    def head_dim(self) -> int:
        if self._head_dim is not None:
            return self._head_dim
        return self.hidden_size // self.num_heads

    @property
    def num_kv_groups(self) -> int:
        assert self.num_kv_heads is not None
        return self.num_heads // self.num_kv_heads


@dataclass
class GPT2Config:
    vocab_size: int = 50257
    n_embd: int = 768
    n_layer: int = 12
    n_head: int = 12
    n_positions: int = 1024
    layer_norm_eps: float = 1e-5

    @property
    def n_inner(self) -> int:
        return 4 * self.n_embd

    def to_transformer_config(self) -> TransformerConfig:
        return TransformerConfig(vocab_size=self.vocab_size, hidden_size=self.n_embd, num_layers=self.n_layer,
                                 num_heads=self.n_head, num_kv_heads=self.n_head, intermediate_size=self.n_inner,
                                 norm_type="layernorm", activation="gelu", use_rope=False, causal=True,
                                 max_position_embeddings=self.n_positions, norm_eps=self.layer_norm_eps)


@dataclass
class LlamaConfig:
    vocab_size: int = 32000
    hidden_size: int = 2048
    intermediate_size: int = 5632
    num_hidden_layers: int = 22
    num_attention_heads: int = 32
    num_key_value_heads: int = 4
    max_position_embeddings: int = 2048
    rms_norm_eps: float = 1e-5
    rope_theta: float = 10000.0

    @property
    def head_dim(self) -> int:
        return self.hidden_size // self.num_attention_heads

    def to_transformer_config(self) -> TransformerConfig:
        return TransformerConfig(vocab_size=self.vocab_size, hidden_size=self.hidden_size,
                                 num_layers=self.num_hidden_layers, num_heads=self.num_attention_heads,
                                 num_kv_heads=self.num_key_value_heads, intermediate_size=self.intermediate_size,
                                 norm_type="rmsnorm", activation="silu", use_rope=True, causal=True,
                                 max_position_embeddings=self.max_position_embeddings, norm_eps=self.rms_norm_eps,
                                 rope_theta=self.rope_theta)


@dataclass
class Qwen3Config:
    vocab_size: int = 151936
    hidden_size: int = 4096
    intermediate_size: int = 12288
    num_hidden_layers: int = 36
    num_attention_heads: int = 32
    num_key_value_heads: int = 8
    head_dim: int = 128
    max_position_embeddings: int = 40960
    rms_norm_eps: float = 1e-6
    rope_theta: float = 1000000.0

    def to_transformer_config(self) -> TransformerConfig:
        return TransformerConfig(vocab_size=self.vocab_size, hidden_size=self.hidden_size,
                                 num_layers=self.num_hidden_layers, num_heads=self.num_attention_heads,
                                 num_kv_heads=self.num_key_value_heads, intermediate_size=self.intermediate_size,
                                 _head_dim=self.head_dim, norm_type="rmsnorm", activation="silu", use_rope=True,
                                 causal=True, max_position_embeddings=self.max_position_embeddings,
                                 norm_eps=self.rms_norm_eps, rope_theta=self.rope_theta)


# ---- the shapes BASELINE.json's configs name (SURVEY.md section 8 header)
def named_config(name: str, max_seq: int = 4096, tp_size: int = 1, tp_rank: int = 0) -> tuple[TransformerConfig, ModelSpec]:
    """(config, spec) of a benchmark model: 'gpt2', 'qwen3-8b', 'llama3-8b', 'llama3-70b', 'qwen3-30b-a3b'
    (and 'tiny-*' variants used by the tests)."""
    kw = dict(max_position_embeddings=max_seq, tp_size=tp_size, tp_rank=tp_rank, tie_word_embeddings=False)
    n = name.lower()
    if n == "gpt2":
        c = GPT2Config().to_transformer_config()
        return c, GPT2_SPEC
    if n == "qwen3-8b":
        return TransformerConfig(vocab_size=151936, hidden_size=4096, num_layers=36, num_heads=32, num_kv_heads=8,
                                 intermediate_size=12288, _head_dim=128, norm_eps=1e-6, rope_theta=1e6, **kw), QWEN3_SPEC
    if n == "llama3-8b":
        return TransformerConfig(vocab_size=128256, hidden_size=4096, num_layers=32, num_heads=32, num_kv_heads=8,
                                 intermediate_size=14336, norm_eps=1e-5, rope_theta=5e5, **kw), LLAMA_SPEC
    if n == "llama3-70b":
        return TransformerConfig(vocab_size=128256, hidden_size=8192, num_layers=80, num_heads=64, num_kv_heads=8,
                                 intermediate_size=28672, norm_eps=1e-5, rope_theta=5e5, **kw), LLAMA_SPEC
    if n == "qwen3-30b-a3b":
        return TransformerConfig(vocab_size=151936, hidden_size=2048, num_layers=48, num_heads=32, num_kv_heads=4,
                                 intermediate_size=6144, _head_dim=128, num_experts=128, num_experts_per_tok=8,
                                 moe_intermediate_size=768, norm_eps=1e-6, rope_theta=1e7, **kw), QWEN3_MOE_SPEC
    if n == "tiny-llama":
        return TransformerConfig(vocab_size=1024, hidden_size=256, num_layers=2, num_heads=4, num_kv_heads=2,
                                 intermediate_size=512, _head_dim=64, norm_eps=1e-5, rope_theta=1e4, **kw), LLAMA_SPEC
    if n == "tiny-qwen3":
        return TransformerConfig(vocab_size=1024, hidden_size=256, num_layers=2, num_heads=4, num_kv_heads=2,
                                 intermediate_size=512, _head_dim=128, norm_eps=1e-6, rope_theta=1e6, **kw), QWEN3_SPEC
    if n == "small-qwen3":      # hidden >= 512: exercises the fused RMSNorm->GEMV prologue
        return TransformerConfig(vocab_size=2048, hidden_size=512, num_layers=2, num_heads=8, num_kv_heads=2,
                                 intermediate_size=1024, _head_dim=128, norm_eps=1e-6, rope_theta=1e6, **kw), QWEN3_SPEC
    if n == "small-llama":
        return TransformerConfig(vocab_size=2048, hidden_size=512, num_layers=2, num_heads=8, num_kv_heads=4,
                                 intermediate_size=1536, _head_dim=64, norm_eps=1e-5, rope_theta=5e5, **kw), LLAMA_SPEC
    if n == "tiny-moe":
        return TransformerConfig(vocab_size=1024, hidden_size=256, num_layers=2, num_heads=4, num_kv_heads=2,
                                 intermediate_size=512, _head_dim=64, num_experts=8, num_experts_per_tok=2,
                                 moe_intermediate_size=128, norm_eps=1e-6, rope_theta=1e6, **kw), QWEN3_MOE_SPEC
    raise ValueError(f"unknown model name {name!r}")


__all__ = ["ModelSpec", "TransformerConfig", "GPT2Config", "LlamaConfig", "Qwen3Config", "GPT2_SPEC", "LLAMA_SPEC",
           "QWEN2_SPEC", "QWEN3_SPEC", "QWEN3_MOE_SPEC", "MIXTRAL_SPEC", "MODEL_SPECS", "detect_model_spec",
           "named_config"]
