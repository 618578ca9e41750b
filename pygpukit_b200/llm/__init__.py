"""pygpukit_b200.llm -- layers, model and decode strategies of the decode hot path (names of
pygpukit/llm/__init__.py, pinned upstream by tests/test_llm_unified.py:198-239).  The tokenizer, chat templates, the lazy / streaming
loaders and the weight-repacking pool are out of scope; the legacy aliases of
models/causal.py:1496-1501 are provided."""
from .buffers import DecodeBuffers, PrefillBuffers
from .config import (GPT2_SPEC, LLAMA_SPEC, MIXTRAL_SPEC, MODEL_SPECS, QWEN2_SPEC, QWEN3_MOE_SPEC, QWEN3_SPEC,
                     GPT2Config, LlamaConfig, ModelSpec, Qwen3Config, TransformerConfig, detect_model_spec,
                     named_config)
from .decode import DecodeBatch, DecodeJacobi, DecodeM1, DecodeM1Graph, DecodeSpeculative, DecodeStrategy
from .layers import (MLP, Attention, Linear, LinearBF16, LinearFP8, LinearNVF4, MoELayer, Norm, TransformerBlock,
                     apply_rotary_pos_emb_numpy, precompute_freqs_cis)
from .models import CausalTransformerModel, GPT2Model, LlamaModel, QwenModel
from .loader import (load_gpt2_from_safetensors, load_llama_from_safetensors, load_mixtral_from_safetensors,
                     load_model_from_safetensors, load_qwen3_from_safetensors, save_model_to_safetensors)
from .random_init import build_random_model
from .safetensors import SafeTensorsFile, TensorInfo, load_safetensors, save_safetensors
from .sampling import sample_token
from .tp import TensorParallel, shard_plan, shard_range

# legacy names (models/causal.py:1496-1501) and the sharded reader (SafeTensorsFile opens *.index.json itself)
RMSNorm = Norm
LayerNorm = Norm
LlamaAttention = Attention
LlamaMLP = MLP
LlamaBlock = TransformerBlock
CausalSelfAttention = Attention
ShardedSafeTensorsFile = SafeTensorsFile

__all__ = [n for n in dir() if not n.startswith("_")]
