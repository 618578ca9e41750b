"""Layer classes of the decode path (mirror of pygpukit/llm/layers/__init__.py)."""
from .attention import Attention
from .block import TransformerBlock
from .linear import Linear, LinearBF16, LinearFP8, LinearNVF4
from .mlp import MLP
from .moe import MoELayer
from .norm import Norm
from .rope import apply_rotary_pos_emb_numpy, precompute_freqs_cis

__all__ = ["Attention", "TransformerBlock", "Linear", "LinearBF16", "LinearFP8", "LinearNVF4", "MLP", "MoELayer",
           "Norm", "apply_rotary_pos_emb_numpy", "precompute_freqs_cis"]
