from .rope import apply_rotary_pos_emb_numpy, precompute_freqs_cis

__all__ = ["precompute_freqs_cis", "apply_rotary_pos_emb_numpy"]
