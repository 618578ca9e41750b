"""Seeded weights of the tiny GPT-2 used by the GPT-2 parity case (numpy only, so the golden generator
and the tests build the SAME model).  Every value is exactly representable in bfloat16, so the
reference's FP32 CPU run and the BF16 GPU run start from identical weights.
GPT-2 layout (llm/loader.py GPT-2 branch, llm/config.py GPT2_SPEC): wte [V,H], wpe [P,H], per layer
ln_1 (g,b), q/k/v/o weights [H,H] + biases, ln_2, fc1 [4H,H] + bias, fc2 [H,4H] + bias, ln_f; tied lm_head."""
import numpy as np

CFG = dict(vocab_size=256, n_embd=128, n_layer=2, n_head=2, n_positions=96)      # head_dim 64, as GPT-2 small
PROMPT = [5, 77, 130, 9, 201, 44, 18, 250, 3, 99, 160, 31]
N_NEW = 24
SEED = 5                       # chosen for a varied greedy sequence (18 distinct tokens in 24)


def _bf16_round(a: np.ndarray) -> np.ndarray:
    u = np.ascontiguousarray(a, np.float32).view(np.uint32)
    u = (u + 0x7FFF + ((u >> 16) & 1)) & 0xFFFF0000
    return u.astype(np.uint32).view(np.float32)


def weights(seed: int = SEED, cfg: dict = CFG) -> dict:
    rng = np.random.default_rng(seed)
    V, H, L, P = cfg["vocab_size"], cfg["n_embd"], cfg["n_layer"], cfg["n_positions"]

    def n(*shape, std=0.08):
        return _bf16_round((rng.standard_normal(shape) * std).astype(np.float32))

    w = {"wte": n(V, H, std=0.2), "wpe": n(P, H, std=0.8), "ln_f.g": _bf16_round(1.0 + n(H, std=0.1)), "ln_f.b": n(H, std=0.05)}
    for i in range(L):
        p = f"h{i}."
        for nm in ("ln_1", "ln_2"):
            w[p + nm + ".g"] = _bf16_round(1.0 + n(H, std=0.1))
            w[p + nm + ".b"] = n(H, std=0.05)
        for nm in ("q", "k", "v", "o"):
            w[p + nm + ".w"] = n(H, H, std=0.15)
            w[p + nm + ".b"] = n(H, std=0.05)
        w[p + "fc1.w"], w[p + "fc1.b"] = n(4 * H, H, std=0.15), n(4 * H, std=0.05)
        w[p + "fc2.w"], w[p + "fc2.b"] = n(H, 4 * H, std=0.08), n(H, std=0.05)
    return w
