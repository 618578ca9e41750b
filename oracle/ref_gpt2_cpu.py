"""TEST / BENCH INFRASTRUCTURE -- BASELINE.json configs[0] on the host cores, through the REFERENCE package itself.

GPT-2 small (124M, random init, FP32), 128-token prompt + 128 greedy tokens through the reference's own
`CausalTransformerModel.generate` (src/pygpukit/llm/models/causal.py:179-255) on its `CPUSimulationBackend`
(src/pygpukit/core/backend.py:281-358).  The package is imported from oracle/_ref/pysrc (staged, unmodified, by
oracle/build_ref.sh; git-ignored) or from /root/reference/src when that tree is mounted.  Nothing of ours is on
this path: no kernel, no model, no engine.

The reference's CPU in-place branches assign `output._data`, an attribute its GPUArray lacks
(ops/nn/linear.py:51-59, ops/nn/norm.py:190-194), so biased models cannot run on the CPU backend as shipped; the
same in-memory shim as tests/golden/make_golden_gpt2.py supplies the attribute (no file of the package is touched).

    python oracle/ref_gpt2_cpu.py [prompt_len] [new_tokens]      -> one JSON line
"""
from __future__ import annotations

import json
import os
import sys
import time

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))


def _import_reference():
    for root in (os.path.join(_HERE, "_ref", "pysrc"), "/root/reference/src"):
        if os.path.isdir(os.path.join(root, "pygpukit")):
            if root not in sys.path:
                sys.path.insert(0, root)
            import pygpukit  # noqa: F401
            return root
    return None


def run(prompt_len: int = 128, new_tokens: int = 128, seed: int = 0) -> dict:
    root = _import_reference()
    if root is None:
        return {"unavailable": "reference package not staged (oracle/_ref/pysrc) and /root/reference absent"}
    import pygpukit as ref
    from pygpukit.core.array import GPUArray as RefArray
    from pygpukit.core.backend import CPUSimulationBackend, get_backend, set_backend
    set_backend(CPUSimulationBackend())
    from pygpukit.llm import GPT2_SPEC, MLP, Attention, CausalTransformerModel, GPT2Config, Norm, TransformerBlock
    RefArray._data = property(lambda self: self,
                              lambda self, other: get_backend().copy_host_to_device(other.to_numpy(), self._device_ptr))
    cfg = GPT2Config().to_transformer_config()           # GPT-2 small: 12 layers, 768 hidden, 12 heads, 50257 vocab
    rng = np.random.default_rng(seed)
    H, I, V, P = cfg.hidden_size, cfg.intermediate_size, cfg.vocab_size, cfg.max_position_embeddings
    g = ref.from_numpy

    def w(*shape, s=0.02):
        return g(rng.standard_normal(shape, dtype=np.float32) * np.float32(s))

    def ln():
        return Norm(g(np.ones(H, np.float32)), w(H, s=0.01), "layernorm", cfg.norm_eps)

    blocks = []
    for _ in range(cfg.num_layers):
        attn = Attention(w(H, H), w(H, H), w(H, H), w(H, H), cfg, q_bias=w(H), k_bias=w(H), v_bias=w(H), o_bias=w(H))
        mlp = MLP(cfg, fc1_weight=w(I, H), fc1_bias=w(I), fc2_weight=w(H, I), fc2_bias=w(H))
        blocks.append(TransformerBlock(ln(), attn, ln(), mlp))
    model = CausalTransformerModel(cfg, w(V, H), blocks, ln(), lm_head=None, position_embed=w(P, H), spec=GPT2_SPEC)
    prompt = [int(t) for t in rng.integers(0, V, prompt_len)]
    # the prompt pass is timed on its own (prompt + 1 token), then prompt + 1 + new_tokens: the difference is the
    # steady-state decode the metric is about (the KV-concat generate() path, causal.py:179-255)
    # (the first call also pays one-time costs -- host copies of the embedding / transposed lm_head the reference
    # caches, causal.py:136-146,167-177 -- so it runs once untimed)
    t0 = time.perf_counter()
    model.generate(prompt, max_new_tokens=1, temperature=0.0, use_cache=True)
    t_first = time.perf_counter() - t0
    t0 = time.perf_counter()
    model.generate(prompt, max_new_tokens=1, temperature=0.0, use_cache=True)
    t_prompt = time.perf_counter() - t0
    t0 = time.perf_counter()
    out = model.generate(prompt, max_new_tokens=1 + new_tokens, temperature=0.0, use_cache=True)
    t_all = time.perf_counter() - t0
    dt = max(t_all - t_prompt, 1e-9)
    threads = None
    try:
        from threadpoolctl import threadpool_info
        threads = max([p.get("num_threads", 1) for p in threadpool_info()] or [1])
    except Exception:
        pass
    return {"value": new_tokens / dt, "unit": "tokens/s", "kind": "reference", "cores": threads or os.cpu_count(),
            "host_cpus": os.cpu_count(), "decode_seconds": dt, "prompt_pass_seconds": t_prompt, "first_call_seconds": t_first,
            "tokens_generated": len(out) - prompt_len - 1,
            "sample": f"GPT-2 small (124M) random-init FP32 through the reference package's "
                      f"CausalTransformerModel.generate on CPUSimulationBackend (NumPy): {new_tokens} greedy decode tokens "
                      f"after a {prompt_len}-token prompt (prompt pass {t_prompt:.1f} s timed separately and excluded; "
                      f"BASELINE.json configs[0] is 128 + 128 tokens, which takes ~260 s on 8 cores, so the bench runs "
                      f"this bounded sample of it); package from {root}"}


if __name__ == "__main__":
    a = [int(x) for x in sys.argv[1:3]]
    print(json.dumps(run(*a)), flush=True)
