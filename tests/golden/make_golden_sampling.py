"""Golden vectors for token sampling, produced by the REFERENCE's own function (run in the build container, where
/root/reference exists; the GPU box only sees the committed .npz):

    python tests/golden/make_golden_sampling.py

pygpukit.llm.sampling.sample_token is called as it is; np.random.choice is patched for the duration of the call so that
(a) the probability vector the reference draws from is recorded and (b) the draw uses a recorded uniform number through
NumPy's own inverse-CDF rule.  Logits are BF16-representable and free of ties among the values a cut can fall on."""
import importlib.util
import os
import sys

import numpy as np

REF = "/root/reference/src/pygpukit/llm/sampling.py"
spec = importlib.util.spec_from_file_location("ref_sampling", REF)
ref = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref)

rng = np.random.default_rng(20261105)


def bf16_round(x):
    b = x.astype(np.float32).view(np.uint32)
    b = (b + 0x7FFF + ((b >> 16) & 1)) & 0xFFFF0000
    return b.view(np.float32)


def distinct_bf16_logits(V, spread):
    """V DISTINCT BF16 values with 2^-6 <= |x| < spread, shuffled: distinct logits that far apart give distinct f32
    probabilities, so no cut of the reference's (unstable) argsort falls on a tie."""
    bits = np.arange(0x10000, dtype=np.uint32)
    vals = (bits << 16).astype(np.uint32).view(np.float32)
    ok = np.isfinite(vals) & (np.abs(vals) >= 2.0 ** -6) & (np.abs(vals) < spread)
    grid = vals[ok]
    assert grid.size >= V, grid.size
    return rng.permutation(rng.choice(grid, V, replace=False)).astype(np.float32)


cases = []
for V, spread in ((1000, 8.0), (2000, 6.0), (2300, 8.0)):
    logits = distinct_bf16_logits(V, spread)
    for temperature, top_k, top_p in ((1.0, 0, 1.0), (0.7, 0, 1.0), (1.0, 50, 1.0), (0.8, 40, 0.9), (1.3, 0, 0.95),
                                      (1.0, 1, 1.0), (0.0, 0, 1.0), (0.0, 5, 0.5), (1.0, 0, 0.3), (0.5, 200, 0.99)):
        for u in (0.0, 0.137, 0.5, 0.731, 0.999):
            rec = {}

            def fake_choice(n, p=None, _u=u, _rec=rec):
                _rec["p"] = np.array(p, np.float64)
                cdf = _rec["p"].cumsum()
                cdf /= cdf[-1]
                return int(cdf.searchsorted(_u, side="right"))

            orig = np.random.choice
            np.random.choice = fake_choice
            try:
                tok = ref.sample_token(logits.copy(), temperature, top_k, top_p)
            finally:
                np.random.choice = orig
            cases.append((V, temperature, top_k, top_p, u, tok))
out = {"n_cases": np.array(len(cases))}
for i, (V, t, k, p, u, tok) in enumerate(cases):
    out[f"c{i}"] = np.array([V, t, k, p, u, tok], np.float64)
rng2 = np.random.default_rng(20261105)          # regenerate the logits in the same order for storage
rng = rng2
for V, spread in ((1000, 8.0), (2000, 6.0), (2300, 8.0)):
    out[f"logits_{V}"] = distinct_bf16_logits(V, spread)
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden_sampling.npz"), **out)
print("cases", len(cases))
