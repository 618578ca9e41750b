"""oracle/sampling.py against the golden vectors of the reference's own sample_token (tests/golden/golden_sampling.npz,
made by tests/golden/make_golden_sampling.py): same token for every (logits, temperature, top_k, top_p, u) case."""
import os

import numpy as np

from oracle import sampling as osm

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_sampling.npz")


def test_oracle_reproduces_reference_sample_token():
    g = np.load(GOLD)
    n = int(g["n_cases"])
    assert n >= 100
    for i in range(n):
        V, t, k, p, u, tok = g[f"c{i}"]
        logits = g[f"logits_{int(V)}"]
        assert osm.sample_token_u(logits, float(t), int(k), float(p), float(u)) == int(tok), (i, V, t, k, p, u)


def test_final_distribution_properties():
    rng = np.random.default_rng(1)
    logits = rng.standard_normal(500).astype(np.float32) * 3
    p = osm.final_distribution(logits, 0.8, 20, 1.0)
    assert np.count_nonzero(p) == 20 and abs(p.sum() - 1) < 1e-5
    p = osm.final_distribution(logits, 1.0, 0, 0.5)
    kept = np.sort(p[p > 0])[::-1]
    full = np.sort(osm.final_distribution(logits, 1.0, 0, 1.0))[::-1]
    assert full[:kept.size].sum() >= 0.5 > full[:kept.size - 1].sum()        # the smallest prefix reaching top_p
    lo, hi = osm.cdf_interval(p, int(np.argmax(p)))
    assert 0 <= lo < hi <= 1
