"""pgk_sample_token_bf16 (temperature / top-k / top-p sampling on the device) against oracle/sampling.py, the
restatement of pygpukit.llm.sampling.sample_token that tests/test_sampling_oracle.py pins to the reference's own outputs.

Exact comparison on the golden cases (tie-free logits: every cut is unambiguous); on vocabulary-sized random logits,
where BF16 logits tie in their thousands, the token must lie in the oracle's support and its cumulative interval must
contain the uniform number up to the probability mass of the ties at the cuts."""
import os

import numpy as np
import pytest

import pygpukit_b200 as gk
from oracle import sampling as osm
from pygpukit_b200 import ops
from tests.util import bf16, f32

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_sampling.npz")


def test_sample_token_matches_reference_golden_cases():
    g = np.load(GOLD)
    dev = {}
    n, exact, near = int(g["n_cases"]), 0, 0
    for i in range(n):
        V, t, k, p, u, tok = g[f"c{i}"]
        V, k, tok = int(V), int(k), int(tok)
        if V not in dev:
            lg = g[f"logits_{V}"]
            assert np.array_equal(f32(bf16(lg)), lg)                 # BF16-representable: both sides see the same values
            dev[V] = (gk.from_numpy(bf16(lg)), lg)
        got = ops.sample_token_gpu(dev[V][0], float(t), k, float(p), random_val=float(u))
        if got == tok:
            exact += 1
            continue
        # f32 summation order differs from NumPy's: a different token is acceptable only when u sits on the boundary
        # between the two tokens' cumulative intervals (or the top-p cut on its threshold)
        probs = osm.final_distribution(dev[V][1], float(t), k, float(p))
        assert t > 0 and probs[got] > 0, (i, got, tok)
        lo, hi = osm.cdf_interval(probs, got)
        assert lo - 1e-5 <= u <= hi + 1e-5, (i, got, tok, lo, hi, u)
        near += 1
    assert exact >= n - 3, (exact, near, n)


@pytest.mark.parametrize("V", [151936, 32000, 1000, 37])
def test_sample_token_vocabulary_sized_logits(V):
    rng = np.random.default_rng(V)
    lg = f32(bf16((rng.standard_normal(V) * 2.5).astype(np.float32)))
    g = gk.from_numpy(bf16(lg))
    # greedy: the first maximum, with and without filters
    assert ops.sample_token_gpu(g, 0.0) == int(np.argmax(lg))
    for t, k, p in ((1.0, 0, 1.0), (0.7, 0, 1.0), (1.0, 50, 1.0), (0.9, 0, 0.9), (0.8, 40, 0.95), (1.0, 1, 1.0), (1.2, 0, 0.2)):
        k_eff = k if 0 < k < V else 0
        probs = osm.final_distribution(lg, t, k_eff, p).astype(np.float64)
        support = probs > 0
        # probability mass of the elements tied with the smallest kept probability (the cut may fall among them)
        pmin = probs[support].min()
        full = osm.final_distribution(lg, t, 0, 1.0).astype(np.float64)
        tie_mass = float(full[np.isclose(full, full[support].min(), rtol=1e-6)].sum() / full[support].sum())
        for u in (0.0, 0.2718, 0.5, 0.8414, 0.99999):
            tok = ops.sample_token_gpu(g, t, k, p, random_val=u)
            assert 0 <= tok < V
            # inside the oracle's support, or tied with its smallest member
            assert support[tok] or np.isclose(full[tok], full[support].min(), rtol=1e-6), (t, k, p, u, tok)
            if support[tok]:
                lo, hi = osm.cdf_interval(probs, tok)
                eps = 1e-4 + 2 * tie_mass
                assert lo - eps <= u <= hi + eps, (t, k, p, u, tok, lo, hi, eps)
        if k_eff:
            kth = np.sort(lg)[-k_eff]
            for u in (0.0, 0.37, 0.93):
                assert lg[ops.sample_token_gpu(g, t, k, 1.0, random_val=u)] >= kth
        assert pmin > 0


def test_sample_token_ties_keep_highest_indices_and_device_random_buffer():
    # eight equal maxima: top_k = 3 keeps the three highest indices (a stable ascending argsort read from its end)
    V = 4096
    lg = np.full(V, -4.0, np.float32)
    tied = np.array([5, 77, 300, 1024, 2047, 2048, 3000, 4095])
    lg[tied] = 2.0
    g = gk.from_numpy(bf16(lg))
    seen = {ops.sample_token_gpu(g, 1.0, 3, 1.0, random_val=u) for u in np.linspace(0.0, 0.999, 41)}
    assert seen == {2048, 3000, 4095}
    assert osm.sample_token_u(lg, 1.0, 3, 1.0, 0.5) in seen
    assert ops.sample_token_gpu(g, 0.0, 3, 1.0) == 2048              # greedy after the cut: lowest kept index
    assert ops.sample_token_gpu(g, 1.0, 0, 1e-6, random_val=0.3) == 4095   # top_p -> 0 keeps one element
    # the uniform number from a device buffer (graph replay form) gives what the scalar gives
    out = gk.zeros((1,), gk.int32)
    lg2 = f32(bf16((np.random.default_rng(3).standard_normal(V) * 2).astype(np.float32)))
    g2 = gk.from_numpy(bf16(lg2))
    for u in (0.05, 0.5, 0.95):
        ub = gk.from_numpy(np.array([u], np.float32))
        ops.sample_token_to_buf(g2, out, 0.9, 64, 0.9, random_buf=ub)
        assert int(out.to_numpy()[0]) == ops.sample_token_gpu(g2, 0.9, 64, 0.9, random_val=float(np.float32(u)))
    with pytest.raises(ValueError):
        ops.sample_token_gpu(gk.zeros((16,), gk.float32), 1.0)
    with pytest.raises(ValueError):
        ops.sample_token_gpu(g, -1.0)


def test_generate_with_gpu_sampling():
    """generate(gpu_sampling=True): the logits stay on the device.  top_k = 1 leaves one candidate, so the sampled
    continuation must equal the greedy one; with a wide nucleus the tokens are valid and reproducible under a seed."""
    from pygpukit_b200 import llm
    model = llm.build_random_model("tiny-qwen3", "fp8", max_seq=96, seed=7, std=0.05)
    model.init_decode(96, max_batch_size=16)
    prompt = [3, 17, 99, 250, 7]
    greedy = model.generate(prompt, 12, temperature=0.0)
    model2 = llm.build_random_model("tiny-qwen3", "fp8", max_seq=96, seed=7, std=0.05)
    model2.init_decode(96, max_batch_size=16)
    assert model2.generate(prompt, 12, temperature=0.8, top_k=1, top_p=1.0, gpu_sampling=True) == greedy
    runs = []
    for _ in range(2):
        m = llm.build_random_model("tiny-qwen3", "fp8", max_seq=96, seed=7, std=0.05)
        m.init_decode(96, max_batch_size=16)
        ops.set_sampling_seed(1234)
        runs.append(m.generate(prompt, 12, temperature=1.0, top_k=50, top_p=0.95, gpu_sampling=True))
    assert runs[0] == runs[1] and all(0 <= t < m.config.vocab_size for t in runs[0])


def test_generate_stream_yields_the_tokens_of_generate():
    """generate_stream (causal.py:257-322): greedy streaming == generate(); with gpu_sampling and one seed, the same run."""
    from pygpukit_b200 import llm

    def fresh():
        m = llm.build_random_model("tiny-qwen3", "fp8", max_seq=96, seed=7, std=0.05)
        m.init_decode(96, max_batch_size=16)
        return m

    prompt = [3, 17, 99, 250, 7]
    greedy = fresh().generate(prompt, 10, temperature=0.0)
    assert list(fresh().generate_stream(prompt, 10, temperature=0.0)) == greedy[len(prompt):]
    ops.set_sampling_seed(99)
    a = list(fresh().generate_stream(prompt, 10, temperature=0.9, top_k=40, top_p=0.9, gpu_sampling=True))
    ops.set_sampling_seed(99)
    bb = fresh().generate(prompt, 10, temperature=0.9, top_k=40, top_p=0.9, gpu_sampling=True)[len(prompt):]
    assert a == bb and len(a) == 10
    gen = fresh().generate_stream(prompt, 10, temperature=0.0, eos_token_id=greedy[len(prompt) + 2])
    assert list(gen) == greedy[len(prompt):len(prompt) + 3]                # stops after yielding the EOS token


def test_reference_named_sampling_ops():
    """sample_multinomial / sample_topk / sample_topp / sample_topk_to_buf_ptr (ops/sampling.py:44-141) are
    sample_token_gpu with the matching filters; seeded, they give what it gives."""
    V = 5000
    lg = f32(bf16((np.random.default_rng(8).standard_normal(V) * 2).astype(np.float32)))
    g = gk.from_numpy(bf16(lg))
    for fn, args, ref in ((ops.sample_multinomial, (0.8,), (0.8, 0, 1.0)), (ops.sample_topk, (30, 0.8), (0.8, 30, 1.0)),
                          (ops.sample_topp, (0.9, 0.8), (0.8, 0, 0.9))):
        ops.set_sampling_seed(5)
        a = [fn(g, *args) for _ in range(6)]
        ops.set_sampling_seed(5)
        assert a == [ops.sample_token_gpu(g, *ref) for _ in range(6)]
    out = gk.zeros((1,), gk.int32)
    ub = gk.from_numpy(np.array([0.625], np.float32))
    ops.sample_topk_to_buf_ptr(g, out, ub, 30, 0.8)
    assert int(out.to_numpy()[0]) == ops.sample_token_gpu(g, 0.8, 30, 1.0, random_val=0.625)
    kth = np.sort(lg)[-30]
    assert lg[int(out.to_numpy()[0])] >= kth
    with pytest.raises(ValueError):
        ops.sample_topk(g, 0, 1.0)
