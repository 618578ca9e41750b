"""End-to-end parity of the decode path against the CPU restatement of the reference's step
(oracle/model.py): logits within 1e-2 relative (BF16 bar of BASELINE.json), greedy tokens identical,
graph replay == eager, for FP8 / NVF4 / BF16 weights, QK-norm on/off, dense and MoE."""
import numpy as np
import pytest

import pygpukit_b200 as gk
from oracle.model import OracleModel
from pygpukit_b200 import llm
from tests.util import f32, rel_err

pytestmark = pytest.mark.gpu

MAX_SEQ = 160
N_NEW = 128            # north_star: greedy tokens identical for the first 128 steps


def _build(name, fmt, **kw):
    m = llm.build_random_model(name, fmt, max_seq=MAX_SEQ, seed=7, std=0.05, **kw)
    m.init_decode(MAX_SEQ, max_batch_size=16)
    return m


@pytest.fixture(params=[0, 1], ids=["pdl0", "pdl1"])
def pdl(request):
    """Programmatic dependent launch off / on: the benchmark runs with it on, so parity is shown in both modes."""
    from pygpukit_b200._lib import api
    api.set_pdl(request.param)
    yield request.param
    api.set_pdl(0)


def _check_tokens(got, exp, margins, n_prompt, tops=None):
    """identical greedy tokens; a divergence is tolerated only at a step whose oracle top-1 / top-2 logits are a genuine
    tie at BF16 resolution, and ends the comparison.  Both logits are BF16-rounded dot products over a hidden state that
    went through an f32 summation order different from the oracle's (split-KV attention, tensor-core accumulation): a
    flipped final rounding moves EITHER of them by one ulp, so two candidates swap only if they are at most TWO BF16
    ulps of the top logit apart -- far inside the 1e-2 relative bar of the floating-point ops themselves (0.06 for a
    logit of 6).  Without the top logits the bar is 2^-5 (two ulps for logits in [2, 4))."""
    for i, (g, e) in enumerate(zip(got[n_prompt:], exp[n_prompt:])):
        if g != e:
            bar = 2.0 ** -5
            if tops is not None:
                bar = max(bar, 2.0 ** (np.floor(np.log2(max(abs(tops[i]), 1e-30))) - 6))
            assert margins[i] <= bar, f"token {i}: got {g}, oracle {e}, margin {margins[i]} (bar {bar})"
            return i
    return len(exp) - n_prompt


_oracle_cache: dict = {}


def _oracle_tokens(key, model, prompt, n_new):
    """The oracle's greedy continuation (cached: the random-init model is a pure function of its name / seed)."""
    if key not in _oracle_cache:
        om = OracleModel(model, MAX_SEQ)
        toks, margins = om.generate(prompt, n_new)
        _oracle_cache[key] = (toks, margins, list(om.top_logits))
    return _oracle_cache[key]


@pytest.mark.parametrize("name,fmt", [("tiny-llama", "fp8"), ("tiny-qwen3", "fp8"), ("tiny-llama", "nvf4"),
                                      ("tiny-qwen3", "bf16"), ("tiny-moe", "fp8"), ("tiny-moe", "nvf4"),
                                      ("small-qwen3", "fp8"), ("small-llama", "nvf4"), ("small-qwen3", "bf16")])
def test_greedy_tokens_match_oracle(name, fmt, pdl):
    model = _build(name, fmt)
    rng = np.random.default_rng(3)
    prompt = [int(t) for t in rng.integers(0, model.config.vocab_size, 12)]
    exp, margins, tops = _oracle_tokens((name, fmt), model, prompt, N_NEW)
    got = model.generate(prompt, N_NEW, temperature=0.0, use_graph=True)
    assert len(got) == len(prompt) + N_NEW
    agreed = _check_tokens(got, exp, margins, len(prompt), tops)
    assert agreed >= N_NEW // 4, f"only {agreed} tokens before a tie"
    # eager (no graph) path gives the same tokens as the graph replay, bit for bit
    model2 = _build(name, fmt)
    got2 = model2.generate(prompt, N_NEW, temperature=0.0, use_graph=False)
    assert got2 == got


def _real_width(family: str):
    """Full hidden size / heads / intermediate size of a BASELINE.json model with 2 layers and a 4096-token vocabulary,
    so the CPU oracle finishes: the shapes every kernel of the step sees are the real ones."""
    from pygpukit_b200.llm.config import named_config
    cfg, spec = named_config(family, max_seq=MAX_SEQ)
    cfg.num_layers, cfg.vocab_size = 2, 4096
    return cfg, spec


@pytest.mark.parametrize("family,fmt,n_new", [("llama3-8b", "fp8", 128), ("qwen3-8b", "fp8", 128),
                                              ("qwen3-30b-a3b", "nvf4", 128), ("qwen3-30b-a3b", "fp8", 64),
                                              ("llama3-70b", "nvf4", 32)])
def test_greedy_tokens_match_oracle_real_width(family, fmt, n_new):
    """One real-width case per model family of BASELINE.json (128 greedy steps; fewer for the widest model, whose
    oracle needs ~1.5 s per token on 8 cores), graph replay with PDL on -- the mode bench.py measures."""
    from pygpukit_b200._lib import api
    cfg, spec = _real_width(family)
    api.set_pdl(1)
    try:
        model = llm.build_random_model(family, fmt, max_seq=MAX_SEQ, seed=11, std=0.02, config=cfg, spec=spec)
        model.init_decode(MAX_SEQ, max_batch_size=8)
        rng = np.random.default_rng(5)
        prompt = [int(t) for t in rng.integers(0, cfg.vocab_size, 8)]
        om = OracleModel(model, MAX_SEQ)
        exp, margins = om.generate(prompt, n_new)
        got = model.generate(prompt, n_new, temperature=0.0, use_graph=True)
    finally:
        api.set_pdl(0)
    agreed = _check_tokens(got, exp, margins, len(prompt), om.top_logits)
    assert agreed >= min(16, n_new // 2), f"only {agreed} tokens before a tie"


@pytest.mark.parametrize("name,fmt", [("tiny-llama", "fp8"), ("tiny-qwen3", "nvf4")])
def test_decode_m1_logits_vs_oracle(name, fmt):
    model = _build(name, fmt)
    orc_model = OracleModel(model, MAX_SEQ)
    strat = llm.DecodeM1()
    strat.bind(model)
    rng = np.random.default_rng(5)
    toks = [int(t) for t in rng.integers(0, model.config.vocab_size, 6)]
    for pos, t in enumerate(toks):
        got = strat.step(t, pos, pos + 1, model.buffers).to_numpy()[0]
        exp, _ = orc_model.step(t, pos, pos + 1)
        assert rel_err(got, exp) < 1e-2, f"pos {pos}"
    # the graph strategy returns the same logits for the next position
    g = llm.DecodeM1Graph()
    g.bind(model)
    g.init_graph(MAX_SEQ)
    # graph capture clobbers cache slot 0: replay the prefix through the graph itself
    for pos, t in enumerate(toks):
        lg = g.step_graph(t, pos, pos + 1).to_numpy()[0]
    exp, tok = orc_model.step(toks[-1], len(toks) - 1, len(toks))
    assert rel_err(lg, exp) < 1e-2
    assert int(model.buffers.next_token.to_numpy()[0]) == int(np.argmax(f32(lg)))


@pytest.mark.parametrize("M", [4, 8])
def test_batch_decode_equals_single_steps(M):
    """DecodeBatch: M consecutive tokens in one call == M single-token steps (same cache contents).  M = 4 stays on
    the CUDA-core GEMV (same arithmetic as M = 1: the layer-0 cache is a bit-exact copy); M = 8 runs the
    projections on the tcgen05 GEMM (BF16 operands), compared at the BF16 bar."""
    model = _build("tiny-qwen3", "fp8")
    rng = np.random.default_rng(6)
    toks = [int(t) for t in rng.integers(0, model.config.vocab_size, M)]
    single = []
    for pos, t in enumerate(toks):
        single.append(model._decode_step_fixed_cache(t, pos, pos + 1).to_numpy()[0].copy())
    model_b = _build("tiny-qwen3", "fp8")
    b = llm.DecodeBatch(M)
    b.bind(model_b)
    h = b.step_batch(toks, 0, len(toks), None).to_numpy()
    for i in range(len(toks)):
        assert rel_err(h[i], single[i]) < 1e-2
    k1 = model.blocks[0].attn._k_cache.to_numpy()
    k2 = model_b.blocks[0].attn._k_cache.to_numpy()
    if M <= 4:
        assert np.array_equal(k1[:, :len(toks)], k2[:, :len(toks)])      # layer-0 cache: bit-exact copy path
    else:
        assert rel_err(k2[:, :len(toks)], k1[:, :len(toks)]) < 1e-2


def test_fp8_device_quantiser_matches_host():
    from pygpukit_b200.llm.quant import quantize_fp8_e4m3_block
    from pygpukit_b200.llm.random_init import normal_bf16, quantize_fp8
    w = normal_bf16((384, 512), 99, 0.02)
    lin = quantize_fp8(w)
    codes, scale = quantize_fp8_e4m3_block(f32(w.to_numpy()))
    assert np.array_equal(lin.scale_inv.to_numpy(), scale)
    assert np.array_equal(lin.weight_fp8.to_numpy(), codes)
    z = f32(w.to_numpy())
    assert abs(z.std() - 0.02) < 1e-3 and abs(z.mean()) < 1e-3      # the device RNG is N(0, std^2)


def test_api_surface_and_errors():
    for n in ("CausalTransformerModel", "GPT2Model", "LlamaModel", "QwenModel", "TransformerConfig", "Attention",
              "MLP", "MoELayer", "Norm", "TransformerBlock", "Linear", "LinearBF16", "LinearFP8", "DecodeM1",
              "DecodeM1Graph", "DecodeBatch", "DecodeBuffers", "precompute_freqs_cis", "sample_token"):
        assert hasattr(llm, n), n
    lin = llm.LinearFP8(gk.zeros((128, 256), gk.uint8), gk.zeros((1, 2), gk.bfloat16))
    with pytest.raises(ValueError):
        lin(gk.zeros((256,), gk.bfloat16))                # 1-D input
    with pytest.raises(ValueError):
        lin(gk.zeros((1, 128), gk.bfloat16))              # wrong in_features
    s = llm.DecodeM1()
    with pytest.raises(RuntimeError):
        _ = s.model                                       # not bound


@pytest.mark.parametrize("fmt", ["fp8", "nvf4", "bf16"])
def test_fused_rmsnorm_gemv_is_bit_identical_to_two_launches(fmt):
    """pgk_rmsnorm_gemv_* == pgk_rmsnorm followed by the plain GEMV, bit for bit (same summation order)."""
    from pygpukit_b200.llm.random_init import normal_bf16, quantize_fp8, quantize_nvf4
    from tests.util import rand_bf16
    rng = np.random.default_rng(17)
    K, N = 1024, 768
    for M in (1, 2, 4, 5):
        x = gk.from_numpy(rand_bf16(rng, (M, K)))
        norm = llm.Norm(gk.from_numpy(rand_bf16(rng, (K,), 0.5)), None, "rmsnorm", 1e-6)
        if fmt == "fp8":
            lin = quantize_fp8(normal_bf16((N, K), 5, 0.05))
        elif fmt == "nvf4":
            lin = quantize_nvf4(normal_bf16((K, N), 5, 0.05))
        else:
            lin = llm.LinearBF16(normal_bf16((N, K), 5, 0.05))
        two = lin.forward(norm(x)).to_numpy()
        one = lin.forward(x, norm=norm).to_numpy()
        assert np.array_equal(one, two), f"M={M}"


@pytest.mark.parametrize("fmt", ["fp8", "nvf4"])
@pytest.mark.parametrize("offset", [0, 2])
@pytest.mark.parametrize("wide", [False, True])
def test_moe_decode_form_equals_generic_pipeline(fmt, offset, wide):
    """T == 1 MoE fast path (route_t1 / gateup_fp8_t1 / down_combine_fp8_t1 / broadcast rows / silu_mul /
    combine_t1) against the 13-launch generic pipeline of the same layer: bit-identical output, with
    and without a residual, also when only part of the experts is local (expert parallel)."""
    from pygpukit_b200.llm.layers import moe as moe_mod

    kw = {}
    if wide:                                    # 64 experts, 8 per token, ragged K (384 = 24 chunks of 16)
        from pygpukit_b200.llm.config import named_config
        cfg, spec = named_config("tiny-moe", max_seq=MAX_SEQ)
        cfg.num_experts, cfg.num_experts_per_tok, cfg.moe_intermediate_size, cfg.num_layers = 64, 8, 384, 1
        kw = {"config": cfg, "spec": spec}
    model = _build("tiny-moe", fmt, **kw)
    layer = next(b.mlp for b in model.blocks if isinstance(b.mlp, llm.MoELayer))
    if offset:                                  # pretend this rank holds experts [offset, offset + local)
        layer.expert_offset = offset
    H = layer.hidden_size
    rng = np.random.default_rng(11)
    for trial in range(4):
        x = gk.from_numpy(rng.standard_normal((1, H)).astype(np.float32)).astype(gk.bfloat16)
        res = gk.from_numpy(rng.standard_normal((1, H)).astype(np.float32)).astype(gk.bfloat16)
        for residual in (None, res):
            outs = []
            for form in (False, True):                       # generic pipeline / decode form
                moe_mod._DECODE_FORM = form
                try:
                    outs.append(layer.forward(x, residual=residual).to_numpy().copy())
                finally:
                    moe_mod._DECODE_FORM = True
            assert np.array_equal(outs[0], outs[1]), f"trial {trial} residual={residual is not None}"
            assert np.any(outs[0] != 0)


# ---------------------------------------------------------------------------- GPT-2 (BASELINE.json configs[0] family)
def _gpt2_tiny_model():
    """The tiny GPT-2 of tests/golden/gpt2_tiny.py assembled through the reference's constructors
    (Attention with biases, GELU MLP, LayerNorm, learned positions, tied lm_head), BF16 on the GPU."""
    from tests.golden import gpt2_tiny as T

    w = T.weights()
    cfg = llm.GPT2Config(**T.CFG).to_transformer_config()

    def g(a):
        return gk.from_numpy(np.ascontiguousarray(a, np.float32)).astype(gk.bfloat16)

    def norm(p):
        return llm.Norm(g(w[p + ".g"]), g(w[p + ".b"]), "layernorm", cfg.norm_eps)

    blocks = []
    for i in range(cfg.num_layers):
        p = f"h{i}."
        attn = llm.Attention(g(w[p + "q.w"]), g(w[p + "k.w"]), g(w[p + "v.w"]), g(w[p + "o.w"]), cfg,
                             q_bias=g(w[p + "q.b"]), k_bias=g(w[p + "k.b"]), v_bias=g(w[p + "v.b"]), o_bias=g(w[p + "o.b"]))
        mlp = llm.MLP(cfg, fc1_weight=g(w[p + "fc1.w"]), fc1_bias=g(w[p + "fc1.b"]), fc2_weight=g(w[p + "fc2.w"]),
                      fc2_bias=g(w[p + "fc2.b"]))
        blocks.append(llm.TransformerBlock(norm(p + "ln_1"), attn, norm(p + "ln_2"), mlp))
    model = llm.GPT2Model(cfg, g(w["wte"]), blocks, norm("ln_f"), lm_head=None, position_embed=g(w["wpe"]),
                          spec=llm.GPT2_SPEC)
    model.init_decode(T.CFG["n_positions"], max_batch_size=16)
    return model, T


def test_gpt2_matches_reference_cpu_run():
    """GPT-2 (LayerNorm, biases, GELU, learned positions, no RoPE, MHA head_dim 64) through the B200 decode path
    in BF16 against the REFERENCE's own FP32 CPU run of the same weights (tests/golden/golden_gpt2.npz).
    Teacher-forced on the reference's tokens: logits within 2e-2 of the logit scale (BF16 end to end over two
    layers; the per-op bar is 1e-2), identical argmax wherever the reference's top-2 margin is clear; then free
    greedy generation (graph replay and eager) identical up to the first near-tie."""
    import os

    model, T = _gpt2_tiny_model()
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "golden_gpt2.npz"))
    toks = [int(t) for t in g["tokens"]]
    P, scale = len(T.PROMPT), float(g["logit_scale"])
    clear = g["margins"] > 0.12                       # ~2 bf16 ulps at |logit| ~ 8
    assert clear.sum() >= T.N_NEW - 2

    # (1) whole sequence in slabs of 16 rows (batch path), logits of the generated positions
    hidden, _ = model(toks[:-1])
    n_last = (len(toks) - 1) % 16 or 16
    assert hidden.shape[0] == n_last
    # (2) single-token steps over the generated positions (M == 1 path), teacher-forced
    model.init_decode(T.CFG["n_positions"], max_batch_size=16)
    model(toks[:P - 1])
    step_logits = []
    for i in range(P - 1, len(toks) - 1):
        h = model._decode_step_fixed_cache(toks[i], i, i + 1)
        step_logits.append(f32(model.get_logits(h).to_numpy())[0])
    step_logits = np.stack(step_logits)
    err = np.abs(step_logits - g["logits_teacher_forced"]).max()
    assert err <= 2e-2 * scale, f"logit error {err} vs scale {scale}"
    am = step_logits.argmax(-1)
    assert np.array_equal(am[clear], g["argmax_teacher_forced"][clear])
    # the batch path's last slab agrees with the single-token path on the same rows
    slab_logits = f32(model.get_logits(hidden).to_numpy())
    assert np.abs(slab_logits - step_logits[-n_last:]).max() <= 2e-2 * scale

    # (3) free greedy generation, device-resident graph replay and eager
    for use_graph in (True, False):
        model.init_decode(T.CFG["n_positions"], max_batch_size=16)
        got = model.generate(list(T.PROMPT), T.N_NEW, temperature=0.0, use_graph=use_graph)
        agreed = 0
        for i in range(T.N_NEW):
            if got[P + i] != toks[P + i]:
                assert not clear[i], f"step {i}: got {got[P + i]}, reference {toks[P + i]}, margin {g['margins'][i]}"
                break
            agreed += 1
        assert agreed >= 3, agreed


def test_gpt2_position_table_bound():
    model, T = _gpt2_tiny_model()
    with pytest.raises(ValueError):
        model._decode_step_fixed_cache_batch([1, 2, 3], T.CFG["n_positions"] - 1, T.CFG["n_positions"])


def test_gpt2_small_full_size_graph_equals_eager():
    """BASELINE.json configs[0] at its full size on the GPU path: GPT-2 small (124M, random init), 128-token prompt +
    128 greedy tokens; the device-resident graph replay and the eager per-op path give the same tokens."""
    model = llm.build_random_model("gpt2", "bf16", max_seq=1024, seed=3, std=0.02)
    assert model.config.num_layers == 12 and model.config.hidden_size == 768 and model.position_embed.shape == (1024, 768)
    rng = np.random.default_rng(9)
    prompt = [int(t) for t in rng.integers(0, model.config.vocab_size, 128)]
    model.init_decode(1024, max_batch_size=64)
    a = model.generate(prompt, 128, temperature=0.0, use_graph=True)
    assert len(a) == 256 and all(0 <= t < model.config.vocab_size for t in a)
    model.init_decode(1024, max_batch_size=64)
    b = model.generate(prompt, 32, temperature=0.0, use_graph=False)
    assert a[:160] == b


@pytest.mark.parametrize("name,fmt,S", [("small-qwen3", "fp8", 3), ("small-qwen3", "nvf4", 3), ("small-qwen3", "fp8", 8)])
def test_multi_sequence_decode_matches_single_sequence_runs(name, fmt, S):
    """Multi-sequence decode (replaces the reference's unused paged_attention.cuh:46-283 /
    continuous_batching.cuh:70-245): S independent sequences with different prompt lengths decode together -- one row
    per sequence through every linear layer, ONE attention launch per layer over the S cache regions with per-sequence
    positions -- and must produce the tokens each sequence produces when it decodes alone (the split-KV fold order and,
    from 5 rows, the tensor-core GEMM differ, so a divergence is accepted only where the lone run's two best logits are
    equal or adjacent BF16 numbers)."""
    n_new = 48
    rng = np.random.default_rng(17)
    model = llm.build_random_model(name, fmt, max_seq=MAX_SEQ, seed=7, std=0.05)
    V = model.config.vocab_size
    prompts = [[int(t) for t in rng.integers(0, V, 5 + 3 * (s % 4))] for s in range(S)]
    # ---- each sequence alone
    alone = []
    for p in prompts:
        model.init_decode(MAX_SEQ, max_batch_size=16)
        alone.append(model.generate(p, n_new, temperature=0.0)[len(p):])
    # ---- together: graph first (its warm-up step writes cache row 0), then the prompts, then S-row steps
    model.init_decode(MAX_SEQ, max_batch_size=16, num_sequences=S)
    model.init_decode_graph()
    firsts = []
    for s, p in enumerate(prompts):
        model.select_sequence(s)
        hidden = model.prefill(p)
        lg = model.get_logits(hidden.row(hidden.shape[0] - 1).view((1, hidden.shape[1]))).to_numpy()[0]
        firsts.append(int(np.argmax(f32(lg))))
    model.set_decode_state(firsts, [len(p) for p in prompts])
    together = [[t] for t in firsts]
    from pygpukit_b200._lib import api
    for _ in range(n_new - 1):
        model.decode_graph_replay()
        api.stream_sync(model._graph_stream.handle)
        for s, t in enumerate(model.buffers.next_token.to_numpy()[:S]):
            together[s].append(int(t))
    # ---- compare
    for s, p in enumerate(prompts):
        same = next((i for i, (a, b) in enumerate(zip(together[s], alone[s])) if a != b), n_new)
        if same < n_new:
            model.init_decode(MAX_SEQ, max_batch_size=16)
            hidden = model.prefill(p + alone[s][:same])
            lg = np.sort(f32(model.get_logits(hidden.row(hidden.shape[0] - 1).view((1, hidden.shape[1]))).to_numpy()[0]))
            ulp = 2.0 ** (np.floor(np.log2(max(abs(lg[-1]), 1e-30))) - 7)
            assert lg[-1] - lg[-2] <= ulp, f"sequence {s}: token {same} differs ({together[s][same]} vs {alone[s][same]}), margin {lg[-1] - lg[-2]}"
        assert same >= 8, f"sequence {s}: only {same} identical tokens"
