"""Self-speculative decode (causal.py:1036-1235, decode/speculative.py): the early layers draft, the full
model verifies in one batched pass.  Whatever the draft proposes, the emitted tokens are the full model's greedy tokens:
they are checked against the oracle's greedy continuation with the tie rule of tests/test_gpu_model.py (the batched
verification pass sums attention in a different order than single steps, so only BF16-resolution ties may differ)."""
import numpy as np
import pytest

from pygpukit_b200 import llm
from pygpukit_b200.core.dtypes import bf16_bits_to_f32
from tests.test_gpu_model import MAX_SEQ, _build, _check_tokens, _oracle_tokens

pytestmark = pytest.mark.gpu
N_NEW = 48


def _speculate(model, prompt, n_new, draft_tokens, draft_layers, via_strategy=False):
    hidden = model.prefill(prompt)
    last = hidden.row(hidden.shape[0] - 1).view((1, hidden.shape[1]))
    tok = int(np.argmax(bf16_bits_to_f32(model.get_logits(last).to_numpy()[0])))
    out, drafted, agreed, steps = [tok], 0, 0, 0
    model.set_lookahead_confirmed_pos(len(prompt))
    strat = llm.DecodeSpeculative(draft_tokens, draft_layers)
    strat.bind(model)
    while len(out) < n_new:
        pos = len(prompt) + len(out) - 1
        if via_strategy:
            accepted, new_pos, stats = strat.step_speculative(tok, pos, pos + 1)
            assert new_pos == pos + len(accepted)
        else:
            assert model.get_lookahead_confirmed_pos() == pos
            accepted, stats = model.decode_step_self_speculative_lookahead(tok, draft_tokens, draft_layers)
        assert 1 <= len(accepted) <= draft_tokens
        assert stats["draft_count"] == draft_tokens and 0 <= stats["accepted_count"] <= len(accepted)
        assert stats["accepted_count"] >= len(accepted) - 1            # only the last token may be a correction
        out.extend(accepted)
        drafted += stats["draft_count"]
        agreed += stats["accepted_count"]
        steps += 1
        tok = accepted[-1]
    return out[:n_new], drafted, agreed, steps


@pytest.mark.parametrize("name,fmt", [("small-qwen3", "fp8"), ("small-llama", "nvf4")])
@pytest.mark.parametrize("via_strategy", [False, True])
def test_speculative_tokens_are_the_full_models_greedy_tokens(name, fmt, via_strategy):
    model = _build(name, fmt)
    rng = np.random.default_rng(3)
    prompt = [int(t) for t in rng.integers(0, model.config.vocab_size, 12)]
    exp, margins, tops = _oracle_tokens((name, fmt), model, prompt, 128)
    n_layers = len(model.blocks)
    got, drafted, agreed, steps = _speculate(model, prompt, N_NEW, 4, max(1, n_layers // 2), via_strategy)
    ok = _check_tokens(list(prompt) + got, exp[:len(prompt) + N_NEW], margins, len(prompt), tops)
    assert ok >= N_NEW // 4, f"only {ok} tokens before a tie"
    assert steps <= N_NEW and drafted == 4 * steps


def test_full_depth_draft_is_always_accepted():
    """With every layer in the draft the draft IS the model: (almost) every drafted token is accepted and a step emits
    max_draft_tokens tokens -- the bookkeeping of positions / committed rows is right."""
    model = _build("small-qwen3", "fp8")
    prompt = [5, 99, 1023, 77, 300, 12, 640, 8]
    got, drafted, agreed, steps = _speculate(model, prompt, 40, 4, len(model.blocks))
    assert agreed >= 0.9 * drafted, (agreed, drafted)
    assert steps <= 13
    with pytest.raises(ValueError):
        model.decode_step_self_speculative_lookahead(got[-1], 64, 1)      # more drafts than max_batch_size


def _jacobi(model, prompt, n_new, n_tokens, max_iter, init, via_strategy=False):
    hidden = model.prefill(prompt)
    last = hidden.row(hidden.shape[0] - 1).view((1, hidden.shape[1]))
    tok = int(np.argmax(bf16_bits_to_f32(model.get_logits(last).to_numpy()[0])))
    out, steps, iters, full = [tok], 0, 0, 0
    model.set_lookahead_confirmed_pos(len(prompt))
    strat = llm.DecodeJacobi(n_tokens, max_iter, init)
    strat.bind(model)
    while len(out) < n_new:
        pos = len(prompt) + len(out) - 1
        if via_strategy:
            accepted, new_pos, stats = strat.step_jacobi(tok, pos, pos + 1)
            assert new_pos == pos + len(accepted)
        else:
            assert model.get_lookahead_confirmed_pos() == pos
            accepted, stats = model.decode_step_jacobi_lookahead(tok, n_tokens, max_iter, init)
        assert 1 <= len(accepted) <= n_tokens and stats["accepted_count"] == len(accepted)
        assert 1 <= stats["iterations"] <= max_iter and (not stats["converged"] or len(accepted) == n_tokens)
        out.extend(accepted)
        steps += 1
        iters += stats["iterations"]
        full += int(len(accepted) == n_tokens)
        tok = accepted[-1]
    return out[:n_new], steps, iters, full


@pytest.mark.parametrize("n_tokens,max_iter,init,via_strategy", [(4, 3, "repeat", False), (8, 3, "repeat", True),
                                                                 (4, 2, "ngram", False), (4, 1, "greedy", False)])
def test_jacobi_tokens_are_the_models_greedy_tokens(n_tokens, max_iter, init, via_strategy):
    """Jacobi decode (causal.py:1357-1476, decode/jacobi.py) emits the greedy continuation whatever the guesses were;
    n_tokens = 8 runs the projections on the tcgen05 GEMM, so, as for every batched pass, only BF16-resolution ties may
    differ from the oracle's single-token steps."""
    name, fmt = "small-qwen3", "fp8"
    model = _build(name, fmt)
    rng = np.random.default_rng(3)
    prompt = [int(t) for t in rng.integers(0, model.config.vocab_size, 12)]
    exp, margins, tops = _oracle_tokens((name, fmt), model, prompt, 128)
    got, steps, iters, full = _jacobi(model, prompt, N_NEW, n_tokens, max_iter, init, via_strategy)
    ok = _check_tokens(list(prompt) + got, exp[:len(prompt) + N_NEW], margins, len(prompt), tops)
    assert ok >= N_NEW // 4, f"only {ok} tokens before a tie"
    assert steps <= N_NEW and iters <= steps * max_iter
    if init == "greedy":                           # the guess is the answer: (almost) every step accepts all n tokens
        assert full >= 0.8 * steps, (full, steps)
    with pytest.raises(ValueError):
        model.decode_step_jacobi_lookahead(got[-1], 4, 3, "nonsense")


def test_kv_cache_snapshot_restore_round_trip():
    """snapshot_kv_cache / restore_kv_cache (causal.py:1003-1034): after steps that overwrite cache rows, restoring the
    snapshot brings back the same bits, in the same device arrays (pointers unchanged), and the same next token."""
    model = _build("tiny-qwen3", "fp8")
    prompt = [5, 99, 1023, 77, 300, 12]
    hidden = model.prefill(prompt)
    ptrs = [(b.attn._k_cache.device_ptr, b.attn._v_cache.device_ptr) for b in model.blocks]
    snap = model.snapshot_kv_cache()
    pos = len(prompt)
    h1 = model._decode_step_fixed_cache(7, pos, pos + 1).to_numpy().copy()
    model._decode_step_fixed_cache(9, pos, pos + 1)                     # another token overwrites row `pos`
    assert any(not np.array_equal(b.attn._k_cache.to_numpy(), s[0]) for b, s in zip(model.blocks, snap))
    model.restore_kv_cache(snap)
    for b, (k, v), (pk, pv) in zip(model.blocks, snap, ptrs):
        assert np.array_equal(b.attn._k_cache.to_numpy(), k) and np.array_equal(b.attn._v_cache.to_numpy(), v)
        assert (b.attn._k_cache.device_ptr, b.attn._v_cache.device_ptr) == (pk, pv)
    assert np.array_equal(model._decode_step_fixed_cache(7, pos, pos + 1).to_numpy(), h1)
    with pytest.raises(ValueError):
        model.restore_kv_cache(snap[:-1])
