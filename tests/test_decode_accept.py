"""Acceptance rules of the self-speculative and Jacobi strategies (pygpukit_b200/llm/decode/accept.py) on a toy
deterministic language model: whatever is drafted or guessed, the emitted tokens are the sequential greedy tokens."""
import numpy as np

from pygpukit_b200.llm.decode.accept import accept_jacobi, accept_speculative

V = 11


def toy_next(history):
    """greedy next token of a toy model: a fixed function of the whole history"""
    h = 1469598103934665603
    for t in history:
        h = ((h ^ (t + 1)) * 1099511628211) % (1 << 61)
    return h % V


def greedy(prompt, n):
    seq = list(prompt)
    for _ in range(n):
        seq.append(toy_next(seq))
    return seq[len(prompt):]


def batched_pass(history, inputs):
    """row argmaxes of one batched pass over `inputs` appended to the committed history"""
    return [toy_next(history + inputs[:i + 1]) for i in range(len(inputs))]


def test_speculative_emits_greedy_tokens_for_any_draft():
    rng = np.random.default_rng(0)
    prompt, want = [3, 1, 4], None
    want = greedy(prompt, 60)
    for trial in range(20):
        hist, out = list(prompt[:-1]), []
        tok = prompt[-1]
        while len(out) < 60:
            truth = greedy(hist + [tok], 4)
            draft = [t if rng.random() < 0.6 else int(rng.integers(0, V)) for t in truth]      # a draft that is often right
            target = batched_pass(hist, [tok] + draft[:-1])
            acc = accept_speculative(draft, target)
            assert 1 <= len(acc) <= 4
            hist += [tok] + acc[:-1]
            tok = acc[-1]
            out += acc
        assert out[:60] == want


def test_jacobi_emits_greedy_tokens_for_any_guess_and_beats_the_reference_rule():
    rng = np.random.default_rng(1)
    prompt = [2, 7, 1]
    want = greedy(prompt, 60)
    emitted_ours = emitted_ref = passes = 0
    for trial in range(20):
        hist, out, tok = list(prompt[:-1]), [], prompt[-1]
        while len(out) < 60:
            n = 5
            truth = greedy(hist + [tok], n)
            guess = [t if rng.random() < 0.5 else int(rng.integers(0, V)) for t in truth]
            prev = None
            for _ in range(3):
                new = batched_pass(hist, [tok] + guess[:-1])
                prev, guess = guess, new
                passes += 1
                if new == prev:
                    break
            acc = accept_jacobi(prev, guess)
            ref = []                                   # the reference's rule: the agreeing prefix only (at least one token)
            for p, g in zip(prev, guess):
                if p != g:
                    break
                ref.append(g)
            ref = ref or [guess[0]]
            assert acc[:len(ref)] == ref and len(acc) >= len(ref)
            emitted_ours += len(acc)
            emitted_ref += len(ref)
            hist += [tok] + acc[:-1]
            tok = acc[-1]
            out += acc
        assert out[:60] == want
    assert emitted_ours > emitted_ref                  # one more token whenever a step does not fully converge
