"""Acceptance rules of the speculative strategies, as pure functions (host logic; unit-tested on CPU with a toy model).

Both rest on one fact about a batched pass over inputs [t, g[0], ..., g[n-2]]: output i is the model's greedy token
after (…, t, g[0..i-1]), hence the TRUE next token whenever g[0..i-1] are the true tokens."""
from __future__ import annotations


def accept_speculative(draft: list[int], target: list[int]) -> list[int]:
    """Self-speculative step (causal.py:1150-1159): `target[i]` is the full model's token after the inputs
    [token, draft[0..i-1]].  Accept drafts while they agree; the first disagreement is replaced by the model's token."""
    accepted = []
    for d, t in zip(draft, target):
        accepted.append(t)
        if t != d:
            break
    return accepted


def accept_jacobi(prev: list[int], new: list[int]) -> list[int]:
    """Jacobi step: `new` = row argmaxes of the pass whose inputs were [token, prev[:-1]].  new[0] is always right;
    new[i] is right if prev[0..i-1] were, which is known as soon as new[j] == prev[j] for all j < i.  So: the prefix on
    which the two agree plus the first token after it (the reference, causal.py:1426-1438, stops one token earlier)."""
    accepted = []
    for p, g in zip(prev, new):
        accepted.append(g)
        if g != p:
            break
    return accepted


__all__ = ["accept_speculative", "accept_jacobi"]
