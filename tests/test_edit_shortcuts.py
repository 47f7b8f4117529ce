"""The two shortcuts of the edit-distance replay (humid_b200/csrc/edit.cu: frame_alive, zero_budget_match),
restated in Python and checked against the recursion they prune — levenshtein_<4,T,false>,
reference lib/trie/src/trieFunctions.tcc:196-233 — on single stored words: no frame with a yield below
it may be called dead, and a live frame without budget yields exactly one word, once."""
import functools
import random

import pytest


def frame_alive(n, t, p, k):
    if p < t:
        return k >= t - p
    if t < n:
        rem = n - t
        subs = min(rem, n - p) if p < n else 0
        return k >= rem - subs
    return True


def yields_below(w, v, d):
    """Number of yields of the recursion below every frame (t, p, k) for one stored word v."""
    n = len(w)

    @functools.lru_cache(None)
    def Y(t, p, k):
        if k < 0:
            return 0
        tot = Y(t, p + 1, k - 1) if p + 1 <= n + d + 1 else 0  # deletion, also past the end of the query (:202-203)
        if t < n:
            c, wl = v[t], (w[p] if p < n else 0)
            if c >= wl:  # asymmetric: children start at word[position]
                if p < n:
                    tot += Y(t + 1, p + 1, k - (c != wl))  # substitution / match
                tot += Y(t + 1, p, k - 1)  # insertion
        if t == n and p >= n:
            tot += 1  # node->leaf
        return tot

    return Y


@pytest.mark.parametrize("n,d", [(5, 2), (6, 3), (8, 2), (7, 3)])
def test_pruning_rules_against_the_recursion(n, d):
    rng = random.Random(100 * n + d)
    for _ in range(400):
        w = [rng.randrange(4) for _ in range(n)]
        v = list(w)
        for _ in range(rng.randrange(0, 3)):
            op, i = rng.randrange(3), rng.randrange(n)
            if op == 0:
                v[i] = rng.randrange(4)
            elif op == 1:
                v = v[:i] + v[i + 1:] + [rng.randrange(4)]
            else:
                v = v[:i] + [rng.randrange(4)] + v[i:-1]
        Y = yields_below(tuple(w), tuple(v), d)
        for t in range(n + 1):
            for p in range(n + d + 2):
                for k in range(d + 1):
                    y = Y(t, p, k)
                    if y:
                        assert frame_alive(n, t, p, k), (w, v, t, p, k, y)
                    if k == 0 and frame_alive(n, t, p, k):
                        if t < n:  # zero_budget_match: the remaining letters are the query's
                            assert p == t and y == (1 if v[t:] == w[t:] else 0)
                        else:
                            assert y == 1
