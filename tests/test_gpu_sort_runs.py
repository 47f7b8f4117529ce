"""sort_words orders the words by their 32 leading key bits and finishes words that share them in the gather behind the
sort (humid_b200/csrc/words.cu: gather_sorted_words_kernel, sort_long_runs_kernel).  Libraries built so that such runs have
every size — pairs, a few dozen (one thread), hundreds to 2048 (bitonic network of one CTA), many thousands (counting
sort in global memory) — must give the oracle's words, counts, first tuples and everything downstream; the 64-bit sort
(HUMID_GPU_SORT=full) and word lengths with at most / exactly / more than 32 key bits are run as well.  The order of
the words is the reference's trie walk order (lib/trie/src/trieFunctions.tcc:129-143): index work, bit-exact."""
import numpy as np
import pytest

import oracle
from humid_b200 import api
from test_gpu_parity import assert_same_as_oracle, run_gpu

pytestmark = pytest.mark.gpu

LETTERS = np.frombuffer(b"ACGT", dtype=np.uint8)


def library_with_runs(n, shared, run_sizes, copies, seed):
    """Words of n letters: for every entry of run_sizes one random stem of `shared` letters with that many distinct
    tails; every word appears 1..copies times; tuple order shuffled."""
    rng = np.random.default_rng(seed)
    words = []
    for size in run_sizes:
        stem = rng.integers(0, 4, shared)
        tails = set()
        while len(tails) < size:
            tails.add(tuple(rng.integers(0, 4, n - shared)))
        for t in tails:
            words.append(np.concatenate([stem, np.array(t)]))
    words = np.array(words, dtype=np.uint8)
    reps = rng.integers(1, copies + 1, len(words))
    rows = np.repeat(words, reps, axis=0)
    rng.shuffle(rows)
    return LETTERS[rows]


@pytest.mark.parametrize("n,shared,run_sizes", [
    (24, 16, [1] * 3000 + [2] * 500 + [3] * 200 + [40] * 30 + [64] * 5 + [65] * 5 + [300] * 6 + [2048] * 2 + [2049, 5000, 20000]),
    (24, 20, [1] * 2000 + [2] * 300 + [17] * 40 + [256] * 3),   # tails of 4 letters: at most 256 words per run
    (17, 16, [1] * 4000 + [2] * 1000 + [4] * 500),               # 34 key bits: two bits below the prefix
    (20, 16, [1] * 1000 + [70] * 20 + [256] * 4),
])
@pytest.mark.parametrize("sort", ["prefix", "full"])
def test_runs_of_every_size(monkeypatch, n, shared, run_sizes, sort):
    if sort == "full":
        monkeypatch.setenv("HUMID_GPU_SORT", "full")
    chars = library_with_runs(n, shared, run_sizes, 3, seed=n * 100 + shared)
    half = n // 2
    seg_take = [half, n - half]
    r = oracle.Result(oracle.pack_words(chars, seg_take), n, 1, False, False)
    dd = run_gpu(chars, seg_take, n, 1, False, False, batches=3)
    assert_same_as_oracle(dd, r, len(chars))
    names = {st["name"] for st in dd.stage_stats()}
    assert {"sort_words", "rank_columns"} <= names


@pytest.mark.parametrize("n", [8, 16, 31])
def test_word_lengths_around_the_prefix(n):
    """16 letters = 32 key bits: the prefix is the whole key; fewer: fewer passes; 31 letters: 30 bits below it."""
    rng = np.random.default_rng(n)
    chars = LETTERS[rng.integers(0, 4, (60000, n))]
    chars[::7] = chars[0]  # a heavy word
    chars[1::11, : n - 1] = chars[1, : n - 1]  # words that differ in the last letter only
    half = n // 2
    r = oracle.Result(oracle.pack_words(chars, [half, n - half]), n, 1, False, True)
    dd = run_gpu(chars, [half, n - half], n, 1, False, True, batches=2)
    assert_same_as_oracle(dd, r, len(chars))
