"""Pins the CPU oracle (oracle/humid_oracle.c) before anything trusts it:

  1. every known-answer vector of the reference's own unit tests for this path
     (tests/test_fastq.cc, tests/test_cluster.cc, lib/trie/tests/test_trie.cc — cited per test);
  2. the survey's probe vectors, checked against the reference binary;
  3. [marker `ref`] random word sets through oracle/_ref/trie_probe (the unmodified trie +
     cluster code) for every mode, and synthetic FastQ through oracle/_ref/humid.
"""
import os

import numpy as np
import pytest

import oracle


# ---------------------------------------------------------------- tests/test_fastq.cc
def test_extract_umi_vectors():  # test_fastq.cc:9-46
    e = oracle.extract_umi
    assert e("header_AATT") == "AATT"
    assert e("Instrument:RunID:FlowCellID:Lane:Tile:X:Y:ATCG") == "ATCG"
    for h in ["header", "header with spaces", "header_with_many_underscores and space",
              "header_ignore_lowercase_umi_aatt", "header space then_underscore",
              "header space then_underscore_AATT",
              "Instrument:RunID:FlowCellID:Lane:Tile:X:Y more stuf"]:
        assert e(h) == "", h
    assert e("header_AATT with spaces") == "AATT"
    assert e("header_with_many_underscores_AATT") == "AATT"
    assert e("header_with_many_underscores_AATT and space") == "AATT"
    assert e("Instrument:RunID:FlowCellID:Lane:Tile:X:Y:ATCG more stuf") == "ATCG"
    assert e("Instrument:RunID:FlowCellID:Lane:Tile:X:Y:ATCG more_underscore") == "ATCG"


def test_valid_umi_vectors():  # test_fastq.cc:168-181 (through extractUMI_)
    for bad in ["", "atcg", "ATCGP", "1234", "ATCGN"]:
        assert oracle.extract_umi("h_" + bad) == ""
    assert oracle.extract_umi("h_A") == "A"
    assert oracle.extract_umi("h_ATCG") == "ATCG"


def test_make_word_vector():  # test_fastq.cc:48-56
    chars = np.frombuffer(b"AAAATTTT", dtype=np.uint8).reshape(1, 8)
    key = oracle.pack_words(chars, [4, 4])[0]
    assert oracle.key_to_word(key, 8) == "AAAATTTT"
    assert [(int(key) >> (2 * (7 - p))) & 3 for p in range(8)] == [0, 0, 0, 0, 3, 3, 3, 3]


def test_get_nucleotides_vectors():  # test_fastq.cc:58-110, :157-166
    g = oracle.get_nucleotides
    seqs = ["TTTT", "GGGG"]
    assert g("header_AAAA", seqs, [4, 4], 4) == "AAAATTTTGGGG"
    assert g("header_AAAA", seqs, [4, 4], 6) == "AAAANNTTTTGGGG"
    assert g("header_AAAA", seqs, [5, 5], 4) == "AAAATTTTNGGGGN"
    assert g("header_AAAA", seqs, [0, 0], 3) == "AAA"
    assert g("header_AAAA", seqs, [2, 2], 0) == "TTGG"
    assert g("header_AAAA", seqs, [1, 3], 0) == "TGGG"
    assert g("header_AAAAAA", ["TTTT"], [0], 4) == "AAAA"


def test_make_string_size_vectors():  # test_fastq.cc:197-202
    for size, want in [(0, ""), (1, "A"), (2, "AA"), (3, "AAN")]:
        assert oracle.get_nucleotides("h", ["AA"], [size], 0) == want


def test_nt_from_file_vectors():  # test_fastq.cc:112-155
    f = oracle.nt_from_file
    assert f(1, 10) == [10]
    assert f(3, 1) == [0, 0, 1]
    assert f(3, 2) == [0, 0, 2]
    assert f(3, 3) == [1, 1, 1]
    assert f(3, 13) == [4, 4, 5]
    assert f(3, 12) == [4, 4, 4]
    assert f(3, 11) == [3, 3, 5]
    assert f(3, 10) == [3, 3, 4]
    assert f(3, 9) == [3, 3, 3]
    assert f(3, 0) == [0, 0, 0]


def test_pre_compute():  # src/humid.cc:38-59
    assert oracle.pre_compute(8, 24, 2) == (8, [8, 8])
    assert oracle.pre_compute(0, 24, 3) == (0, [8, 8, 8])
    assert oracle.pre_compute(10, 24, 2) == (10, [7, 7])
    assert oracle.pre_compute(30, 24, 2) == (24, [0, 0])
    assert oracle.pre_compute(0, 25, 2) == (0, [12, 13])


def test_filtering_rules():  # src/fastq.cc:146-161
    chars = np.frombuffer(b"ACGTACGN" b"ACGTacgt" b"ACGTACGT", dtype=np.uint8).reshape(3, 8)
    k = oracle.pack_words(chars, [4, 4])
    assert k[0] == oracle.INVALID and k[1] == oracle.INVALID
    assert oracle.key_to_word(k[2], 8) == "ACGTACGT"
    seg_len = np.array([[4, 3]], dtype=np.uint16)  # short second read => 'N' pad => filtered
    assert oracle.pack_words(chars[2:3], [4, 4], seg_len)[0] == oracle.INVALID


# ---------------------------------------------------------- lib/trie/tests/test_trie.cc
W5 = [[0, 1, 2], [0, 1, 3], [0, 2, 2], [0, 2, 3], [1, 1, 2]]
W9 = [[0, 0, 1, 2], [0, 0, 2], [0, 0, 3, 2], [0, 1], [0, 1, 2], [0, 1, 2, 3], [0, 2], [0, 2, 2], [0, 3]]


def _q(words, query, d, kind):
    return [words[i] for i in oracle.trie_query(words, query, d, kind)]


def test_trie_hamming_vectors():  # test_trie.cc:101-126
    assert _q(W5, [0, 1, 3], 1, "hamming") == [[0, 1, 2], [0, 1, 3], [0, 2, 3]]
    assert _q(W5, [0, 1, 3], 1, "asymmetricHamming") == [[0, 1, 3], [0, 2, 3]]


def test_trie_levenshtein_vectors():  # test_trie.cc:155-191 (order AND multiplicity)
    assert _q(W9, [0, 1, 2], 1, "levenshtein") == [
        [0, 2], [0, 0, 2], [0, 0, 1, 2], [0, 1], [0, 1, 2], [0, 1, 2, 3], [0, 1, 2], [0, 2, 2],
        [0, 0, 1, 2]]
    assert _q(W9, [0, 1, 2], 1, "asymmetricLevenshtein") == [
        [0, 2], [0, 1], [0, 1, 2], [0, 1, 2, 3], [0, 1, 2], [0, 2, 2], [0, 0, 1, 2]]


def test_walk_order_and_counts():  # test_trie.cc:17-40, :79-99
    words = ["ACG", "ACT", "AGG", "AGT", "CCG"]
    keys = np.array([oracle.word_to_key(w) for w in reversed(words)] * 2, dtype=np.uint64)
    r = oracle.Result(keys, 3, 1, False, True)
    assert [oracle.key_to_word(k, 3) for k in r.keys] == words
    assert list(r.count) == [2] * 5


# ------------------------------------------------------------- tests/test_cluster.cc
def test_at_least_double():  # test_cluster.cc:17-21
    assert oracle.at_least_double(1, 0) and oracle.at_least_double(2, 1)
    assert not oracle.at_least_double(3, 2)


def test_max_neighbour_vectors():  # test_cluster.cc:23-71
    assert oracle.max_neighbour([0], [[]], [0], 0) == 0
    assert oracle.max_neighbour([1, 2], [[1], [0]], [0, 1], 0) == 0
    assert oracle.max_neighbour([1, 2], [[1], [0]], [0, 0], 0) == 1
    assert oracle.max_neighbour([1, 2, 4], [[1], [0, 2], [1]], [0, 0, 0], 0) == 2
    assert oracle.max_neighbour([1, 2, 4, 7], [[1], [0, 2], [1, 3], [2]], [0] * 4, 0) == 2


def test_directional_chain():  # test_cluster.cc:73-137
    count = [2, 4, 8, 10, 3]
    adj = [[1], [0, 2], [1, 3], [2, 4], [3]]
    cl, size, mx = oracle.cluster_graph(count, adj, False, [0, 3])
    assert list(cl) == [1, 1, 1, 2, 2]
    assert list(size) == [14, 13]
    assert list(mx) == [2, 3]


# ------------------------------------------------------------ SURVEY §4.4 probe vectors
def _run_words(words_counts, n, d, edit, maximum):
    keys = []
    for w, c in words_counts:
        keys += [oracle.word_to_key(w)] * c
    return oracle.Result(np.array(keys, dtype=np.uint64), n, d, edit, maximum)


def test_dominance_rule():  # AC/CA at Hamming 2 are never linked
    r = _run_words([("AA", 1), ("AC", 1), ("CA", 1), ("CC", 1)], 2, 2, False, True)
    nb = [list(r.neighbours(i)) for i in range(4)]
    assert nb == [[1, 2, 3], [0, 3], [0, 3], [0, 1, 2]]


def test_maxleaf_dfs_tiebreak():  # AA(1) AC(1) CA(5) CC(5), -n 2 -x keeps CC, not CA
    r = _run_words([("AA", 1), ("AC", 1), ("CA", 5), ("CC", 5)], 2, 1, False, True)
    assert r.clusters == 1 and list(r.cluster_max) == [3] and list(r.cluster_size) == [12]
    r = _run_words([("AA", 1), ("AC", 1), ("CA", 5), ("CC", 5)], 2, 1, False, False)
    assert r.clusters == 2
    assert sorted(r.cluster_size) == [6, 6]
    assert list(r.cluster_of) == [1, 2, 1, 2]  # {AA,CA}, {AC,CC}
    assert list(r.cluster_max) == [2, 3]


# --------------------------------------------------- against the unmodified reference
ref = pytest.mark.skipif(not oracle.have_ref(), reason="oracle/_ref not built")


def _random_words(rng, n, m, alphabet=4):
    base = rng.integers(0, alphabet, size=(max(2, m // 6), n))
    rows = base[rng.integers(0, base.shape[0], size=m)].copy()
    mut = rng.random(rows.shape) < 0.08
    rows[mut] = rng.integers(0, alphabet, size=int(mut.sum()))
    return ["".join("ACGT"[c] for c in r) for r in rows]


@ref
@pytest.mark.parametrize("edit,d", [(False, 1), (False, 2), (False, 3), (True, 1), (True, 2), (True, 3)])
@pytest.mark.parametrize("maximum", [True, False])
def test_against_reference_trie(edit, d, maximum):
    rng = np.random.default_rng(100 * d + 10 * edit + maximum)
    for n, m, alpha in [(6, 300, 4), (8, 600, 4), (10, 500, 2), (12, 800, 4)]:
        words = _random_words(rng, n, m, alpha)
        leaves, clusters, _ = oracle.run_ref_probe(words, [1] * len(words), edit, d, maximum)
        keys = np.array([oracle.word_to_key(w) for w in words], dtype=np.uint64)
        r = oracle.Result(keys, n, d, edit, maximum)
        assert r.unique == len(leaves)
        for lf in leaves:
            i = lf["rank"]
            assert oracle.key_to_word(r.keys[i], n) == lf["word"]
            assert int(r.count[i]) == lf["count"]
            assert list(r.neighbours(i)) == lf["neighbours"], (n, i)
            assert int(r.cluster_of[i]) == lf["cluster"]
        assert r.clusters == len(clusters)
        for c in clusters:
            assert int(r.cluster_size[c["id"] - 1]) == c["size"]
            assert int(r.cluster_max[c["id"] - 1]) == c["max_leaf"]


@ref
def test_reference_query_vectors():
    """The golden vectors above, re-derived from the reference code itself."""
    words = ["ACG", "ACT", "AGG", "AGT", "CCG"]
    _, _, q = oracle.run_ref_probe(words, [1] * 5, False, 1, True, [("h", 1, "ACT"), ("H", 1, "ACT")])
    assert q == [["ACG", "ACT", "AGT"], ["ACT", "AGT"]]


def _wide_keys(words):
    ks = [oracle.word_to_key(w) for w in words]
    lo = np.array([k & 0xFFFFFFFFFFFFFFFF for k in ks], dtype=np.uint64)
    hi = np.array([k >> 64 for k in ks], dtype=np.uint64)
    return lo, hi


@ref
@pytest.mark.parametrize("edit,d", [(False, 1), (False, 2), (True, 1), (True, 2)])
@pytest.mark.parametrize("maximum", [True, False])
def test_wide_words_against_reference_trie(edit, d, maximum):
    """Words of 32..63 letters (two 64-bit halves, `-n 36` of docs/usage.rst:18-20 and beyond)."""
    rng = np.random.default_rng(7000 + 100 * d + 10 * edit + maximum)
    for n, m, alpha in [(32, 300, 4), (36, 500, 4), (40, 400, 2), (63, 300, 4)]:
        if edit and d == 2 and n > 40:
            continue  # the reference's recursion takes minutes there
        words = _random_words(rng, n, m, alpha)
        leaves, clusters, _ = oracle.run_ref_probe(words, [1] * len(words), edit, d, maximum)
        r = oracle.Result(_wide_keys(words), n, d, edit, maximum)
        assert r.unique == len(leaves)
        for lf in leaves:
            i = lf["rank"]
            assert oracle.key_to_word((int(r.keys_hi[i]) << 64) | int(r.keys[i]), n) == lf["word"]
            assert int(r.count[i]) == lf["count"]
            assert list(r.neighbours(i)) == lf["neighbours"], (n, i)
            assert int(r.cluster_of[i]) == lf["cluster"]
        assert r.clusters == len(clusters)
        for c in clusters:
            assert int(r.cluster_size[c["id"] - 1]) == c["size"]
            assert int(r.cluster_max[c["id"] - 1]) == c["max_leaf"]


def test_wide_pack_words():
    """makeWord on 36 letters: halves split 4 | 32, filtered tuples are all-ones in both halves."""
    w = "ACGT" * 9
    chars = np.frombuffer((w + w[:35] + "N").encode(), dtype=np.uint8).reshape(2, 36)
    lo, hi = oracle.pack_words(chars, [12, 12, 12])
    k = oracle.word_to_key(w)
    assert int(lo[0]) == k & 0xFFFFFFFFFFFFFFFF and int(hi[0]) == k >> 64
    assert lo[1] == oracle.INVALID and hi[1] == oracle.INVALID
