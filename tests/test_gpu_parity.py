"""Parity of the CUDA path (through the C ABI) with the reference.

  * golden fixtures = outputs of the unmodified reference binary (tests/golden/*.npz);
  * the CPU oracle on seeded synthetic inputs, including what the binary never prints
    (rank-ordered words, neighbour lists with order and multiplicity, cluster sizes and
    maxLeaf);
  * the reference's own edge cases: empty input, all tuples filtered, short segments,
    count ties in maximum mode, the dominance rule at distance 2.
Everything here is integer work: the bar is bit-exact.
"""
import glob
import os

import numpy as np
import pytest

import oracle
from humid_b200 import api
from humid_b200.synth import Synth, baseline_config

pytestmark = pytest.mark.gpu

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))


def load_golden(path):
    z = np.load(path)
    g = {k: z[k] for k in z.files}
    for k in ("n", "d", "edit", "maximum"):
        g[k] = int(g[k])
    for k in ("counts", "neigh", "clusters", "stats"):
        g[k] = bytes(g[k]).decode()
    return g


def run_gpu(chars, seg_take, n, d, edit, maximum, seg_len=None, batches=1):
    dd = api.Dedup(word_length=n, distance=d, seg_take=seg_take, edit=edit, maximum=maximum)
    R = chars.shape[0]
    step = max(1, (R + batches - 1) // batches)
    for lo in range(0, R, step):
        dd.add_batch(chars[lo:lo + step], None if seg_len is None else seg_len[lo:lo + step])
    dd.finalize()
    return dd


def assert_same_as_oracle(dd, r, R):
    c = dd.counts()
    assert (c["total"], c["usable"], c["unique"], c["clusters"]) == (r.total, r.usable, r.unique, r.clusters)
    keys, cnt, first, cid = dd.words()
    assert np.array_equal(keys, r.keys)
    assert np.array_equal(dd.words_hi(), r.keys_hi)
    assert np.array_equal(cnt.astype(np.uint64), r.count)
    assert np.array_equal(first.astype(np.uint64), r.first)
    row_ptr, col = dd.neighbours()
    assert np.array_equal(row_ptr, r.row_ptr)
    assert np.array_equal(col, r.col)
    assert np.array_equal(cid, r.cluster_of)
    size, mx = dd.clusters()
    assert np.array_equal(size, r.cluster_size)
    assert np.array_equal(mx, r.cluster_max)
    assert np.array_equal(dd.keep_mask(), r.keep.astype(bool))
    assert np.array_equal(dd.cluster_ids(), r.cid)
    assert dd.stats_text() == r.stats_text()
    assert int(dd.keep_mask().sum()) == r.clusters


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_golden_reference_outputs(path):
    g = load_golden(path)
    dd = run_gpu(g["chars"], g["seg_take"], g["n"], g["d"], bool(g["edit"]), bool(g["maximum"]), batches=3)
    txt = dd.stats_text()
    assert txt["stats.dat"] == g["stats"]
    assert txt["counts.dat"] == g["counts"]
    assert txt["neigh.dat"] == g["neigh"]
    assert txt["clusters.dat"] == g["clusters"]
    assert np.array_equal(dd.keep_mask(), g["keep"].astype(bool))
    assert np.array_equal(dd.cluster_ids(), g["cid"])


CASES = [  # (config index, tuples, overrides)
    (0, 40000, {}),
    (0, 40000, dict(maximum=False)),
    (1, 30000, {}),
    (2, 30000, {}),
    (2, 30000, dict(maximum=True)),
    (2, 4000, dict(distance=3)),
    (4, 50000, {}),
    (0, 20000, dict(word_length=8, genome_len=2000)),
    (0, 20000, dict(word_length=8, genome_len=2000, maximum=False)),
    (2, 20000, dict(word_length=10, genome_len=3000, maximum=True)),
    (2, 20000, dict(word_length=10, genome_len=3000)),
    (3, 8000, dict(distance=1)),
    (3, 8000, {}),
    (3, 8000, dict(maximum=True)),
    (3, 4000, dict(word_length=10, genome_len=3000)),
    (3, 1500, dict(distance=3, word_length=12, genome_len=3000)),
]


@pytest.mark.parametrize("ci,R,kw", CASES, ids=[f"c{c + 1}_{R}_{'_'.join(f'{k}{v}' for k, v in kw.items())}" for c, R, kw in CASES])
def test_against_oracle(ci, R, kw):
    cfg = baseline_config(ci, R, **kw)
    s = Synth(cfg)
    chars = s.word_chars()
    hdr, take = s.layout()
    seg_take = ([hdr] if hdr else []) + take
    keys = oracle.pack_words(chars, seg_take)
    r = oracle.Result(keys, cfg.word_length, cfg.distance, cfg.edit, cfg.maximum)
    dd = run_gpu(chars, seg_take, cfg.word_length, cfg.distance, cfg.edit, cfg.maximum, batches=2)
    assert_same_as_oracle(dd, r, R)


WIDE_CASES = [  # words of 32..63 letters: two key halves (config 2 at `-n 36`)
    (1, 30000, dict(word_length=36)),
    (1, 30000, dict(word_length=36, maximum=True)),
    (1, 20000, dict(word_length=36, distance=2)),
    (1, 20000, dict(word_length=32)),
    (1, 20000, dict(word_length=33, distance=2, maximum=True)),
    (2, 20000, dict(word_length=48, distance=3)),
    (2, 20000, dict(word_length=63, distance=2)),
    (0, 20000, dict(word_length=40, maximum=False)),
    (3, 8000, dict(word_length=36, distance=1)),  # edit distance 1 == Hamming 1
    (2, 20000, dict(word_length=34, genome_len=3000, distance=2)),  # few distinct fragments: long groups
]


@pytest.mark.parametrize("ci,R,kw", WIDE_CASES,
                         ids=[f"c{c + 1}_{R}_{'_'.join(f'{k}{v}' for k, v in kw.items())}" for c, R, kw in WIDE_CASES])
def test_wide_words_against_oracle(ci, R, kw):
    cfg = baseline_config(ci, R, **kw)
    s = Synth(cfg)
    chars = s.word_chars()
    hdr, take = s.layout()
    seg_take = ([hdr] if hdr else []) + take
    keys = oracle.pack_words(chars, seg_take)
    assert isinstance(keys, tuple)
    r = oracle.Result(keys, cfg.word_length, cfg.distance, cfg.edit, cfg.maximum)
    dd = run_gpu(chars, seg_take, cfg.word_length, cfg.distance, cfg.edit, cfg.maximum, batches=3)
    assert_same_as_oracle(dd, r, R)


def test_wide_words_edge_cases():
    """Empty input, everything filtered, short segments, and a handle reused after reset."""
    n, take = 36, [12, 12, 12]
    dd = api.Dedup(word_length=n, distance=1, seg_take=take)
    dd.finalize()
    assert dd.counts() == dict(total=0, usable=0, unique=0, clusters=0)
    dd.reset()
    chars = np.full((100, n), ord("N"), dtype=np.uint8)
    dd.add_batch(chars)
    dd.finalize()
    assert dd.counts() == dict(total=100, usable=0, unique=0, clusters=0)
    assert not dd.keep_mask().any() and not dd.cluster_ids().any()
    dd.reset()
    rng = np.random.default_rng(5)
    chars = np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, size=(5000, n))]
    chars[rng.integers(0, 5000, 60), rng.integers(0, n, 60)] = ord("N")
    chars[1000:2000] = chars[:1000]  # duplicates
    seg_len = np.tile(np.array(take, dtype=np.uint16), (5000, 1))
    seg_len[rng.integers(0, 5000, 50), rng.integers(0, 3, 50)] = 11  # short segment: padded with N => filtered
    dd.add_batch(chars[:3000], seg_len[:3000])
    dd.add_batch(chars[3000:], seg_len[3000:])
    dd.finalize()
    r = oracle.Result(oracle.pack_words(chars, take, seg_len), n, 1, False, False)
    assert_same_as_oracle(dd, r, 5000)
    with pytest.raises(api.HumidGpuError):
        api.Dedup(word_length=36, distance=2, seg_take=take, edit=True)
    with pytest.raises(api.HumidGpuError):
        api.Dedup(word_length=64, distance=1, seg_take=[32, 32])


def test_wide_words_million():
    """1 M tuples at -n 36 (configs[1] shape): the two-sort word construction over many tiles."""
    cfg = baseline_config(1, 1_000_000, word_length=36)
    s = Synth(cfg)
    chars = s.word_chars()
    hdr, take = s.layout()
    seg_take = ([hdr] if hdr else []) + take
    r = oracle.Result(oracle.pack_words(chars, seg_take), cfg.word_length, cfg.distance, cfg.edit, cfg.maximum)
    dd = run_gpu(chars, seg_take, cfg.word_length, cfg.distance, cfg.edit, cfg.maximum, batches=3)
    assert_same_as_oracle(dd, r, 1_000_000)


def test_against_oracle_millions():
    """Enough words that the sorts, scans and tables span more tiles than are resident at once."""
    cfg = baseline_config(1, 4_000_000)
    s = Synth(cfg)
    chars = s.word_chars()
    hdr, take = s.layout()
    seg_take = ([hdr] if hdr else []) + take
    r = oracle.Result(oracle.pack_words(chars, seg_take), cfg.word_length, cfg.distance, cfg.edit, cfg.maximum)
    dd = run_gpu(chars, seg_take, cfg.word_length, cfg.distance, cfg.edit, cfg.maximum, batches=3)
    assert_same_as_oracle(dd, r, chars.shape[0])


@pytest.mark.parametrize("hint", [30000, 100000, 17000, 1])
def test_words_counted_while_batches_arrive(hint):
    """expected_tuples sizes the table up front and add_batch counts into it; a job that outgrows
    the hint (17000, 1) is recounted at finalize.  Same result either way, also after a reset."""
    cfg = baseline_config(1, 30000)
    s = Synth(cfg)
    chars = s.word_chars()
    hdr, take = s.layout()
    seg_take = ([hdr] if hdr else []) + take
    r = oracle.Result(oracle.pack_words(chars, seg_take), cfg.word_length, cfg.distance, cfg.edit, cfg.maximum)
    dd = api.Dedup(word_length=cfg.word_length, distance=cfg.distance, seg_take=seg_take, edit=cfg.edit,
                   maximum=cfg.maximum, expected_tuples=hint)
    for rounds in range(2):
        for lo in range(0, 30000, 7000):
            dd.add_batch(chars[lo:lo + 7000])
        dd.finalize()
        assert_same_as_oracle(dd, r, 30000)
        dd.reset()


JOIN_CASES = [  # (config index, tuples, overrides) — dense enough that small caps split the work both ways
    (0, 20000, dict(word_length=8, genome_len=2000)),
    (2, 20000, dict(word_length=10, genome_len=3000)),
    (2, 20000, dict(word_length=10, genome_len=3000, maximum=True)),
    (2, 3000, dict(distance=3, word_length=9, genome_len=1500)),
    (1, 30000, {}),
]


@pytest.mark.parametrize("kernel", ["walk", "tile"])
@pytest.mark.parametrize("group_cap", [1, 3, 16, 100000])
@pytest.mark.parametrize("ci,R,kw", JOIN_CASES, ids=[f"c{c + 1}_{R}_{'_'.join(f'{k}{v}' for k, v in kw.items())}" for c, R, kw in JOIN_CASES])
def test_oversized_groups_take_the_probe_path(monkeypatch, ci, R, kw, group_cap, kernel):
    """The block join hands groups above the cap to variant probing; with tiny caps most or all pairs
    come from that path, with middling ones the two paths share the pairs.  Same lists either way."""
    monkeypatch.setenv("HUMID_GPU_JOIN_CAP", str(group_cap))
    monkeypatch.setenv("HUMID_GPU_JOIN_KERNEL", kernel)  # both group-walk kernels, whatever the group sizes
    cfg = baseline_config(ci, R, **kw)
    s = Synth(cfg)
    chars = s.word_chars()
    hdr, take = s.layout()
    seg_take = ([hdr] if hdr else []) + take
    r = oracle.Result(oracle.pack_words(chars, seg_take), cfg.word_length, cfg.distance, cfg.edit, cfg.maximum)
    dd = run_gpu(chars, seg_take, cfg.word_length, cfg.distance, cfg.edit, cfg.maximum, batches=2)
    assert_same_as_oracle(dd, r, R)
    names = {st["name"] for st in dd.stage_stats()}
    assert "join_hamming_b0" in names
    if group_cap == 1:
        assert "probe_flagged" in names


SCHEME_CASES = [
    (2, 30000, {}),
    (2, 30000, dict(maximum=True)),
    (2, 20000, dict(word_length=10, genome_len=3000)),           # long groups
    (2, 20000, dict(word_length=9, genome_len=3000, maximum=True)),  # blocks of 2-3 letters
    (1, 30000, dict(distance=2)),
    (0, 20000, dict(distance=2, word_length=31)),
]


@pytest.mark.parametrize("group_cap", [2, 100000])
@pytest.mark.parametrize("scheme", ["blocks", "combos"])
@pytest.mark.parametrize("ci,R,kw", SCHEME_CASES, ids=[f"c{c + 1}_{R}_{'_'.join(f'{k}{v}' for k, v in kw.items())}" for c, R, kw in SCHEME_CASES])
def test_distance_two_by_blocks_and_by_block_pairs(monkeypatch, ci, R, kw, scheme, group_cap):
    """Hamming distance 2: three blocks (agree on one) or four blocks taken in pairs (agree on two) — the
    same pairs either way, also when oversized groups go through variant probing."""
    monkeypatch.setenv("HUMID_GPU_JOIN_SCHEME", scheme)
    monkeypatch.setenv("HUMID_GPU_JOIN_CAP", str(group_cap))
    cfg = baseline_config(ci, R, **kw)
    s = Synth(cfg)
    chars = s.word_chars()
    hdr, take = s.layout()
    seg_take = ([hdr] if hdr else []) + take
    r = oracle.Result(oracle.pack_words(chars, seg_take), cfg.word_length, cfg.distance, cfg.edit, cfg.maximum)
    dd = run_gpu(chars, seg_take, cfg.word_length, cfg.distance, cfg.edit, cfg.maximum, batches=2)
    assert_same_as_oracle(dd, r, R)
    names = {st["name"] for st in dd.stage_stats()}
    assert ("join_hamming_b5" in names) == (scheme == "combos")


EDIT_REPLAY_CASES = [
    (3, 8000, {}),
    (3, 8000, dict(maximum=True)),
    (3, 4000, dict(word_length=10, genome_len=3000)),  # many candidates per query
    (3, 1500, dict(distance=3, word_length=12, genome_len=3000)),
]


@pytest.mark.parametrize("shifts", ["probe", "join"])
@pytest.mark.parametrize("ci,R,kw", EDIT_REPLAY_CASES,
                         ids=[f"c{c + 1}_{R}_{'_'.join(f'{k}{v}' for k, v in kw.items())}" for c, R, kw in EDIT_REPLAY_CASES])
def test_edit_shift_candidates_by_probing_and_by_deletion_join(monkeypatch, ci, R, kw, shifts):
    """One deletion + one insertion: variants probed through the word table, or (distance 2) words joined on
    their deletion variants.  Same candidate sets, same lists."""
    monkeypatch.setenv("HUMID_GPU_EDIT_SHIFTS", shifts)
    cfg = baseline_config(ci, R, **kw)
    s = Synth(cfg)
    chars = s.word_chars()
    hdr, take = s.layout()
    seg_take = ([hdr] if hdr else []) + take
    r = oracle.Result(oracle.pack_words(chars, seg_take), cfg.word_length, cfg.distance, cfg.edit, cfg.maximum)
    dd = run_gpu(chars, seg_take, cfg.word_length, cfg.distance, cfg.edit, cfg.maximum, batches=2)
    assert_same_as_oracle(dd, r, R)
    names = {st["name"] for st in dd.stage_stats()}
    assert ("edit_deletion_join" in names) == (shifts == "join" and cfg.distance == 2)
    assert ("edit_candidates" in names) == (shifts == "probe" or cfg.distance != 2)


@pytest.mark.parametrize("heavy_from", [3, 16])
@pytest.mark.parametrize("task_cap", [0, 2, 7, 96])
@pytest.mark.parametrize("ci,R,kw", EDIT_REPLAY_CASES,
                         ids=[f"c{c + 1}_{R}_{'_'.join(f'{k}{v}' for k, v in kw.items())}" for c, R, kw in EDIT_REPLAY_CASES])
def test_edit_replay_by_counting_and_by_walking(monkeypatch, ci, R, kw, task_cap, heavy_from):
    """Queries with several candidates: subtrees holding one word are counted (task list), queries whose
    list overflows or that have `heavy_from` candidates or more are walked.  0 walks everything, tiny lists
    mix the kernels.  Same lists, same order."""
    monkeypatch.setenv("HUMID_GPU_EDIT_TASKS", str(task_cap))
    monkeypatch.setenv("HUMID_GPU_EDIT_HEAVY", str(heavy_from))
    cfg = baseline_config(ci, R, **kw)
    s = Synth(cfg)
    chars = s.word_chars()
    hdr, take = s.layout()
    seg_take = ([hdr] if hdr else []) + take
    r = oracle.Result(oracle.pack_words(chars, seg_take), cfg.word_length, cfg.distance, cfg.edit, cfg.maximum)
    dd = run_gpu(chars, seg_take, cfg.word_length, cfg.distance, cfg.edit, cfg.maximum, batches=2)
    assert_same_as_oracle(dd, r, R)
    names = {st["name"] for st in dd.stage_stats()}
    assert ("edit_replay_count" in names) == (task_cap > 0)
    if task_cap == 0 or (task_cap == 2 and heavy_from == 16):
        assert "edit_replay" in names


def _chars(words):
    return np.frombuffer("".join(words).encode(), dtype=np.uint8).reshape(len(words), -1).copy()


def test_histogram_values_beyond_the_shared_memory_bins():
    """Words counted 4096 times and more, and a cluster of > 9000 reads, take the listed-value path of the histograms."""
    rng = np.random.default_rng(5)
    words = ["ACGTACGTAC"] * 5000 + ["ACGTACGTAG"] * 4500 + ["TTTTTTTTTT"] * 4096 + ["GGGGGGGGGG"] * 4095
    words += ["".join(rng.choice(list("ACGT"), 10)) for _ in range(3000)]
    order = rng.permutation(len(words))
    chars = _chars([words[i] for i in order])
    r = oracle.Result(oracle.pack_words(chars, [5, 5]), 10, 1, False, False)
    dd = run_gpu(chars, [5, 5], 10, 1, False, False, batches=2)
    assert_same_as_oracle(dd, r, chars.shape[0])
    assert max(int(x.split()[0]) for x in dd.stats_text()["counts.dat"].splitlines()) == 5000


def test_empty_input():
    dd = api.Dedup(word_length=8, distance=1, seg_take=[4, 4])
    dd.finalize()
    assert dd.counts() == dict(total=0, usable=0, unique=0, clusters=0)
    assert dd.stats_text() == {"counts.dat": "", "neigh.dat": "", "clusters.dat": "",
                               "stats.dat": "total: 0\nusable: 0\nunique: 0\nclusters: 0\n"}


def test_all_filtered():
    dd = run_gpu(_chars(["ACGN", "NNNN", "acgt"]), [2, 2], 4, 1, False, True)
    assert dd.counts() == dict(total=3, usable=0, unique=0, clusters=0)
    assert not dd.keep_mask().any()
    assert list(dd.cluster_ids()) == [0, 0, 0]


def test_short_segments_filter():  # makeStringSize_ pads with 'N' (src/fastq.cc:57-66)
    chars = _chars(["ACGTACGT", "ACGTACGT", "ACGTACGT"])
    seg_len = np.array([[4, 4], [4, 3], [2, 4]], dtype=np.uint16)
    keys = oracle.pack_words(chars, [4, 4], seg_len)
    r = oracle.Result(keys, 8, 1, False, False)
    dd = run_gpu(chars, [4, 4], 8, 1, False, False, seg_len=seg_len)
    assert_same_as_oracle(dd, r, 3)
    assert dd.counts()["usable"] == 1


def test_dominance_rule_distance2():  # AC / CA are never linked
    chars = _chars(["AA", "AC", "CA", "CC"])
    dd = run_gpu(chars, [2], 2, 2, False, True)
    row_ptr, col = dd.neighbours()
    nb = [list(col[int(row_ptr[i]):int(row_ptr[i + 1])]) for i in range(4)]
    assert nb == [[1, 2, 3], [0, 3], [0, 3], [0, 1, 2]]


def test_maxleaf_dfs_tiebreak():  # AA(1) AC(1) CA(5) CC(5) keeps the first CC read
    words = ["AA", "AC"] + ["CA"] * 5 + ["CC"] * 5
    dd = run_gpu(_chars(words), [2], 2, 1, False, True)
    assert dd.counts()["clusters"] == 1
    assert list(np.nonzero(dd.keep_mask())[0]) == [7]
    dd = run_gpu(_chars(words), [2], 2, 1, False, False)
    assert dd.counts()["clusters"] == 2
    assert list(np.nonzero(dd.keep_mask())[0]) == [2, 7]
    assert dd.stats_text()["clusters.dat"] == "6 2\n"


def test_reset_reuses_handle():
    s = Synth(baseline_config(0, 5000))
    chars = s.word_chars()
    dd = api.Dedup(word_length=24, distance=1, seg_take=[8, 8, 8], maximum=True)
    dd.add_batch(chars)
    dd.finalize()
    a = (dd.counts(), dd.keep_mask().copy(), dd.stats_text())
    dd.reset()
    dd.add_batch(chars[:2500])
    dd.add_batch(chars[2500:])
    dd.finalize()
    assert a[0] == dd.counts() and np.array_equal(a[1], dd.keep_mask()) and a[2] == dd.stats_text()


def test_call_order_errors():
    dd = api.Dedup(word_length=4, distance=1, seg_take=[4])
    with pytest.raises(api.HumidGpuError):
        dd.counts()
    dd.add_batch(_chars(["ACGT"]))
    dd.finalize()
    with pytest.raises(api.HumidGpuError):
        dd.add_batch(_chars(["ACGT"]))
    with pytest.raises(api.HumidGpuError):
        api.Dedup(word_length=64, distance=1, seg_take=[64])  # longer than two 64-bit keys
