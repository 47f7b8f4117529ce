"""BASELINE.json sizes (configs[1] — also at -n 36 — and configs[2] at 50 M tuples, configs[3] at 20 M): too
large for the oracle, so the result is checked through properties that hold at any size."""
import numpy as np
import pytest
import torch

from humid_b200 import api
from humid_b200.synth import HashSynth, baseline_config

pytestmark = pytest.mark.gpu

CASES = [(1, 50_000_000, {}), (2, 50_000_000, {}), (1, 50_000_000, dict(word_length=36)), (3, 20_000_000, {})]


def run(cfg, chars, batches, expected=0):
    hdr, take = HashSynth(cfg).layout()
    seg_take = ([hdr] if hdr else []) + take
    dd = api.Dedup(word_length=cfg.word_length, distance=cfg.distance, seg_take=seg_take, edit=cfg.edit,
                   maximum=cfg.maximum, expected_tuples=expected)
    n = chars.shape[0]
    step = (n + batches - 1) // batches
    for lo in range(0, n, step):
        m = min(step, n - lo)
        dd.add_batch_device(chars[lo:lo + m].data_ptr(), m)
    dd.finalize()
    return dd


def hist_dict(dd, which):
    k, v = dd.histogram(which)
    return k.astype(np.int64), v.astype(np.int64)


@pytest.mark.parametrize("ci,R,kw", CASES, ids=["c2", "c3", "c2_n36", "c4_edit2"])
def test_properties_at_full_size(ci, R, kw):
    cfg = baseline_config(ci, R, **kw)
    chars = HashSynth(cfg).word_chars(0, R, device=torch.device("cuda", 0))
    dd = run(cfg, chars, batches=1, expected=R)
    c = dd.counts()
    assert c["total"] == R and 0 < c["clusters"] <= c["unique"] <= c["usable"] <= R
    # counts.dat: sizes of the words
    k, v = hist_dict(dd, "counts")
    assert int(v.sum()) == c["unique"] and int((k * v).sum()) == c["usable"] and k.min() >= 1
    # neigh.dat: list lengths; every undirected pair is in two lists
    k, v = hist_dict(dd, "neigh")
    assert int(v.sum()) == c["unique"] and int((k * v).sum()) % 2 == 0
    # clusters.dat: the clusters partition the usable reads
    k, v = hist_dict(dd, "clusters")
    assert int(v.sum()) == c["clusters"] and int((k * v).sum()) == c["usable"] and k.min() >= 1  # size = reads in the cluster
    # one tuple is kept per cluster, and it is a usable one: its cluster id is set and distinct
    keep = dd.keep_mask()
    assert int(keep.sum()) == c["clusters"]
    ids = dd.cluster_ids()
    kept = ids[keep]
    assert kept.min() >= 1 and kept.max() == c["clusters"] and np.unique(kept).size == c["clusters"]
    assert int((ids == 0).sum()) == R - c["usable"]
    # ... and the first tuple of its word (nothing with the same word comes earlier): spot-check by
    # recounting words per cluster id == counts of a second, differently batched run
    dd2 = run(cfg, chars, batches=7)  # table sized at finalize instead of up front, uneven batches
    assert dd2.counts() == c
    assert np.array_equal(dd2.keep_mask(), keep)
    assert np.array_equal(dd2.cluster_ids(), ids)
    assert dd2.stats_text() == dd.stats_text()
    dd.close()
    dd2.close()


ALTERNATIVES = [  # (config, tuples, environment of the second run)
    (2, 50_000_000, dict(HUMID_GPU_JOIN_SCHEME="blocks")),                       # three blocks instead of block pairs
    (3, 20_000_000, dict(HUMID_GPU_EDIT_SHIFTS="probe", HUMID_GPU_EDIT_TASKS="0",  # variant probing + plain walk of the
                         HUMID_GPU_JOIN_SCHEME="blocks")),                          # recursion instead of joins + counting
]


@pytest.mark.parametrize("ci,R,env", ALTERNATIVES, ids=["c3_blocks_vs_pairs", "c4_probe_walk_vs_join_count"])
def test_independent_paths_agree_at_full_size(monkeypatch, ci, R, env):
    """Every search has two implementations (forced through the testing knobs); at BASELINE.json's sizes, where
    the oracle cannot follow, they must produce the same reads, ids and statistics files."""
    cfg = baseline_config(ci, R)
    chars = HashSynth(cfg).word_chars(0, R, device=torch.device("cuda", 0))
    dd = run(cfg, chars, batches=1, expected=R)
    ref = (dd.counts(), dd.keep_mask().copy(), dd.cluster_ids().copy(), dd.stats_text())
    dd.close()
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    dd = run(cfg, chars, batches=3)
    assert dd.counts() == ref[0]
    assert np.array_equal(dd.keep_mask(), ref[1])
    assert np.array_equal(dd.cluster_ids(), ref[2])
    assert dd.stats_text() == ref[3]
    dd.close()
