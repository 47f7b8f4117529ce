"""The oracle against the committed golden fixtures (outputs of the unmodified reference
binary, see tests/golden/make_golden.py).  Runs without /root/reference."""
import glob
import os

import numpy as np
import pytest

import oracle

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))


def load_golden(path):
    z = np.load(path)
    g = {k: z[k] for k in z.files}
    for k in ("n", "d", "edit", "maximum"):
        g[k] = int(g[k])
    for k in ("counts", "neigh", "clusters", "stats"):
        g[k] = bytes(g[k]).decode()
    return g


def test_fixtures_present():
    assert len(GOLDEN) >= 14


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_oracle_matches_reference_outputs(path):
    g = load_golden(path)
    keys = oracle.pack_words(g["chars"], g["seg_take"])
    r = oracle.Result(keys, g["n"], g["d"], bool(g["edit"]), bool(g["maximum"]))
    txt = r.stats_text()
    assert txt["stats.dat"] == g["stats"]
    assert txt["counts.dat"] == g["counts"]
    assert txt["neigh.dat"] == g["neigh"]
    assert txt["clusters.dat"] == g["clusters"]
    assert np.array_equal(r.keep, g["keep"])
    assert np.array_equal(r.cid, g["cid"])
