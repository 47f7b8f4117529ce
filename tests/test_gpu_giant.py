"""Components of more than 32 words go through the data-parallel clustering of humid_b200/csrc/cluster_big.cu
(the reference's own failure mode is recursion depth on such components, docs/troubleshooting.rst:6-18):
  * every fixture-sized case again with the limit lowered to 1, so that EVERY component takes that path;
  * one component of 10^5..10^6 words (dense low-complexity 12-letter words), both clustering modes, against the
    oracle, with the time the clustering stages took."""
import numpy as np
import pytest

import oracle
from humid_b200 import api
from humid_b200.synth import Synth, baseline_config
from test_gpu_parity import assert_same_as_oracle, run_gpu

pytestmark = pytest.mark.gpu

CASES = [
    (0, 40000, {}), (0, 40000, dict(maximum=False)), (1, 30000, {}), (2, 30000, {}), (2, 30000, dict(maximum=True)),
    (0, 20000, dict(word_length=8, genome_len=2000)), (0, 20000, dict(word_length=8, genome_len=2000, maximum=False)),
    (2, 20000, dict(word_length=10, genome_len=3000, maximum=True)), (2, 20000, dict(word_length=10, genome_len=3000)),
    (3, 8000, {}), (3, 8000, dict(maximum=True)), (3, 4000, dict(word_length=10, genome_len=3000)),
    (1, 30000, dict(word_length=36)), (1, 30000, dict(word_length=36, maximum=True)),
]


@pytest.mark.parametrize("limit", [1, 4])
@pytest.mark.parametrize("ci,R,kw", CASES, ids=[f"c{c + 1}_{R}_{'_'.join(f'{k}{v}' for k, v in kw.items())}" for c, R, kw in CASES])
def test_every_component_through_the_parallel_path(monkeypatch, ci, R, kw, limit):
    monkeypatch.setenv("HUMID_GPU_SMALL_COMP", str(limit))
    cfg = baseline_config(ci, R, **kw)
    s = Synth(cfg)
    chars = s.word_chars()
    hdr, take = s.layout()
    seg_take = ([hdr] if hdr else []) + take
    r = oracle.Result(oracle.pack_words(chars, seg_take), cfg.word_length, cfg.distance, cfg.edit, cfg.maximum)
    dd = run_gpu(chars, seg_take, cfg.word_length, cfg.distance, cfg.edit, cfg.maximum, batches=2)
    assert_same_as_oracle(dd, r, R)
    names = {st["name"] for st in dd.stage_stats()}
    assert ("cluster_big_maximum" if cfg.maximum else "cluster_big_directional") in names


def low_complexity_words(R, n, seed, p_g=0.4):
    """R words of n letters, G-rich, with a heavy head of repeated words (counts from 1 to thousands)."""
    rng = np.random.default_rng(seed)
    p = np.array([(1 - p_g) / 3] * 2 + [p_g] + [(1 - p_g) / 3])  # A C G T
    distinct = rng.choice(4, size=(R // 2, n), p=p).astype(np.uint8)
    pick = np.minimum((rng.pareto(1.2, size=R) * 50).astype(np.int64), R // 2 - 1)  # Zipf-like reuse of early words
    fresh = rng.random(R) < 0.5
    idx = np.where(fresh, rng.integers(0, R // 2, size=R), pick)
    return np.frombuffer(b"ACGT", dtype=np.uint8)[distinct[idx]]


@pytest.mark.parametrize("maximum", [False, True])
def test_giant_component(maximum):
    n, R = 12, 3_000_000
    chars = low_complexity_words(R, n, 7)
    r = oracle.Result(oracle.pack_words(chars, [6, 6]), n, 1, False, maximum)
    # the point of the case: one component holding a large share of all words
    comp_sizes = np.bincount(r.cluster_of) if maximum else None
    with api.Dedup(word_length=n, distance=1, seg_take=[6, 6], maximum=maximum) as dd:
        for rep in range(2):  # second pass: allocations in place, times are the steady state
            dd.reset()
            dd.add_batch(chars)
            dd.finalize()
        assert_same_as_oracle(dd, r, R)
        st = {}
        for x in dd.stage_stats():
            st[x["name"]] = st.get(x["name"], 0.0) + x["ms"]
    cluster_ms = sum(v for k, v in st.items() if k in ("union_find", "classify_components", "group_big_components",
                                                        "replay_directional", "replay_maximum", "cluster_big_directional",
                                                        "cluster_big_maximum"))
    big = max(comp_sizes[1:]) if maximum else None
    print(f"\ngiant component, maximum={maximum}: U={r.unique} clusters={r.clusters} largest cluster (words)={big} "
          f"clustering stages {cluster_ms:.2f} ms: " + ", ".join(f"{k}={v:.2f}" for k, v in st.items()))
    if maximum:
        assert big >= 100_000, "the generator no longer yields a giant component"
    assert cluster_ms < 50.0
