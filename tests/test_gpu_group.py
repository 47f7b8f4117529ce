"""The library-run multi-GPU path (include/humid_gpu_shard.h "Groups", humid_b200/csrc/group.cu) on ONE GPU: W ranks
= W threads of this process, each with its own handle on device 0, exchanging through each other's device buffers.
Every rank must end with exactly the oracle's whole-job counts / statistics and its own slice of the keep mask and
cluster ids; twice through the same handles (reset keeps the group's buffers)."""
import threading

import numpy as np
import pytest

import oracle
from humid_b200 import sharded
from humid_b200.synth import Synth, baseline_config

pytestmark = pytest.mark.gpu

CASES = [  # (ranks, config index, tuples, overrides)
    (2, 1, 30000, {}),
    (3, 0, 30000, {}),
    (4, 4, 40000, {}),
    (2, 2, 20000, {}),                                        # Hamming 2
    (4, 2, 12000, dict(word_length=10, genome_len=3000)),     # dense graph, many cross-shard edges
    (3, 0, 20000, dict(word_length=8, genome_len=2000, maximum=False)),
    (2, 3, 8000, dict(distance=1)),                           # edit distance 1 == Hamming 1
    (5, 0, 3, {}),                                            # more ranks than words
    (8, 1, 60000, {}),
]


def run_group(W, chars, seg_take, cfg, passes=2):
    R = chars.shape[0]
    cuts = [0] + [int(R * (r + 1) / W * (0.9 if r % 2 == 0 else 1.0)) for r in range(W - 1)] + [R]
    grp = sharded.new_group(W)
    out, err = [None] * W, []

    def work(r):
        try:
            dd = sharded.GroupDedup(r, W, grp, None, device=0, word_length=cfg.word_length, distance=cfg.distance,
                                    seg_take=seg_take, edit=cfg.edit, maximum=cfg.maximum)
            part = chars[cuts[r]:cuts[r + 1]]
            res = []
            for _ in range(passes):
                dd.reset()
                half = part.shape[0] // 2
                dd.add_batch(part[:half])
                dd.add_batch(part[half:])
                dd.finalize()
                res.append(dict(counts=dd.counts(), keep=dd.keep_mask(), ids=dd.cluster_ids(), stats=dd.stats_text()))
            out[r] = (dd, res)
        except BaseException as e:  # noqa: BLE001
            err.append(e)

    th = [threading.Thread(target=work, args=(r,)) for r in range(W)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    for o in out:
        if o is not None:
            o[0].close()
    sharded.free_group(grp)
    if err:
        raise err[0]
    return [o[1] for o in out], cuts


@pytest.mark.parametrize("W,ci,R,kw", CASES, ids=[f"w{w}_c{c + 1}_{R}_{'_'.join(f'{k}{v}' for k, v in kw.items())}"
                                                   for w, c, R, kw in CASES])
def test_group_ranks_match_oracle(W, ci, R, kw):
    cfg = baseline_config(ci, R, **kw)
    s = Synth(cfg)
    chars = s.word_chars()
    hdr, take = s.layout()
    seg_take = ([hdr] if hdr else []) + take
    ref = oracle.Result(oracle.pack_words(chars, seg_take), cfg.word_length, cfg.distance, cfg.edit, cfg.maximum)
    results, cuts = run_group(W, chars, seg_take, cfg)
    for r, passes in enumerate(results):
        lo, hi = cuts[r], cuts[r + 1]
        for res in passes:
            c = res["counts"]
            assert (c["total"], c["usable"], c["unique"], c["clusters"]) == (ref.total, ref.usable, ref.unique, ref.clusters)
            assert np.array_equal(res["keep"], ref.keep[lo:hi].astype(bool))
            assert np.array_equal(res["ids"], ref.cid[lo:hi])
            assert res["stats"] == ref.stats_text()


def test_group_rejects_what_is_not_sharded():
    grp = sharded.new_group(1)
    try:
        with pytest.raises(Exception):
            sharded.GroupDedup(0, 1, grp, None, device=0, word_length=8, distance=2, seg_take=[4, 4], edit=True)
        with pytest.raises(Exception):
            sharded.GroupDedup(0, 1, grp, None, device=0, word_length=36, distance=1, seg_take=[18, 18])
    finally:
        sharded.free_group(grp)
