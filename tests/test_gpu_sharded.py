"""The multi-GPU phases (include/humid_gpu_shard.h) on ONE GPU: W logical shards, each with its
own handle, run the orchestration of humid_b200/sharded.py in W threads with loop-back
collectives; every shard must end with exactly the oracle's whole-job result."""
import threading

import numpy as np
import pytest

import oracle
from humid_b200 import sharded
from humid_b200.synth import Synth, baseline_config

pytestmark = pytest.mark.gpu

CASES = [  # (shards, config index, tuples, overrides)
    (2, 1, 30000, {}),
    (3, 0, 30000, {}),
    (4, 4, 40000, {}),
    (2, 2, 20000, {}),                                        # Hamming 2
    (4, 2, 12000, dict(word_length=10, genome_len=3000)),     # dense graph, many cross-shard edges
    (3, 0, 20000, dict(word_length=8, genome_len=2000, maximum=False)),
    (2, 3, 8000, dict(distance=1)),                           # edit distance 1 == Hamming 1
    (5, 0, 3, {}),                                            # more shards than words
]


def run_sharded(W, chars, seg_take, cfg):
    R = chars.shape[0]
    cuts = [0] + [int(R * (r + 1) / W * (0.9 if r % 2 == 0 else 1.0)) for r in range(W - 1)] + [R]
    shared = sharded.ThreadComm.Shared(W)
    out, err = [None] * W, []

    def work(r):
        try:
            dd = sharded.ShardedDedup(comm=sharded.ThreadComm(shared, r), word_length=cfg.word_length,
                                      distance=cfg.distance, seg_take=seg_take, edit=cfg.edit, maximum=cfg.maximum,
                                      device=0)
            part = chars[cuts[r]:cuts[r + 1]]
            half = part.shape[0] // 2
            dd.add_batch(part[:half])
            dd.add_batch(part[half:])
            dd.finalize()
            # every result accessor is a collective: all shards call them in the same order
            dd.res = dict(counts=dd.counts(), words=dd.words(), neighbours=dd.neighbours(), clusters=dd.clusters(),
                          keep=dd.keep_mask(), ids=dd.cluster_ids(), stats=dd.stats_text())
            out[r] = dd
        except BaseException as e:  # noqa: BLE001
            err.append(e)
            shared.barrier.abort()

    th = [threading.Thread(target=work, args=(r,)) for r in range(W)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    if err:
        raise err[0]
    return out, cuts


@pytest.mark.parametrize("W,ci,R,kw", CASES, ids=[f"w{w}_c{c + 1}_{R}_{'_'.join(f'{k}{v}' for k, v in kw.items())}"
                                                   for w, c, R, kw in CASES])
def test_logical_shards_match_oracle(W, ci, R, kw):
    cfg = baseline_config(ci, R, **kw)
    s = Synth(cfg)
    chars = s.word_chars()
    hdr, take = s.layout()
    seg_take = ([hdr] if hdr else []) + take
    ref = oracle.Result(oracle.pack_words(chars, seg_take), cfg.word_length, cfg.distance, cfg.edit, cfg.maximum)
    dds, cuts = run_sharded(W, chars, seg_take, cfg)
    for r, dd in enumerate(dds):
        lo, hi = cuts[r], cuts[r + 1]
        res = dd.res
        c = res["counts"]
        assert (c["total"], c["usable"], c["unique"], c["clusters"]) == (ref.total, ref.usable, ref.unique, ref.clusters)
        keys, cnt, first, cid = res["words"]
        assert np.array_equal(keys, ref.keys)
        assert np.array_equal(cnt.astype(np.uint64), ref.count)
        assert np.array_equal(first.astype(np.uint64), ref.first)
        assert np.array_equal(cid, ref.cluster_of)
        row, col = res["neighbours"]
        assert np.array_equal(row, ref.row_ptr) and np.array_equal(col, ref.col)
        size, mx = res["clusters"]
        assert np.array_equal(size, ref.cluster_size) and np.array_equal(mx, ref.cluster_max)
        assert np.array_equal(res["keep"], ref.keep[lo:hi].astype(bool))
        assert np.array_equal(res["ids"], ref.cid[lo:hi])
        assert res["stats"] == ref.stats_text()
    if W > 1 and R > 100:
        assert sum(dd.words_exchanged for dd in dds) > 0
    for dd in dds:
        dd.close()


def test_sharded_rejects_edit2():
    shared = sharded.ThreadComm.Shared(1)
    dd = sharded.ShardedDedup(comm=sharded.ThreadComm(shared, 0), word_length=8, distance=2, seg_take=[4, 4], edit=True,
                              device=0)
    dd.add_batch(np.frombuffer(b"ACGTACGT", dtype=np.uint8).reshape(1, 8).copy())
    with pytest.raises(Exception):
        dd.finalize()
