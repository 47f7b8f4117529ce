"""world_size-2 (and 3) runs of the multi-GPU orchestration (humid_b200/sharded.py) over gloo on
the CPU, with the numpy shard engine of tests/shard_cpu_engine.py standing in for the CUDA
library.  The result on every rank must equal the oracle's for the concatenated input."""
import os
import sys

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle
from humid_b200.synth import Synth, baseline_config

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, ci, R, kw, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from shard_cpu_engine import CpuShardEngine

    from humid_b200 import sharded

    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        cfg = baseline_config(ci, R, **kw)
        s = Synth(cfg)
        chars = s.word_chars()
        hdr, take = s.layout()
        seg_take = ([hdr] if hdr else []) + take
        # uneven slices, rank order == input order
        cuts = [0] + [int(R * (r + 1) / world * (0.9 if r % 2 == 0 else 1.0)) for r in range(world - 1)] + [R]
        lo, hi = cuts[rank], cuts[rank + 1]
        eng = CpuShardEngine(cfg.word_length, cfg.distance, seg_take, cfg.edit, cfg.maximum)
        dd = sharded.ShardedDedup(engine=eng)
        dd.add_batch(chars[lo:hi])
        dd.finalize()
        ref = oracle.Result(oracle.pack_words(chars, seg_take), cfg.word_length, cfg.distance, cfg.edit, cfg.maximum)
        c = dd.counts()
        assert (c["total"], c["usable"], c["unique"], c["clusters"]) == (ref.total, ref.usable, ref.unique, ref.clusters)
        keys, cnt, first, cid = dd.words()
        assert np.array_equal(keys, ref.keys) and np.array_equal(cnt.astype(np.uint64), ref.count)
        assert np.array_equal(first.astype(np.uint64), ref.first) and np.array_equal(cid, ref.cluster_of)
        row, col = dd.neighbours()
        assert np.array_equal(row, ref.row_ptr) and np.array_equal(col, ref.col)
        assert np.array_equal(dd.keep_mask(), ref.keep[lo:hi].astype(bool))
        assert np.array_equal(dd.cluster_ids(), ref.cid[lo:hi])
        assert dd.stats_text() == ref.stats_text()
        q.put((rank, "ok", dd.words_exchanged))
    except Exception as e:  # noqa: BLE001
        import traceback

        q.put((rank, "fail", traceback.format_exc()))
        raise e
    finally:
        dist.destroy_process_group()


CASES = [(2, 0, 6000, {}), (2, 1, 6000, {}), (3, 2, 3000, dict(word_length=10, genome_len=3000)),
         (2, 0, 4000, dict(word_length=8, genome_len=2000, maximum=False))]


@pytest.mark.parametrize("world,ci,R,kw", CASES)
def test_sharded_orchestration_gloo(world, ci, R, kw):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() + ci * 7 + world) % 2000
    procs = [ctx.Process(target=_worker, args=(r, world, port, ci, R, kw, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=300)
    res = [q.get(timeout=5) for _ in range(world)]
    for r in res:
        assert r[1] == "ok", r[2]
    assert all(p.exitcode == 0 for p in procs)
    assert sum(r[2] for r in res) > 0  # some variant queries really crossed shards
