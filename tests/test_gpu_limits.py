"""The library's hard limits, at their edges, through the C ABI: every one of them must come back as an error code with a
message BEFORE any memory is read or allocated for the oversized request — and leave the handle usable.

  * fewer than 2^32 - 1 read tuples per handle (32-bit tuple indices: first-of-word, cluster_first, the wire of the
    sharded exchange)                                                         humid_gpu_add_batch / _packed4 / _device
  * fewer than 2^32 - 1 tuples over ALL shards (global 32-bit tuple indices)  humid_gpu_shard_begin
  * at most 16 shards, fewer than 2^32 - 1 distinct words over all shards      humid_gpu_shard_partition / _set_layout
  * sharded runs: no edit distance >= 2, no words of more than 31 letters      humid_gpu_shard_begin
  * words of at most 63 letters, edit distance >= 2 only up to 31 letters      humid_gpu_create
The reference has none of these limits (Trie::add takes any word, size_t counters: lib/trie/src/trieFunctions.tcc:45-59);
the executable states them in usage() and refuses such a run before it reads a file (tests/test_cli_cpu.py)."""
import ctypes as C

import numpy as np
import pytest

from humid_b200 import api, sharded
from humid_b200.synth import Synth, baseline_config

pytestmark = pytest.mark.gpu

UNSUPPORTED, ARG = "(-4)", "(-1)"


def small_job(cfg_index=1, R=3000):
    cfg = baseline_config(cfg_index, R)
    s = Synth(cfg)
    hdr, take = s.layout()
    return cfg, s.word_chars(), ([hdr] if hdr else []) + take


def test_tuple_count_limit_on_every_way_in():
    cfg, chars, seg_take = small_job()
    with api.Dedup(word_length=cfg.word_length, distance=cfg.distance, seg_take=seg_take, edit=cfg.edit, maximum=cfg.maximum) as dd:
        dd.add_batch(chars[:1000])
        L, h = dd._L, dd._h
        too_many = (1 << 32) - 1 - 1000  # 1000 + too_many == 2^32 - 1: one more than the handle can number
        p = chars[1000:].ctypes.data_as(C.c_void_p)
        for call in (lambda n: L.humid_gpu_add_batch(h, p, None, 1000, n),
                     lambda n: L.humid_gpu_add_batch_packed4(h, p, None, 1000, n),
                     lambda n: L.humid_gpu_add_batch_device(h, p, None, 1000, n)):
            with pytest.raises(api.HumidGpuError) as e:
                dd._check(call(too_many), "add_batch")
            assert UNSUPPORTED in str(e.value) and "2^32" in str(e.value)
        # (one fewer would be accepted as far as the numbering goes; it is not tried: 100 GB of characters)
        # the handle is as it was: the rest of the job goes in and gives the oracle's counts
        dd.add_batch(chars[1000:])
        dd.finalize()
        whole = api.Dedup(word_length=cfg.word_length, distance=cfg.distance, seg_take=seg_take, edit=cfg.edit, maximum=cfg.maximum)
        whole.add_batch(chars)
        whole.finalize()
        assert dd.counts() == whole.counts() and dd.counts()["total"] == len(chars)
        whole.close()


def test_shard_limits():
    cfg, chars, seg_take = small_job()
    eng = sharded.CudaShardEngine(device=0, word_length=cfg.word_length, distance=cfg.distance, seg_take=seg_take, edit=cfg.edit,
                                  maximum=cfg.maximum)
    eng.add_batch(chars)
    with pytest.raises(api.HumidGpuError) as e:  # this shard's tuples would be numbered past 2^32 - 2
        eng.begin((1 << 32) - 1 - len(chars))
    assert UNSUPPORTED in str(e.value) and "2^32" in str(e.value)
    assert eng.begin((1 << 32) - 2 - len(chars)) > 0  # the last offset that fits
    with pytest.raises(api.HumidGpuError) as e:
        eng.partition(np.arange(16, dtype=np.uint64), 17)
    assert ARG in str(e.value)
    with pytest.raises(api.HumidGpuError) as e:
        eng.set_layout(0, 17, np.arange(16, dtype=np.uint64), np.arange(18, dtype=np.uint64))
    assert ARG in str(e.value)
    bases = np.array([0, 5, (1 << 32) - 1], dtype=np.uint64)  # 2^32 - 1 distinct words over both shards
    with pytest.raises(api.HumidGpuError) as e:
        eng.set_layout(0, 2, np.array([7], dtype=np.uint64), bases)
    assert UNSUPPORTED in str(e.value) and "2^32" in str(e.value)
    eng.close()


@pytest.mark.parametrize("kw,what", [
    (dict(word_length=24, distance=2, seg_take=[12, 12], edit=True), "edit distance >= 2"),
    (dict(word_length=36, distance=1, seg_take=[12, 12, 12]), "more than 31 letters"),
])
def test_configurations_that_are_not_sharded(kw, what):
    eng = sharded.CudaShardEngine(device=0, **kw)
    with pytest.raises(api.HumidGpuError) as e:
        eng.begin(0)
    assert UNSUPPORTED in str(e.value) and what in str(e.value)
    eng.close()


@pytest.mark.parametrize("kw", [
    dict(word_length=64, distance=1, seg_take=[32, 32]),             # -n > 63
    dict(word_length=36, distance=2, seg_take=[12, 12, 12], edit=True),  # edit distance 2 on two-key words
    dict(word_length=24, distance=9, seg_take=[12, 12]),             # Hamming distance > 8
    dict(word_length=24, distance=4, seg_take=[12, 12], edit=True),  # edit distance > 3
])
def test_configurations_refused_at_create_or_finalize(kw):
    cfg, chars, _ = small_job(1, 500)
    try:
        dd = api.Dedup(**kw)
    except api.HumidGpuError as e:
        assert UNSUPPORTED in str(e) or ARG in str(e)
        return
    with pytest.raises(api.HumidGpuError) as e:
        wl = kw["word_length"]
        rows = np.frombuffer(np.random.default_rng(1).choice(list(b"ACGT"), (500, wl)).astype(np.uint8).tobytes(), np.uint8).reshape(500, wl)
        dd.add_batch(rows)
        dd.finalize()
    assert UNSUPPORTED in str(e.value) or "not supported" in str(e.value)
    dd.close()
