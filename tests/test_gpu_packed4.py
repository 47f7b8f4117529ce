"""The 4-bit wire format (humid_gpu_add_batch_packed4) against the oracle and against the ASCII entry point: same
words, counts, first tuples, neighbours, clusters, keep mask, ids and statistics files, for even and odd word lengths,
words of more than 31 letters, short segments, and bytes that are not nucleotides (N, lower case, padding)."""
import numpy as np
import pytest

import oracle
from humid_b200 import api
from humid_b200.synth import Synth, baseline_config
from test_gpu_parity import assert_same_as_oracle

pytestmark = pytest.mark.gpu

CASES = [
    (1, 30000, {}), (0, 30000, {}), (2, 20000, {}), (4, 40000, {}),                # 8/8/8, 8+8/8, 12/12, 10+7/7
    (0, 20000, dict(word_length=9, genome_len=2000)), (2, 20000, dict(word_length=11, genome_len=3000)),
    (1, 20000, dict(word_length=31)), (1, 20000, dict(word_length=36)), (2, 20000, dict(word_length=63, distance=2)),
    (1, 20000, dict(word_length=33, maximum=True)), (3, 6000, {}),
]


def run(chars, seg_take, cfg, packed, seg_len=None, batches=3):
    dd = api.Dedup(word_length=cfg.word_length, distance=cfg.distance, seg_take=seg_take, edit=cfg.edit, maximum=cfg.maximum)
    R = chars.shape[0]
    step = max(1, (R + batches - 1) // batches)
    for lo in range(0, R, step):
        sl = None if seg_len is None else seg_len[lo:lo + step]
        if packed:
            dd.add_batch_packed4(api.pack_nibbles(chars[lo:lo + step]), sl)
        else:
            dd.add_batch(chars[lo:lo + step], sl)
    dd.finalize()
    return dd


@pytest.mark.parametrize("ci,R,kw", CASES, ids=[f"c{c + 1}_{R}_{'_'.join(f'{k}{v}' for k, v in kw.items())}" for c, R, kw in CASES])
def test_packed4_matches_oracle(ci, R, kw):
    cfg = baseline_config(ci, R, **kw)
    s = Synth(cfg)
    chars = s.word_chars()
    rng = np.random.default_rng(R + cfg.word_length)
    rows = rng.integers(0, R, size=60)
    chars[rows[:20], rng.integers(0, cfg.word_length, size=20)] = ord("a")   # lower case is not a nucleotide
    chars[rows[20:40], rng.integers(0, cfg.word_length, size=20)] = 0        # nor is a NUL
    chars[rows[40:], cfg.word_length - 1] = ord("N")                         # the last character of the word
    hdr, take = s.layout()
    seg_take = ([hdr] if hdr else []) + take
    r = oracle.Result(oracle.pack_words(chars, seg_take), cfg.word_length, cfg.distance, cfg.edit, cfg.maximum)
    with run(chars, seg_take, cfg, True) as dd:
        assert_same_as_oracle(dd, r, R)


def test_packed4_short_segments_filter():
    """makeStringSize_ (src/fastq.cc:57-66): a segment shorter than its share is padded with N => the tuple is filtered,
    whatever the bytes behind the segment's end are."""
    cfg = baseline_config(1, 5000)
    s = Synth(cfg)
    chars = s.word_chars()
    hdr, take = s.layout()
    seg_take = ([hdr] if hdr else []) + take
    R = chars.shape[0]
    seg_len = np.tile(np.array(seg_take, dtype=np.uint16), (R, 1))
    rng = np.random.default_rng(3)
    short = rng.random(R) < 0.05
    which = rng.integers(0, len(seg_take), size=R)
    seg_len[short, which[short]] -= 1
    r = oracle.Result(oracle.pack_words(chars, seg_take, seg_len), cfg.word_length, cfg.distance, cfg.edit, cfg.maximum)
    assert r.usable < R - 100
    with run(chars, seg_take, cfg, True, seg_len) as a, run(chars, seg_take, cfg, False, seg_len) as b:
        assert_same_as_oracle(a, r, R)
        assert_same_as_oracle(b, r, R)


def test_pack_nibbles_layout():
    x = np.frombuffer(b"ACGTNacgt", dtype=np.uint8).reshape(1, 9)
    assert api.pack_nibbles(x).tolist() == [[0x01, 0x23, 0x44, 0x44, 0x40]]


@pytest.mark.parametrize("ci,R,kw", [(1, 30000, {}), (0, 20000, dict(word_length=9, genome_len=2000)), (1, 20000, dict(word_length=36)),
                                     (4, 5000, {}), (1, 1025, {}), (1, 1, {})])
def test_pack_tiles_by_bulk_copy(monkeypatch, ci, R, kw):
    """pack_words_bulk_kernel (cp.async.bulk + mbarrier tile loads, HUMID_GPU_PACK=bulk) against the oracle: whole
    tiles, a last partial tile, sizes that are not a multiple of 16 bytes, words of 36 letters."""
    monkeypatch.setenv("HUMID_GPU_PACK", "bulk")
    cfg = baseline_config(ci, R, **kw)
    s = Synth(cfg)
    chars = s.word_chars()
    hdr, take = s.layout()
    seg_take = ([hdr] if hdr else []) + take
    r = oracle.Result(oracle.pack_words(chars, seg_take), cfg.word_length, cfg.distance, cfg.edit, cfg.maximum)
    with run(chars, seg_take, cfg, False, batches=2) as dd:
        assert_same_as_oracle(dd, r, R)
