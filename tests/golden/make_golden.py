"""Generate tests/golden/*.npz from the UNMODIFIED reference binary (oracle/_ref/humid).

Run in the build container (needs /root/reference to have been built into oracle/_ref):

    python tests/golden/make_golden.py

For every case a small seeded synthetic FastQ set is written to a temp dir, the reference
CLI is run with `-s -a` on it, and what it wrote is stored next to the inputs:

    chars      uint8[R, n]   the word characters of every tuple (the C-ABI input)
    seg_take   uint32[S]     header-UMI size (if > 0) followed by nt per file (preCompute)
    keep       uint8[R]      1 iff the tuple appears in *_dedup.fastq
    cid        uint32[R]     the cluster id appended by -a in *_annotated.fastq
    counts/neigh/clusters/stats   the four .dat files, verbatim bytes
    n, d, edit, maximum      the options

The fixtures are what the `-m gpu` parity tests compare the CUDA path with on the GPU box,
where /root/reference does not exist.
"""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import oracle  # noqa: E402
from humid_b200.synth import Synth, baseline_config  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))

# (file stem, BASELINE config index, tuples, overrides)
CASES = [
    ("c1_ham1_max", 0, 6000, {}),
    ("c1_ham1_dir", 0, 6000, dict(maximum=False)),
    ("c2_umifile_ham1_dir", 1, 6000, {}),
    ("c2_umifile_n30", 1, 4000, dict(word_length=30)),
    ("c3_noumi_ham2_dir", 2, 6000, {}),
    ("c3_noumi_ham2_max", 2, 6000, dict(maximum=True)),
    ("c3_noumi_ham3_dir", 2, 3000, dict(distance=3)),
    ("c4_indel_edit1", 3, 5000, dict(distance=1)),
    ("c4_indel_edit2_dir", 3, 5000, {}),
    ("c4_indel_edit2_max", 3, 5000, dict(maximum=True)),
    ("c5_umi10_ham1_dir", 4, 6000, {}),
    ("short_n8_ham1_dir", 0, 5000, dict(word_length=8, genome_len=2000)),
    ("short_n10_edit2_dir", 3, 3000, dict(word_length=10, genome_len=3000)),
    ("short_n9_ham2_max", 2, 4000, dict(word_length=9, maximum=True, genome_len=3000)),
]


def read_fastq_headers(path):
    with open(path, "rb") as f:
        return [ln.rstrip(b"\n") for i, ln in enumerate(f) if i % 4 == 0]


def make(stem, ci, R, kw):
    cfg = baseline_config(ci, R, **kw)  # a tiny genome_len => dense neighbourhoods, count ties
    s = Synth(cfg)
    with tempfile.TemporaryDirectory() as tmp:
        files = s.write_fastq(tmp, prefix="g")
        out = os.path.join(tmp, "out")
        oracle.run_ref_humid(files, out, cfg.cli_flags(), stats=True, annotate=True)
        in_hdr = read_fastq_headers(files[0])
        dedup = os.path.join(out, "g_R1_dedup.fastq")
        kept = set(read_fastq_headers(dedup))
        assert len(kept) == len(read_fastq_headers(dedup)), "synthetic headers must be unique"
        keep = np.array([h in kept for h in in_hdr], dtype=np.uint8)
        ann = read_fastq_headers(os.path.join(out, "g_R1_annotated.fastq"))
        cid = np.array([int(h.rsplit(b":", 1)[1]) for h in ann], dtype=np.uint32)
        assert all(a.rsplit(b":", 1)[0] == h for a, h in zip(ann, in_hdr))
        dat = {k: open(os.path.join(out, k + ".dat"), "rb").read() for k in ("counts", "neigh", "clusters", "stats")}
    hdr, take = s.layout()
    seg_take = np.array(([hdr] if hdr else []) + take, dtype=np.uint32)
    np.savez_compressed(
        os.path.join(HERE, stem + ".npz"), chars=s.word_chars(), seg_take=seg_take, keep=keep, cid=cid,
        n=cfg.word_length, d=cfg.distance, edit=int(cfg.edit), maximum=int(cfg.maximum),
        **{k: np.frombuffer(v, dtype=np.uint8) for k, v in dat.items()})
    print(stem, "R", R, "kept", int(keep.sum()), dat["stats"].decode().replace("\n", " "))


if __name__ == "__main__":
    assert oracle.have_ref(), "build oracle/_ref first (make -C oracle ref)"
    only = sys.argv[1:]
    for stem, ci, R, kw in CASES:
        if not only or stem in only:
            make(stem, ci, R, kw)
