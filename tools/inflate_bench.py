"""Inflate throughput of the host's .gz readers on real-entropy FastQ: zlib's gzread (what the
executable used before), host/inflate_fast.h on one thread, host/inflate_parallel.h on N threads — all through
tests/inflate_check.cc, which also checks size and CRC-32 of what comes out.

    python tools/inflate_bench.py [--reads 300000] [--level 6] [--threads 2,4,8] [--runs 5]

The FastQ is synthetic but not compressible beyond what real files are: Illumina-style names, 150 random bases,
qualities from the binned alphabet with runs.  Prints one markdown table; times are the best of --runs (the readers
compute the CRC-32 the gzip trailer asks for, the tool's own CRC is not in the time)."""
import argparse
import gzip
import json
import os
import subprocess
import sys
import tempfile
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def fastq(n, rng):
    L = 150
    seq = rng.choice(np.frombuffer(b"ACGT", np.uint8), (n, L), p=[.28, .22, .22, .28])
    # qualities: runs of a binned alphabet (NovaSeq-like), mostly 'F'
    q = rng.choice(np.frombuffer(b"FFFFFF:,#", np.uint8), (n, L // 5)).repeat(5, axis=1)
    noise = rng.random((n, L)) < 0.08
    q = np.where(noise, rng.choice(np.frombuffer(b"F:,#", np.uint8), (n, L)), q)
    parts = []
    for i in range(n):
        parts.append(b"@A00123:45:HXXXXXXXX:1:%d:%d:%d 1:N:0:ACGTACGT\n" % (1101 + i // 50000, 1000 + i % 30000, 2000 + (i * 7) % 30000))
        parts.append(seq[i].tobytes())
        parts.append(b"\n+\n")
        parts.append(q[i].tobytes())
        parts.append(b"\n")
    return b"".join(parts)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reads", type=int, default=300000)
    ap.add_argument("--level", type=int, default=6)
    ap.add_argument("--threads", default="2,4,8")
    ap.add_argument("--runs", type=int, default=5)
    a = ap.parse_args()
    with tempfile.TemporaryDirectory() as d:
        exe = os.path.join(d, "inflate_check")
        subprocess.run(["g++", "-O2", "-std=c++17", "-pthread", "-o", exe, os.path.join(ROOT, "tests", "inflate_check.cc"), "-lz"], check=True)
        data = fastq(a.reads, np.random.default_rng(1))
        path = os.path.join(d, "reads.fastq.gz")
        with open(path, "wb") as f:
            f.write(gzip.compress(data, a.level))
        gz_bytes = os.path.getsize(path)
        want = (len(data), zlib.crc32(data))
        rows = []
        for mode in ["zlib", "fast"] + ["par%s" % t for t in a.threads.split(",") if t]:
            best = None
            for _ in range(a.runs):
                out = subprocess.run([exe, mode, path], capture_output=True, text=True, check=True).stdout.split()
                assert (int(out[0]), int(out[1], 16)) == want, (mode, out)
                best = float(out[2]) if best is None else min(best, float(out[2]))
            rows.append((mode, best))
        base = rows[0][1]
        print(json.dumps({"fastq_bytes": len(data), "gz_bytes": gz_bytes, "level": a.level, "host_cores": os.cpu_count(),
                          "seconds": {m: s for m, s in rows}}))
        print("| reader | seconds | MB/s inflated | vs zlib |\n|---|---|---|---|")
        names = {"zlib": "zlib gzread", "fast": "host/inflate_fast.h, 1 thread"}
        for m, s in rows:
            print("| %s | %.3f | %.0f | %.2fx |" % (names.get(m, "host/inflate_parallel.h, %s threads" % m[3:]), s, len(data) / s / 1e6, base / s))


if __name__ == "__main__":
    sys.exit(main())
