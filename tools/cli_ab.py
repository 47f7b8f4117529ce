"""A/B of the executable's host side on one box: the whole program on bench.py's configs[1] leg (4 M tuples, plain FastQ),
alternating the default with one of
    --ab rewrite   HUMID_REWRITE_PARSE=1  (every chunk of the output passes parsed again, round 1's pass 2)
    --ab ingest    HUMID_INGEST_THREADS=1 (one parser thread per input file instead of host/ingest_parallel.h), and that
                   with HUMID_NO_MMAP=1 (the files read in blocks instead of parsed out of a mapping: round 1's reader)
Prints every run's phases; the box, the files and the page cache are the same for both.

    python tools/cli_ab.py [--ab ingest] [--tuples 4000000] [--runs 5]"""
import argparse
import os
import shutil
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tuples", type=int, default=4_000_000)
    ap.add_argument("--runs", type=int, default=5)
    ap.add_argument("--config", type=int, default=1)
    ap.add_argument("--gz", action="store_true", help="gzip -1 copies of the files in, .gz files out")
    ap.add_argument("--ab", default="ingest", choices=["ingest", "rewrite", "pool", "inflate"])
    a = ap.parse_args()
    cfg, files, d, _ = bench.fastq_files(a.config, a.tuples, device="cuda")
    if a.gz:
        files, _ = bench.gzip_copies(files, d)
    out = os.path.join(d, "out_ab")
    others = {"rewrite": [("parse_again", dict(HUMID_REWRITE_PARSE="1"))],
              "inflate": [("four_inflate_threads_per_file", dict(HUMID_INFLATE_THREADS="4")),
                          ("one_inflate_thread_per_file", dict(HUMID_INFLATE_THREADS="1"))],
              "pool": [("inflate_buffers_freed_and_allocated", dict(HUMID_INFLATE_NO_POOL="1")),
                       ("one_inflate_thread_per_file", dict(HUMID_INFLATE_THREADS="1"))],
              "ingest": [("one_parser_per_file", dict(HUMID_INGEST_THREADS="1")),
                         ("one_parser_per_file_read_in_blocks", dict(HUMID_INGEST_THREADS="1", HUMID_NO_MMAP="1"))]}[a.ab]
    envs = dict(others)
    rows = {"default": []}
    rows.update({name: [] for name, _ in others})
    for run in range(a.runs):
        for mode in rows:
            shutil.rmtree(out, ignore_errors=True)
            os.makedirs(out)
            env = dict(os.environ, HUMID_TIMING="1")
            if mode != "default":
                env.update(envs[mode])
            t0 = time.monotonic()
            p = subprocess.run([bench.HUMID_EXE, *cfg.cli_flags(), "-s", "-d", out, "-l", os.path.join(out, "log"), *files], check=True,
                               env=env, capture_output=True, text=True)
            total = time.monotonic() - t0
            ph = {}
            for ln in p.stderr.splitlines():
                nm, sep, rest = ln.rpartition(": ")
                if sep and rest.endswith(" ms"):
                    ph[nm.strip()] = float(rest[:-3]) / 1e3
            rows[mode].append((total, ph.get("Reading data", 0), ph.get("Writing filtered results", 0),
                               ph.get("humid_gpu_create (CUDA context, beside the readers)", 0)))
    for mode, rr in rows.items():
        print(mode, "total / reading / writing / context creation (s):")
        for r in rr:
            print("   %.3f  %.3f  %.3f  %.3f" % r)
        med = lambda i: sorted(r[i] for r in rr)[len(rr) // 2]
        print("   medians: total %.3f  reading %.3f  writing %.3f" % (med(0), med(1), med(2)))
    shutil.rmtree(out, ignore_errors=True)


if __name__ == "__main__":
    main()
