"""Wall time of the `humid` executable (host/ + libhumid_gpu) on synthetic FastQ files:
python tools/cli_bench.py <config index> <tuples> [extra humid args...]"""
import os
import subprocess
import sys

sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))))
import tempfile
import time

from humid_b200.synth import HashSynth, baseline_config

ci, R = int(sys.argv[1]), int(sys.argv[2])
extra = sys.argv[3:]
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
exe = os.path.join(root, "humid_b200", "humid")
with tempfile.TemporaryDirectory(dir="/tmp") as d:
    t0 = time.time()
    cfg = baseline_config(ci, R)
    files = HashSynth(cfg).write_fastq(d)
    size = sum(os.path.getsize(f) for f in files)
    print(f"{cfg.name}: {R} tuples, {len(files)} files, {size / 1e9:.2f} GB written in {time.time() - t0:.1f} s", flush=True)
    args = [exe, "-n", str(cfg.word_length), "-m", str(cfg.distance), "-s", "-d", os.path.join(d, "out"), "-l", os.path.join(d, "log")]
    args += (["-e"] if cfg.edit else []) + (["-x"] if cfg.maximum else []) + extra + files
    for rep in range(2):
        t0 = time.time()
        subprocess.run(args, check=True)
        dt = time.time() - t0
        print(f"run {rep}: {dt:.2f} s wall = {R / dt / 1e6:.2f} M read tuples/s ({size / dt / 1e9:.2f} GB/s of FastQ)", flush=True)
    print(open(os.path.join(d, "log")).read())
    print(open(os.path.join(d, "out", "stats.dat")).read())
