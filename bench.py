#!/usr/bin/env python
"""Benchmark of the deduplication hot path (BASELINE.json: read pairs deduplicated / s).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--tuples R]

One "step" = one complete pass of the hot path over one batch of synthetic read tuples:
word construction (pack) -> distinct-word table with counts -> Hamming neighbour search ->
clustering -> keep mask + the three histograms (all inside humid_gpu_finalize).  The workload is
BASELINE.json configs[1] (50 M read pairs, 12 bp UMI in a separate FastQ, word length 24 = 8/8/8 nt,
Hamming 1, directional clustering), `--tuples` read tuples PER GPU (weak scaling: N GPUs process
N x that many tuples of one common library; the word space is range-partitioned over the GPUs).

  value         whole-job read pairs/s with the word characters already resident in HBM
  e2e           the same through the C ABI with HOST buffers: per step the H2D copy of the word characters
                (pinned memory) and the D2H read of the keep mask, the histograms and the stats scalars are
                inside the timed region.  Its like-for-like CPU figure is the reference arm's `value`: the
                reference's own functions driven from host memory (oracle/_ref/hotpath_probe), same scope —
                word characters in, keep decision + statistics out, no FastQ parsing, no files.
  cli_e2e       whole program against whole program: `humid_b200/humid` here, `oracle/_ref/humid` in the reference
                arm, on IDENTICAL FastQ files (generated once per box under /tmp and shared by both arms), FastQ in ->
                deduplicated FastQ + the four statistics files out, with digests of the outputs so that the two arms
                can be compared byte for byte
  roofline      the stage with the largest device time: `frac` by SURVEY.md 8(d)'s algorithmic bytes of the
                stage's kernel family, `pass_frac` by the bytes its launches actually have to move (one read + one
                write per radix pass), `traffic_ratio` = measured DRAM bytes / 8(d) bytes; `kernels_8d` and
                `stages` give the same for every kernel family / stage
  cpu_baseline  the reference hot path from host memory on a bounded sample of the same workload (1 core:
                the reference is single-threaded by construction — one global trie), checked against the GPU
                result on the same sample
  parity        (N > 1) rank 0 also runs the SAME total job on one GPU outside the timed region; counts, the
                four statistics files and the keep mask of every rank's slice must be identical, else exit 1

`--impl reference` times only the CPU arm and prints its own line (same `config`).
"""
from __future__ import annotations

import argparse
import hashlib
import json
import os
import shutil
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "read pairs deduplicated/sec"
UNIT = "read pairs/s"
WORKLOAD = "configs[1]: synthetic read pairs, 12bp UMI in separate FastQ, -n 24 (8/8/8), Hamming 1, directional"
REF_SAMPLE_TUPLES = 250_000
# whole-program legs (config index, read tuples): configs[0] at its own size (BASELINE.md's 19.4 s row) and
# configs[1] at the largest size the reference finishes in about two minutes
CLI_LEGS = ((0, 1_000_000), (1, 4_000_000))
CLI_CACHE = os.environ.get("HUMID_BENCH_CACHE", "/tmp/humid_bench_fastq")
HUMID_EXE = os.path.join(ROOT, "humid_b200", "humid")


# stage (humid_gpu_stage_stats) -> the kernel that does its work
KERNEL_OF = {
    "sort_words": "radix_pass_kernel<true, ShiftDigit, unsigned int> (1 histogram + 4 digit passes over 32-bit prefixes)",
    "sort_buckets": "bucket_sort_kernel",
    "join_sort": "radix_pass_kernel<true, ShiftDigit> (1 histogram + 3 digit passes)",
    "table_partition": "radix_pass_kernel<true, PrefixDigit>", "table_insert": "table_insert_kernel",
    "pack": "pack_words_kernel", "rank_columns": "gather_sorted_words_kernel", "join_hamming_b0": "hamming_join_kernel",
    "join_hamming_b1": "hamming_join_kernel", "join_hamming_b2": "hamming_join_kernel", "csr_build": "csr_fill_kernel",
    "union_find": "uf_union_pairs_kernel", "replay_directional": "replay_small_kernel", "replay_maximum": "replay_small_kernel",
}

# kernel family of every stage; the family's algorithmic bytes are a formula of the job's counts
FAMILY_OF = {
    "pack": "K1", "table_clear": "K1", "table_partition": "K1", "table_insert": "K1", "prefix_hist": "K1",
    "table_compact": "K2", "sort_words": "K2", "sort_buckets": "K2", "rank_columns": "K2", "word_runs": "K2",
    "sort_tuples": "K2",
    "join_sort": "K3", "join_hamming_b0": "K3", "join_hamming_b1": "K3", "join_hamming_b2": "K3",
    "join_hamming_b3": "K3", "join_hamming_b4": "K3", "join_hamming_b5": "K3",
    "table_ranks": "K3", "filter_build": "K3", "probe_flagged": "K3",
    "csr_build": "K3'", "csr_sort": "K3'",
    "union_find": "K4", "group_big_components": "K4", "replay_directional": "K4", "replay_maximum": "K4",
    "replay_roots": "K4", "replay_large": "K4", "cluster_ids_scan": "K4", "compact_clusters": "K4",
    "keep": "K5", "cluster_ids": "K5",
    "hist_counts": "a12", "hist_neigh": "a12", "hist_clusters": "a12",
}


def family_bytes(cfg, counts, entries):
    """SURVEY.md 8(d) 'Algorithmic bytes per unit' x the units of one step (K = 1 key word; n, d from cfg)."""
    n, d = cfg.word_length, cfg.distance
    K = 1 if n <= 31 else 2
    usable, U, C, R = counts["usable"], counts["unique"], counts["clusters"], counts["total"]
    variants = 1.5 * n + (2.25 * n * (n - 1) / 2.0 if d >= 2 else 0.0)
    return {
        "K1": usable * (n + 8 * K + 8 + 8 + 4),
        "K2": U * (2 * (8 * K + 4) + 4),
        "K3": U * 8 * K * (1 + variants) + 16 * (entries // 2),
        "K3'": entries * 16,
        "K4": entries * 16 + U * 12,
        "K5": R * 24,
        "a12": U * 12 + C * 8,  # not in 8(d)'s table: count (4) + row pointer (8) per word, size (8) per cluster
    }


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def like_for_like_config(cfg, args):
    """The `config` object: identical in both arms (what differs — sizes, samples — sits beside it)."""
    return {"workload": WORKLOAD if args.config == 1 and not args.word_length else cfg.name,
            "word_length": cfg.word_length, "distance": cfg.distance, "edit": bool(cfg.edit),
            "clustering": "maximum" if cfg.maximum else "directional",
            "tuples_per_gpu": args.tuples,
            "l2": "inputs larger than L2 (word characters %.2f GB per GPU, every array of a step is larger than the "
                  "126 MB L2)" % (args.tuples * cfg.word_length / 1e9)}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs.  The process is
    started before the warm-up steps (its start-up takes the driver's locks for a while, which stretched the
    first timed steps of host-synchronising paths several times over); mark() drops the samples taken so far."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.rows = []
        self.start = 0
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.rows.append([x.strip() for x in ln.split(",")])

    def mark(self):
        """The timed region starts now (waits for the first sample if nvidia-smi is still starting up)."""
        t0 = time.time()
        while self.proc and not self.rows and time.time() - t0 < 5.0:
            time.sleep(0.02)
        self.start = len(self.rows)

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows[self.start:]:
            try:
                sm.append(float(r[0]))
                mx = max(mx, float(r[1]))
            except (ValueError, IndexError):
                continue
            for nm, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


# ----------------------------------------------------------------------------------------
# whole program against whole program: identical FastQ files for both executables
# ----------------------------------------------------------------------------------------
def sha_file(path, h=None):
    h = h or hashlib.sha256()
    with open(path, "rb") as f:
        for blk in iter(lambda: f.read(1 << 22), b""):
            h.update(blk)
    return h


def fastq_files(ci, R, device="cpu"):
    """FastQ files of BASELINE.json configs[ci] at R tuples under CLI_CACHE: written by whichever arm runs first
    on this box, found by the other (the generator is a pure function of (seed, tuple index): a regenerated copy
    holds the same bytes)."""
    from humid_b200.synth import HashSynth, baseline_config

    cfg = baseline_config(ci, R)
    d = os.path.join(CLI_CACHE, f"{cfg.name}_{R}")
    marker = os.path.join(d, "files.json")
    if os.path.exists(marker):
        try:
            meta = json.load(open(marker))
            if all(os.path.getsize(f) == s for f, s in meta["files"]):
                return cfg, [f for f, _ in meta["files"]], d, 0.0
        except Exception:
            pass
    shutil.rmtree(d, ignore_errors=True)
    os.makedirs(d, exist_ok=True)
    t0 = time.monotonic()
    files = HashSynth(cfg).write_fastq(d, prefix="in", device=device)
    dt = time.monotonic() - t0
    json.dump({"files": [(f, os.path.getsize(f)) for f in files]}, open(marker, "w"))
    return cfg, files, d, dt


def gzip_copies(files, d):
    """`<file>.gz` next to every plain FastQ file (gzip -1, the files side by side); made once per box like the files."""
    out = [f + ".gz" for f in files]
    marker = os.path.join(d, "files_gz.json")
    if os.path.exists(marker) and all(os.path.exists(f) for f in out):
        return out, 0.0
    t0 = time.monotonic()
    procs = [subprocess.Popen(["gzip", "-1", "-k", "-f", f]) for f in files]
    for p in procs:
        if p.wait() != 0:
            raise RuntimeError("gzip failed")
    json.dump({"files": [(f, os.path.getsize(f)) for f in out]}, open(marker, "w"))
    return out, time.monotonic() - t0


def sha_inflated(path):
    import gzip

    h = hashlib.sha256()
    with gzip.open(path, "rb") as f:
        for blk in iter(lambda: f.read(1 << 22), b""):
            h.update(blk)
    return h


def output_digest(out_dir, names):
    """sha256 over the deduplicated FastQ files (a .gz by its inflated bytes: the deflate libraries of the two arms
    differ) and the four statistics files, in a fixed order."""
    h = hashlib.sha256()
    per = {}
    for nm in names:
        p = os.path.join(out_dir, nm)
        hh = sha_inflated(p) if nm.endswith(".gz") else sha_file(p)
        per[nm] = hh.hexdigest()[:16]
        h.update(hh.digest())
    return h.hexdigest(), per


def cli_leg(kind, ci, R, device="cpu", reps=1, gz=False):
    """One whole-program run (`kind` = "humid_b200" | "reference") of configs[ci] at R tuples; gz: the same FastQ as
    .gz files in, .gz files out."""
    import oracle

    cfg, files, d, gen_s = fastq_files(ci, R, device)
    plain_bytes = sum(os.path.getsize(f) for f in files)
    if gz:
        files, gz_s = gzip_copies(files, d)
        gen_s += gz_s
    size = sum(os.path.getsize(f) for f in files)
    out_names = [os.path.basename(f).replace(".fastq", "_dedup.fastq") for f in files]
    out_names += ["stats.dat", "counts.dat", "neigh.dat", "clusters.dat"]
    flags = cfg.cli_flags()
    times, breakdown = [], {}
    out = os.path.join(d, "out_" + kind + ("_gz" if gz else ""))
    for rep in range(reps):
        shutil.rmtree(out, ignore_errors=True)
        if kind == "reference":
            t = oracle.run_ref_humid(files, out, flags, stats=True)["wall_s"]
            try:  # the reference's own phase lines (whole seconds)
                for ln in open(os.path.join(out, "humid.log")):
                    if "... done. (" in ln:
                        nm, _, rest = ln.partition("... done. (")
                        m, _, s = rest.partition("m")
                        breakdown[nm.strip()] = int(m) * 60 + int(s.split("s")[0])
            except Exception:
                pass
        else:
            os.makedirs(out, exist_ok=True)
            cmd = [HUMID_EXE, *flags, "-s", "-d", out, "-l", os.path.join(out, "humid.log"), *files]
            env = dict(os.environ, HUMID_TIMING="1")
            t0 = time.monotonic()
            p = subprocess.run(cmd, check=True, env=env, capture_output=True, text=True)
            t = time.monotonic() - t0
            breakdown = {}
            for ln in p.stderr.splitlines():
                nm, sep, rest = ln.rpartition(": ")
                if sep and rest.endswith(" ms"):
                    try:
                        breakdown[nm.strip()] = round(float(rest[:-3]) / 1e3, 3)
                    except ValueError:
                        pass
        times.append((t, breakdown))
    times.sort(key=lambda x: x[0])
    sec, breakdown = times[len(times) // 2]
    digest, per = output_digest(out, out_names)
    shutil.rmtree(out, ignore_errors=True)
    leg = {"config": cfg.name, "input": "gzip -1 files in, .gz files out" if gz else "plain FastQ files",
           "flags": " ".join(flags + ["-s"]), "tuples": R, "fastq_bytes": plain_bytes, "input_bytes": size, "seconds": round(sec, 3),
           "runs": len(times), "pairs_per_s": R / sec, "breakdown_s": breakdown, "outputs_sha256": digest,
           "outputs": per, "fastq_generated_s": round(gen_s, 1)}
    # the other arm's digest of the same job, if it ran on this box before: the outputs must agree byte for byte
    tag = "_gz" if gz else ""
    note = os.path.join(d, "digest_" + kind + tag + ".json")
    json.dump({"outputs_sha256": digest}, open(note, "w"))
    other = os.path.join(d, "digest_" + ("humid_b200" if kind == "reference" else "reference") + tag + ".json")
    if os.path.exists(other):
        try:
            leg["identical_to_other_arm"] = json.load(open(other))["outputs_sha256"] == digest
        except Exception:
            pass
    return leg


def cli_legs(kind, args, device="cpu"):
    legs = []
    for ci, R in CLI_LEGS:
        R = int(R * args.cli_scale)
        if R <= 0:
            continue
        legs.append(cli_leg(kind, ci, R, device, reps=1 if kind == "reference" else 3))
    # configs[0] once more with compressed files on both sides: the usual way FastQ is stored; the reference reads it
    # through isa-l and writes through libdeflate, this repo through host/inflate_parallel.h and zlib
    R = int(CLI_LEGS[0][1] * args.cli_scale)
    if R > 0 and shutil.which("gzip"):
        legs.append(cli_leg(kind, CLI_LEGS[0][0], R, device, reps=1 if kind == "reference" else 3, gz=True))
    return legs


# ----------------------------------------------------------------------------------------
# CPU arm: the reference's hot path from host memory on a bounded sample
# ----------------------------------------------------------------------------------------
def sample_chars(args, sample):
    from humid_b200.synth import HashSynth, baseline_config

    cfg = baseline_config(args.config, sample, **({"word_length": args.word_length} if args.word_length else {}))
    syn = HashSynth(cfg)
    hdr, take = syn.layout()
    return cfg, syn.word_chars(0, sample).numpy(), hdr, take


def cpu_reference_run(args, steps: int, warmup: int, sample: int):
    """Returns (pairs/s, ms per step, kind, cores, sample description, last result)."""
    import oracle

    cfg, chars, hdr, take = sample_chars(args, sample)
    tmp = tempfile.mkdtemp(prefix="humid_ref_")
    res = None
    try:
        if os.access(oracle.REF_HOTPATH, os.X_OK):
            times = []
            for i in range(warmup + steps):
                res = oracle.run_ref_hotpath(chars, hdr, take, cfg.distance, cfg.edit, cfg.maximum, tmp)
                if i >= warmup:
                    times.append(res["seconds"]["hot_path"])
            kind = "reference"
            what = (f"{sample} read tuples of the same library model; the reference's own makeWord / Trie::add / "
                    f"find*Neighbours / findClusters / trie.find + visited test / runStatistics / clusterStats "
                    f"(oracle/_ref/hotpath_probe: unmodified src/humid.cc, src/cluster.cc, lib/trie) driven from word "
                    f"characters in host memory — no FastQ parsing, no file output; sum of the phase times")
        else:
            import numpy as np

            seg = ([hdr] if hdr else []) + take
            times = []
            for i in range(warmup + steps):
                t0 = time.monotonic()
                keys = oracle.pack_words(chars, seg)
                o = oracle.Result(keys, cfg.word_length, cfg.distance, cfg.edit, cfg.maximum)
                if i >= warmup:
                    times.append(time.monotonic() - t0)
            res = {"total": o.total, "usable": o.usable, "unique": o.unique, "clusters": o.clusters,
                   "keep": o.keep.astype(bool), "stats": o.stats_text(), "seconds": {"hot_path": times[-1]}}
            kind = "port"
            what = f"{sample} read tuples, C restatement (oracle/humid_oracle.c) from word characters in memory"
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    t = sum(times) / len(times)
    return sample / t, t * 1e3, kind, 1, what, res


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    from humid_b200.synth import baseline_config

    cfg = baseline_config(args.config, args.tuples, **({"word_length": args.word_length} if args.word_length else {}))
    val, ms, kind, cores, what, res = cpu_reference_run(args, args.steps, args.warmup, args.ref_sample)
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u64", "data": "synthetic",
        "config": like_for_like_config(cfg, args),
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": kind, "sample": what,
                         "sample_tuples": args.ref_sample, "host_cores": os.cpu_count(),
                         "phase_seconds": res["seconds"], "result": {k: res[k] for k in ("total", "usable", "unique", "clusters")}},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    if world == 1 and not args.no_cli:
        line["cli_e2e"] = cli_legs("reference", args)
    print(json.dumps(line), flush=True)


def bind_near_gpu(local):
    """Run this rank's host threads on the cores next to its GPU (the box's own PCIe / NUMA layout from
    sysfs), so that the pinned upload buffer lives in that socket's memory.  Returns the NUMA node or None."""
    try:
        import torch

        p = torch.cuda.get_device_properties(local)
        bdf = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        base = "/sys/bus/pci/devices/" + bdf
        cpus = set()
        for part in open(base + "/local_cpulist").read().strip().split(","):
            if part:
                a, _, b = part.partition("-")
                cpus.update(range(int(a), int(b or a) + 1))
        if cpus:
            os.sched_setaffinity(0, cpus)
        return int(open(base + "/numa_node").read())
    except Exception:
        return None


# ----------------------------------------------------------------------------------------
# GPU arm
# ----------------------------------------------------------------------------------------
def stats_digests(stats):
    return {k: hashlib.sha256(v.encode()).hexdigest()[:16] for k, v in sorted(stats.items())}


def run_gpu_arm(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    from humid_b200 import api
    from humid_b200.synth import HashSynth, baseline_config

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the deduplication hot path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa = bind_near_gpu(local)  # before the pinned buffer is allocated: first touch decides the node
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    if world != args.gpus:
        if rank == 0:
            sys.stderr.write(f"warning: --gpus {args.gpus} but WORLD_SIZE={world}; using {world}\n")

    R_per = args.tuples
    R_all = R_per * world
    cfg = baseline_config(args.config, R_all, **({"word_length": args.word_length} if args.word_length else {}))
    syn = HashSynth(cfg)
    hdr, take = syn.layout()
    seg_take = ([hdr] if hdr else []) + take
    lo = rank * R_per
    d_chars = syn.word_chars(lo, lo + R_per, device=dev)
    h_chars = torch.empty(d_chars.shape, dtype=torch.uint8, pin_memory=True)
    h_chars.copy_(d_chars)
    torch.cuda.synchronize()
    h_np = h_chars.numpy()
    # the 4-bit wire format of humid_gpu_add_batch_packed4 (two characters per byte), packed once outside the timed
    # region the way the executable's parser writes it as it goes; page-locked like the ASCII buffer
    wire = args.wire
    row_bytes = (cfg.word_length + 1) // 2
    h_nib = torch.empty((R_per, row_bytes), dtype=torch.uint8, pin_memory=True)
    h_nib_np = h_nib.numpy()
    for a in range(0, R_per, 4_000_000):
        h_nib_np[a:a + 4_000_000] = api.pack_nibbles(h_np[a:a + 4_000_000])

    kw = dict(word_length=cfg.word_length, distance=cfg.distance, seg_take=seg_take, edit=cfg.edit, maximum=cfg.maximum,
              device=local)
    transport = None
    if world > 1:
        from humid_b200 import sharded
        dd = sharded.make_sharded(expected_tuples=R_per, **kw)
        transport = getattr(dd, "transport", "nccl")
    else:
        dd = api.Dedup(expected_tuples=R_per, **kw)
    ext = torch.cuda.ExternalStream(dd.stream, device=dev)
    BATCH = 8_000_000

    def step_device():
        dd.reset()
        dd.add_batch_device(d_chars.data_ptr(), R_per)
        dd.finalize()  # ... through the keep bitmap and the three histograms
        return dd.counts()

    d2h = [0]

    def step_e2e(packed=None):
        packed = (wire == "packed4") if packed is None else packed
        dd.reset()
        for a in range(0, R_per, BATCH):
            if packed:
                dd.add_batch_packed4(h_nib_np[a:a + BATCH])
            else:
                dd.add_batch(h_np[a:a + BATCH])
        dd.finalize()
        bits = dd.keep_bits()
        n = bits.nbytes
        for w in ("counts", "neigh", "clusters"):
            k, v = dd.histogram(w)
            n += k.nbytes + v.nbytes
        c = dd.counts()
        d2h[0] = n + 32
        return c, bits

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record(ext)
        for _ in range(steps):
            out = fn()
        e1.record(ext)
        barrier()
        wall = time.perf_counter() - t0
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms, wall * 1e3], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms, wall = float(t[0]), float(t[1]) / 1e3
        return ms, wall, out

    sampler = ClockSampler(local) if rank == 0 else None
    for _ in range(args.warmup):
        step_device()
    if sampler:
        sampler.mark()
    ms_dev, wall_dev, counts = timed(step_device, args.steps)
    stages = dd.stage_stats()
    launches_per_step = sum(s["launches"] for s in stages)
    entries = int(getattr(dd, "n_entries", lambda: 0)())
    for _ in range(max(1, min(args.warmup, 2))):
        step_e2e()
    ms_e2e, wall_e2e, (counts_e2e, _) = timed(step_e2e, args.steps)
    # the other wire format beside it (a few steps): what the bytes on the host link cost
    other = (lambda: step_e2e(wire != "packed4"))
    other()
    ms_alt, wall_alt, (counts_alt, _) = timed(other, max(2, args.steps // 2))
    t_alt = max(ms_alt / 1e3, wall_alt) / max(2, args.steps // 2)
    assert counts_alt == counts_e2e, (counts_alt, counts_e2e)
    clocks = sampler.stop() if sampler else None
    assert counts == counts_e2e, (counts, counts_e2e)
    phases = [(n, round(t, 3)) for n, t in (getattr(dd, "phases", None) or [])] if world > 1 else None
    if world > 1:  # every rank's phases of its last step: what the slowest rank spends where (diagnostic, untimed)
        allp = [None] * world
        dist.all_gather_object(allp, phases)
        if rank == 0 and phases and all(a and len(a) == len(phases) for a in allp):
            spread = []
            for i, (nm, _) in enumerate(phases):
                vals = [a[i][1] for a in allp]
                spread.append((nm, min(vals), max(vals), int(np.argmax(vals))))
            phases = [{"phase": nm, "min_ms": lo, "max_ms": hi, "slowest_rank": r} for nm, lo, hi, r in spread]
    if phases and rank == 0:
        sys.stderr.write("phases (ms, last step): " + json.dumps(phases) + "\n")

    # ---- N > 1: the same total job on ONE GPU (outside the timed region) must give identical results ----
    parity = None
    if world > 1:
        keep_local = dd.keep_bits()  # this rank's slice of the job's keep mask (collective)
        stats_sharded = dd.stats_text()  # collective
        mine = hashlib.sha256(np.ascontiguousarray(keep_local).tobytes()).hexdigest()
        gathered = [None] * world
        dist.all_gather_object(gathered, mine)
        if rank == 0:
            with api.Dedup(expected_tuples=R_all, **kw) as one:
                for r in range(world):
                    if r == 0:
                        one.add_batch_device(d_chars.data_ptr(), R_per)
                    else:
                        ch = syn.word_chars(r * R_per, (r + 1) * R_per, device=dev)
                        torch.cuda.synchronize()
                        one.add_batch_device(ch.data_ptr(), R_per)
                        torch.cuda.synchronize()
                        del ch
                one.finalize()
                c1 = one.counts()
                s1 = one.stats_text()
                k1 = [hashlib.sha256(np.ascontiguousarray(one.keep_bits(r * R_per, R_per)).tobytes()).hexdigest()
                      for r in range(world)] if R_per % 64 == 0 else None
            keep_sha = hashlib.sha256("".join(gathered).encode()).hexdigest()
            equal = c1 == counts and s1 == stats_sharded and (k1 is None or k1 == gathered)
            parity = {"against": f"the same {R_all} tuples on one GPU (api.Dedup), outside the timed region",
                      "counts": counts, "counts_one_gpu": c1, "keep_sha256": keep_sha,
                      "keep_slices_equal": None if k1 is None else k1 == gathered,
                      "stats_digests": stats_digests(stats_sharded), "stats_digests_one_gpu": stats_digests(s1),
                      "equal": bool(equal)}

    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    # the device-side clock brackets exactly the K steps; host-side syncs inside finalize make
    # the wall clock agree with it — report the larger of the two
    t_dev = max(ms_dev / 1e3, wall_dev) / args.steps
    t_e2e = max(ms_e2e / 1e3, wall_e2e) / args.steps
    value = R_all / t_dev
    e2e = R_all / t_e2e

    # ---- roofline: every stage, every 8(d) kernel family, and the dominant stage --------------------------
    peak, peak_src = peaks()
    agg = {}
    for s in stages:
        a = agg.setdefault(s["name"], dict(ms=0.0, launches=0, algo_bytes=0, calls=0, units=0))
        a["ms"] += s["ms"]
        a["launches"] += s["launches"]
        a["algo_bytes"] += s["algo_bytes"]
        a["units"] += s["units"]
        a["calls"] += 1
    local_counts = dict(counts)
    if world > 1:  # per-GPU share of the job for the per-GPU stage times
        local_counts = {k: v // world for k, v in counts.items()}
        entries = entries or 0
    fam_bytes = family_bytes(cfg, local_counts, entries)
    fam = {}
    for nm, a in agg.items():
        f = fam.setdefault(FAMILY_OF.get(nm, "other"), dict(ms=0.0, launches=0, stages=[]))
        f["ms"] += a["ms"]
        f["launches"] += a["launches"]
        f["stages"].append(nm)
    kernels_8d = {}
    for k, f in sorted(fam.items()):
        b = fam_bytes.get(k)
        kernels_8d[k] = {"ms": round(f["ms"], 4), "launches": f["launches"], "algo_bytes": b,
                         "frac": (b / (f["ms"] / 1e3) / 1e9 / peak) if b and f["ms"] > 0 else None, "stages": f["stages"]}
    stage_table = {nm: {"ms": round(a["ms"], 4), "launches": a["launches"], "moved_bytes_model": a["algo_bytes"],
                        "frac_of_copy_bandwidth": (a["algo_bytes"] / (a["ms"] / 1e3) / 1e9 / peak) if a["ms"] > 0 else None,
                        "family": FAMILY_OF.get(nm, "other")} for nm, a in agg.items()}
    stage_sum_ms = sum(a["ms"] for a in agg.values())
    moved_sum = sum(a["algo_bytes"] for a in agg.values())
    name, a = max(agg.items(), key=lambda kv: kv[1]["ms"])
    f_name = FAMILY_OF.get(name, "other")
    b8d = fam_bytes.get(f_name) or a["algo_bytes"]
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        try:
            traffic = json.load(open(tp)).get(name)
        except Exception:
            traffic = None
    achieved = b8d / (a["ms"] / 1e3) / 1e9 if a["ms"] > 0 else 0.0
    pass_achieved = a["algo_bytes"] / (a["ms"] / 1e3) / 1e9 if a["ms"] > 0 else 0.0
    roofline = {"bound": "hbm", "kernel": KERNEL_OF.get(name, name), "stage": name, "family": f_name,
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "frac_definition": f"SURVEY 8(d) algorithmic bytes of family {f_name} ({b8d:.0f} B per step) / device time "
                                   f"of stage {name}",
                "family_frac": (kernels_8d.get(f_name) or {}).get("frac"), "family_ms": (kernels_8d.get(f_name) or {}).get("ms"),
                "family_frac_definition": "the same bytes / device time of ALL stages of the family (the stricter reading: "
                                          "the family's bytes are not the dominant stage's alone)",
                "pass_achieved": pass_achieved, "pass_frac": pass_achieved / peak,
                "pass_frac_definition": "bytes the stage's launches have to move by the method used (e.g. one read + one "
                                        "write per radix pass) / the same time",
                "traffic": traffic, "traffic_ratio": (traffic / b8d) if traffic and b8d else None,
                "traffic_source": "ncu dram__bytes_read.sum + dram__bytes_write.sum of the stage's launches, profiles/traffic.json",
                "peak_source": peak_src,
                "kernel_ms": a["ms"] / max(1, a["calls"]), "launches": a["launches"],
                "algo_bytes_per_launch": b8d / max(1, a["launches"]),
                "step": {"stage_sum_ms": round(stage_sum_ms, 4), "moved_bytes_model": moved_sum,
                         "streaming_bound_ms": round(moved_sum / (peak * 1e9) * 1e3, 4),
                         "frac_of_streaming_bound": (moved_sum / (peak * 1e9) * 1e3) / (t_dev * 1e3) if t_dev > 0 else None},
                "kernels_8d": kernels_8d, "stages": stage_table,
                "stages_ms": {k: round(v["ms"], 3) for k, v in agg.items()}}

    # BASELINE.json's second figure, "neighbour-search GB/s", by SURVEY.md §8(d)'s definition: the bytes of the
    # variant-probe model (8 B per word and per upward variant, 296 B per word at n = 24, d = 1) over the
    # time of whatever finds the neighbour lists here (block join: sorts + joins; probe stages if any)
    ns_ms = sum(v["ms"] for k, v in agg.items()
                if k.startswith(("join_", "probe_", "shard_join", "edit_candidates", "edit_replay")))
    n_, d_ = cfg.word_length, cfg.distance
    variants = 1.5 * n_ + (2.25 * n_ * (n_ - 1) / 2.0 if d_ >= 2 else 0.0)
    ns_bytes = 8.0 * (1.0 + variants) * counts["unique"] / max(world, 1)
    neighbour_search = {"ms": round(ns_ms, 3), "model_bytes": ns_bytes, "unit": "GB/s",
                        "value": ns_bytes / (ns_ms / 1e3) / 1e9 if ns_ms > 0 else None,
                        "definition": "SURVEY 8(d): 8 B x (1 + upward variants) x distinct words / neighbour-search time; "
                                      "the block join does not probe, so this is an equivalent rate, not traffic"}

    # ---- CPU baseline beside it (N = 1): the reference hot path on a bounded sample, checked against this library ----
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        v, ms, kind, cores, what, res = cpu_reference_run(args, 1, 0, args.ref_sample)
        _, chars_s, _, _ = sample_chars(args, args.ref_sample)
        with api.Dedup(**kw) as small:
            small.add_batch(chars_s)
            small.finalize()
            same = (small.counts() == {k: res[k] for k in ("total", "usable", "unique", "clusters")}
                    and np.array_equal(small.keep_mask(), res["keep"]) and small.stats_text() == res["stats"])
        cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": kind, "sample": what, "sample_tuples": args.ref_sample,
               "host_cores": os.cpu_count(), "phase_seconds": res["seconds"], "gpu_result_identical_on_sample": bool(same)}
        if not same:
            sys.stderr.write("PARITY FAILURE: GPU result differs from the reference on the CPU-baseline sample\n")

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": t_dev * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u64", "data": "synthetic",
        "config": like_for_like_config(cfg, args),
        "detail": {"tuples_total": R_all, "result": counts, "host_numa_node": numa,
                   "parallelism": "1 GPU" if world == 1 else
                   f"word space range-partitioned over {world} GPUs, one process per GPU; exchanges: {transport}"},
        "e2e": {"value": e2e, "unit": UNIT, "ms_per_step": t_e2e * 1e3,
                "h2d_bytes_per_step": int(R_all * (row_bytes if wire == "packed4" else cfg.word_length)),
                "d2h_bytes_per_step": int(d2h[0]) * world,
                "wire_format": "4 bits per character (humid_gpu_add_batch_packed4)" if wire == "packed4" else "ASCII (humid_gpu_add_batch)",
                "other_wire_format": {"format": "ASCII" if wire == "packed4" else "4 bits per character", "ms_per_step": t_alt * 1e3,
                                      "value": R_all / t_alt,
                                      "h2d_bytes_per_step": int(R_all * (cfg.word_length if wire == "packed4" else row_bytes))},
                "scope": "C ABI with host buffers: word characters in pinned host memory in, keep mask + histograms + "
                         "counts on the host out; compare with the reference arm's `value` (same scope, bounded sample)"},
        "gpu_launches": int(launches_per_step * args.steps),
        "roofline": roofline, "neighbour_search": neighbour_search, "clocks": clocks,
    }
    if cpu:
        line["cpu_baseline"] = cpu
    if parity:
        line["parity"] = parity
    if phases:
        line["detail"]["phases_ms_last_e2e_step"] = phases
    if world == 1 and not args.no_cli:
        dd.close()
        del d_chars
        torch.cuda.empty_cache()
        line["cli_e2e"] = cli_legs("humid_b200", args, device=dev)
    print(json.dumps(line), flush=True)
    ok = (parity is None or parity["equal"]) and (cpu is None or cpu["gpu_result_identical_on_sample"])
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if not ok:
        raise SystemExit(1)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="humid_b200", choices=["humid_b200", "reference"])
    ap.add_argument("--tuples", type=int, default=50_000_000, help="read tuples per GPU")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-cli", action="store_true", help="skip the whole-program legs (cli_e2e)")
    ap.add_argument("--cli-scale", type=float, default=float(os.environ.get("HUMID_BENCH_CLI_SCALE", "1.0")),
                    help="scale the tuple counts of the whole-program legs")
    ap.add_argument("--wire", default="packed4", choices=["packed4", "ascii"],
                    help="host -> device format of the e2e leg: 4 bits per character (default) or ASCII")
    ap.add_argument("--ref-sample", type=int, default=REF_SAMPLE_TUPLES, help="tuples of the CPU arm's bounded sample")
    ap.add_argument("--config", type=int, default=1, help="BASELINE.json configs[i] (0-based); the bench line is configs[1]")
    ap.add_argument("--word-length", type=int, default=0, help="override -n (e.g. 36: all 12 UMI letters of configs[1])")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        if args.warmup < 3:
            sys.stderr.write("note: timing rules ask for >= 3 warm-up steps\n")
        run_gpu_arm(args)


if __name__ == "__main__":
    main()
