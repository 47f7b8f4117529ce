#!/usr/bin/env python
"""Benchmark of the deduplication hot path (BASELINE.json: read pairs deduplicated / s).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--tuples R]

One "step" = one complete pass of the hot path over one batch of synthetic read tuples:
word construction (pack) -> distinct-word table with counts -> Hamming neighbour search ->
clustering -> keep mask + the three histograms.  The workload is BASELINE.json configs[1]
(50 M read pairs, 12 bp UMI in a separate FastQ, word length 24 = 8/8/8 nt, Hamming 1,
directional clustering), `--tuples` read tuples PER GPU (weak scaling: N GPUs process
N x that many tuples of one common library; the word space is range-partitioned over the
GPUs, humid_b200/sharded.py).

  value  whole-job read pairs/s with the word characters already resident in HBM
  e2e    the same through the C ABI with HOST buffers: per step the H2D copy of the word
         characters (pinned memory) and the D2H read of the keep mask, the histograms and
         the stats scalars are inside the timed region
  roofline      the stage with the largest device time against the measured HBM copy bandwidth:
                algorithmic bytes of its launches / CUDA-event time of the stage
  cpu_baseline  the reference `humid` binary (oracle/_ref/humid, built from the reference's
                own sources) — or the C restatement when that binary is absent — timed on a
                bounded sample of the same workload on the host cores

`--impl reference` times only that CPU arm and prints its own line.
"""
from __future__ import annotations

import argparse
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


# stage (humid_gpu_stage_stats) -> the kernel that does its work
KERNEL_OF = {
    "sort_words": "radix_pass_kernel<true, ShiftDigit> (1 histogram + 6 digit passes)",
    "join_sort": "radix_pass_kernel<true, ShiftDigit> (1 histogram + 3 digit passes)",
    "table_partition": "radix_pass_kernel<true, SlotDigit>", "table_insert": "table_insert_kernel",
    "pack": "pack_words_kernel", "rank_columns": "gather_columns_kernel", "join_hamming_b0": "hamming_join_kernel",
    "join_hamming_b1": "hamming_join_kernel", "join_hamming_b2": "hamming_join_kernel", "csr_build": "csr_fill_kernel",
    "union_find": "uf_union_kernel", "replay_directional": "replay_kernel<false>", "replay_maximum": "replay_kernel<true>",
}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


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
# CPU arm: the reference binary on a bounded sample
# ----------------------------------------------------------------------------------------
def cpu_reference_run(steps: int, warmup: int, sample: int, word_length: int = 0):
    """Returns (pairs/s, ms per step, kind, cores, sample description)."""
    import oracle
    from humid_b200.synth import HashSynth, baseline_config

    cfg = baseline_config(1, sample, **({"word_length": word_length} if word_length else {}))
    syn = HashSynth(cfg)
    if os.access(oracle.REF_HUMID, os.X_OK):
        tmp = tempfile.mkdtemp(prefix="humid_ref_")
        try:
            files = syn.write_fastq(tmp, prefix="s")
            times = []
            for i in range(warmup + steps):
                out = os.path.join(tmp, f"out{i}")
                t = oracle.run_ref_humid(files, out, cfg.cli_flags(), stats=True)["wall_s"]
                shutil.rmtree(out, ignore_errors=True)
                if i >= warmup:
                    times.append(t)
        finally:
            shutil.rmtree(tmp, ignore_errors=True)
        kind = "reference"
        what = (f"{sample} read tuples of the same library model (3 FastQ files, 2x150 nt + 12 nt UMI), whole "
                f"`humid -n {cfg.word_length} -m 1 -s` process wall time incl. FastQ parsing and writing the dedup files")
    else:
        import numpy as np

        chars = syn.word_chars(0, sample).numpy()
        hdr, take = syn.layout()
        seg = ([hdr] if hdr else []) + take
        times = []
        for i in range(warmup + steps):
            t0 = time.monotonic()
            keys = oracle.pack_words(chars, seg)
            oracle.Result(keys, cfg.word_length, cfg.distance, cfg.edit, cfg.maximum)
            if i >= warmup:
                times.append(time.monotonic() - t0)
        kind = "port"
        what = f"{sample} read tuples, C restatement (oracle/humid_oracle.c) from word characters in memory"
    t = sum(times) / len(times)
    return sample / t, t * 1e3, kind, 1, what


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    val, ms, kind, cores, what = cpu_reference_run(args.steps, args.warmup, REF_SAMPLE_TUPLES, args.word_length)
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample_tuples": REF_SAMPLE_TUPLES, "host_cores": os.cpu_count()},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": kind, "sample": what},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
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

    if world > 1:
        from humid_b200 import sharded
        dd = sharded.ShardedDedup(word_length=cfg.word_length, distance=cfg.distance, seg_take=seg_take, edit=cfg.edit,
                                  maximum=cfg.maximum, device=local, expected_tuples=R_per)
    else:
        dd = api.Dedup(word_length=cfg.word_length, distance=cfg.distance, seg_take=seg_take, edit=cfg.edit,
                       maximum=cfg.maximum, device=local, expected_tuples=R_per)
    ext = torch.cuda.ExternalStream(dd.stream, device=dev)
    BATCH = 8_000_000

    def step_device():
        dd.reset()
        dd.add_batch_device(d_chars.data_ptr(), R_per)
        dd.finalize()
        return dd.counts()

    d2h = [0]

    def step_e2e():
        dd.reset()
        for a in range(0, R_per, BATCH):
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
    for _ in range(max(1, min(args.warmup, 2))):
        step_e2e()
    ms_e2e, wall_e2e, (counts_e2e, _) = timed(step_e2e, args.steps)
    clocks = sampler.stop() if sampler else None
    assert counts == counts_e2e, (counts, counts_e2e)
    if world > 1 and rank == 0 and getattr(dd, "phases", None):
        sys.stderr.write("phases (ms, last step, rank 0): " + json.dumps([(n, round(t, 3)) for n, t in dd.phases]) + "\n")

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # the device-side clock brackets exactly the K steps; host-side syncs inside finalize make
    # the wall clock agree with it — report the larger of the two
    t_dev = max(ms_dev / 1e3, wall_dev) / args.steps
    t_e2e = max(ms_e2e / 1e3, wall_e2e) / args.steps
    value = R_all / t_dev
    e2e = R_all / t_e2e

    # dominant kernel of the step
    peak, peak_src = peaks()
    agg = {}
    for s in stages:
        a = agg.setdefault(s["name"], dict(ms=0.0, launches=0, algo_bytes=0, calls=0, units=0))
        a["ms"] += s["ms"]
        a["launches"] += s["launches"]
        a["algo_bytes"] += s["algo_bytes"]
        a["units"] += s["units"]
        a["calls"] += 1
    top = max(agg.items(), key=lambda kv: kv[1]["ms"])
    name, a = top
    achieved = a["algo_bytes"] / (a["ms"] / 1e3) / 1e9 if a["ms"] > 0 else 0.0
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        try:
            traffic = json.load(open(tp)).get(name)
        except Exception:
            traffic = None
    roofline = {"bound": "hbm", "kernel": KERNEL_OF.get(name, name), "stage": name, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                "kernel_ms": a["ms"] / max(1, a["calls"]), "launches": a["launches"],
                "algo_bytes_per_launch": a["algo_bytes"] / max(1, a["calls"]),
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

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        v, ms, kind, cores, what = cpu_reference_run(1, 0, REF_SAMPLE_TUPLES, args.word_length)
        cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": kind, "sample": what, "host_cores": os.cpu_count()}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": t_dev * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u64", "data": "synthetic",
        "config": {"workload": WORKLOAD if args.config == 1 and not args.word_length else cfg.name, "tuples_per_gpu": R_per, "tuples_total": R_all, "word_length": cfg.word_length,
                   "distance": cfg.distance, "clustering": "maximum" if cfg.maximum else "directional",
                   "l2": "inputs larger than L2 (word characters %.2f GB per GPU)" % (R_per * cfg.word_length / 1e9),
                   "parallelism": "1 GPU" if world == 1 else f"word space range-partitioned over {world} GPUs; NCCL all-to-all-v (tuples, variant queries), all-gather (components that cross shards)",
                   "host_numa_node": numa, "result": counts},
        "e2e": {"value": e2e, "unit": UNIT, "ms_per_step": t_e2e * 1e3,
                "h2d_bytes_per_step": int(R_all * cfg.word_length), "d2h_bytes_per_step": int(d2h[0]) * world},
        "gpu_launches": int(launches_per_step * args.steps),
        "roofline": roofline, "neighbour_search": neighbour_search, "clocks": clocks,
    }
    if cpu:
        line["cpu_baseline"] = cpu
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="humid_b200", choices=["humid_b200", "reference"])
    ap.add_argument("--tuples", type=int, default=50_000_000, help="read tuples per GPU")
    ap.add_argument("--no-cpu-baseline", action="store_true")
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
