"""One job of `--tuples` read tuples (BASELINE.json configs[--config]) on however many GPUs the launcher gives
(plain python: 1 GPU through api.Dedup; torchrun: one rank per GPU through sharded.ShardedDedup).  Prints one JSON
line with the counts, the three statistics files' digests and a digest of the keep mask: the lines of a 1-GPU and
an N-GPU run of the same job must be identical (the 1-GPU path is the parity reference)."""
import argparse
import hashlib
import json
import os
import time

import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

ap = argparse.ArgumentParser()
ap.add_argument("--config", type=int, default=4)
ap.add_argument("--tuples", type=int, default=500_000_000)
args = ap.parse_args()

from humid_b200 import api, sharded  # noqa: E402
from humid_b200.synth import HashSynth, baseline_config  # noqa: E402

rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    torch.distributed.init_process_group("nccl", device_id=dev)
cfg = baseline_config(args.config, args.tuples)
syn = HashSynth(cfg)
hdr, take = syn.layout()
seg_take = ([hdr] if hdr else []) + take
lo, hi = args.tuples * rank // world, args.tuples * (rank + 1) // world
kw = dict(word_length=cfg.word_length, distance=cfg.distance, seg_take=seg_take, edit=cfg.edit, maximum=cfg.maximum,
          device=local, expected_tuples=hi - lo)
dd = sharded.make_sharded(**kw) if world > 1 else api.Dedup(**kw)  # peer-memory group (HUMID_SHARD_TRANSPORT=nccl: the NCCL path)
CH = 50_000_000
t_gen = 0.0
torch.cuda.synchronize()
t0 = time.perf_counter()
for a in range(lo, hi, CH):
    b = min(a + CH, hi)
    g0 = time.perf_counter()
    ch = syn.word_chars(a, b, device=dev)
    torch.cuda.synchronize()
    t_gen += time.perf_counter() - g0
    dd.add_batch_device(ch.data_ptr(), b - a)
    del ch
dd.finalize()
torch.cuda.synchronize()
t_job = time.perf_counter() - t0 - t_gen
counts = dd.counts()
stats = dd.stats_text()
if world > 1:
    keep = dd.keep_mask()
    h = hashlib.sha256(np.packbits(keep, bitorder="little").tobytes()).hexdigest() if (hi - lo) % 8 == 0 else None
    gathered = [None] * world
    torch.distributed.all_gather_object(gathered, (int(keep.sum()), h))
    kept = sum(g[0] for g in gathered)
    digest = hashlib.sha256("".join(g[1] or "" for g in gathered).encode()).hexdigest()
else:
    keep = dd.keep_mask()
    kept = int(keep.sum())
    parts = []
    for r in range(int(os.environ.get("SPLIT", "8"))):
        s0, s1 = args.tuples * r // int(os.environ.get("SPLIT", "8")), args.tuples * (r + 1) // int(os.environ.get("SPLIT", "8"))
        parts.append(hashlib.sha256(np.packbits(keep[s0:s1], bitorder="little").tobytes()).hexdigest())
    digest = hashlib.sha256("".join(parts).encode()).hexdigest()
if rank == 0:
    print(json.dumps({"config": cfg.name, "tuples": args.tuples, "gpus": world, "counts": counts, "kept": kept,
                      "keep_digest_8_slices": digest,
                      "stats_digest": {k: hashlib.sha256(v.encode()).hexdigest()[:16] for k, v in stats.items()},
                      "job_seconds_excl_generation": round(t_job, 3)}), flush=True)
if world > 1:
    torch.distributed.destroy_process_group()
