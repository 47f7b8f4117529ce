"""Worker of tests/test_gpu_multiproc.py: one rank of a real multi-process run (torchrun, NCCL, one GPU per rank).
Every rank deduplicates its slice of a small job through humid_b200.sharded; rank 0 compares the whole-job result
with the CPU oracle (test infrastructure) and writes `ok` / the mismatch to the file named by argv[1]."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import oracle  # noqa: E402
from humid_b200 import sharded  # noqa: E402
from humid_b200.synth import Synth, baseline_config  # noqa: E402

out_path = sys.argv[1]
cases = json.loads(sys.argv[2])
rank, world, local = (int(os.environ[k]) for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"))
# HUMID_TEST_ONE_DEVICE=1: all ranks on device 0 (CUDA IPC between processes works on one device too; NCCL does not
# take two ranks on one GPU, so the rendezvous is gloo and only the library's own peer-memory transport can run)
one_device = os.environ.get("HUMID_TEST_ONE_DEVICE") == "1"
if one_device:
    local = 0
torch.cuda.set_device(local)
if one_device:
    dist.init_process_group("gloo")
else:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
problems = []
for ci, R, kw in cases:
    cfg = baseline_config(ci, R, **kw)
    s = Synth(cfg)
    chars = s.word_chars()
    hdr, take = s.layout()
    seg_take = ([hdr] if hdr else []) + take
    lo, hi = R * rank // world, R * (rank + 1) // world
    dd = sharded.make_sharded(word_length=cfg.word_length, distance=cfg.distance, seg_take=seg_take, edit=cfg.edit,
                              maximum=cfg.maximum, device=local)
    for step in range(2):  # twice through one handle: reset must leave nothing behind
        if one_device:
            assert isinstance(dd, sharded.GroupDedup), "the peer-memory group could not be formed"
        dd.reset()
        half = (hi - lo) // 2
        dd.add_batch(chars[lo:lo + half])
        dd.add_batch(chars[lo + half:hi])
        dd.finalize()
        counts = dd.counts()
        keep = dd.keep_mask()
        ids = dd.cluster_ids()
        stats = dd.stats_text()
        ref = oracle.Result(oracle.pack_words(chars, seg_take), cfg.word_length, cfg.distance, cfg.edit, cfg.maximum)
        tag = f"{cfg.name} R={R} rank {rank} step {step}"
        if (counts["total"], counts["usable"], counts["unique"], counts["clusters"]) != (ref.total, ref.usable, ref.unique, ref.clusters):
            problems.append(f"{tag}: counts {counts}")
        if not np.array_equal(keep, ref.keep[lo:hi].astype(bool)):
            problems.append(f"{tag}: keep mask")
        if not np.array_equal(ids, ref.cid[lo:hi]):
            problems.append(f"{tag}: cluster ids")
        if stats != ref.stats_text():
            problems.append(f"{tag}: statistics files")
    dd.close()
gathered = [None] * world
dist.all_gather_object(gathered, problems)
if rank == 0:
    allp = [p for g in gathered for p in g]
    with open(out_path, "w") as f:
        f.write("ok\n" if not allp else "\n".join(allp) + "\n")
dist.barrier()
dist.destroy_process_group()
