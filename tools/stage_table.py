"""Per-stage device times of one configuration: python tools/stage_table.py <config index> <tuples> [steps] [word length]."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from humid_b200 import api
from humid_b200.synth import HashSynth, baseline_config

ci, R = int(sys.argv[1]), int(sys.argv[2])
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
cfg = baseline_config(ci, R, **({"word_length": int(sys.argv[4])} if len(sys.argv) > 4 else {}))
syn = HashSynth(cfg)
hdr, take = syn.layout()
seg_take = ([hdr] if hdr else []) + take
ch = syn.word_chars(0, R, device=torch.device("cuda", 0))
dd = api.Dedup(word_length=cfg.word_length, distance=cfg.distance, seg_take=seg_take, edit=cfg.edit, maximum=cfg.maximum,
               expected_tuples=R)
for _ in range(steps):
    dd.reset()
    dd.add_batch_device(ch.data_ptr(), R)
    dd.finalize()
print(cfg.name, dd.counts())
tot = 0.0
for st in dd.stage_stats():
    tot += st["ms"]
    gbs = st["algo_bytes"] / st["ms"] / 1e6 if st["ms"] > 0 else 0
    print(f"{st['name']:26s} {st['ms']:9.3f} ms  units {st['units']:>12d}  launches {st['launches']:3d}  {gbs:8.0f} GB/s (algorithmic)")
print(f"{'sum':26s} {tot:9.3f} ms")
