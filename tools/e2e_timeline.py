"""Where the end-to-end step goes: python tools/e2e_timeline.py [tuples] [word length].
Host clock around the calls of one step (configs[1] shape, pinned host buffer), medians over 5 steps."""
import os
import statistics
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from humid_b200 import api
from humid_b200.synth import HashSynth, baseline_config

R = int(sys.argv[1]) if len(sys.argv) > 1 else 50_000_000
kw = {"word_length": int(sys.argv[2])} if len(sys.argv) > 2 else {}
cfg = baseline_config(1, R, **kw)
syn = HashSynth(cfg)
hdr, take = syn.layout()
seg_take = ([hdr] if hdr else []) + take
d_chars = syn.word_chars(0, R, device=torch.device("cuda", 0))
h_chars = torch.empty(d_chars.shape, dtype=torch.uint8, pin_memory=True)
h_chars.copy_(d_chars)
torch.cuda.synchronize()
h_np = h_chars.numpy()
dd = api.Dedup(word_length=cfg.word_length, distance=cfg.distance, seg_take=seg_take, edit=cfg.edit, maximum=cfg.maximum,
               expected_tuples=R)
BATCH = 8_000_000
rows = []
for it in range(8):
    t = [time.perf_counter()]
    dd.reset()
    t.append(time.perf_counter())
    for a in range(0, R, BATCH):
        dd.add_batch(h_np[a:a + BATCH])
        t.append(time.perf_counter())
    dd.finalize()
    t.append(time.perf_counter())
    dd.keep_bits()
    t.append(time.perf_counter())
    if it >= 3:
        rows.append([(b - a) * 1e3 for a, b in zip(t, t[1:])])
med = [statistics.median(c) for c in zip(*rows)]
nb = (R + BATCH - 1) // BATCH
print("reset %.2f ms" % med[0])
print("add_batch", " ".join("%.2f" % x for x in med[1:1 + nb]), "ms (sum %.2f)" % sum(med[1:1 + nb]))
print("finalize %.2f ms" % med[1 + nb])
print("keep_bits %.2f ms" % med[2 + nb])
print("total %.2f ms; upload at the rate of the add_batch calls: %.1f GB/s" %
      (sum(med), R * cfg.word_length / (sum(med[1:1 + nb]) / 1e3) / 1e9))
tail = {s["name"]: s["ms"] for s in dd.stage_stats()}
print("device stages of the last step (ms):", " ".join(f"{k}={v:.2f}" for k, v in tail.items()))
