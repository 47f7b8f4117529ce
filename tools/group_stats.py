"""Group-size statistics of the Hamming block join on the bench workload (diagnostic)."""
import sys

sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))))
import numpy as np
import torch
from humid_b200 import api
from humid_b200.synth import HashSynth, baseline_config

ci = int(sys.argv[1]) if len(sys.argv) > 1 else 1
R = int(sys.argv[2]) if len(sys.argv) > 2 else 50_000_000
cfg = baseline_config(ci, R)
syn = HashSynth(cfg)
hdr, take = syn.layout()
seg_take = ([hdr] if hdr else []) + take
dev = torch.device("cuda", 0)
ch = syn.word_chars(0, R, device=dev)
dd = api.Dedup(word_length=cfg.word_length, distance=cfg.distance, seg_take=seg_take, edit=cfg.edit, maximum=cfg.maximum,
               expected_tuples=R)
dd.add_batch_device(ch.data_ptr(), R)
dd.finalize()
keys = dd.words()[0]
n, d = cfg.word_length, cfg.distance
print("U", keys.size, "pairs", dd.neighbours()[1].size // 2)
for b in range(d + 1):
    p0, p1 = b * n // (d + 1), (b + 1) * n // (d + 1)
    lo = 2 * (n - p1)
    blk = (keys >> np.uint64(lo)) & np.uint64((1 << (2 * (p1 - p0))) - 1)
    _, cnt = np.unique(blk, return_counts=True)
    c = cnt.astype(np.float64)
    print(f"block {b}: letters [{p0},{p1}) groups {cnt.size} mean {c.mean():.2f} max {cnt.max()} "
          f"sum s^2/U {float((c * c).sum()) / keys.size:.1f} words in groups >64: {int(cnt[cnt > 64].sum())} >2048: {int(cnt[cnt > 2048].sum())}")
for st in dd.stage_stats():
    print(st["name"], round(st["ms"], 3))
