"""Pinned host -> device copy bandwidth of this box (what bounds the end-to-end step)."""
import time
import torch

n = 1_200_000_000
h = torch.empty(n, dtype=torch.uint8, pin_memory=True)
h.fill_(65)
d = torch.empty(n, dtype=torch.uint8, device="cuda")
for chunk in (n, 192_000_000, 48_000_000):
    for _ in range(2):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for a in range(0, n, chunk):
            d[a:a + chunk].copy_(h[a:a + chunk], non_blocking=True)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
    print(f"chunk {chunk / 1e6:7.0f} MB: {n / dt / 1e9:6.1f} GB/s ({dt * 1e3:.1f} ms for 1.2 GB)")
