#!/usr/bin/env python
"""PCIe floor of the end-to-end leg: pinned H2D of one cfg2 step's input (1.97 GB), pinned D2H of its output (1.57 GB),
alone and concurrently on two streams.  Prints ms and GB/s; bench.py's e2e.ms_per_step is to be read against the
concurrent figure."""
import time
import torch

B, L, M, T = 1024, 480000, 128, 3000
xh = torch.empty((B, L), dtype=torch.float32, pin_memory=True).normal_()
oh = torch.empty((B, M, T), dtype=torch.float32, pin_memory=True)
xd = torch.empty((B, L), dtype=torch.float32, device="cuda")
od = torch.zeros((B, M, T), dtype=torch.float32, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def run(h2d, d2h, reps=5):
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        t0 = time.perf_counter()
        if h2d:
            with torch.cuda.stream(s1):
                xd.copy_(xh, non_blocking=True)
        if d2h:
            with torch.cuda.stream(s2):
                oh.copy_(od, non_blocking=True)
        torch.cuda.synchronize()
        best = min(best, time.perf_counter() - t0)
    return best * 1e3


for name, a, b in (("h2d", 1, 0), ("d2h", 0, 1), ("both", 1, 1)):
    ms = run(a, b)
    gb = (a * xh.numel() + b * oh.numel()) * 4 / 1e9
    print(f"{name:5s} {ms:7.2f} ms  {gb / ms * 1e3:6.1f} GB/s")
