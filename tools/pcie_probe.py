#!/usr/bin/env python
"""PCIe ceiling of the box for the inference e2e leg: pinned H2D, D2H and both at once (8 x 24.9 MB)."""
import time
import torch

n = 8 * 1920 * 1080 * 3
h_in = torch.empty(n, dtype=torch.float32).pin_memory()
h_out = torch.empty(n, dtype=torch.float32).pin_memory()
d_in = torch.empty(n, dtype=torch.float32, device="cuda")
d_out = torch.empty(n, dtype=torch.float32, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def run(up, down, reps=10):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        if up:
            with torch.cuda.stream(s1):
                d_in.copy_(h_in, non_blocking=True)
        if down:
            with torch.cuda.stream(s2):
                h_out.copy_(d_out, non_blocking=True)
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps


for _ in range(2):
    run(True, True, 2)
gb = n * 4 / 1e9
print(f"H2D alone  {gb / run(True, False):6.1f} GB/s")
print(f"D2H alone  {gb / run(False, True):6.1f} GB/s")
t = run(True, True)
print(f"both       {gb / t:6.1f} GB/s each way, {t * 1e3:.2f} ms per 8 images -> ceiling {8 / t:.0f} images/s")
