#!/usr/bin/env python
"""HBM floor of the conv layers: pure-write, pure-read and copy bandwidth (torch kernels, CUDA events)."""
import torch

n = 1 << 29                     # 2 GiB of fp32
a = torch.empty(n, dtype=torch.float32, device="cuda")
b = torch.empty(n, dtype=torch.float32, device="cuda")


def timed(fn, reps=10):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e-3


gb = n * 4 / 1e9
print(f"write only (fill_)   {gb / timed(lambda: a.fill_(1.0)):7.0f} GB/s")
print(f"write only (memset)  {gb / timed(lambda: a.zero_()):7.0f} GB/s")
print(f"read only (sum)      {gb / timed(lambda: a.sum()):7.0f} GB/s")
print(f"copy (read + write)  {2 * gb / timed(lambda: b.copy_(a)):7.0f} GB/s")
