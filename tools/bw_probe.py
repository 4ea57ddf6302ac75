#!/usr/bin/env python
"""HBM bandwidth probes on the box (read-only, write-only, copy) with torch kernels, CUDA events, best of 10.
Gives the denominators for write-dominated kernels (COLLECT) beside MEASURED_PEAKS.json's copy figure."""
import torch

dev = torch.device("cuda:0")
n = 1 << 30  # 8 GiB of fp64
x = torch.empty(n, dtype=torch.float64, device=dev)
y = torch.empty(n, dtype=torch.float64, device=dev)


def best(f, reps=10):
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        f()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return min(ts)


x.fill_(1.0)
y.fill_(2.0)
torch.cuda.synchronize()
t = best(lambda: x.fill_(3.0))
print("write-only  fill_        %.1f GB/s" % (n * 8 / t / 1e6))
t = best(lambda: x.zero_())
print("write-only  zero_ (memset) %.1f GB/s" % (n * 8 / t / 1e6))
t = best(lambda: x.sum())
print("read-only   sum          %.1f GB/s" % (n * 8 / t / 1e6))
t = best(lambda: y.copy_(x))
print("copy        copy_ (r+w)  %.1f GB/s" % (2 * n * 8 / t / 1e6))
t = best(lambda: torch.add(x, y, out=y))
print("2r+1w       add          %.1f GB/s" % (3 * n * 8 / t / 1e6))
