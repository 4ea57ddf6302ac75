#!/usr/bin/env python
"""e2e (pinned host buffers through fbn_run) throughput of ALARM 1M cases for several pipeline chunk sizes (FBN_CHUNK_MB),
plus a raw pinned D2H/H2D copy bandwidth probe of the box for reference."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

import bench  # noqa: E402
from fastbn_b200 import capi, synth  # noqa: E402

n = 1_000_000
tree = bench.load_tree("alarm")
off, ev, ew = synth.uniform_evidence(tree["var_dims"], n, 7, seed=1)
off_p, ev_p, ew_p = [torch.from_numpy(x).pin_memory() for x in (off, ev, ew)]
plan = capi.Plan(tree, device=0)
sd = plan.sum_dims
lab = torch.empty(n, dtype=torch.int32).pin_memory()
post = torch.empty((n, sd), dtype=torch.float64).pin_memory()
d = torch.empty((n, sd), dtype=torch.float64, device="cuda")
for _ in range(2):
    post.copy_(d, non_blocking=True)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(5):
    post.copy_(d, non_blocking=True)
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / 5
print("raw pinned D2H of %d MB: %.2f ms = %.1f GB/s" % (post.numel() * 8 / 1e6, dt * 1e3, post.numel() * 8 / dt / 1e9))
for mb in sys.argv[1:] or ["128"]:
    if mb == "default":
        os.environ.pop("FBN_CHUNK_MB", None)
    else:
        os.environ["FBN_CHUNK_MB"] = mb
    for _ in range(2):
        plan.run(off_p.numpy(), ev_p.numpy(), ew_p.numpy(), labels_out=lab.numpy(), post_out=post.numpy())
    t0 = time.perf_counter()
    for _ in range(5):
        _, _, st = plan.run(off_p.numpy(), ev_p.numpy(), ew_p.numpy(), labels_out=lab.numpy(), post_out=post.numpy())
    dt = (time.perf_counter() - t0) / 5
    print("chunk %s MB: %.2f ms/step -> %.1f M cases/s (launches %d)" % (mb, dt * 1e3, n / dt / 1e6, st.n_launches))
plan.close()
