#!/usr/bin/env python
"""One device-resident launch set of a workload, for ncu captures (numbers printed under ncu are never bench values).
  python tools/prof_run.py <workload> <cases> [team] [reps] [mode 0 auto|1 fused|2 level] [max_resident]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402
from fastbn_b200 import capi, synth  # noqa: E402

w = sys.argv[1]
n = int(sys.argv[2])
team = int(sys.argv[3]) if len(sys.argv) > 3 else 0
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 2
tree = bench.load_tree(w)
mode = int(sys.argv[5]) if len(sys.argv) > 5 else 0
max_res = int(sys.argv[6]) if len(sys.argv) > 6 else 0
plan = capi.Plan(tree, device=0, team_warps=team, mode=mode, max_resident=max_res)
dev = torch.device("cuda:0")
off, ev, ew = synth.uniform_evidence(tree["var_dims"], n, bench.WORKLOADS[w]["n_evidence"], seed=1)
off_d, ev_d, ew_d = [torch.from_numpy(x).to(dev) for x in (off, ev, ew)]
lab = torch.empty(n, dtype=torch.int32, device=dev)
post = torch.empty((n, plan.sum_dims), dtype=torch.float64, device=dev)
for _ in range(reps):
    st = plan.run_device(n, off_d.data_ptr(), ev_d.data_ptr(), ew_d.data_ptr(), lab.data_ptr(), post.data_ptr())
torch.cuda.synchronize()
i = plan.info
print("workload %s cases %d mode %d team %d resident %d launches %d kernel_ms %.3f  -> %.0f cases/s, alg %.1f GB/s, moved %.1f GB/s" % (
    w, n, i.mode, i.team_warps, i.resident_tiles, st.n_launches, st.kernel_ms, n / st.kernel_ms * 1e3, i.alg_bytes_per_case * n / st.kernel_ms / 1e6,
    i.kernel_bytes_per_case * n / st.kernel_ms / 1e6))
plan.close()
