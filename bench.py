#!/usr/bin/env python
"""Benchmark of the junction-tree hot path (BASELINE.json metric: test cases/sec of exact JT inference; % of HBM roofline).

  python bench.py --gpus N --steps K --warmup W            our arm (one process per GPU; torchrun for N > 1)
  python bench.py --impl reference --gpus N --steps K ...  the reference's own CPU implementation on the host cores

A "step" = one pass of the hot path over one batch of synthetic test cases (config[1]: ALARM, 1M cases, 20% evidence,
i.e. 7 of 37 variables; weak scaling: every GPU gets its own 1M-case slice, no data-path collective).
`value`  : whole-job cases/s with evidence already resident in HBM and results left in HBM (CUDA events).
`e2e`    : same metric through fbn_run with pinned HOST buffers, H2D/D2H copies inside the timed region.
One JSON line is printed by rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")

WORKLOADS = {
    # name: tree source, #evidence vars, default cases per GPU per step
    "alarm": dict(ref_cases=60000, tree="alarm_tree_ref.npz", xml=os.path.join("alarm", "alarm.xml"), n_evidence=7, cases=1_000_000,
                  desc="ALARM (37 nodes), 1M synthetic test cases, 20% evidence (7 vars), batched"),
    # BASELINE.json configs 3-5 (parity-test shapes; benchable with --workload): trees built by the C++ host mirror
    "hepar2": dict(ref_cases=1200, shape="hepar2", n_evidence=14, cases=1_000_000, desc="synthetic HEPAR2-shape BN (70 nodes, card 2-4), 20% evidence"),
    "hepar2_50": dict(ref_cases=2400, shape="hepar2", n_evidence=35, cases=1_000_000, desc="synthetic HEPAR2-shape BN (70 nodes, card 2-4), 50% evidence"),
    "pigs": dict(ref_cases=200, shape="pigs", n_evidence=88, cases=65_536, desc="synthetic PIGS-shape BN (441 nodes, card 3, max clique 3^10), 20% evidence"),
    "munin_small": dict(ref_cases=8, shape="munin_small", n_evidence=0, cases=2048, desc="synthetic MUNIN-shape stand-in (max clique 5^9), 0% evidence"),
    "munin": dict(ref_cases=1, shape="munin", n_evidence=0, cases=512, desc="synthetic MUNIN-shape BN (max clique 5^10 ~ 1e7), 0% evidence"),
}


def workload_xml(w):
    """Path of the workload's XMLBIF (committed for alarm/hepar2/pigs, generated from the recorded seed otherwise)."""
    from fastbn_b200 import synth
    cfg = WORKLOADS[w]
    if "xml" in cfg:
        return os.path.join(GOLDEN, cfg["xml"])
    p = os.path.join(GOLDEN, "synth", cfg["shape"] + ".xml")
    if not os.path.exists(p):
        os.makedirs(os.path.join(ROOT, "build", "synth"), exist_ok=True)
        p = os.path.join(ROOT, "build", "synth", "%s_%d.xml" % (cfg["shape"], os.getpid()))
        synth.make_shape(cfg["shape"], p)
    return p


def load_tree(w):
    cfg = WORKLOADS[w]
    if "tree" in cfg:
        z = np.load(os.path.join(GOLDEN, cfg["tree"]))
        return {k: z[k] for k in z.files}
    from fastbn_b200 import hostapi
    m = hostapi.HostModel(workload_xml(w))
    t = m.tree()
    m.close()
    return t


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, gpu):
        super().__init__(daemon=True)
        self.gpu = gpu
        self.rows = []
        self.stop_flag = False

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        sm = sorted(int(r[0]) for r in self.rows if r[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in self.rows)]
        mx = [int(r[1]) for r in self.rows if r[1].isdigit()]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.rows)}


# ---------------------------------------------------------------------------------------------------------------
# reference arm / cpu_baseline: the UNMODIFIED reference (oracle/_ref) on the host cores, one independent `-t 1`
# instance per core on its own slice of the cases (SURVEY.md 8(d) "Baseline B"; the reference's own -t N only
# parallelises inside one case and is slower than -t 1 on ALARM-size trees).
def _ref_worker(args):
    xml, cases_file, off, ev, ew, kind, tree = args
    if kind == "reference":
        from oracle import refdrv
        r = refdrv.Reference(xml, cases_file)
        _, _, secs = r.run_batch(off, ev, ew, want_post=True)
        r.close()
        return secs
    from oracle import jtoracle
    o = jtoracle.Oracle(tree)
    t0 = time.perf_counter()
    o.run_batch(off, ev, ew)
    secs = time.perf_counter() - t0
    o.close()
    return secs


def cpu_reference_throughput(workload, cases_per_core, cores, seed):
    """Returns (cases/s over all cores, kind, sample description)."""
    import multiprocessing as mp
    from fastbn_b200 import synth
    from oracle import refdrv
    w = WORKLOADS[workload]
    tree = load_tree(workload)
    kind = "reference" if refdrv.available() else "port"
    xml = workload_xml(workload)
    cases_file = os.path.join(ROOT, "build", "ref_cases_%d" % os.getpid())  # the reference's ctor wants a LIBSVM file
    os.makedirs(os.path.dirname(cases_file), exist_ok=True)
    synth.write_libsvm(cases_file, *synth.uniform_evidence(tree["var_dims"], 20, 0, seed=1))
    jobs = []
    for c in range(cores):
        off, ev, ew = synth.uniform_evidence(tree["var_dims"], cases_per_core, w["n_evidence"], seed=seed + c)
        jobs.append((xml, cases_file, off, ev, ew, kind, tree))
    ctx = mp.get_context("fork")
    t0 = time.perf_counter()
    with ctx.Pool(cores) as pool:
        secs = pool.map(_ref_worker, jobs)
    wall = time.perf_counter() - t0
    # throughput = total cases / slowest worker's time inside the per-case loop (construction/parsing excluded, as for the GPU arm)
    val = cases_per_core * cores / max(secs)
    return val, kind, "%d cases/core x %d cores of the same synthetic workload (max worker loop %.2f s, wall %.2f s)" % (
        cases_per_core, cores, max(secs), wall)


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    w = WORKLOADS[args.workload]
    # bounded sample per step: a quarter of the cpu_baseline sample (a few seconds of CPU work per step)
    per_core = args.ref_cases_per_core or max(1, w["ref_cases"] // 4)
    for _ in range(min(args.warmup, 1)):
        cpu_reference_throughput(args.workload, max(1, per_core // 10), cores, 777)
    vals = []
    t0 = time.perf_counter()
    for s in range(args.steps):
        v, kind, sample = cpu_reference_throughput(args.workload, per_core, cores, 1000 + s)
        vals.append(v)
    wall = time.perf_counter() - t0
    value = float(np.mean(vals))
    line = {"impl": "reference", "metric": "test cases/sec, exact JT inference", "value": value, "unit": "cases/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": wall / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": w["desc"], "cases_per_step": per_core * cores, "n_evidence": w["n_evidence"]},
            "cpu_baseline": {"value": value, "unit": "cases/s", "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": "cases/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="alarm", choices=sorted(WORKLOADS))
    ap.add_argument("--cases", type=int, default=0, help="cases per GPU per step (default: the workload's)")
    ap.add_argument("--team", type=int, default=0)
    ap.add_argument("--mode", type=int, default=0, help="0 auto, 1 fused kernel, 2 level kernels")
    ap.add_argument("--ref-cases-per-core", type=int, default=0, help="CPU sample per core (default: the workload's, ~10-20 s)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--e2e-steps", type=int, default=0, help="timed e2e steps (default: min(steps, 5))")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)

    import torch
    import torch.distributed as dist
    from fastbn_b200 import capi, synth

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    w = WORKLOADS[args.workload]
    tree = load_tree(args.workload)
    plan = capi.Plan(tree, device=local, team_warps=args.team, mode=args.mode)
    info = plan.info
    sd = info.sum_dims
    n = args.cases or w["cases"]
    if not args.cases and args.workload != "alarm":
        wave = info.resident_tiles * 32  # whole waves of resident tiles (the ALARM config is fixed at 1M cases)
        n = max(wave, n // wave * wave)
    waves = -(-((n + 31) // 32) // info.resident_tiles)
    launches_per_step = info.launches_per_wave * waves if info.mode == capi.MODE_LEVEL else 1
    # every rank gets its own slice of the synthetic case stream (weak scaling, no data-path collective)
    off_h, ev_h, ew_h = synth.uniform_evidence(tree["var_dims"], n, w["n_evidence"], seed=20260101 + rank)

    # ---- device-resident arm -------------------------------------------------------------------------------------
    off_d = torch.from_numpy(off_h).to(dev)
    ev_d = torch.from_numpy(ev_h).to(dev)
    ew_d = torch.from_numpy(ew_h).to(dev)
    labels_d = torch.empty(n, dtype=torch.int32, device=dev)
    post_d = torch.empty((n, sd), dtype=torch.float64, device=dev)
    stream = torch.cuda.current_stream().cuda_stream

    def step_device():
        plan.run_device(n, off_d.data_ptr(), ev_d.data_ptr(), ew_d.data_ptr(), labels_d.data_ptr(), post_d.data_ptr(),
                        stream=stream, timed=False)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step_device()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    e0 = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    e1 = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    barrier()
    t_all0 = torch.cuda.Event(enable_timing=True)
    t_all1 = torch.cuda.Event(enable_timing=True)
    t_all0.record()
    for s in range(args.steps):
        e0[s].record()
        step_device()
        e1[s].record()
    t_all1.record()
    barrier()
    total_ms = t_all0.elapsed_time(t_all1)
    kern_ms = float(np.mean([a.elapsed_time(b) for a, b in zip(e0, e1)]))  # all kernel launches of one step
    if world > 1:
        t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t.item())
    ms_per_step = total_ms / args.steps
    value = world * n / (ms_per_step * 1e-3)

    # sanity inside the bench: accuracy-style count reduction over ranks (the path's only exchange step)
    counts = torch.tensor([float((labels_d >= 0).sum().item()), float(n)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(counts)

    # ---- end-to-end arm: pinned host buffers through fbn_run ---------------------------------------------------------
    off_p = torch.from_numpy(off_h).pin_memory()
    ev_p = torch.from_numpy(ev_h).pin_memory()
    ew_p = torch.from_numpy(ew_h).pin_memory()
    labels_p = torch.empty(n, dtype=torch.int32).pin_memory()
    post_p = torch.empty((n, sd), dtype=torch.float64).pin_memory()
    e2e_steps = args.e2e_steps or min(args.steps, 5)
    st = None
    for _ in range(2):
        _, _, st = plan.run(off_p.numpy(), ev_p.numpy(), ew_p.numpy(), labels_out=labels_p.numpy(), post_out=post_p.numpy())
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        _, _, st = plan.run(off_p.numpy(), ev_p.numpy(), ew_p.numpy(), labels_out=labels_p.numpy(), post_out=post_p.numpy())
    barrier()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    e2e_value = world * n * e2e_steps / e2e_s
    sampler.stop_flag = True
    sampler.join(timeout=3)
    same = bool(np.array_equal(labels_p.numpy(), labels_d.cpu().numpy()))

    if rank == 0:
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(peaks_path):
            peak = float(json.load(open(peaks_path))["hbm_gbs"])
            peak_src = "measured (MEASURED_PEAKS.json)"
        else:
            peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
        achieved = info.alg_bytes_per_case * n / (kern_ms * 1e-3) / 1e9
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "traffic.json")  # ncu dram__bytes_read+write per case, by workload
        if os.path.exists(tpath):
            tj = json.load(open(tpath)).get(args.workload)
            if tj:
                traffic = tj["dram_bytes_per_case"] * n
        line = {
            "metric": "test cases/sec, exact JT inference", "value": value, "unit": "cases/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": w["desc"], "cases_per_gpu": n, "n_evidence": w["n_evidence"], "team_warps": info.team_warps,
                       "mode": "level" if info.mode == capi.MODE_LEVEL else "fused", "launches_per_step": launches_per_step,
                       "resident_tiles": info.resident_tiles, "l2": "per-step working set (tile scratch + evidence + posteriors) %.0f MB > 126 MB L2"
                       % ((info.scratch_bytes + off_h.nbytes + ev_h.nbytes + ew_h.nbytes + n * sd * 8 + n * 4) / 1e6),
                       "labels_e2e_equal_device": same, "labelled_cases": int(counts[0].item()), "total_cases": int(counts[1].item())},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "peak_source": peak_src,
                         "moved_gbs": info.kernel_bytes_per_case * n / (kern_ms * 1e-3) / 1e9, "alg_bytes_per_case": info.alg_bytes_per_case,
                         "kernel_bytes_per_case": info.kernel_bytes_per_case, "kernel_ms": kern_ms,
                         "kernel": ("jt_flat_kernel" if info.flat_records else "jt_fused_kernel") if info.mode == capi.MODE_FUSED else "lvl_collect_kernel+lvl_pull_kernel (all launches of a step)"},
            "e2e": {"value": e2e_value, "unit": "cases/s", "h2d_bytes_per_step": int(st.h2d_bytes), "d2h_bytes_per_step": int(st.d2h_bytes),
                    "steps": e2e_steps, "launches_per_step": int(st.n_launches)},
            "gpu_launches": args.steps * launches_per_step,
            "clocks": sampler.summary(),
        }
        if world == 1 and not args.no_cpu_baseline:
            cores = os.cpu_count() or 1
            v, kind, sample = cpu_reference_throughput(args.workload, args.ref_cases_per_core or w["ref_cases"], cores, 555)
            line["cpu_baseline"] = {"value": v, "unit": "cases/s", "cores": cores, "kind": kind, "sample": sample}
        print(json.dumps(line), flush=True)
    plan.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
