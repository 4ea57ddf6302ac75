#!/usr/bin/env python
"""Benchmark of the junction-tree hot path (BASELINE.json metric: test cases/sec of exact JT inference; % of HBM roofline).

  python bench.py --gpus N --steps K --warmup W            our arm (one process per GPU; torchrun for N > 1)
  python bench.py --impl reference --gpus N --steps K ...  the reference's own CPU implementation on the host cores

A "step" = one pass of the hot path over one batch of synthetic test cases. The headline (`value`, `e2e`, `roofline`,
`cpu_baseline`) is BASELINE.json configs[1]: ALARM, 1M cases, 20% evidence (7 of 37 variables). The same line carries
`workloads`: configs[2..4] at the sizes BASELINE.json states (HEPAR2-shape 1M cases at 20% and 50% evidence, PIGS-shape 1M,
MUNIN-shape 10k), each with its own device-resident rate, e2e rate, DRAM-based roofline fraction and a parity check of a
golden prefix of the TIMED batch against the reference's fp64 posteriors (tests/golden/*_ref.npz); config[0] (ALARM, the
shipped 1000 cases, reference default run) is in `cpu_baseline.config1`. Weak scaling: every GPU gets its own batch of the
same size, no data-path collective; the only exchange is the final reduction of the four scores (fbn_reduce, NCCL).
`value`  : whole-job cases/s with evidence already resident in HBM and results left in HBM (CUDA events).
`e2e`    : same metric through fbn_run with pinned HOST buffers, H2D/D2H copies inside the timed region: evidence in, labels out
           (what the reference's PredictUseJTInfer returns); `e2e.full_posteriors`: the all-node posteriors shipped to the host too.
One JSON line is printed by rank 0; the exit code is non-zero when a parity prefix differs by more than 1e-9.
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
TOL = 1e-9  # north_star: posterior marginals in fp64 within max-abs 1e-9, labels bit-identical

WORKLOADS = {
    # name: tree source, #evidence vars, cases per GPU per step (BASELINE.json sizes), seed, golden prefix
    "alarm": dict(tree="alarm_tree_ref.npz", xml=os.path.join("alarm", "alarm.xml"), n_evidence=7, cases=1_000_000, seed=20260101,
                  golden=("alarm_synth_4k_ref", 0, 4000), ref_cases=60000, ref_t1=20000, ref_tn=1000, steps=(10, 3, 5),
                  desc="ALARM (37 nodes), 1M synthetic test cases, 20% evidence (7 vars), batched"),
    "hepar2": dict(shape="hepar2", n_evidence=14, cases=1_000_000, seed=20260103, golden=(os.path.join("synth", "hepar2_ref"), 0, 256),
                   ref_cases=1200, ref_t1=600, ref_tn=150, steps=(4, 3, 2),
                   desc="synthetic HEPAR2-shape BN (70 nodes, card 2-4), 1M cases, 20% evidence"),
    "hepar2_50": dict(shape="hepar2", n_evidence=35, cases=1_000_000, seed=20260104, golden=(os.path.join("synth", "hepar2_ref"), 256, 512),
                      ref_cases=2400, ref_t1=1200, ref_tn=300, steps=(4, 3, 2),
                      desc="synthetic HEPAR2-shape BN (70 nodes, card 2-4), 1M cases, 50% evidence"),
    "pigs": dict(shape="pigs", n_evidence=88, cases=1_000_000, seed=20260105, golden=(os.path.join("synth", "pigs_ref"), 0, 48),
                 ref_cases=200, ref_t1=100, ref_tn=40, steps=(3, 3, 1),
                 desc="synthetic PIGS-shape BN (441 nodes, card 3, max clique 3^10), 1M cases, 20% evidence"),
    "munin_small": dict(shape="munin_small", n_evidence=0, cases=40_000, seed=20260107, golden=(os.path.join("synth", "munin_small_ref"), 0, 1),
                        ref_cases=8, ref_t1=8, ref_tn=8, steps=(3, 3, 1),
                        desc="synthetic MUNIN-shape stand-in (max clique 5^9), 0% evidence"),
    "munin": dict(shape="munin", n_evidence=0, cases=10_000, seed=20260106, golden=(os.path.join("synth", "munin_ref"), 0, 1),
                  ref_cases=1, ref_t1=2, ref_tn=2, steps=(3, 3, 1),
                  desc="synthetic MUNIN-shape BN (max clique 5^10 ~ 1e7), 10k cases, 0% evidence, intra-table multi-CTA split"),
}
SECONDARY = ["hepar2", "hepar2_50", "pigs", "munin"]  # BASELINE.json configs[2..4]


def workload_xml(w):
    """Path of the workload's XMLBIF (committed under tests/golden)."""
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


_TREES = {}


def load_tree(w):
    cfg = WORKLOADS[w]
    key = cfg.get("tree") or cfg["shape"]
    if key not in _TREES:
        if "tree" in cfg:
            z = np.load(os.path.join(GOLDEN, cfg["tree"]))
            _TREES[key] = {k: z[k] for k in z.files}
        else:
            from fastbn_b200 import hostapi
            m = hostapi.HostModel(workload_xml(w))
            _TREES[key] = m.tree()
            m.close()
    return _TREES[key]


def make_batch(w, n, rank):
    """The synthetic batch of workload w for one rank: seeded uniform evidence, its first cases replaced by the golden cases
    the reference's posteriors are committed for (so the parity prefix is part of the timed batch).
    Returns (ev_off, ev_vars, ev_vals, golden_post, golden_labels)."""
    from fastbn_b200 import synth
    cfg = WORKLOADS[w]
    tree = load_tree(w)
    off, ev, ew = synth.uniform_evidence(tree["var_dims"], n, cfg["n_evidence"], seed=cfg["seed"] + 1000 * rank)
    name, g0, g1 = cfg["golden"]
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    k = min(g1 - g0, n)
    goff = z["ev_off"][g0:g0 + k + 1].astype(np.int64)
    gev, gew = z["ev_vars"][goff[0]:goff[-1]], z["ev_vals"][goff[0]:goff[-1]]
    goff = goff - goff[0]
    off = np.concatenate([goff, goff[-1] + (off[k + 1:] - off[k])]).astype(np.int64)
    ev = np.concatenate([gev, ev[cfg["n_evidence"] * k:]]).astype(np.int32)
    ew = np.concatenate([gew, ew[cfg["n_evidence"] * k:]]).astype(np.int32)
    return off, ev, ew, z["post"][g0:g0 + k], z["labels"][g0:g0 + k]


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons sampled DURING the timed regions (B200_PROFILING.md recipe)."""

    def __init__(self, gpu):
        super().__init__(daemon=True)
        self.gpu = gpu
        self.rows = []  # (time, fields)
        self.stop_flag = False

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append((time.time(), [x.strip() for x in out.split(",")]))
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self, t0=None, t1=None):
        rows = [r for t, r in self.rows if (t0 is None or t >= t0) and (t1 is None or t <= t1)]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"], "samples": 0}
        sm = sorted(int(r[0]) for r in rows if r[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in rows)]
        mx = [int(r[1]) for r in rows if r[1].isdigit()]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(rows)}


# ---------------------------------------------------------------------------------------------------------------
# CPU baselines (rank 0 only): the UNMODIFIED reference (oracle/_ref) on the host cores.
#   value : N independent `-t 1` instances, one per core, each driving the stock PredictUseJTInfer loop over its own
#           slice of the cases (SURVEY.md 8(d) "baseline B", the throughput-oriented number)
#   t1    : ONE instance, `-t 1` (the reference default), cases / its `jt` span
#   tN    : ONE instance, `-t N` (the reference's OpenMP path: parallel INSIDE a case, cases stay sequential)
def _cpu_model():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def _write_cases(path, w, n, seed):
    from fastbn_b200 import synth
    tree = load_tree(w)
    off, ev, ew = synth.uniform_evidence(tree["var_dims"], n, WORKLOADS[w]["n_evidence"], seed=seed)
    synth.write_libsvm(path, off, ev, ew)
    return off, ev, ew


def _ref_worker(args):
    """One `-t 1` instance on its own slice. Returns seconds inside the per-case loop."""
    xml, w, n, seed, kind, stock, idx = args
    from fastbn_b200 import synth
    tmp = os.path.join(ROOT, "build", "ref_%d_%d" % (os.getppid(), idx))
    if kind == "reference":
        from oracle import refdrv
        if stock:  # the stock driver loop (PredictUseJTInfer incl. MSE/HD per case) over a LIBSVM file, needs >= 20 cases
            _write_cases(tmp, w, n, seed)
            synth.write_empty_pt(tmp + "_pt")
            r = refdrv.Reference(xml, tmp)
            _, secs = r.predict_timed(tmp + "_pt", 1)
        else:
            off, ev, ew = _write_cases(tmp, w, 20, 1)
            r = refdrv.Reference(xml, tmp)
            off, ev, ew = synth.uniform_evidence(load_tree(w)["var_dims"], n, WORKLOADS[w]["n_evidence"], seed=seed)
            _, _, secs = r.run_batch(off, ev, ew, want_post=True)
        r.close()
        for p in (tmp, tmp + "_pt"):
            if os.path.exists(p):
                os.remove(p)
        return secs
    from oracle import jtoracle
    o = jtoracle.Oracle(load_tree(w))
    off, ev, ew = synth.uniform_evidence(load_tree(w)["var_dims"], n, WORKLOADS[w]["n_evidence"], seed=seed)
    t0 = time.perf_counter()
    o.run_batch(off, ev, ew)
    secs = time.perf_counter() - t0
    o.close()
    return secs


def cpu_reference_throughput(workload, cases_per_core, cores, seed):
    """N independent `-t 1` instances. Returns (cases/s over all cores, kind, sample description)."""
    import multiprocessing as mp
    from oracle import refdrv
    kind = "reference" if refdrv.available() else "port"
    if kind == "reference":
        refdrv.lib()  # loaded in the parent too, so that the process running this arm shows which library does the work
    os.makedirs(os.path.join(ROOT, "build"), exist_ok=True)
    load_tree(workload)
    xml = workload_xml(workload)
    stock = cases_per_core >= 20
    jobs = [(xml, workload, cases_per_core, seed + c, kind, stock, c) for c in range(cores)]
    ctx = mp.get_context("fork")
    t0 = time.perf_counter()
    with ctx.Pool(cores) as pool:
        secs = pool.map(_ref_worker, jobs)
    wall = time.perf_counter() - t0
    # throughput = total cases / slowest worker's time inside the per-case loop (construction/parsing excluded, as for the GPU arm)
    val = cases_per_core * cores / max(secs)
    return val, kind, "%d cases/core x %d independent -t 1 instances (%s; max worker loop %.2f s, wall %.2f s)" % (
        cases_per_core, cores, "stock PredictUseJTInfer loop" if stock and kind == "reference" else "per-case sequence", max(secs), wall)


def cpu_single_instance(workload, n_t1, n_tn, cores):
    """ONE reference instance: `-t 1` and `-t N` (N = cores). Returns dict(t1, tN, samples) in cases/s, or None without oracle/_ref."""
    from oracle import refdrv
    from fastbn_b200 import synth
    if not refdrv.available():
        return None
    os.makedirs(os.path.join(ROOT, "build"), exist_ok=True)
    xml = workload_xml(workload)
    tmp = os.path.join(ROOT, "build", "ref_single_%d" % os.getpid())
    out = {}
    for key, n, threads in (("t1", n_t1, 1), ("tN", n_tn, cores)):
        if n >= 20:
            _write_cases(tmp, workload, n, 4242)
            synth.write_empty_pt(tmp + "_pt")
            r = refdrv.Reference(xml, tmp)
            _, secs = r.predict_timed(tmp + "_pt", threads)
            how = "stock PredictUseJTInfer loop"
        else:
            _write_cases(tmp, workload, 20, 1)
            r = refdrv.Reference(xml, tmp)
            off, ev, ew = synth.uniform_evidence(load_tree(workload)["var_dims"], n, WORKLOADS[workload]["n_evidence"], seed=4242)
            _, _, secs = r.run_batch_threads(off, ev, ew, threads)
            how = "per-case sequence"
        r.close()
        out[key] = n / secs
        out[key + "_sample"] = "%d cases, %d thread(s), %s, %.2f s" % (n, threads, how, secs)
    for p in (tmp, tmp + "_pt"):
        if os.path.exists(p):
            os.remove(p)
    return out


def config1_reference_default_run(cores):
    """BASELINE.json configs[0]: ALARM, the shipped 1000 cases (testing_alarm_1k_p20 + alarm_1k_pt), the reference's default run
    (`-t 1`) and `-t N`: cases / `jt` span of the stock loop; accuracy as printed by the reference."""
    from oracle import refdrv
    if not refdrv.available():
        return None
    a = os.path.join(GOLDEN, "alarm")
    r = refdrv.Reference(os.path.join(a, "alarm.xml"), os.path.join(a, "testing_alarm_1k_p20"))
    out = {"workload": "ALARM, shipped testing_alarm_1k_p20 (1000 cases, 20% evidence), alarm_1k_pt"}
    for key, threads in (("t1", 1), ("tN", cores)):
        best = None
        for _ in range(2):
            acc, secs = r.predict_timed(os.path.join(a, "alarm_1k_pt"), threads)
            best = secs if best is None else min(best, secs)
        out[key] = 1000.0 / best
        out[key + "_jt_seconds"] = best
        out["accuracy"] = acc
    r.close()
    return out


def cpu_baseline_block(workload, cores, per_core=0, full=True):
    w = WORKLOADS[workload]
    v, kind, sample = cpu_reference_throughput(workload, per_core or w["ref_cases"], cores, 555)
    blk = {"value": v, "unit": "cases/s", "cores": cores, "kind": kind, "sample": sample, "n_procs": cores, "cpu_model": _cpu_model()}
    if full:
        s = cpu_single_instance(workload, w["ref_t1"], w["ref_tn"], cores)
        if s:
            blk.update(s)
    return blk


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    w = WORKLOADS[args.workload]
    # bounded sample per step: a quarter of the cpu_baseline sample (a few seconds of CPU work per step)
    per_core = args.ref_cases_per_core or max(1, w["ref_cases"] // 4)
    for _ in range(min(args.warmup, 1)):
        cpu_reference_throughput(args.workload, max(20, per_core // 10) if per_core >= 20 else per_core, cores, 777)
    vals = []
    t0 = time.perf_counter()
    for s in range(args.steps):
        v, kind, sample = cpu_reference_throughput(args.workload, per_core, cores, 1000 + s)
        vals.append(v)
    wall = time.perf_counter() - t0
    value = float(np.mean(vals))
    blk = {"value": value, "unit": "cases/s", "cores": cores, "kind": kind, "sample": sample, "n_procs": cores, "cpu_model": _cpu_model()}
    s1 = cpu_single_instance(args.workload, w["ref_t1"], w["ref_tn"], cores)
    if s1:
        blk.update(s1)
    line = {"impl": "reference", "metric": "test cases/sec, exact JT inference", "value": value, "unit": "cases/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": wall / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": w["desc"], "cases_per_step": per_core * cores, "n_evidence": w["n_evidence"],
                       "note": "bounded sample of the same synthetic workload (cases/s does not depend on the batch size); the GPU arm runs %d cases per step" % w["cases"]},
            "cpu_baseline": blk,
            "e2e": {"value": value, "unit": "cases/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------------------
class Ctx:
    pass


class PinnedPool:
    """One pinned host arena reused by every workload (the PIGS-shape posteriors of 1M cases are 10.6 GB)."""

    def __init__(self, torch):
        self.torch = torch
        self.buf = None

    def arrays(self, specs):
        """specs: list of (shape, numpy dtype). Returns numpy views into the arena (256-byte aligned)."""
        sizes = [int(np.prod(s)) * np.dtype(d).itemsize for s, d in specs]
        offs, tot = [], 0
        for sz in sizes:
            offs.append(tot)
            tot += (sz + 255) // 256 * 256
        if self.buf is None or self.buf.numel() < tot:
            self.buf = None
            self.buf = self.torch.empty(max(tot, 1), dtype=self.torch.uint8).pin_memory()
        base = self.buf.numpy()
        return [base[o:o + sz].view(d).reshape(s) for (s, d), o, sz in zip(specs, offs, sizes)]


def traffic_entry(workload, info, kernel):
    """ncu-measured DRAM bytes per case of this workload's kernels (profiles/traffic.json); None when the entry was
    profiled with another kernel revision / mode / team than the one running now."""
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(tpath):
        return None, None
    tj = json.load(open(tpath)).get(workload)
    if not tj:
        return None, None
    same_team = info.mode == 2 or tj.get("team_warps") == info.team_warps  # level mode: the nominal gang size follows the free memory
    if tj.get("kernel_rev") != info.kernel_rev or tj.get("mode") != info.mode or not same_team:
        return None, "stale: %s was profiled at kernel_rev %s mode %s team %s" % (tj.get("source"), tj.get("kernel_rev"), tj.get("mode"), tj.get("team_warps"))
    return float(tj["dram_bytes_per_case"]), tj.get("source")


def smem_entry(workload, info):
    """ncu-measured shared-memory traffic per case (l1tex__data_pipe_lsu_wavefronts_mem_shared x 128 B, one --set full capture;
    profiles/traffic.json, same kernel revision): the on-chip roofline of jt_resident_kernel, whose tile state never leaves
    shared memory. (None, None) when absent or stale."""
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(tpath):
        return None, None
    tj = json.load(open(tpath)).get(workload) or {}
    if "smem_bytes_per_case" not in tj or tj.get("smem_kernel_rev") != info.kernel_rev or tj.get("smem_team_warps") != info.team_warps:
        return None, None
    return float(tj["smem_bytes_per_case"]), tj.get("smem_source")


def run_workload(cx, name, steps, warmup, e2e_steps, cases=0, team=0, mode=0, max_resident=0):
    """Device-resident arm, parity prefix, e2e arm of one workload on this rank. Returns the workload's result dict (complete on rank 0)."""
    torch, dist, capi = cx.torch, cx.dist, cx.capi
    w = WORKLOADS[name]
    tree = load_tree(name)
    plan = capi.Plan(tree, device=cx.local, team_warps=team, mode=mode, max_resident=max_resident)
    info = plan.info
    sd = info.sum_dims
    n = cases or w["cases"]
    tile_cases = info.tile_cases
    waves = -(-((n + tile_cases - 1) // tile_cases) // info.resident_tiles)
    launches_per_step = info.launches_per_wave * waves if info.mode == capi.MODE_LEVEL else 1
    off_h, ev_h, ew_h, gpost, glab = make_batch(name, n, cx.rank)
    k = len(glab)

    def barrier():
        if cx.world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident arm -------------------------------------------------------------------------------------
    dev = cx.dev
    off_d = torch.from_numpy(off_h).to(dev)
    ev_d = torch.from_numpy(ev_h).to(dev)
    ew_d = torch.from_numpy(ew_h).to(dev)
    labels_d = torch.empty(n, dtype=torch.int32, device=dev)
    post_d = torch.empty((n, sd), dtype=torch.float64, device=dev)
    stream = torch.cuda.current_stream().cuda_stream

    def step_device():
        plan.run_device(n, off_d.data_ptr(), ev_d.data_ptr(), ew_d.data_ptr(), labels_d.data_ptr(), post_d.data_ptr(),
                        stream=stream, timed=False)

    for _ in range(warmup):
        step_device()
    barrier()
    t_wall0 = time.time()
    e0 = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
    e1 = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
    t_all0 = torch.cuda.Event(enable_timing=True)
    t_all1 = torch.cuda.Event(enable_timing=True)
    barrier()
    t_all0.record()
    for s in range(steps):
        e0[s].record()
        step_device()
        e1[s].record()
    t_all1.record()
    barrier()
    total_ms = t_all0.elapsed_time(t_all1)
    kern_ms = float(np.mean([a.elapsed_time(b) for a, b in zip(e0, e1)]))  # all kernel launches of one step
    if cx.world > 1:
        t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t.item())
    ms_per_step = total_ms / steps
    value = cx.world * n / (ms_per_step * 1e-3)

    # ---- parity of the golden prefix of the timed batch (reference fp64 posteriors, tests/golden) ----------------
    p_dev = post_d[:k].cpu().numpy()
    l_dev = labels_d[:k].cpu().numpy()
    max_abs = float(np.max(np.abs(p_dev - gpost))) if k else 0.0
    labels_equal = bool(np.array_equal(l_dev, glab))
    # accuracy-style counts of this rank (the path's only exchange step is their reduction)
    n_labelled = float((labels_d >= 0).sum().item())
    row_sum = float(post_d.sum().item())
    counts = cx.reduce4([n_labelled, float(n), row_sum, 0.0])

    # ---- end-to-end arm: pinned host buffers through fbn_run ---------------------------------------------------------
    off_p, ev_p, ew_p, labels_p, post_p = cx.pool.arrays([(off_h.shape, np.int64), (ev_h.shape, np.int32), (ew_h.shape, np.int32),
                                                          ((n,), np.int32), ((n, sd), np.float64)])
    off_p[:] = off_h
    ev_p[:] = ev_h
    ew_p[:] = ew_h
    st = None
    for _ in range(1 if steps < 5 else 2):
        _, _, st = plan.run(off_p, ev_p, ew_p, labels_out=labels_p, post_out=post_p)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        _, _, st = plan.run(off_p, ev_p, ew_p, labels_out=labels_p, post_out=post_p)
    barrier()
    e2e_s = time.perf_counter() - t0
    t_wall1 = time.time()
    if cx.world > 1:
        t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    e2e_value = cx.world * n * e2e_steps / e2e_s
    same = bool(np.array_equal(labels_p, labels_d.cpu().numpy()))
    # the same call when the caller wants what PredictUseJTInfer returns (labels): posteriors never leave the device, 4 bytes per case back
    lab_only = np.empty(n, np.int32)
    plan.run(off_p, ev_p, ew_p, want_post=False, labels_out=labels_p)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        _, _, st_l = plan.run(off_p, ev_p, ew_p, want_post=False, labels_out=labels_p)
    barrier()
    e2e_l_s = time.perf_counter() - t0
    if cx.world > 1:
        t = torch.tensor([e2e_l_s], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_l_s = float(t.item())
    e2e_labels_value = cx.world * n * e2e_steps / e2e_l_s
    del lab_only
    e2e_max_abs = float(np.max(np.abs(post_p[:k] - gpost))) if k else 0.0
    ok = max_abs <= TOL and e2e_max_abs <= TOL and labels_equal and same
    if cx.world > 1:
        t = torch.tensor([0.0 if ok else 1.0], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ok = float(t.item()) == 0.0

    if info.mode == capi.MODE_LEVEL:
        kernel = "lvl_sum_kernel (all launches of a step)"
    else:
        kernel = plan.kernel_name
    traffic, tsrc = traffic_entry(name, info, kernel)
    peak, peak_src = cx.peak
    alg_gbs = info.alg_bytes_per_case * n / (kern_ms * 1e-3) / 1e9
    dram_gbs = traffic * n / (kern_ms * 1e-3) / 1e9 if traffic else None
    smem_b, smem_src = smem_entry(name, info)
    on_chip = None
    if smem_b:
        # shared-memory bandwidth: 128 B per clock and SM (B300_MICROARCH.md: smem crossbar) at the SM clock sampled during the run
        clk = cx.sampler.summary(t_wall0, t_wall1)
        smem_peak = torch.cuda.get_device_properties(dev).multi_processor_count * 128.0 * (clk.get("sm_mhz") or clk.get("sm_max_mhz") or 0) * 1e6 / 1e9
        smem_gbs = smem_b * n / (kern_ms * 1e-3) / 1e9
        on_chip = {"bound": "shared-memory bandwidth", "smem_bytes_per_case": smem_b, "achieved": smem_gbs, "peak": smem_peak, "unit": "GB/s",
                   "frac": smem_gbs / smem_peak if smem_peak else None, "source": smem_src,
                   "peak_source": "n_sms x 128 B/clk x SM clock sampled during the run"}
    res = {
        "workload": w["desc"], "cases": n, "n_evidence": w["n_evidence"], "steps": steps, "warmup": warmup,
        "value": value, "unit": "cases/s", "ms_per_step": ms_per_step,
        # e2e = the reference-facing call: PredictUseJTInfer returns the labels (src/JunctionTree.cpp:1508-1534); the posteriors of all
        # nodes are computed for every case and stay on the device (fbn_run with posteriors_out = NULL; fbn_evaluate scores them
        # there). `full_posteriors`: the same call also shipping 8*sum_dims bytes of posteriors per case to the host: bound by the
        # box's pinned device->host bandwidth (PCIe), not by the GPU
        "e2e": {"value": e2e_labels_value, "unit": "cases/s", "steps": e2e_steps, "h2d_bytes_per_step": int(st_l.h2d_bytes),
                "d2h_bytes_per_step": int(st_l.d2h_bytes), "launches_per_step": int(st_l.n_launches),
                "returns": "labels of all cases (what PredictUseJTInfer returns); all-node posteriors computed, left on the device",
                "full_posteriors": {"value": e2e_value, "unit": "cases/s", "h2d_bytes_per_step": int(st.h2d_bytes),
                                    "d2h_bytes_per_step": int(st.d2h_bytes),
                                    "note": "posteriors_out = host buffer: + 8*sum_dims bytes per case device->host; PCIe-bound"}},
        "kernel": kernel, "kernel_ms": kern_ms, "mode": "level" if info.mode == capi.MODE_LEVEL else "fused",
        "team_warps": info.team_warps, "resident_tiles": info.resident_tiles, "tile_cases": tile_cases, "launches_per_step": launches_per_step,
        "kernel_rev": info.kernel_rev,
        # DRAM-based roofline: ncu dram__bytes_read+write per case (profiles/traffic.json, same kernel revision) x cases / device time
        "dram_bytes_per_case": traffic, "dram_source": tsrc, "dram_gbs": dram_gbs, "dram_frac": dram_gbs / peak if dram_gbs else None,
        # algorithmic-equivalent figure of SURVEY.md 8(d): bytes the reference's table passes would move; NOT a utilisation (> 1: the kernels fuse passes)
        "alg_bytes_per_case": info.alg_bytes_per_case, "alg_gbs": alg_gbs, "alg_frac": alg_gbs / peak,
        "kernel_bytes_per_case": info.kernel_bytes_per_case,
        "on_chip": on_chip,
        "parity_prefix": {"cases": k, "max_abs": max_abs, "e2e_max_abs": e2e_max_abs, "labels_equal": labels_equal,
                          "labels_e2e_equal_device": same, "tol": TOL, "golden": WORKLOADS[name]["golden"][0] + ".npz", "ok": ok},
        "labelled_cases": int(round(counts[0])), "total_cases": int(round(counts[1])),
        "l2": "per-step working set (tile scratch in HBM + evidence + posteriors) %.0f MB > 126 MB L2"
              % ((info.scratch_bytes + off_h.nbytes + ev_h.nbytes + ew_h.nbytes + n * sd * 8 + n * 4) / 1e6),
        "clocks": cx.sampler.summary(t_wall0, t_wall1),
        "_window": (t_wall0, t_wall1), "_ok": ok,
    }
    plan.close()
    del off_d, ev_d, ew_d, labels_d, post_d
    torch.cuda.empty_cache()
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="alarm", choices=sorted(WORKLOADS), help="the headline workload")
    ap.add_argument("--workloads", default=None, help="secondary workloads in the same line: 'all' (default with the alarm headline), 'none', or a comma list")
    ap.add_argument("--cases", type=int, default=0, help="cases per GPU per step of the headline workload (default: BASELINE.json's)")
    ap.add_argument("--team", type=int, default=0)
    ap.add_argument("--mode", type=int, default=0, help="0 auto, 1 fused kernel, 2 level kernels")
    ap.add_argument("--max-resident", type=int, default=0)
    ap.add_argument("--ref-cases-per-core", type=int, default=0, help="CPU sample per core (default: the workload's, ~10-20 s)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--e2e-steps", type=int, default=0, help="timed e2e steps (default: min(steps, 5))")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)

    import torch
    import torch.distributed as dist
    from fastbn_b200 import capi

    # stdout carries ONE JSON line: whatever libraries print on their own (NCCL writes "NCCL version ..." to stdout when the box
    # sets NCCL_DEBUG=VERSION) goes to stderr; the saved descriptor gets the line at the end
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)

    cx = Ctx()
    cx.torch, cx.dist, cx.capi = torch, dist, capi
    cx.world = int(os.environ.get("WORLD_SIZE", "1"))
    cx.rank = int(os.environ.get("RANK", "0"))
    cx.local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    torch.cuda.set_device(cx.local)
    cx.dev = torch.device("cuda", cx.local)
    comm = None
    if cx.world > 1:
        dist.init_process_group("nccl", device_id=cx.dev)
        # the path's one exchange step goes through the C ABI (fbn_reduce = one ncclAllReduce of the four scores)
        from fastbn_b200 import shard
        comm = shard.comm_from_process_group(cx.local)

    def reduce4(v):
        if comm is None:
            return list(v)
        return comm.reduce(v)
    cx.reduce4 = reduce4
    cx.pool = PinnedPool(torch)
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        cx.peak = (float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)")
    else:
        cx.peak = (6650.0, "fallback (B200_PROFILING.md)")
    cx.sampler = ClockSampler(cx.local)
    cx.sampler.start()

    warm = max(args.warmup, 3)
    head = run_workload(cx, args.workload, args.steps, warm, args.e2e_steps or min(args.steps, 5), cases=args.cases, team=args.team,
                        mode=args.mode, max_resident=args.max_resident)
    sel = args.workloads if args.workloads is not None else ("all" if args.workload == "alarm" and not args.cases else "none")
    names = [] if sel == "none" else (SECONDARY if sel == "all" else [x for x in sel.split(",") if x])
    others = {}
    for nm in names:
        k_steps, k_warm, k_e2e = WORKLOADS[nm]["steps"]
        others[nm] = run_workload(cx, nm, k_steps, k_warm, k_e2e)

    # strong scaling of the headline beside the weak-scaling value: ONE batch of the headline size cut N ways
    strong = None
    if cx.world > 1:
        n_total = args.cases or WORKLOADS[args.workload]["cases"]
        s = run_workload(cx, args.workload, args.steps, warm, 1, cases=max(1, n_total // cx.world), team=args.team, mode=args.mode)
        strong = {"cases_total": (n_total // cx.world) * cx.world, "value": s["value"], "ms_per_step": s["ms_per_step"], "e2e": s["e2e"]["value"], "e2e_full_posteriors": s["e2e"]["full_posteriors"]["value"]}
    cx.sampler.stop_flag = True
    cx.sampler.join(timeout=3)

    all_ok = head["_ok"] and all(o["_ok"] for o in others.values())
    if cx.rank == 0:
        peak, peak_src = cx.peak
        w = WORKLOADS[args.workload]
        for r in [head] + list(others.values()):
            r.pop("_window", None)
            r.pop("_ok", None)
        launches = args.steps * head["launches_per_step"] + sum(o["steps"] * o["launches_per_step"] for o in others.values())
        line = {
            "metric": "test cases/sec, exact JT inference", "value": head["value"], "unit": "cases/s", "n_gpus": cx.world,
            "steps": args.steps, "warmup": warm, "ms_per_step": head["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": w["desc"], "cases_per_gpu": head["cases"], "n_evidence": w["n_evidence"], "team_warps": head["team_warps"],
                       "mode": head["mode"], "launches_per_step": head["launches_per_step"], "resident_tiles": head["resident_tiles"],
                       "tile_cases": head["tile_cases"], "l2": head["l2"], "labels_e2e_equal_device": head["parity_prefix"]["labels_e2e_equal_device"],
                       "labelled_cases": head["labelled_cases"], "total_cases": head["total_cases"]},
            # frac = DRAM traffic (ncu, per case) x cases / device time / measured peak: a utilisation, <= ~1
            "roofline": {"bound": "hbm", "achieved": head["dram_gbs"], "peak": peak, "unit": "GB/s", "frac": head["dram_frac"],
                         "traffic": head["dram_bytes_per_case"] * head["cases"] if head["dram_bytes_per_case"] else None,
                         "traffic_source": head["dram_source"], "peak_source": peak_src, "kernel": head["kernel"], "kernel_ms": head["kernel_ms"],
                         "dram_bytes_per_case": head["dram_bytes_per_case"],
                         "alg_bytes_per_case": head["alg_bytes_per_case"], "alg_equiv_gbs": head["alg_gbs"], "alg_frac": head["alg_frac"],
                         "kernel_bytes_per_case": head["kernel_bytes_per_case"],
                         # the resident-tile kernel moves only evidence + posteriors through HBM; what bounds it is on the chip
                         "on_chip": head["on_chip"]},
            "e2e": head["e2e"],
            "parity_prefix": head["parity_prefix"],
            "workloads": others,
            "gpu_launches": launches,
            "clocks": cx.sampler.summary(),
        }
        if strong:
            line["strong_scaling"] = strong
        if cx.world == 1 and not args.no_cpu_baseline:
            cores = os.cpu_count() or 1
            line["cpu_baseline"] = cpu_baseline_block(args.workload, cores, args.ref_cases_per_core)
            if args.workload == "alarm":
                c1 = config1_reference_default_run(cores)
                if c1:
                    line["cpu_baseline"]["config1"] = c1
            for nm, o in others.items():
                o["cpu_baseline"] = cpu_baseline_block(nm, cores)
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    if comm is not None:
        comm.close()
    if cx.world > 1:
        dist.destroy_process_group()
    if not all_ok:
        sys.stderr.write("bench.py: PARITY FAILURE — a golden prefix of a timed batch differs from the reference by more than %g\n" % TOL)
        sys.exit(3)


if __name__ == "__main__":
    main()
