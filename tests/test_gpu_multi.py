"""Multi-GPU correctness on real GPUs (run with -m gpu on a box with >= 2 devices; skipped on one): the case slices of
several devices + the one exchange step (fbn_reduce = ncclAllReduce of the four scores behind the C ABI) must give the
single-GPU answer. Also: two live plans that use the same kernel instantiation with different shared-memory needs."""
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import GOLDEN, ROOT, load_golden
from fastbn_b200 import capi, hostapi

pytestmark = pytest.mark.gpu
A = os.path.join(GOLDEN, "alarm")


def _ndev():
    return capi.load_library().fbn_device_count()


def test_two_devices_through_the_host_mirror_match_one_device():
    """FASTBN_DEVICES=0,1 path of the C++ JunctionTree: slices on two devices, scores summed by fbn_reduce."""
    if _ndev() < 2:
        pytest.skip("needs 2 GPUs")
    m = hostapi.HostModel(os.path.join(A, "alarm.xml"), os.path.join(A, "alarm_1k_p20"))
    r1 = m.evaluate(os.path.join(A, "alarm_1k_pt"), devices=[0], want_post=True)
    r2 = m.evaluate(os.path.join(A, "alarm_1k_pt"), devices=[0, 1], want_post=True)
    m.close()
    assert r2["accuracy"] == r1["accuracy"] and abs(r1["accuracy"] - 0.513) < 1e-12
    assert abs(r2["average_mse"] - r1["average_mse"]) <= 1e-12 * abs(r1["average_mse"])
    assert abs(r2["average_hd"] - r1["average_hd"]) <= 1e-12 * abs(r1["average_hd"])
    assert np.array_equal(r1["labels"], r2["labels"]) and np.array_equal(r1["post"], r2["post"])


def test_two_ranks_sharded_evaluate_matches_one_gpu(tmp_path):
    """one process per GPU (torchrun, 2 ranks): fastbn_b200.shard prints the reference's result lines from the reduced scores"""
    if _ndev() < 2:
        pytest.skip("needs 2 GPUs")
    args = ["--xml", os.path.join(A, "alarm.xml"), "--cases", os.path.join(A, "alarm_1k_p20"), "--pt", os.path.join(A, "alarm_1k_pt")]
    env = dict(os.environ, PYTHONPATH=ROOT)
    one = subprocess.run([sys.executable, "-m", "fastbn_b200.shard"] + args, capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert one.returncode == 0, one.stderr
    two = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                          "--master-port", "29611", "-m", "fastbn_b200.shard"] + args, capture_output=True, text=True, timeout=900, env=env, cwd=ROOT)
    assert two.returncode == 0, two.stderr

    def lines(out):
        return {l.split(" = ")[0]: float(l.split(" = ")[1]) for l in out.splitlines() if " = " in l and l.split(" = ")[0] in ("average MSE", "average HD", "accuracy")}
    a, b = lines(one.stdout), lines(two.stdout)
    assert a["accuracy"] == b["accuracy"] == 0.513
    assert abs(a["average MSE"] - b["average MSE"]) <= 1e-5 * abs(a["average MSE"])  # printed with 6 significant digits
    assert abs(a["average HD"] - b["average HD"]) <= 1e-5 * abs(a["average HD"])


def test_single_rank_communicator_roundtrip():
    """fbn_comm_unique_id / fbn_comm_create / fbn_reduce / fbn_comm_info / fbn_comm_destroy with world size 1 (any GPU box)"""
    c = capi.Comm(0, 0, 1, capi.comm_unique_id())
    assert c.reduce([513.0, 1000.0, 6.5e-14, 4.7e-14]) == [513.0, 1000.0, 6.5e-14, 4.7e-14]
    assert c.nccl_version() >= 20000
    c.close()


def test_two_live_plans_of_one_kernel_instantiation():
    """cudaFuncAttributeMaxDynamicSharedMemorySize is per function, not per plan: a plan that needs > 48 KB of shared memory
    must still launch after a smaller plan of the same <team> instantiation was created."""
    big = hostapi.HostModel(os.path.join(GOLDEN, "synth", "pigs.xml"))
    tb = big.tree()
    big.close()
    g = load_golden(os.path.join("synth", "pigs_ref"))
    p_big = capi.Plan(tb, device=0, team_warps=16, max_resident=4)
    small = hostapi.HostModel(os.path.join(GOLDEN, "synth", "hepar2.xml"))
    ts = small.tree()
    small.close()
    p_small = capi.Plan(ts, device=0, team_warps=16, max_resident=4)
    labels, post, _ = p_big.run(g["ev_off"], g["ev_vars"], g["ev_vals"])
    assert np.nanmax(np.abs(post - g["post"])) <= 1e-9 and np.array_equal(labels, g["labels"])
    gs = load_golden(os.path.join("synth", "hepar2_ref"))
    labels, post, _ = p_small.run(gs["ev_off"], gs["ev_vars"], gs["ev_vals"])
    assert np.nanmax(np.abs(post - gs["post"])) <= 1e-9 and np.array_equal(labels, gs["labels"])
    p_big.close()
    p_small.close()
