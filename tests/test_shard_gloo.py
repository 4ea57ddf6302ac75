"""N > 1 path on the CPU: world_size-2 gloo. Each rank takes its contiguous slice of the shipped 1000 ALARM cases, scores
it (the CPU oracle stands in for the GPU here — tests may use it as the checker), and the all-reduce of
{n_correct, n_total, sum_mse, sum_hd} must reproduce the reference's printout for the whole set."""
import os
import socket

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import GOLDEN, load_golden
from fastbn_b200 import hostapi, shard


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _round7(x):
    ip = np.trunc(x)
    f = x - ip
    for _ in range(7):
        f = f * 10
    f = np.trunc(f + 0.5)
    for _ in range(7):
        f = f / 10
    return ip + f


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import jtoracle
    m = hostapi.HostModel(os.path.join(GOLDEN, "alarm", "alarm.xml"), os.path.join(GOLDEN, "alarm", "testing_alarm_1k_p20"))
    tree = m.tree()
    off, ev, ew, truths = m.cases()
    golden = m.golden(os.path.join(GOLDEN, "alarm", "alarm_1k_pt"))
    m.close()
    c0, c1 = shard.case_slice(len(off) - 1, rank, world)
    o, v, w = shard.slice_csr(off, ev, ew, c0, c1)
    orc = jtoracle.Oracle(tree)
    post, lab = orc.run_batch(o, v, w)
    orc.close()
    g = golden[c0:c1]
    d = tree["var_dims"]
    voff = np.concatenate([[0], np.cumsum(d)])
    mse = hd = 0.0
    for i in range(c1 - c0):  # Inference::CalculateMSE / CalculateHellingerDistance, src/Inference.cpp:153-192
        use = np.repeat(g[i, voff[:-1]] > 0, d)
        a, b = _round7(post[i, use]), g[i, use]
        mse += np.sqrt(np.sum((a - b) ** 2) / use.sum())
        hd += np.sqrt(np.sum((np.sqrt(a) - np.sqrt(b)) ** 2) / use.sum())
    sc = shard.reduce_scores(dict(n_correct=int(np.sum(lab == truths[c0:c1])), n_total=c1 - c0, sum_mse=mse, sum_hd=hd))
    q.put((rank, c0, c1, sc))
    dist.destroy_process_group()


def test_two_rank_gloo_reduction_reproduces_reference_printout():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in ps:
        p.start()
    res = sorted(q.get(timeout=240) for _ in range(world))
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert [(r[1], r[2]) for r in res] == [(0, 500), (500, 1000)]
    for _, _, _, sc in res:  # every rank holds the global result
        assert sc["n_correct"] == 1000 and sc["n_total"] == 1000
        assert abs(sc["sum_mse"] / 1000 - 6.50867e-17) < 1e-19  # reference: average MSE = 6.50867e-17
        assert abs(sc["sum_hd"] / 1000 - 4.6999e-17) < 1e-19    # reference: average HD = 4.6999e-17


def test_case_slices_cover_everything_once():
    for n in (0, 1, 7, 1000, 1_000_003):
        for world in (1, 2, 3, 8):
            cuts = [shard.case_slice(n, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n
            assert all(cuts[i][1] == cuts[i + 1][0] for i in range(world - 1))
            assert max(c1 - c0 for c0, c1 in cuts) - min(c1 - c0 for c0, c1 in cuts) <= 1
    off = np.array([0, 2, 2, 5, 9], np.int64)
    o, v, w = shard.slice_csr(off, np.arange(9, dtype=np.int32), np.arange(9, dtype=np.int32) * 2, 1, 3)
    assert o.tolist() == [0, 0, 3] and v.tolist() == [2, 3, 4] and w.tolist() == [4, 6, 8]
