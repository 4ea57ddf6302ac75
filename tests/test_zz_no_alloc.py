"""fbn_score_device_static (include/fastbn_b200.h): the scoring entry point that allocates nothing — its per-case scratch is a
static __device__ array (fastbn_b200/csrc/noalloc.cu). Checked through the C ABI against fbn_score_device on two calls in a
row (the second smaller and with other inputs, so left-over scratch of the first would show). Collected last on purpose
(file name). Passes alone in a fresh process and after the other tests."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, load_golden
from fastbn_b200 import capi

pytestmark = pytest.mark.gpu
RTOL = 1e-12  # same kernels' arithmetic in the same order: the sums agree to rounding (the other GPU tests allow 1e-9 on posteriors)


def _score_both(plan, dev, n, voff_d, post, labels, gold, truths):
    import torch
    ref = plan.score_device(n, post.data_ptr(), labels.data_ptr(), gold.data_ptr(), truths.data_ptr())
    out4 = torch.full((4,), -1.0, dtype=torch.float64, device=dev)
    capi.score_device_static(0, n, voff_d.numel() - 1, voff_d.data_ptr(), post.data_ptr(), labels.data_ptr(), gold.data_ptr(),
                             truths.data_ptr(), out4.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    got = out4.cpu().numpy()
    assert got[0] == ref["n_correct"] and got[1] == ref["n_total"] == n
    assert abs(got[2] - ref["sum_mse"]) <= RTOL * max(abs(ref["sum_mse"]), 1e-300)
    assert abs(got[3] - ref["sum_hd"]) <= RTOL * max(abs(ref["sum_hd"]), 1e-300)
    return ref


def test_static_scoring_matches_allocating_entry_point_twice(alarm_tree):
    import torch
    from test_oracle import read_pt
    dev = torch.device("cuda:0")
    g = load_golden("alarm_1k_ref")
    dims = np.asarray(alarm_tree["var_dims"], np.int64)
    voff_d = torch.from_numpy(np.concatenate([[0], np.cumsum(dims)]).astype(np.int32)).to(dev)
    n = 1000
    plan = capi.Plan(alarm_tree, device=0)
    try:
        labels = torch.empty(n, dtype=torch.int32, device=dev)
        post = torch.empty((n, int(dims.sum())), dtype=torch.float64, device=dev)
        off, ev, ew = [torch.from_numpy(g[k]).to(dev) for k in ("ev_off", "ev_vars", "ev_vals")]
        plan.run_device(n, off.data_ptr(), ev.data_ptr(), ew.data_ptr(), labels.data_ptr(), post.data_ptr())
        torch.cuda.synchronize()
        gold = torch.from_numpy(read_pt(os.path.join(GOLDEN, "alarm", "alarm_1k_pt"), alarm_tree["var_dims"])).to(dev)
        truths = torch.from_numpy(g["truths"]).to(dev)
        # call 1: the shipped golden file (reference prints accuracy 1, MSE 6.50867e-17, HD 4.6999e-17)
        r1 = _score_both(plan, dev, n, voff_d, post, labels, gold, truths)
        assert r1["n_correct"] == 1000
        # call 2, right after: fewer cases, perturbed posteriors and labels (scores of another magnitude)
        n2 = 617
        rng = np.random.default_rng(3)
        post2 = (post[:n2] * torch.from_numpy(rng.uniform(0.5, 1.0, (n2, post.shape[1]))).to(dev)).contiguous()
        labels2 = labels[:n2].clone()
        labels2[::5] += 1
        r2 = _score_both(plan, dev, n2, voff_d, post2, labels2, gold[:n2].contiguous(), truths[:n2].contiguous())
        assert r2["n_correct"] < n2 and r2["sum_mse"] > 1e-3
        # an empty batch is four zeros from both
        _score_both(plan, dev, 0, voff_d, post, labels, gold, truths)
    finally:
        plan.close()


def test_static_scoring_rejects_batches_beyond_the_static_scratch():
    import torch
    dev = torch.device("cuda:0")
    z = torch.zeros(8, dtype=torch.float64, device=dev)
    with pytest.raises(capi.FastBNError, match="exceeds the static scratch"):
        capi.score_device_static(0, capi.STATIC_SCORE_MAX_CASES + 1, 1, z.data_ptr(), z.data_ptr(), z.data_ptr(), z.data_ptr(),
                                 z.data_ptr(), z.data_ptr())
    with pytest.raises(capi.FastBNError, match="device ordinal"):
        capi.score_device_static(4096, 1, 1, z.data_ptr(), z.data_ptr(), z.data_ptr(), z.data_ptr(), z.data_ptr(), z.data_ptr())
