"""GPU parity tests (run on the B200 with -m gpu): the CUDA path, called through the C ABI, against the committed
reference fp64 fixtures and against the CPU oracle on fresh seeded cases. Tolerance from BASELINE.json's north_star:
posteriors max-abs <= 1e-9 in fp64, labels bit-identical."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, load_golden
from fastbn_b200 import capi, synth
from oracle import jtoracle

pytestmark = pytest.mark.gpu
TOL = 1e-9


@pytest.fixture(scope="module")
def plan(alarm_tree):
    p = capi.Plan(alarm_tree, device=0)
    yield p
    p.close()


@pytest.mark.parametrize("name", ["alarm_1k_ref", "alarm_synth_4k_ref", "alarm_edge_ref"])
def test_matches_reference_fixtures(plan, name):
    g = load_golden(name)
    labels, post, st = plan.run(g["ev_off"], g["ev_vars"], g["ev_vals"])
    assert np.array_equal(np.isnan(post), np.isnan(g["post"]))
    assert np.nanmax(np.abs(post - g["post"])) <= TOL
    assert np.array_equal(labels, g["labels"])
    assert st.n_launches >= 1


@pytest.mark.parametrize("team", [1, 2, 4, 8, 16, 32])
def test_every_team_size(alarm_tree, team):
    g = load_golden("alarm_1k_ref")
    p = capi.Plan(alarm_tree, device=0, team_warps=team, max_resident=37)
    labels, post, _ = p.run(g["ev_off"], g["ev_vars"], g["ev_vals"])
    p.close()
    assert np.nanmax(np.abs(post - g["post"])) <= TOL
    assert np.array_equal(labels, g["labels"])


def test_fresh_cases_vs_oracle_ragged(plan, alarm_tree):
    """ragged evidence (0..20 pairs per case), unsorted, with ignored out-of-range variables; not a multiple of 32"""
    rng = np.random.default_rng(5)
    dims = alarm_tree["var_dims"]
    n = 3001
    vl, wl = [], []
    for i in range(n):
        k = int(rng.integers(0, 21))
        v = rng.permutation(np.arange(1, 37))[:k].astype(np.int32)
        w = (rng.integers(0, 1 << 30, k) % dims[v]).astype(np.int32)
        if i % 97 == 0:
            v = np.concatenate([v, [37, 1000, -3]]).astype(np.int32)
            w = np.concatenate([w, [0, 1, 0]]).astype(np.int32)
        vl.append(v)
        wl.append(w)
    off = np.concatenate([[0], np.cumsum([len(v) for v in vl])]).astype(np.int64)
    ev = np.concatenate(vl)
    ew = np.concatenate(wl)
    labels, post, _ = plan.run(off, ev, ew)
    o = jtoracle.Oracle(alarm_tree)
    opost, olab = o.run_batch(off, ev, ew)
    o.close()
    assert np.max(np.abs(post - opost)) <= TOL
    assert np.array_equal(labels, olab)


def test_empty_batch_and_labels_only(plan):
    z = np.zeros(1, np.int64)
    e = np.zeros(0, np.int32)
    labels, post, _ = plan.run(z, e, e)
    assert labels.shape == (0,)
    g = load_golden("alarm_1k_ref")
    labels, post, _ = plan.run(g["ev_off"], g["ev_vars"], g["ev_vals"], want_post=False)
    assert post is None and np.array_equal(labels, g["labels"])


def test_decreasing_offsets_are_rejected_in_any_chunk(plan, monkeypatch):
    """ev_offsets are checked chunk by chunk inside the pipeline: a bad entry in the first or in a later chunk fails
    the call with FBN_ERR_INVALID, and the plan stays usable."""
    g = load_golden("alarm_synth_4k_ref")
    monkeypatch.setenv("FBN_CHUNK_MB", "1")  # ~1.2k ALARM cases per chunk -> several chunks
    for where in (3, len(g["ev_off"]) - 5):
        off = g["ev_off"].copy()
        off[where] = off[where + 1] + 1
        with pytest.raises(capi.FastBNError, match="non-decreasing"):
            plan.run(off, g["ev_vars"], g["ev_vals"])
    labels, post, _ = plan.run(g["ev_off"], g["ev_vars"], g["ev_vals"])
    assert np.array_equal(labels, g["labels"]) and np.nanmax(np.abs(post - g["post"])) <= TOL


def test_impossible_evidence_gives_nan_like_reference(plan, alarm_tree):
    """value >= dim wipes every table: the reference's Normalize divides 0/0 -> NaN (no zero guard)."""
    off = np.array([0, 1], np.int64)
    labels, post, _ = plan.run(off, np.array([5], np.int32), np.array([9], np.int32))
    d = alarm_tree["var_dims"]
    o5 = int(np.sum(d[:5]))
    assert np.all(post[0, o5:o5 + d[5]] == 0)  # the evidence variable itself is reported as zeros
    assert np.all(np.isnan(np.delete(post[0], np.arange(o5, o5 + d[5]))))
    assert labels[0] == -1


def test_device_resident_path_and_scoring(plan, alarm_tree):
    import torch
    g = load_golden("alarm_1k_ref")
    dev = torch.device("cuda:0")
    off = torch.from_numpy(g["ev_off"]).to(dev)
    ev = torch.from_numpy(g["ev_vars"]).to(dev)
    ew = torch.from_numpy(g["ev_vals"]).to(dev)
    n = 1000
    labels = torch.empty(n, dtype=torch.int32, device=dev)
    post = torch.empty((n, 105), dtype=torch.float64, device=dev)
    torch.cuda.synchronize()
    st = plan.run_device(n, off.data_ptr(), ev.data_ptr(), ew.data_ptr(), labels.data_ptr(), post.data_ptr())
    torch.cuda.synchronize()
    assert st.kernel_ms > 0
    assert np.max(np.abs(post.cpu().numpy() - g["post"])) <= TOL
    assert np.array_equal(labels.cpu().numpy(), g["labels"])
    # on-device MSE / HD / accuracy against the shipped 7-decimal golden file (reference prints 6.50867e-17, 4.6999e-17, 1)
    from test_oracle import read_pt
    gold = torch.from_numpy(read_pt(os.path.join(GOLDEN, "alarm", "alarm_1k_pt"), alarm_tree["var_dims"])).to(dev)
    truths = torch.from_numpy(g["truths"]).to(dev)
    sc = plan.score_device(n, post.data_ptr(), labels.data_ptr(), gold.data_ptr(), truths.data_ptr())
    assert sc["n_correct"] == 1000 and sc["n_total"] == 1000
    assert abs(sc["sum_mse"] / n - 6.50867e-17) < 1e-19
    assert abs(sc["sum_hd"] / n - 4.6999e-17) < 1e-19


def test_large_batch_properties(plan, alarm_tree):
    """200k synthetic cases (several waves of tiles): rows sum to 1 per non-evidence variable, evidence rows are 0,
    permuting the cases permutes the results, and the first 2000 match the oracle."""
    n = 200_000
    off, ev, ew = synth.uniform_evidence(alarm_tree["var_dims"], n, 7, seed=20260101)
    labels, post, _ = plan.run(off, ev, ew)
    d = alarm_tree["var_dims"]
    voff = np.concatenate([[0], np.cumsum(d)])
    sums = np.add.reduceat(post, voff[:-1], axis=1)
    is_ev = np.zeros((n, len(d)), bool)
    is_ev[np.repeat(np.arange(n), 7), ev] = True
    assert np.all(np.abs(sums[~is_ev] - 1.0) < 1e-12)
    assert np.all(sums[is_ev] == 0.0)
    o = jtoracle.Oracle(alarm_tree)
    opost, olab = o.run_batch(off[:2001], ev[:2000 * 7], ew[:2000 * 7])
    o.close()
    assert np.max(np.abs(post[:2000] - opost)) <= TOL and np.array_equal(labels[:2000], olab)
    perm = np.random.default_rng(1).permutation(n)
    ev2 = ev.reshape(n, 7)[perm].reshape(-1)
    ew2 = ew.reshape(n, 7)[perm].reshape(-1)
    labels2, post2, _ = plan.run(off, ev2, ew2)
    assert np.array_equal(post2, post[perm]) and np.array_equal(labels2, labels[perm])


@pytest.mark.parametrize("want_post", [False, True])
def test_evaluate_pipeline_matches_reference_printout(plan, alarm_tree, want_post):
    """fbn_evaluate (host buffers, inference + scoring pipelined): reference prints accuracy 1, MSE 6.50867e-17, HD 4.6999e-17"""
    from test_oracle import read_pt
    g = load_golden("alarm_1k_ref")
    gold = read_pt(os.path.join(GOLDEN, "alarm", "alarm_1k_pt"), alarm_tree["var_dims"])
    labels, post, sc, st = plan.evaluate(g["ev_off"], g["ev_vars"], g["ev_vals"], gold, g["truths"], want_post=want_post)
    assert np.array_equal(labels, g["labels"])
    assert sc["n_correct"] == 1000 and sc["n_total"] == 1000
    assert abs(sc["sum_mse"] / 1000 - 6.50867e-17) < 1e-19
    assert abs(sc["sum_hd"] / 1000 - 4.6999e-17) < 1e-19
    if want_post:
        assert np.nanmax(np.abs(post - g["post"])) <= TOL
    other = np.array([int(l.split(" ")[0]) for l in open(os.path.join(GOLDEN, "alarm", "alarm_1k_p20")).read().strip().split("\n")], np.int32)
    _, _, sc2, _ = plan.evaluate(g["ev_off"], g["ev_vars"], g["ev_vals"], gold, other)
    assert sc2["n_correct"] == 513


@pytest.mark.parametrize("team", [0, 4])
def test_duplicate_conflicting_and_invalid_evidence(alarm_tree, team):
    """same pair twice = once; two values for one variable, a value >= dim or < 0 = nothing consistent (NaN, label -1): the
    CUDA path must give what the CPU oracle gives (the reference's DiscreteConfig is a set of pairs)"""
    from test_plan_host import _odd_evidence
    off, ev, ew = _odd_evidence()
    o = jtoracle.Oracle(alarm_tree)
    want, wlab = o.run_batch(off, ev, ew)
    o.close()
    p = capi.Plan(alarm_tree, device=0, team_warps=team)
    labels, post, _ = p.run(off, ev, ew)
    p.close()
    assert np.array_equal(np.isnan(post), np.isnan(want))
    assert np.nanmax(np.abs(post - want)) <= TOL and np.array_equal(labels, wlab)


@pytest.mark.parametrize("max_fx", ["1", "8"])
@pytest.mark.parametrize("team,mode", [(0, 0), (4, 0), (0, 2)])
def test_lambda_gather_path(alarm_tree, team, mode, max_fx, monkeypatch):
    """FBN_MAX_FX caps the evidence patterns per entry: uncovered variables beyond the cap are gathered as lambda rows (the path
    big pattern tables take). Flat, fused and level kernels against the reference fixtures."""
    monkeypatch.setenv("FBN_MAX_FX", max_fx)
    p = capi.Plan(alarm_tree, device=0, team_warps=team, mode=mode)
    for name in ["alarm_1k_ref", "alarm_edge_ref"]:
        g = load_golden(name)
        labels, post, _ = p.run(g["ev_off"], g["ev_vars"], g["ev_vals"])
        assert np.array_equal(np.isnan(post), np.isnan(g["post"]))
        assert np.nanmax(np.abs(post - g["post"])) <= TOL and np.array_equal(labels, g["labels"])
    p.close()


@pytest.mark.parametrize("warps", ["8", "16", "24", "32"])
def test_resident_kernel_many_waves_against_oracle(alarm_tree, warps, monkeypatch):
    """jt_resident_kernel (tile state in shared memory, constants by cp.async.bulk, one CTA per SM): 30 001 cases = more than three
    waves of tiles per CTA plus a ragged last tile, ragged evidence (0..12 pairs per case), EVERY case against the CPU oracle, so
    that the pattern double buffer, the next-tile evidence item and the mbarrier phases of later tiles are checked too"""
    monkeypatch.setenv("FBN_RES_WARPS", warps)
    p = capi.Plan(alarm_tree, device=0)
    assert p.info.flat_records == 2 and p.info.team_warps == int(warps)
    n = 30001
    rng = np.random.default_rng(5)
    counts = rng.integers(0, 13, n)
    off = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
    nv = len(alarm_tree["var_dims"])
    ev = rng.integers(0, nv, off[-1]).astype(np.int32)
    ew = (rng.random(off[-1]) * alarm_tree["var_dims"][ev]).astype(np.int32)
    labels, post, _ = p.run(off, ev, ew)
    p.close()
    o = jtoracle.Oracle(alarm_tree)
    want, wlab = o.run_batch(off, ev, ew)
    o.close()
    assert np.array_equal(np.isnan(post), np.isnan(want))
    assert np.nanmax(np.abs(post - want)) <= TOL and np.array_equal(labels, wlab)


def test_resident_and_flat_kernels_agree(alarm_tree, monkeypatch):
    """FBN_NO_RESIDENT=1 keeps the one-warp-per-tile kernel (tile state in HBM / L2): same fixtures, and both kernels within 1e-12
    of each other (they differ only in where Normalize is applied)"""
    g = load_golden("alarm_synth_4k_ref")
    p = capi.Plan(alarm_tree, device=0)
    assert p.info.flat_records == 2
    l1, p1, _ = p.run(g["ev_off"], g["ev_vars"], g["ev_vals"])
    p.close()
    monkeypatch.setenv("FBN_NO_RESIDENT", "1")
    p = capi.Plan(alarm_tree, device=0)
    assert p.info.flat_records == 1
    l2, p2, _ = p.run(g["ev_off"], g["ev_vars"], g["ev_vals"])
    p.close()
    for lab, post in ((l1, p1), (l2, p2)):
        assert np.nanmax(np.abs(post - g["post"])) <= TOL and np.array_equal(lab, g["labels"])
    assert np.nanmax(np.abs(p1 - p2)) <= 1e-12


def test_deep_chain_runs_the_flat_kernel(tmp_path, monkeypatch):
    """a small tree the resident planner declines (a chain of 150 small cliques: one CTA barrier per level): AUTO falls back to
    jt_flat_kernel; results against the CPU oracle"""
    from fastbn_b200 import hostapi
    monkeypatch.delenv("FBN_RES_MIN_WORK", raising=False)
    var_dims, parents, cpts = synth.window_network(150, (2, 3), 1, 1, 21)
    xml = str(tmp_path / "chain.xml")
    synth.write_xmlbif(xml, var_dims, parents, cpts)
    m = hostapi.HostModel(xml)
    tree = m.tree()
    m.close()
    off, ev, ew = synth.uniform_evidence(tree["var_dims"], 3000, 20, seed=3)
    p = capi.Plan(tree, device=0)
    assert p.info.flat_records == 1
    labels, post, _ = p.run(off, ev, ew)
    p.close()
    o = jtoracle.Oracle(tree)
    want, wlab = o.run_batch(off, ev, ew)
    o.close()
    assert np.array_equal(np.isnan(post), np.isnan(want))
    assert np.nanmax(np.abs(post - want)) <= TOL and np.array_equal(labels, wlab)


def test_forward_sampled_set_10k_against_oracle(alarm_tree):
    """the second evidence set (values forward-sampled from the network, seed 20260102) and the parity
    subset of the first 10 000 cases through the oracle: labels bit-identical, posteriors <= 1e-9"""
    off, ev, ew = synth.forward_sampled_evidence(os.path.join(GOLDEN, "alarm", "alarm.xml"), 10000, 7, 20260102)
    p = capi.Plan(alarm_tree, device=0)
    labels, post, _ = p.run(off, ev, ew)
    p.close()
    o = jtoracle.Oracle(alarm_tree)
    want, wlab = o.run_batch(off, ev, ew)
    o.close()
    assert not np.isnan(post).any()   # a forward sample never has probability zero
    assert np.max(np.abs(post - want)) <= TOL and np.array_equal(labels, wlab)
    off, ev, ew = synth.uniform_evidence(alarm_tree["var_dims"], 10000, 7, seed=20260101)   # the first set
    p = capi.Plan(alarm_tree, device=0)
    labels, post, _ = p.run(off, ev, ew)
    p.close()
    o = jtoracle.Oracle(alarm_tree)
    want, wlab = o.run_batch(off, ev, ew)
    o.close()
    assert np.nanmax(np.abs(post - want)) <= TOL and np.array_equal(labels, wlab)
