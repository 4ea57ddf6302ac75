"""CPU tests of the host side: the planner (fastbn_b200/csrc/plan.cpp) interpreted by oracle/plan_emu.cpp must
reproduce the reference's posteriors for every team size, and must reject malformed trees; the C-ABI library must
load and export every symbol include/fastbn_b200.h declares (no compute without a GPU)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import emu
from conftest import ROOT, load_golden
from fastbn_b200 import capi
from oracle import jtoracle


@pytest.mark.parametrize("team", [1, 2, 8, 32])
@pytest.mark.parametrize("name", ["alarm_1k_ref", "alarm_edge_ref"])
def test_plan_reproduces_reference(alarm_tree, team, name):
    g = load_golden(name)
    post, lab, info = emu.run(alarm_tree, team, g["ev_off"], g["ev_vars"], g["ev_vals"])
    assert np.array_equal(np.isnan(post), np.isnan(g["post"]))
    assert np.nanmax(np.abs(post - g["post"])) <= 1e-9
    assert np.array_equal(lab, g["labels"])
    assert info[6] == 15  # reference prints "maximum level of the junction tree = 15"
    assert info[7] == 1517  # sum over edges of |parent clique|


@pytest.mark.parametrize("gang", [8, 24, 504])
def test_level_mode_plan_reproduces_reference(alarm_tree, gang):
    """level mode (one launch per tree level, `gang` warps over several CTAs per tile): negative team selects it in the emulator"""
    g = load_golden("alarm_1k_ref")
    post, lab, info = emu.run(alarm_tree, -gang, g["ev_off"][:201], g["ev_vars"][:g["ev_off"][200]], g["ev_vals"][:g["ev_off"][200]])
    assert np.nanmax(np.abs(post - g["post"][:200])) <= 1e-9
    assert np.array_equal(lab, g["labels"][:200])
    g = load_golden("alarm_edge_ref")
    post, lab, _ = emu.run(alarm_tree, -gang, g["ev_off"], g["ev_vars"], g["ev_vals"])
    assert np.array_equal(np.isnan(post), np.isnan(g["post"]))
    assert np.nanmax(np.abs(post - g["post"])) <= 1e-9


def test_planner_rejects_malformed_trees(alarm_tree):
    z = np.zeros(1, np.int64)
    e = np.zeros(0, np.int32)
    t = dict(alarm_tree)
    t["root"] = np.int32(3)  # a non-root clique as root: it has an upstream separator
    with pytest.raises(ValueError):
        emu.run(t, 1, z[:1].repeat(2), e, e)
    t = dict(alarm_tree)
    tv = t["table_vars"].copy()
    tv[0], tv[1] = tv[1], tv[0]  # breaks "child ends with its separator's variables" or the scope check
    t["table_vars"] = tv
    with pytest.raises(ValueError):
        emu.run(t, 1, z[:1].repeat(2), e, e)
    t = dict(alarm_tree)
    sp = t["sep_child"].copy()
    sp[0] = sp[1]
    t["sep_child"] = sp
    with pytest.raises(ValueError):
        emu.run(t, 1, z[:1].repeat(2), e, e)
    with pytest.raises(ValueError):
        emu.run(alarm_tree, 3, z[:1].repeat(2), e, e)  # team must be a power of two


def test_abi_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "fastbn_b200.h")).read()
    declared = set(re.findall(r"\b(fbn_[a-z_]+)\s*\(", hdr))
    assert declared == set(capi.ABI_SYMBOLS)
    L = capi.load_library()
    for s in declared:
        assert hasattr(L, s), s


def test_no_cpu_fallback_without_device(alarm_tree):
    L = capi.load_library()
    if L.fbn_device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(capi.FastBNError) as ei:
        capi.Plan(alarm_tree)
    assert "no CUDA device" in str(ei.value)


@pytest.mark.parametrize("team", [32, -64])
def test_plan_with_permuted_block_order_on_pigs_shape(team):
    """PIGS-shape (max clique 3^10): the planner visits the column blocks of big cliques in a locality order
    (SumOp::perm_off); the emulated plan must still reproduce the unmodified reference (fused team 32 and level mode)."""
    from fastbn_b200 import hostapi
    m = hostapi.HostModel(os.path.join(ROOT, "tests", "golden", "synth", "pigs.xml"))
    t = m.tree()
    m.close()
    g = load_golden(os.path.join("synth", "pigs_ref"))
    n = 6
    k = g["ev_off"][n]
    post, lab, info = emu.run(t, team, g["ev_off"][:n + 1], g["ev_vars"][:k], g["ev_vals"][:k])
    assert np.nanmax(np.abs(post - g["post"][:n])) <= 1e-9
    assert np.array_equal(lab, g["labels"][:n])


@pytest.mark.parametrize("no_flat", [False, True])
def test_flat_records_of_small_trees(alarm_tree, no_flat, monkeypatch):
    """team 1 plans of small trees carry flat per-entry records (jt_flat_kernel); the emulator interprets them when present.
    Both forms of the plan must reproduce the reference."""
    if no_flat:
        monkeypatch.setenv("FBN_NO_FLAT", "1")
    for name in ["alarm_1k_ref", "alarm_edge_ref"]:
        g = load_golden(name)
        post, lab, _ = emu.run(alarm_tree, 1, g["ev_off"], g["ev_vars"], g["ev_vals"])
        assert np.array_equal(np.isnan(post), np.isnan(g["post"]))
        assert np.nanmax(np.abs(post - g["post"])) <= 1e-9
        assert np.array_equal(lab, g["labels"])


@pytest.mark.parametrize("team", [8, 16])
def test_unit_order_of_a_phase_does_not_change_results(team, monkeypatch):
    """Fused plans deal the most expensive units of a phase first (longest-processing-time order, plan.cpp emit_phase) and keep
    the per-warp shares of the clique norms in max_normslots x team scratch rows appended to the tile. Ops of one phase are
    independent; only the grouping of a clique norm's per-warp shares follows the rotation, so the plan with the ops in storage
    order (FBN_NO_LPT) must give the same posteriors up to fp64 re-association (1e-13), identical labels, and both must
    reproduce the unmodified reference (PIGS shape, 16 warps per tile is what AUTO picks)."""
    from fastbn_b200 import hostapi
    m = hostapi.HostModel(os.path.join(ROOT, "tests", "golden", "synth", "pigs.xml"))
    t = m.tree()
    m.close()
    g = load_golden(os.path.join("synth", "pigs_ref"))
    n = 4
    k = g["ev_off"][n]
    args = (g["ev_off"][:n + 1], g["ev_vars"][:k], g["ev_vals"][:k])
    post, lab, info = emu.run(t, team, *args)
    monkeypatch.setenv("FBN_NO_LPT", "1")
    post0, lab0, info0 = emu.run(t, team, *args)
    assert np.nanmax(np.abs(post - post0)) <= 1e-13 and np.array_equal(lab, lab0)
    assert np.nanmax(np.abs(post - g["post"][:n])) <= 1e-9
    assert np.array_equal(lab, g["labels"][:n])
    assert info[1] == info0[1] and info[3] == info0[3]  # same units and partial slots, only their order differs
    t1 = emu.run(t, 1, *args)[2]
    assert info[5] >= t1[5] + info[4] * team  # tile rows: tables + messages + one norm row per (clique of a phase, warp)


def _odd_evidence():
    """duplicate pair (same value twice), conflicting values for one variable, a value >= dim, a negative value, a plain case"""
    vl = [[5, 5, 9], [5, 5, 9], [7], [7], [3, 9]]
    wl = [[1, 1, 0], [0, 1, 0], [9], [-1], [1, 2]]
    off = np.concatenate([[0], np.cumsum([len(v) for v in vl])]).astype(np.int64)
    return off, np.concatenate(vl).astype(np.int32), np.concatenate(wl).astype(np.int32)


@pytest.mark.parametrize("team", [1, 4, -16])
def test_evidence_patterns_duplicates_conflicts_and_invalid_values(alarm_tree, team):
    """Evidence is a set of (variable, value) pairs in the reference (DiscreteConfig): the same pair twice is harmless, two
    different values for one variable or a value outside [0, dim) leave nothing consistent (NaN posteriors, label -1). The
    planner's evidence patterns (0 / 1 + state / dim + 1) must give exactly what the CPU oracle gives."""
    from oracle import jtoracle
    off, ev, ew = _odd_evidence()
    o = jtoracle.Oracle(alarm_tree)
    want, wlab = o.run_batch(off, ev, ew)
    o.close()
    post, lab, _ = emu.run(alarm_tree, team, off, ev, ew)
    assert np.array_equal(np.isnan(post), np.isnan(want)) and np.isnan(want[1:4]).any(axis=1).all() and not np.isnan(want[[0, 4]]).any()
    assert np.nanmax(np.abs(post - want)) <= 1e-12 and np.array_equal(lab, wlab)


@pytest.mark.parametrize("max_fx", ["1", "8", "1024"])
@pytest.mark.parametrize("team", [1, 8])
def test_folded_and_gathered_evidence_agree(alarm_tree, team, max_fx, monkeypatch):
    """An uncovered variable of a view is either folded into per-pattern potentials or gathered as lambda rows
    (FBN_MAX_FX caps the patterns per entry: 1 = gather everything, 1024 = fold everything). Every split must reproduce the
    reference, on ALARM and on the HEPAR2 shape."""
    from fastbn_b200 import hostapi
    monkeypatch.setenv("FBN_MAX_FX", max_fx)
    g = load_golden("alarm_1k_ref")
    n = 300
    k = g["ev_off"][n]
    post, lab, info = emu.run(alarm_tree, team, g["ev_off"][:n + 1], g["ev_vars"][:k], g["ev_vals"][:k])
    assert np.nanmax(np.abs(post - g["post"][:n])) <= 1e-9 and np.array_equal(lab, g["labels"][:n])
    assert (info[13] > 0) == (max_fx != "1024") and (info[12] > 0) == (max_fx != "1")  # lambda gathers / folded variables
    g = load_golden("alarm_edge_ref")
    post, lab, _ = emu.run(alarm_tree, team, g["ev_off"], g["ev_vars"], g["ev_vals"])
    assert np.array_equal(np.isnan(post), np.isnan(g["post"])) and np.nanmax(np.abs(post - g["post"])) <= 1e-9
    m = hostapi.HostModel(os.path.join(ROOT, "tests", "golden", "synth", "hepar2.xml"))
    t = m.tree()
    m.close()
    g = load_golden(os.path.join("synth", "hepar2_ref"))
    idx = list(range(8)) + list(range(256, 264))
    off = np.concatenate([[0], np.cumsum([g["ev_off"][i + 1] - g["ev_off"][i] for i in idx])]).astype(np.int64)
    ev = np.concatenate([g["ev_vars"][g["ev_off"][i]:g["ev_off"][i + 1]] for i in idx])
    ew = np.concatenate([g["ev_vals"][g["ev_off"][i]:g["ev_off"][i + 1]] for i in idx])
    post, lab, _ = emu.run(t, team, off, ev, ew)
    assert np.nanmax(np.abs(post - g["post"][idx])) <= 1e-9 and np.array_equal(lab, g["labels"][idx])


# ---- resident-tile plan (fastbn_b200/csrc/resident.cpp), interpreted block by block like jt_resident_kernel ----
@pytest.mark.parametrize("warps", [1, 8, 16, 32])
def test_resident_plan_reproduces_reference(alarm_tree, warps):
    """live-range row allocation (the emulator's state has only the plan's live rows and is poisoned per case), static per-warp
    schedule, Normalize left to the final marginals, per-target factor tables: reference fixtures incl. the NaN patterns"""
    for name in ["alarm_1k_ref", "alarm_edge_ref"]:
        g = load_golden(name)
        post, lab, info = emu.run_resident(alarm_tree, warps, g["ev_off"], g["ev_vars"], g["ev_vals"])
        assert np.array_equal(np.isnan(post), np.isnan(g["post"]))
        assert np.nanmax(np.abs(post - g["post"])) <= 1e-9
        assert np.array_equal(lab, g["labels"])
        assert info[0] == 1 and info[1] < info[7] // 2   # ALARM: 775 scratch rows -> fewer than half live at once
        assert info[3] <= 232448   # fits one SM's shared memory


@pytest.mark.parametrize("max_fx", ["1", "8"])
def test_resident_plan_with_lambda_rows(alarm_tree, max_fx, monkeypatch):
    """views that gather evidence indicators (pattern tables capped by FBN_MAX_FX): lambda rows keep state rows for good"""
    monkeypatch.setenv("FBN_MAX_FX", max_fx)
    for name in ["alarm_1k_ref", "alarm_edge_ref"]:
        g = load_golden(name)
        # with nothing folded (cap 1) ALARM's state + constant blocks are 320 bytes over one SM's shared memory (the flat kernel
        # would run it): the schedule is emulated all the same
        post, lab, info = emu.run_resident(alarm_tree, 16, g["ev_off"], g["ev_vars"], g["ev_vals"], smem_limit=240000)
        assert np.array_equal(np.isnan(post), np.isnan(g["post"]))
        assert np.nanmax(np.abs(post - g["post"])) <= 1e-9 and np.array_equal(lab, g["labels"])


def test_resident_plan_odd_evidence_and_eligibility(alarm_tree):
    off, ev, ew = _odd_evidence()
    o = jtoracle.Oracle(alarm_tree)
    want, wlab = o.run_batch(off, ev, ew)
    o.close()
    post, lab, _ = emu.run_resident(alarm_tree, 16, off, ev, ew)
    assert np.array_equal(np.isnan(post), np.isnan(want))
    assert np.nanmax(np.abs(post - want)) <= 1e-9 and np.array_equal(lab, wlab)
    with pytest.raises(ValueError, match="shared memory"):   # a device with less shared memory: not eligible, the flat kernel runs the tree
        emu.run_resident(alarm_tree, 16, off, ev, ew, smem_limit=100000)
    t = dict(alarm_tree)
    t["clique_potentials"] = alarm_tree["clique_potentials"] * 1e-40   # unnormalised messages could underflow: not eligible
    with pytest.raises(ValueError, match="range of a double"):
        emu.run_resident(t, 16, off, ev, ew)


def test_resident_plan_declines_deep_chains(tmp_path, monkeypatch):
    """a chain of 150 small cliques: one CTA barrier per tree level would dominate -> not eligible, jt_flat_kernel runs such trees"""
    from fastbn_b200 import hostapi, synth
    monkeypatch.delenv("FBN_RES_MIN_WORK", raising=False)
    var_dims, parents, cpts = synth.window_network(150, (2, 3), 1, 1, 21)
    xml = str(tmp_path / "chain.xml")
    synth.write_xmlbif(xml, var_dims, parents, cpts)
    m = hostapi.HostModel(xml)
    tree = m.tree()
    m.close()
    off, ev, ew = synth.uniform_evidence(tree["var_dims"], 8, 20, seed=3)
    with pytest.raises(ValueError, match="too little work per phase"):
        emu.run_resident(tree, 24, off, ev, ew)
    post, lab, _ = emu.run(tree, 1, off, ev, ew)   # the flat plan of the same tree against the oracle
    o = jtoracle.Oracle(tree)
    want, wlab = o.run_batch(off, ev, ew)
    o.close()
    assert np.nanmax(np.abs(post - want)) <= 1e-9 and np.array_equal(lab, wlab)


def test_forward_sampled_evidence_set(alarm_tree):
    """the second evidence set of SURVEY.md 8(d) config 2 (forward samples of the network): 7 distinct non-query variables per case,
    states inside the domains, typical evidence (no case of probability zero); flat and resident plans against the oracle"""
    from fastbn_b200 import synth
    from conftest import GOLDEN
    import os
    off, ev, ew = synth.forward_sampled_evidence(os.path.join(GOLDEN, "alarm", "alarm.xml"), 600, 7, 20260102)
    assert np.array_equal(np.diff(off), np.full(600, 7))
    e2 = ev.reshape(600, 7)
    assert (np.diff(e2, axis=1) > 0).all() and e2.min() >= 1 and e2.max() <= 36
    assert (ew >= 0).all() and (ew < alarm_tree["var_dims"][ev]).all()
    o = jtoracle.Oracle(alarm_tree)
    want, wlab = o.run_batch(off, ev, ew)
    o.close()
    assert not np.isnan(want).any()
    for post, lab in (emu.run(alarm_tree, 1, off, ev, ew)[:2], emu.run_resident(alarm_tree, 24, off, ev, ew)[:2]):
        assert np.max(np.abs(post - want)) <= 1e-9 and np.array_equal(lab, wlab)
