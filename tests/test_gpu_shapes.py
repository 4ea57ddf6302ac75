"""GPU parity on the synthetic shapes of BASELINE.json configs 3-5 and on the host mirror / CLI (run with -m gpu).
Trees are built by the C++ host mirror from the committed XMLBIF files; expected posteriors come from the unmodified
reference (tests/golden/synth/*_ref.npz). Tolerance: posteriors max-abs <= 1e-9, labels bit-identical."""
import os
import subprocess

import numpy as np
import pytest

from conftest import GOLDEN, ROOT, load_golden
from fastbn_b200 import capi, hostapi, synth
from oracle import jtoracle

pytestmark = pytest.mark.gpu
TOL = 1e-9


def _shape(name):
    m = hostapi.HostModel(os.path.join(GOLDEN, "synth", name + ".xml"))
    t = m.tree()
    m.close()
    return t, load_golden(os.path.join("synth", name + "_ref"))


@pytest.mark.parametrize("name,team", [("hepar2", 0), ("hepar2", 1), ("hepar2", 4), ("hepar2", 8), ("hepar2", 16), ("pigs", 0), ("pigs", 2), ("pigs", 8), ("pigs", 16), ("pigs", 32)])
def test_shape_matches_reference(name, team):
    t, g = _shape(name)
    p = capi.Plan(t, device=0, team_warps=team)
    labels, post, st = p.run(g["ev_off"], g["ev_vars"], g["ev_vals"])
    p.close()
    assert np.nanmax(np.abs(post - g["post"])) <= TOL
    assert np.array_equal(labels, g["labels"])


@pytest.mark.parametrize("name,team,max_resident", [("hepar2", 0, 0), ("hepar2", 8, 3), ("hepar2", 200, 5), ("pigs", 0, 0), ("pigs", 64, 2),
                                                    ("pigs", 1000, 1)])
def test_level_mode_matches_reference(name, team, max_resident):
    """one launch per tree level, many CTAs per tile (the MUNIN-shape path), incl. several waves of few resident tiles"""
    t, g = _shape(name)
    p = capi.Plan(t, device=0, team_warps=team, max_resident=max_resident, mode=capi.MODE_LEVEL)
    assert p.info.mode == capi.MODE_LEVEL and p.info.launches_per_wave > p.info.n_phases
    labels, post, st = p.run(g["ev_off"], g["ev_vars"], g["ev_vals"])
    p.close()
    assert np.nanmax(np.abs(post - g["post"])) <= TOL
    assert np.array_equal(labels, g["labels"])
    assert st.n_launches >= p.info.launches_per_wave


@pytest.mark.parametrize("name", ["munin_small", "munin"])
def test_munin_shape_auto_level_mode(name):
    """BASELINE.json config 5: max clique 5^9 / 5^10 entries, intra-table multi-CTA split. Expected values: the unmodified
    reference's posteriors for the committed cases (0 % evidence + cases with 24 evidence variables)."""
    t, g = _shape(name)
    p = capi.Plan(t, device=0)
    assert p.info.mode == capi.MODE_LEVEL
    labels, post, st = p.run(g["ev_off"], g["ev_vars"], g["ev_vals"])
    assert np.nanmax(np.abs(post - g["post"])) <= TOL
    assert np.array_equal(labels, g["labels"])
    # 0 % evidence: many identical cases (several tiles) must give identical rows, equal to the reference's
    n = 100
    labels2, post2, _ = p.run(np.zeros(n + 1, np.int64), np.zeros(0, np.int32), np.zeros(0, np.int32))
    p.close()
    assert np.all(post2 == post2[0]) and np.all(labels2 == labels2[0])
    assert np.max(np.abs(post2[0] - g["post"][0])) <= TOL


def test_level_mode_alarm_fixtures_and_nan(alarm_tree):
    p = capi.Plan(alarm_tree, device=0, mode=capi.MODE_LEVEL)
    for name in ["alarm_1k_ref", "alarm_edge_ref"]:
        g = load_golden(name)
        labels, post, _ = p.run(g["ev_off"], g["ev_vars"], g["ev_vals"])
        assert np.array_equal(np.isnan(post), np.isnan(g["post"]))
        assert np.nanmax(np.abs(post - g["post"])) <= TOL
        assert np.array_equal(labels, g["labels"])
    p.close()


@pytest.mark.parametrize("name,n,nev", [("hepar2", 20000, 14), ("hepar2", 5000, 35), ("pigs", 3000, 88)])
def test_shape_batch_properties(name, n, nev):
    """many tiles: every non-evidence variable sums to 1, evidence variables are zero rows, repeated cases give repeated
    results, and a seeded prefix matches the CPU oracle."""
    t, _ = _shape(name)
    d = t["var_dims"]
    off, ev, ew = synth.uniform_evidence(d, n, nev, seed=4242)
    ev[-nev:] = ev[:nev]      # last case == first case
    ew[-nev:] = ew[:nev]
    p = capi.Plan(t, device=0)
    labels, post, _ = p.run(off, ev, ew)
    p.close()
    voff = np.concatenate([[0], np.cumsum(d)])
    sums = np.add.reduceat(post, voff[:-1], axis=1)
    is_ev = np.zeros((n, len(d)), bool)
    is_ev[np.repeat(np.arange(n), nev), ev] = True
    assert np.all(np.abs(sums[~is_ev] - 1.0) < 1e-10)
    assert np.all(sums[is_ev] == 0.0)
    assert np.array_equal(post[0], post[-1]) and labels[0] == labels[-1]
    k = 256  # seeded prefix through the CPU oracle (pinned bit-exact to the reference, tests/test_oracle.py)
    o = jtoracle.Oracle(t)
    opost, olab = o.run_batch(off[:k + 1], ev[:k * nev], ew[:k * nev])
    o.close()
    assert np.max(np.abs(post[:k] - opost)) <= TOL and np.array_equal(labels[:k], olab)


def test_host_mirror_evaluate_alarm():
    """JunctionTree::EvaluateAccuracy of the C++ mirror on the reference's default run: accuracy 1, MSE 6.50867e-17, HD 4.6999e-17"""
    g = load_golden("alarm_1k_ref")
    m = hostapi.HostModel(os.path.join(GOLDEN, "alarm", "alarm.xml"), os.path.join(GOLDEN, "alarm", "testing_alarm_1k_p20"))
    r = m.evaluate(os.path.join(GOLDEN, "alarm", "alarm_1k_pt"), want_post=True)
    assert r["accuracy"] == 1.0
    assert abs(r["average_mse"] - 6.50867e-17) < 1e-19 and abs(r["average_hd"] - 4.6999e-17) < 1e-19
    assert np.array_equal(r["labels"], g["labels"])
    assert np.nanmax(np.abs(r["post"] - g["post"])) <= TOL
    m.close()
    m = hostapi.HostModel(os.path.join(GOLDEN, "alarm", "alarm.xml"), os.path.join(GOLDEN, "alarm", "alarm_1k_p20"))
    r = m.evaluate(os.path.join(GOLDEN, "alarm", "alarm_1k_pt"))
    assert abs(r["accuracy"] - 0.513) < 1e-12
    m.close()


def test_cli_default_run_prints_reference_lines(tmp_path):
    """`BayesianNetwork` with no flags = `-a 2 -t 1` on ALARM (reference README / SURVEY.md 8(c) fixtures); -o writes a pt
    file that round-trips through the loader to MSE 0 against itself and reproduces the shipped golden file to 7 decimals."""
    exe = os.path.join(ROOT, "fastbn_b200", "bin", "BayesianNetwork")
    ds = tmp_path / "dataset"
    bd = tmp_path / "build"
    ds.mkdir()
    bd.mkdir()
    os.symlink(os.path.join(GOLDEN, "alarm"), ds / "alarm")
    r = subprocess.run([exe], cwd=bd, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    for line in ["number of cliques = 27, number of separators = 26", "ave clique size = 41.1111, max size = 144",
                 "maximum level of the junction tree = 15", "average MSE = 6.50867e-17", "average HD = 4.6999e-17", "accuracy = 1"]:
        assert line in r.stdout, line
    assert "jt: " in r.stdout
    # the reference's progress lines (one per n/20 cases) and timer block names (src/JunctionTree.cpp:76-84, :1509-1524)
    for line in ["Progress indicator: ", "5.000000%...", "100.000000%...", "load evidence: ", "msg passing: ", "reset: ", "predict: "]:
        assert line in r.stdout, line
    assert r.stdout.count("%...") == 20
    r = subprocess.run([exe, "-a", "2", "-t", "4", "-f3", "alarm/alarm_1k_p20"], cwd=bd, capture_output=True, text=True, timeout=300)
    assert "accuracy = 0.513" in r.stdout
    out = tmp_path / "out_pt"
    r = subprocess.run([exe, "-D", str(ds), "-o", "../out_pt"], cwd=bd, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert open(out).read() == open(os.path.join(GOLDEN, "alarm", "alarm_1k_pt")).read()
