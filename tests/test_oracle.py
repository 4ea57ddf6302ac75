"""CPU tests: the oracle (oracle/jt_oracle.c) is pinned against the reference's own fixtures and against fp64
outputs of the unmodified reference (tests/golden/*_ref.npz, made by tools/make_golden.py)."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, load_golden
from oracle import jtoracle, refdrv


@pytest.mark.parametrize("name", ["alarm_1k_ref", "alarm_synth_4k_ref", "alarm_edge_ref"])
def test_oracle_matches_reference_fp64(alarm_tree, name):
    g = load_golden(name)
    o = jtoracle.Oracle(alarm_tree)
    post, lab = o.run_batch(g["ev_off"], g["ev_vars"], g["ev_vals"])
    o.close()
    assert np.array_equal(np.isnan(post), np.isnan(g["post"]))
    assert np.nanmax(np.abs(post - g["post"])) <= 1e-12  # observed: bit-identical (0.0)
    assert np.array_equal(lab, g["labels"])


def read_pt(path, var_dims):
    """dataset/alarm/alarm_1k_pt: one line per (case, node), '%.7f ' per state, empty line = evidence node."""
    off = np.concatenate([[0], np.cumsum(var_dims)])
    lines = open(path).read().split("\n")
    n_vars = len(var_dims)
    n = len(lines) // n_vars
    out = np.full((n, off[-1]), -1.0)
    for i in range(n):
        for v in range(n_vars):
            ln = lines[i * n_vars + v].strip()
            if ln:
                out[i, off[v]:off[v + 1]] = [float(x) for x in ln.split(" ")]
    return out


def test_oracle_matches_shipped_golden_pt(alarm_tree):
    """The reference's only self-test: posteriors vs alarm_1k_pt after rounding to 7 decimals, accuracy 1 / 0.513."""
    g = load_golden("alarm_1k_ref")
    gold = read_pt(os.path.join(GOLDEN, "alarm", "alarm_1k_pt"), alarm_tree["var_dims"])
    assert gold.shape == (1000, 105)
    o = jtoracle.Oracle(alarm_tree)
    post, lab = o.run_batch(g["ev_off"], g["ev_vars"], g["ev_vals"])
    o.close()
    m = gold >= 0
    assert np.max(np.abs(post[m] - gold[m])) <= 5.1e-8  # pure 7-decimal rounding
    assert np.all(post[~m] == 0)  # evidence nodes are printed as empty lines / returned as zeros
    assert np.mean(lab == g["truths"]) == 1.0
    other = [int(l.split(" ")[0]) for l in open(os.path.join(GOLDEN, "alarm", "alarm_1k_p20")).read().strip().split("\n")]
    assert abs(np.mean(lab == np.array(other)) - 0.513) < 1e-12


@pytest.mark.skipif(not refdrv.available(), reason="oracle/_ref not built (needs /root/reference)")
def test_oracle_matches_live_reference_on_fresh_cases(alarm_tree):
    from fastbn_b200 import synth
    ref = refdrv.Reference(os.path.join(GOLDEN, "alarm", "alarm.xml"), os.path.join(GOLDEN, "alarm", "testing_alarm_1k_p20"))
    tree = ref.tree()
    # The reference breaks Prim ties by POINTER order (src/JunctionTreeStructure.cpp:264-277), so which clique becomes a
    # separator's parent can change between processes; the cliques themselves and the CPT products may not.
    for k in ("var_dims", "n_cliques", "n_separators"):
        assert np.array_equal(tree[k], alarm_tree[k]), "reference tree differs from the committed fixture: " + k

    def clique_sets(t):
        pos = np.concatenate([[0], np.cumsum(t["table_nvars"])])
        return [frozenset(t["table_vars"][pos[i]:pos[i + 1]].tolist()) for i in range(int(t["n_cliques"]))]
    assert clique_sets(tree) == clique_sets(alarm_tree)
    off, ev, ew = synth.uniform_evidence(tree["var_dims"], 300, 11, seed=99)
    rpost, rlab, _ = ref.run_batch(off, ev, ew)
    assert abs(ref.evaluate_accuracy(os.path.join(GOLDEN, "alarm", "alarm_1k_pt"), 1) - 1.0) < 1e-15
    ref.close()
    o = jtoracle.Oracle(tree)
    post, lab = o.run_batch(off, ev, ew)
    o.close()
    assert np.nanmax(np.abs(post - rpost)) <= 1e-12
    assert np.array_equal(lab, rlab)
    # calibrated posteriors do not depend on the tree shape: the committed fixture tree gives the same answers
    o = jtoracle.Oracle(alarm_tree)
    post2, lab2 = o.run_batch(off, ev, ew)
    o.close()
    assert np.nanmax(np.abs(post2 - rpost)) <= 1e-12
    assert np.array_equal(lab2, rlab)
