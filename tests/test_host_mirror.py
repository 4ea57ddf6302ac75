"""CPU tests of the C++ host mirror (fastbn_b200/host/): XMLBIF loader, LIBSVM loader, junction-tree construction.
The tree it builds for ALARM must be the reference's tree (fixture read from the live reference objects), and on the
synthetic shapes the oracle run on OUR tree must reproduce the reference's fp64 posteriors (tests/golden/synth/*_ref.npz,
made by tools/make_synth_golden.py with the unmodified reference)."""
import os
import re
import subprocess

import numpy as np
import pytest

import emu
from conftest import GOLDEN, ROOT, load_golden
from fastbn_b200 import capi, hostapi, synth
from oracle import jtoracle

ALARM_XML = os.path.join(GOLDEN, "alarm", "alarm.xml")
ALARM_CASES = os.path.join(GOLDEN, "alarm", "testing_alarm_1k_p20")


def test_alarm_tree_is_the_reference_tree(alarm_tree):
    m = hostapi.HostModel(ALARM_XML, ALARM_CASES)
    t = m.tree()
    assert (m.n_cliques, m.n_separators, m.max_level, m.max_clique) == (27, 26, 15, 144)  # the reference's printout
    for k in ("var_dims", "table_nvars", "table_vars", "root", "query_var"):
        assert np.array_equal(t[k], alarm_tree[k]), k
    # CPT products: same counts (int)(p*10000), same Laplace smoothing, same multiplication order -> bit-identical
    assert np.array_equal(t["clique_potentials"], alarm_tree["clique_potentials"])
    # separators: same set of (variables, {two cliques}); the orientation of an edge may differ (pointer-order tie-breaks)
    def edges(tr):
        pos = np.concatenate([[0], np.cumsum(tr["table_nvars"])])
        nc = int(tr["n_cliques"])
        return sorted((tuple(tr["table_vars"][pos[nc + s]:pos[nc + s + 1]]), frozenset((int(tr["sep_parent"][s]), int(tr["sep_child"][s]))))
                      for s in range(int(tr["n_separators"])))
    assert [e[0] for e in edges(t)] == [e[0] for e in edges(alarm_tree)]
    m.close()


def test_alarm_cases_and_posteriors_through_native_tree():
    g = load_golden("alarm_1k_ref")
    m = hostapi.HostModel(ALARM_XML, ALARM_CASES)
    off, ev, ew, truths = m.cases()
    assert m.n_cases == 1000 and np.array_equal(off, g["ev_off"]) and np.array_equal(truths, g["truths"])
    assert np.array_equal(ev, g["ev_vars"]) and np.array_equal(ew, g["ev_vals"])  # set order: ascending variable
    o = jtoracle.Oracle(m.tree())
    post, lab = o.run_batch(off, ev, ew)
    o.close()
    m.close()
    assert np.nanmax(np.abs(post - g["post"])) <= 1e-12
    assert np.array_equal(lab, g["labels"])


@pytest.mark.parametrize("name", ["hepar2", "pigs", "munin_small"])
def test_synthetic_shapes_match_reference(name, tmp_path):
    xml = os.path.join(GOLDEN, "synth", name + ".xml")
    g = load_golden(os.path.join("synth", name + "_ref"))
    # the committed XML is what the generator writes for the recorded seed
    regen = str(tmp_path / "regen.xml")
    synth.make_shape(name, regen)
    assert open(regen).read() == open(xml).read()
    m = hostapi.HostModel(xml)
    assert [m.n_cliques, m.n_separators, m.max_level, m.max_clique, m.clique_entries] == g["ref_tree_stats"].tolist()
    assert m.max_clique == synth.SHAPES[name]["max_clique"]
    t = m.tree()
    m.close()
    n = {"hepar2": 64, "pigs": 16, "munin_small": 2}[name]
    e = int(g["ev_off"][n])
    o = jtoracle.Oracle(t)
    post, lab = o.run_batch(g["ev_off"][:n + 1], g["ev_vars"][:e], g["ev_vals"][:e])
    o.close()
    assert np.nanmax(np.abs(post - g["post"][:n])) <= 1e-12
    assert np.array_equal(lab, g["labels"][:n])
    # and the device plan built from that tree, interpreted on the CPU
    team = -120 if name == "munin_small" else 4  # MUNIN-shape runs in level mode (many CTAs per tile)
    post2, lab2, _ = emu.run(t, team, g["ev_off"][:n + 1], g["ev_vars"][:e], g["ev_vals"][:e])
    assert np.nanmax(np.abs(post2 - g["post"][:n])) <= 1e-9
    assert np.array_equal(lab2, g["labels"][:n])


def test_libsvm_loader_quirks(tmp_path):
    """last line without a newline is dropped; label-only line = no evidence; unsorted / duplicate pairs become a set"""
    p = str(tmp_path / "cases")
    with open(p, "w") as f:
        f.write("1 3:1 2:0 \n0\n2 5:1 5:1 4:2   \n1 7:0")
    m = hostapi.HostModel(ALARM_XML, p)
    off, ev, ew, truths = m.cases()
    m.close()
    assert truths.tolist() == [1, 0, 2]
    assert off.tolist() == [0, 2, 2, 4]
    assert ev.tolist() == [2, 3, 4, 5] and ew.tolist() == [0, 1, 2, 1]


def test_xmlbif_count_truncation_and_laplace():
    """(int)(p*10000) truncates (0.69 -> 6899) and P = (count+1)/(total+dim): no zero entries, SURVEY.md 8(c) quirk 1"""
    m = hostapi.HostModel(ALARM_XML)
    t = m.tree()
    m.close()
    assert np.all(t["clique_potentials"] > 0)
    # clique C0 = [0 5] holds P(HISTORY | LVFAILURE) * (nothing else containing var 0): 0.9 0.01 / 0.1 0.99
    pos = np.concatenate([[0], np.cumsum(t["table_nvars"])])
    assert t["table_vars"][pos[0]:pos[1]].tolist() == [0, 5]


def test_host_abi_symbols_exported():
    hdr = open(os.path.join(ROOT, "include", "fastbn_b200_host.h")).read()
    declared = set(re.findall(r"\b(fbn_host_[a-z_]+)\s*\(", hdr))
    assert declared == set(hostapi.HOST_ABI_SYMBOLS)
    L = capi.load_library()
    for s in declared:
        assert hasattr(L, s), s


def test_cli_fails_loudly_without_gpu(tmp_path):
    L = capi.load_library()
    if L.fbn_device_count() > 0:
        pytest.skip("a CUDA device is present")
    exe = os.path.join(ROOT, "fastbn_b200", "bin", "BayesianNetwork")
    r = subprocess.run([exe, "-a", "2", "-D", GOLDEN], capture_output=True, text=True, timeout=120)
    assert r.returncode == 1
    assert "no CUDA device" in r.stderr
    assert "number of cliques = 27, number of separators = 26" in r.stdout
    assert "maximum level of the junction tree = 15" in r.stdout
