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


def test_libsvm_loader_parallel_pieces_and_binary_cache(tmp_path, monkeypatch):
    """A file of several MB is mmap'ed and parsed in pieces of whole lines by several threads: the result must be the
    sequential one (incl. the dropped unterminated last line). With FASTBN_EVIDENCE_CACHE=1 the parsed arrays are kept in
    <file>.fbncache, reused while the text file's size and mtime match, and rebuilt when the file changes."""
    n = 400_000
    tree = hostapi.HostModel(ALARM_XML)
    dims = tree.tree()["var_dims"]
    tree.close()
    off, ev, ew = synth.uniform_evidence(dims, n, 7, seed=3)
    labels = (np.arange(n) % 2).astype(np.int32)
    p = str(tmp_path / "cases")
    synth.write_libsvm(p, off, ev, ew, labels=labels)
    with open(p, "a") as f:
        f.write("1 3:1")  # unterminated last line: dropped (src/Dataset.cpp:179-227)
    assert os.path.getsize(p) > (12 << 20)  # several 4 MB pieces
    m = hostapi.HostModel(ALARM_XML, p)
    o1, v1, w1, t1 = m.cases()
    m.close()
    assert np.array_equal(o1, off) and np.array_equal(v1, ev) and np.array_equal(w1, ew) and np.array_equal(t1, labels)
    assert not os.path.exists(p + ".fbncache")  # opt-in
    monkeypatch.setenv("FASTBN_EVIDENCE_CACHE", "1")
    m = hostapi.HostModel(ALARM_XML, p)
    m.close()
    assert os.path.exists(p + ".fbncache")
    stamp = os.path.getmtime(p + ".fbncache")
    m = hostapi.HostModel(ALARM_XML, p)   # served from the cache
    o2, v2, w2, t2 = m.cases()
    m.close()
    assert np.array_equal(o2, off) and np.array_equal(v2, ev) and np.array_equal(w2, ew) and np.array_equal(t2, labels)
    assert os.path.getmtime(p + ".fbncache") == stamp
    with open(p, "a") as f:
        f.write(" 4:0\n0 2:1 \n")       # the text file changes: the cache no longer matches and is rebuilt
    m = hostapi.HostModel(ALARM_XML, p)
    o3, v3, w3, t3 = m.cases()
    m.close()
    assert len(t3) == n + 2 and t3[-2:].tolist() == [1, 0] and v3[-3:].tolist() == [3, 4, 2] and w3[-3:].tolist() == [1, 0, 1]
    m = hostapi.HostModel(ALARM_XML, p)
    o4, v4, w4, t4 = m.cases()
    m.close()
    assert np.array_equal(o4, o3) and np.array_equal(v4, v3) and np.array_equal(t4, t3)
    with open(p + ".fbncache", "r+b") as f:  # a damaged cache is ignored, not trusted
        f.seek(0)
        f.write(b"garbage!")
    m = hostapi.HostModel(ALARM_XML, p)
    o5, v5, w5, t5 = m.cases()
    m.close()
    assert np.array_equal(o5, o3) and np.array_equal(v5, v3) and np.array_equal(w5, w3)


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


def test_text_fast_paths_are_bit_identical_to_libc():
    """"%.7f" writer and decimal parser of the potential-table files (fastbn_b200/host/textio.cpp) against Python's
    correctly rounded formatting / parsing (== glibc's) on random values, exact ties, neighbours of ties and odd tokens."""
    import ctypes as C
    import struct
    L = capi.load_library()
    L.fbn_host_format_fixed.argtypes = [C.c_double, C.c_char_p]
    L.fbn_host_parse_double.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.c_double)]
    buf = C.create_string_buffer(400)
    rng = np.random.default_rng(11)
    xs = list(rng.random(60000)) + list(rng.random(20000) * 1e-6) + list(rng.random(2000) * 1000)
    xs += [0.0, 1.0, 0.5, 5e-8, 1.5e-7, 2.5e-7, 1e-8, 4.9999999e-8, 5.0000001e-8, 0.99999995, 0.99999994999, 1.5, 123.4567891,
           float("nan"), -0.0, -1.25, 1e300, float("inf"), 1048575.99999995, 1048576.0]
    ties = [(k + 0.5) / 1e7 for k in range(3000)]
    xs += ties + [float(np.nextafter(t, 1)) for t in ties] + [float(np.nextafter(t, 0)) for t in ties]
    for x in xs:
        n = L.fbn_host_format_fixed(float(x), buf)
        assert buf.raw[:n].decode() == "%.7f" % x, repr(x)
    out = C.c_double()
    toks = ["%.7f" % x for x in xs[:40000] if x == x and abs(x) < 1e200]
    toks += ["0.1", "1", "1.", "-.5", "+3.25", "1e-3", "1.5E2", "nan", "inf", "123456789012345678", "0.123456789012345678",
             "0.00000000000000000000000012", "12345678901234567890123", "1.7976931348623157e308", "4.9e-324", "-0.0000000"]
    for t in toks:
        b = t.encode()
        n = L.fbn_host_parse_double(b, len(b), C.byref(out))
        exp = float(t)
        assert n == len(b), t
        assert struct.pack("d", out.value) == struct.pack("d", exp) or (exp != exp and out.value != out.value), t


def _parse_pt_py(path, n_cases, var_dims):
    """straightforward restatement of Inference::LoadGroundTruthProbabilityTable (src/Inference.cpp:108-146)"""
    lines = open(path).read().split("\n")
    if lines and lines[-1] == "":
        lines.pop()
    offs = np.concatenate([[0], np.cumsum(var_dims)])
    out = np.zeros((n_cases, offs[-1]))
    nv = len(var_dims)
    for ln in range(n_cases * nv):
        i, j = divmod(ln, nv)
        toks = lines[ln].split() if ln < len(lines) else []
        if not toks:
            out[i, offs[j]] = -1
            continue
        for k, t in enumerate(toks[:var_dims[j]]):
            out[i, offs[j] + k] = float(t)
    return out


def test_potential_table_loader_matches_plain_parse(tmp_path):
    """the parallel -f4 loader (pieces read with pread, line index built per piece, exact fast decimal parser) against a
    line-by-line Python parse: the shipped 1000-case file, a truncated copy and a copy without the final newline"""
    xml = os.path.join(GOLDEN, "alarm", "alarm.xml")
    cases = os.path.join(GOLDEN, "alarm", "testing_alarm_1k_p20")
    pt = os.path.join(GOLDEN, "alarm", "alarm_1k_pt")
    m = hostapi.HostModel(xml, cases)
    dims = m.tree()["var_dims"]
    n = len(m.cases()[0]) - 1
    raw = open(pt).read()
    variants = {"full": raw, "truncated": raw[:len(raw) // 3], "no_final_newline": raw.rstrip("\n")}
    for name, text in variants.items():
        p = str(tmp_path / name)
        open(p, "w").write(text)
        got = m.golden(p)
        exp = _parse_pt_py(p, n, dims)
        assert np.array_equal(got, exp), name
    m.close()
