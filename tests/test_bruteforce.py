"""Secondary cross-check oracle: brute-force enumeration of the full joint of a SMALL network.
Independent of the junction-tree algorithm altogether -- it only shares the reference's CPT semantics
((int)(p*10000) counts, Laplace (count+1)/(total+dim), src/XMLBIFParser.cpp:176, src/DiscreteNode.cpp:152-161) -- so it
checks that what the oracle (CPU) and the CUDA path (GPU) print really is P(X_v | evidence) for every variable.
The reference has the same kind of check in its BruteForce class; here it is 30 lines of numpy."""
import numpy as np
import pytest

import emu
from fastbn_b200 import capi, hostapi, synth
from oracle import jtoracle

NETS = [  # (n, dims, max_par, window, seed)
    (9, (2, 3), 3, 5, 11),
    (12, (2, 3), 2, 6, 12),
    (10, (2, 4), 3, 9, 13),
    (7, (5,), 2, 3, 14),
]


def joint_table(var_dims, parents, cpts):
    """Full joint, axis v = variable v, with the reference's smoothed CPTs."""
    n = len(var_dims)
    joint = np.ones([int(d) for d in var_dims], np.float64)
    for v in range(n):
        text = np.array([float("%.4f" % x) for x in cpts[v].reshape(-1)]).reshape(cpts[v].shape)  # what the XML holds
        counts = np.floor(text * 10000.0)  # (int)(p*10000): truncation of the double product
        prob = (counts + 1.0) / (counts.sum(axis=0, keepdims=True) + int(var_dims[v]))
        pa = parents[v]
        prob = prob.reshape([int(var_dims[v])] + [int(var_dims[p]) for p in pa])  # first GIVEN most significant
        axes = [v] + list(pa)
        shape = [1] * n
        order = np.argsort(axes)
        prob = np.transpose(prob, order)
        for a in axes:
            shape[a] = int(var_dims[a])
        joint = joint * prob.reshape(shape)
    return joint


def brute_posteriors(joint, var_dims, ev_off, ev_vars, ev_vals):
    n = len(var_dims)
    pos = np.concatenate([[0], np.cumsum(var_dims)])
    out = np.zeros((len(ev_off) - 1, int(pos[-1])), np.float64)
    for c in range(len(ev_off) - 1):
        sl = [slice(None)] * n
        evs = ev_vars[ev_off[c]:ev_off[c + 1]]
        for v, w in zip(evs, ev_vals[ev_off[c]:ev_off[c + 1]]):
            sl[v] = slice(int(w), int(w) + 1)
        sub = joint[tuple(sl)]
        z = sub.sum()
        for v in range(n):
            if v in evs:
                continue  # the reference prints nothing for evidence variables (zeros in the posterior row)
            out[c, pos[v]:pos[v + 1]] = sub.sum(axis=tuple(a for a in range(n) if a != v)) / z
    return out


def small_case(tmp_path, spec, n_cases=48):
    n, dims, max_par, window, seed = spec
    var_dims, parents, cpts = synth.window_network(n, dims, max_par, window, seed)
    xml = str(tmp_path / "net.xml")
    synth.write_xmlbif(xml, var_dims, parents, cpts)
    m = hostapi.HostModel(xml)
    tree = m.tree()
    m.close()
    rng = np.random.default_rng(seed)
    counts = rng.integers(0, max(2, n // 2), n_cases)
    counts[0] = 0  # a case without evidence: prior marginals
    ev_off = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
    ev_vars = np.concatenate([np.sort(rng.choice(n, k, replace=False)) for k in counts] + [np.zeros(0, np.int64)]).astype(np.int32)
    ev_vals = np.array([rng.integers(0, var_dims[v]) for v in ev_vars], np.int32)
    want = brute_posteriors(joint_table(var_dims, parents, cpts), var_dims, ev_off, ev_vars, ev_vals)
    return tree, ev_off, ev_vars, ev_vals, want


@pytest.mark.parametrize("spec", NETS)
def test_oracle_agrees_with_joint_enumeration(tmp_path, spec):
    tree, ev_off, ev_vars, ev_vals, want = small_case(tmp_path, spec)
    o = jtoracle.Oracle(tree)
    post, lab = o.run_batch(ev_off, ev_vars, ev_vals)
    o.close()
    assert np.max(np.abs(post - want)) <= 1e-12
    q = int(tree["query_var"])
    pos = np.concatenate([[0], np.cumsum(tree["var_dims"])])
    free = np.array([q not in ev_vars[ev_off[c]:ev_off[c + 1]] for c in range(len(lab))])
    assert np.array_equal(lab[free], np.argmax(post[free, pos[q]:pos[q + 1]], axis=1))
    # and the device plans built from that tree (flat records for team 1, fused units, level mode), interpreted on the CPU
    for team in (1, 4, -8):
        post2, lab2, _ = emu.run(tree, team, ev_off, ev_vars, ev_vals)
        assert np.max(np.abs(post2 - want)) <= 1e-12 and np.array_equal(lab2, lab)
    post2, lab2, _ = emu.run_resident(tree, 16, ev_off, ev_vars, ev_vals)  # the resident-tile plan (shared-memory state)
    assert np.max(np.abs(post2 - want)) <= 1e-12 and np.array_equal(lab2, lab)


@pytest.mark.gpu
@pytest.mark.parametrize("spec", NETS)
def test_cuda_path_agrees_with_joint_enumeration(tmp_path, spec):
    tree, ev_off, ev_vars, ev_vals, want = small_case(tmp_path, spec, n_cases=200)
    import os
    for kw in (dict(), dict(flat=True), dict(team_warps=4), dict(team_warps=8, mode=capi.MODE_LEVEL)):  # resident, flat, fused and level kernels
        if kw.pop("flat", False):
            os.environ["FBN_NO_RESIDENT"] = "1"
        try:
            p = capi.Plan(tree, device=0, **kw)
        finally:
            os.environ.pop("FBN_NO_RESIDENT", None)
        labels, post, _ = p.run(ev_off, ev_vars, ev_vals)
        p.close()
        assert np.max(np.abs(post - want)) <= 1e-12, kw


def star_case(tmp_path, n_cases=64, seed=3):
    """Two cliques whose three binary parents appear nowhere else: the collect view of each folds THREE variables into its pattern
    tables (4^3 = 64 patterns per entry, the default cap) — the resident kernel then reads the fold list from the op's
    descriptor instead of the two pairs an item carries (csrc/resident.h)."""
    rng = np.random.default_rng(seed)
    var_dims = np.array([2, 2, 2, 3, 2, 3, 2, 2, 2, 2], np.int32)
    parents = [[], [], [], [0, 1, 2], [3], [4], [], [], [5, 6, 7], [8]]
    cpts = []
    for i, pa in enumerate(parents):
        ncfg = int(np.prod(var_dims[pa])) if pa else 1
        cpts.append(np.round(rng.dirichlet(np.ones(int(var_dims[i])), size=ncfg).T, 4))
    xml = str(tmp_path / "star.xml")
    synth.write_xmlbif(xml, var_dims, parents, cpts)
    m = hostapi.HostModel(xml)
    tree = m.tree()
    m.close()
    n = len(var_dims)
    counts = rng.integers(0, 7, n_cases)
    counts[0] = 0
    ev_off = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
    ev_vars = np.concatenate([np.sort(rng.choice(n, k, replace=False)) for k in counts] + [np.zeros(0, np.int64)]).astype(np.int32)
    ev_vals = np.array([rng.integers(0, var_dims[v]) for v in ev_vars], np.int32)
    want = brute_posteriors(joint_table(var_dims, parents, cpts), var_dims, ev_off, ev_vars, ev_vals)
    return tree, ev_off, ev_vars, ev_vals, want


def test_three_folded_variables_against_joint_enumeration(tmp_path):
    tree, ev_off, ev_vars, ev_vals, want = star_case(tmp_path)
    o = jtoracle.Oracle(tree)
    post, lab = o.run_batch(ev_off, ev_vars, ev_vals)
    o.close()
    ok = ~np.isnan(want)   # evidence of probability zero: NaN on both sides
    assert np.array_equal(np.isnan(post), np.isnan(want)) and np.max(np.abs(post[ok] - want[ok])) <= 1e-12
    for team in (1, 4, -8):
        post2, lab2, _ = emu.run(tree, team, ev_off, ev_vars, ev_vals)
        assert np.array_equal(np.isnan(post2), np.isnan(want)) and np.max(np.abs(post2[ok] - want[ok])) <= 1e-12 and np.array_equal(lab2, lab)
    post2, lab2, info = emu.run_resident(tree, 16, ev_off, ev_vars, ev_vals)
    assert np.array_equal(np.isnan(post2), np.isnan(want)) and np.max(np.abs(post2[ok] - want[ok])) <= 1e-12 and np.array_equal(lab2, lab)
