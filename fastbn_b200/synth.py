"""Synthetic workloads for the junction-tree hot path (BASELINE.json configs 2-5): evidence batches in the CSR layout
of fbn_run, following the recipe of the reference authors' generator comment (src/XMLBIFParser.cpp:198-215) and the
shipped dataset/alarm/testing_alarm_1k_p20: variable 0 is the query and never evidence, every case observes the same
number of distinct variables, values uniform in [0, dim)."""
import numpy as np


def uniform_evidence(var_dims, n_cases, n_evidence, seed, sort_vars=True):
    """Returns (ev_off int64[n+1], ev_vars int32, ev_vals int32)."""
    var_dims = np.asarray(var_dims, np.int64)
    n_vars = len(var_dims)
    n_evidence = int(min(n_evidence, n_vars - 1))
    rng = np.random.Generator(np.random.MT19937(seed))
    ev_vars = np.empty((n_cases, n_evidence), np.int32)
    step = max(1, (1 << 22) // max(1, n_vars))
    for s in range(0, n_cases, step):
        e = min(n_cases, s + step)
        keys = rng.random((e - s, n_vars - 1))
        if n_evidence > 0:
            idx = np.argpartition(keys, n_evidence - 1, axis=1)[:, :n_evidence]
            if sort_vars:
                idx = np.sort(idx, axis=1)
            ev_vars[s:e] = idx.astype(np.int32) + 1
    u = rng.random((n_cases, n_evidence))
    ev_vals = np.minimum((u * var_dims[ev_vars]).astype(np.int32), (var_dims[ev_vars] - 1).astype(np.int32))
    ev_off = (np.arange(n_cases + 1, dtype=np.int64) * n_evidence)
    return ev_off, ev_vars.reshape(-1), ev_vals.reshape(-1).astype(np.int32)


def read_xmlbif(path):
    """(var_dims, parents, cpts) of an XMLBIF file as the reference reads it: variables in file order, <TABLE> with the FOR variable
    most significant, then the <GIVEN>s in file order (src/XMLBIFParser.cpp). Probabilities as printed (no count truncation /
    smoothing: used for sampling only)."""
    import xml.etree.ElementTree as ET
    net = ET.parse(path).getroot().find("NETWORK")
    names, dims = [], []
    for v in net.findall("VARIABLE"):
        names.append(v.find("NAME").text.strip())
        dims.append(len(v.findall("VALUE")) or len(v.findall("OUTCOME")))
    idx = {nm: i for i, nm in enumerate(names)}
    parents, cpts = [None] * len(names), [None] * len(names)
    for pr in list(net.findall("PROBABILITY")) + list(net.findall("DEFINITION")):
        v = idx[pr.find("FOR").text.strip()]
        pa = [idx[g.text.strip()] for g in pr.findall("GIVEN")]
        tab = np.array([float(x) for x in pr.find("TABLE").text.split()], np.float64)
        parents[v] = pa
        cpts[v] = tab.reshape(dims[v], -1)
    return np.array(dims, np.int32), parents, cpts


def forward_sampled_evidence(xml_path, n_cases, n_evidence, seed):
    """SURVEY.md 8(d) config 2, second set: every case is a forward sample of the network (ancestral order), of which n_evidence
    distinct non-query variables (uniform from 1..n-1, sorted) are observed. Returns (ev_off, ev_vars, ev_vals)."""
    var_dims, parents, cpts = read_xmlbif(xml_path)
    n = len(var_dims)
    rng = np.random.Generator(np.random.MT19937(seed))
    state = np.full((n_cases, n), -1, np.int64)
    done = np.zeros(n, bool)
    while not done.all():
        progressed = False
        for v in range(n):
            if done[v] or not all(done[p] for p in parents[v]):
                continue
            cfg = np.zeros(n_cases, np.int64)
            for p in parents[v]:  # first GIVEN most significant
                cfg = cfg * int(var_dims[p]) + state[:, p]
            cdf = np.cumsum(cpts[v][:, cfg], axis=0)  # (dim, n_cases)
            u = rng.random(n_cases) * cdf[-1]
            state[:, v] = np.minimum((u[None, :] >= cdf).sum(axis=0), int(var_dims[v]) - 1)
            done[v] = progressed = True
        if not progressed:
            raise ValueError("cycle in the network")
    n_evidence = int(min(n_evidence, n - 1))
    keys = rng.random((n_cases, n - 1))
    ev_vars = (np.sort(np.argpartition(keys, n_evidence - 1, axis=1)[:, :n_evidence], axis=1) + 1).astype(np.int32)
    ev_vals = np.take_along_axis(state, ev_vars.astype(np.int64), axis=1).astype(np.int32)
    ev_off = np.arange(n_cases + 1, dtype=np.int64) * n_evidence
    return ev_off, ev_vars.reshape(-1), ev_vals.reshape(-1)


def write_libsvm(path, ev_off, ev_vars, ev_vals, labels=None):
    """'label idx:val idx:val ... \\n' per case (trailing space like the shipped files; src/Dataset.cpp:162-262)."""
    n = len(ev_off) - 1
    with open(path, "w") as f:
        for i in range(n):
            lab = 0 if labels is None else int(labels[i])
            parts = [str(lab)] + ["%d:%d" % (ev_vars[k], ev_vals[k]) for k in range(ev_off[i], ev_off[i + 1])]
            f.write(" ".join(parts) + " \n")


# ---------------------------------------------------------------------------------------------------------------
# Synthetic networks of the named shapes (BASELINE.json configs 3-5; recipe of SURVEY.md Appendix B): node i draws
# 1..max_par parents uniformly from the previous `window` nodes (bounded treewidth), CPT rows are Dirichlet(1) samples
# printed with 4 decimals in the reference's <TABLE> order (FOR variable most significant, then <GIVEN>s in file order).
SHAPES = {
    # name: generator parameters; the seeds were found by tools/make_synth_golden.py --sweep so that the junction tree the
    # reference (and our host mirror) builds has the named maximum clique size. n_evidence = int(frac * n).
    # hepar2: 52 cliques, 29 levels, max clique 6 912, sum |clique| 67 286
    "hepar2": dict(n=70, dims=(2, 3, 4), max_par=6, window=8, par_geom=0.55, seed=20260103, max_clique=6912, evidence=(14, 35),
                   desc="synthetic HEPAR2-shape BN (70 nodes, card 2-4)"),
    # pigs: 380 cliques, 75 levels, max clique 3^10, sum |clique| 264 690
    "pigs": dict(n=441, dims=(3,), max_par=2, window=10, seed=20260126, max_clique=59049, evidence=(88,),
                 desc="synthetic PIGS-shape BN (441 nodes, card 3, max clique 3^10)"),
    # munin_small: 102 cliques, 37 levels, max clique 5^9, sum |clique| 3 185 150 (oracle-friendly stand-in)
    "munin_small": dict(n=120, dims=(5,), max_par=2, window=9, seed=20260148, max_clique=1953125, evidence=(0,),
                        desc="synthetic MUNIN-shape stand-in (120 nodes, card 5, max clique 5^9)"),
    # munin: 105 cliques, 29 levels, max clique 5^10 = 9.8e6, sum |clique| 14 551 925
    "munin": dict(n=120, dims=(5,), max_par=2, window=12, seed=20260127, max_clique=9765625, evidence=(0,),
                  desc="synthetic MUNIN-shape BN (120 nodes, card 5, max clique 5^10 ~ 1e7)"),
}


def window_network(n, dims, max_par, window, seed, par_geom=0.0):
    """Returns (var_dims int32[n], parents list[list[int]] in GIVEN order, cpts list[np.ndarray (dim, n_parent_cfgs)])."""
    rng = np.random.Generator(np.random.MT19937(seed))
    var_dims = np.array([dims[int(rng.integers(0, len(dims)))] for _ in range(n)], np.int32)
    parents, cpts = [], []
    for i in range(n):
        lo = max(0, i - window)
        if par_geom:  # most nodes get 1-2 parents, a few up to max_par (HEPAR2-like in-degree profile)
            k = 0 if i == 0 else int(min(i - lo, max_par, rng.geometric(par_geom)))
        else:
            k = 0 if i == 0 else int(min(i - lo, rng.integers(1, max_par + 1)))
        pa = sorted(int(x) for x in rng.choice(np.arange(lo, i), size=k, replace=False)) if k else []
        ncfg = int(np.prod(var_dims[pa])) if pa else 1
        p = rng.dirichlet(np.ones(int(var_dims[i])), size=ncfg).T  # (dim, ncfg): FOR variable most significant
        parents.append(pa)
        cpts.append(np.round(p, 4))
    return var_dims, parents, cpts


def write_xmlbif(path, var_dims, parents, cpts, name="synthetic"):
    with open(path, "w") as f:
        f.write('<?xml version="1.0" encoding="UTF-8"?>\n<BIF>\n    <NETWORK>\n        <NAME>%s</NAME>\n' % name)
        for i, d in enumerate(var_dims):
            f.write("        <VARIABLE>\n            <NAME>V%d</NAME>\n            <TYPE>discrete</TYPE>\n" % i)
            for s in range(int(d)):
                f.write("            <VALUE>s%d</VALUE>\n" % s)
            f.write("        </VARIABLE>\n")
        for i in range(len(var_dims)):
            f.write("        <PROBABILITY>\n            <FOR>V%d</FOR>\n" % i)
            for p in parents[i]:
                f.write("            <GIVEN>V%d</GIVEN>\n" % p)
            f.write("            <TABLE>" + " ".join("%.4f" % x for x in cpts[i].reshape(-1)) + " </TABLE>\n        </PROBABILITY>\n")
        f.write("    </NETWORK>\n</BIF>\n")


def make_shape(name, path, seed=None, **override):
    """Writes the XMLBIF of a named shape; returns var_dims."""
    cfg = dict(SHAPES[name])
    cfg.update(override)
    if seed is not None:
        cfg["seed"] = seed
    var_dims, parents, cpts = window_network(cfg["n"], cfg["dims"], cfg["max_par"], cfg["window"], cfg["seed"], cfg.get("par_geom", 0.0))
    write_xmlbif(path, var_dims, parents, cpts, name=name)
    return var_dims


def write_empty_pt(path):
    """The reference only needs the -f4 file to OPEN (src/Inference.cpp:120-124); MSE/HD then print as nan."""
    with open(path, "w") as f:
        f.write("\n")
