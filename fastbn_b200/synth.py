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


def write_libsvm(path, ev_off, ev_vars, ev_vals, labels=None):
    """'label idx:val idx:val ... \\n' per case (trailing space like the shipped files; src/Dataset.cpp:162-262)."""
    n = len(ev_off) - 1
    with open(path, "w") as f:
        for i in range(n):
            lab = 0 if labels is None else int(labels[i])
            parts = [str(lab)] + ["%d:%d" % (ev_vars[k], ev_vals[k]) for k in range(ev_off[i], ev_off[i + 1])]
            f.write(" ".join(parts) + " \n")
