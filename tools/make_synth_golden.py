#!/usr/bin/env python
"""Fixtures for the synthetic shapes (BASELINE.json configs 3-5), made in the build container with the UNMODIFIED
reference (oracle/_ref):

  tests/golden/synth/<shape>.xml        the generated network (fastbn_b200.synth.make_shape)
  tests/golden/synth/<shape>_ref.npz    seeded test cases + the reference's fp64 posteriors and labels for them, and the
                                        reference's own tree statistics (cliques, levels, max clique) for the same XML

  python tools/make_synth_golden.py --sweep       find seeds that give the target maximum clique (uses our host mirror)
  python tools/make_synth_golden.py hepar2 pigs   write fixtures
"""
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fastbn_b200 import synth, hostapi  # noqa: E402
from oracle import refdrv  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "synth")
N_ORACLE = {"hepar2": 256, "pigs": 48, "munin_small": 1, "munin": 1}
# MUNIN shapes: config 5 is 0 % evidence (every case identical -> one reference case pins it); parity of the multi-CTA split,
# the evidence masks and PULL2 on these big tables rests on extra cases WITH 24 evidence variables
N_ORACLE_EVIDENCE = {"munin_small": 32, "munin": 8}


def sweep():
    d = tempfile.mkdtemp()
    for name, cfg in synth.SHAPES.items():
        for seed in range(cfg["seed"], cfg["seed"] + 400):
            p = os.path.join(d, "x.xml")
            synth.make_shape(name, p, seed=seed)
            m = hostapi.HostModel(p)
            r = (m.n_cliques, m.max_level, m.max_clique, m.clique_entries)
            m.close()
            if r[2] == cfg["max_clique"]:
                print(name, "seed", seed, "cliques %d levels %d max %d sum %d" % r)
                break


def make(name):
    os.makedirs(OUT, exist_ok=True)
    cfg = synth.SHAPES[name]
    xml = os.path.join(OUT, name + ".xml")
    var_dims = synth.make_shape(name, xml)
    parts = []
    for k, nev in enumerate(cfg["evidence"]):
        parts.append(synth.uniform_evidence(var_dims, N_ORACLE[name], nev, seed=cfg["seed"] + 1000 + k))
    if name.startswith("munin"):  # beyond the 0 % evidence of config 5: a few cases WITH evidence for parity
        parts.append(synth.uniform_evidence(var_dims, N_ORACLE_EVIDENCE[name], 24, seed=cfg["seed"] + 2000))
    off = np.concatenate([[0]] + [p[0][1:] + sum(len(q[1]) for q in parts[:i]) for i, p in enumerate(parts)]).astype(np.int64)
    ev = np.concatenate([p[1] for p in parts]).astype(np.int32)
    ew = np.concatenate([p[2] for p in parts]).astype(np.int32)
    with tempfile.TemporaryDirectory() as d:
        cases = os.path.join(d, "cases")
        o20, v20, w20 = synth.uniform_evidence(var_dims, 20, 0, seed=1)  # the reference's ctor needs a LIBSVM file
        synth.write_libsvm(cases, o20, v20, w20)
        t0 = time.time()
        ref = refdrv.Reference(xml, cases)
        tree = ref.tree()
        t1 = time.time()
        post, labels, secs = ref.run_batch(off, ev, ew)
        ref.close()
    pos = np.concatenate([[0], np.cumsum(tree["table_nvars"])])
    sizes = [int(np.prod(tree["var_dims"][tree["table_vars"][pos[i]:pos[i + 1]]])) for i in range(int(tree["n_cliques"]))]
    stats = np.array([int(tree["n_cliques"]), int(tree["n_separators"]), int(tree["max_level"]), max(sizes), sum(sizes)], np.int64)
    np.savez_compressed(os.path.join(OUT, name + "_ref.npz"), ev_off=off, ev_vars=ev, ev_vals=ew, post=post, labels=labels,
                        ref_tree_stats=stats, ref_seconds=np.array([t1 - t0, secs]))
    print("%s: reference tree %s, construct %.2f s, %d cases in %.2f s (%.2f cases/s)" %
          (name, stats.tolist(), t1 - t0, len(off) - 1, secs, (len(off) - 1) / max(secs, 1e-9)))


if __name__ == "__main__":
    if "--sweep" in sys.argv:
        sweep()
    else:
        for nm in sys.argv[1:] or ["hepar2", "pigs", "munin_small"]:
            make(nm)
