#!/usr/bin/env python
"""Generates the committed golden fixtures under tests/golden/ by running the UNMODIFIED reference
(oracle/_ref/libfastbn_ref.so, built by oracle/build_ref.sh from /root/reference). Run in the build container only;
the GPU box just reads the fixtures.

  tests/golden/alarm_tree_ref.npz      flat tree read from the live reference objects (fbn_tree_desc fields)
  tests/golden/alarm_1k_ref.npz        the 1000 shipped cases: evidence CSR, truths, reference fp64 posteriors + labels
  tests/golden/alarm_synth_4k_ref.npz  4000 synthetic cases (BASELINE.json config 2 recipe, seed 20260101): same fields
  tests/golden/alarm_edge_ref.npz      edge cases: no evidence, single evidence, many evidence, out-of-range variable
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import refdrv  # noqa: E402
from fastbn_b200 import synth  # noqa: E402

G = os.path.join(ROOT, "tests", "golden")


def csr_from_lists(vars_list, vals_list):
    off = np.zeros(len(vars_list) + 1, np.int64)
    for i, v in enumerate(vars_list):
        off[i + 1] = off[i] + len(v)
    return off, np.concatenate(vars_list + [np.zeros(0, np.int32)]).astype(np.int32), \
        np.concatenate(vals_list + [np.zeros(0, np.int32)]).astype(np.int32)


def main():
    xml = os.path.join(G, "alarm", "alarm.xml")
    cases = os.path.join(G, "alarm", "testing_alarm_1k_p20")
    ref = refdrv.Reference(xml, cases)
    tree = ref.tree()
    np.savez_compressed(os.path.join(G, "alarm_tree_ref.npz"), **tree)

    vl, wl, truths = [], [], []
    for i in range(ref.n_cases):
        v, w, t = ref.case_evidence(i)
        vl.append(v)
        wl.append(w)
        truths.append(t)
    off, ev, ew = csr_from_lists(vl, wl)
    post, labels, _ = ref.run_batch(off, ev, ew)
    np.savez_compressed(os.path.join(G, "alarm_1k_ref.npz"), ev_off=off, ev_vars=ev, ev_vals=ew,
                        truths=np.array(truths, np.int32), post=post, labels=labels)

    off, ev, ew = synth.uniform_evidence(tree["var_dims"], 4000, 7, seed=20260101)
    post, labels, _ = ref.run_batch(off, ev, ew)
    np.savez_compressed(os.path.join(G, "alarm_synth_4k_ref.npz"), ev_off=off, ev_vars=ev, ev_vals=ew, post=post, labels=labels)

    dims = tree["var_dims"]
    vl = [np.zeros(0, np.int32), np.array([5], np.int32), np.arange(1, 37, dtype=np.int32),
          np.array([3, 99], np.int32), np.array([36], np.int32), np.arange(1, 20, dtype=np.int32)]
    wl = [np.zeros(0, np.int32), np.array([1], np.int32), (np.arange(1, 37) % dims[1:37]).astype(np.int32),
          np.array([1, 0], np.int32), np.array([dims[36] - 1], np.int32), np.zeros(19, np.int32)]
    off, ev, ew = csr_from_lists(vl, wl)
    post, labels, _ = ref.run_batch(off, ev, ew)
    np.savez_compressed(os.path.join(G, "alarm_edge_ref.npz"), ev_off=off, ev_vars=ev, ev_vals=ew, post=post, labels=labels)
    ref.close()
    print("golden fixtures written to", G)


if __name__ == "__main__":
    main()
