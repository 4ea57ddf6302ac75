"""ctypes driver for oracle/plan_emu.cpp (test infrastructure: scalar CPU interpreter of the device plan)."""
import ctypes as C
import os

import numpy as np

from fastbn_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_L = None


def _lib():
    global _L
    if _L is None:
        _L = C.CDLL(os.path.join(ROOT, "oracle", "_build", "libplan_emu.so"))
        _L.emu_last_error.restype = C.c_char_p
    return _L


def tree_desc(t):
    keep = {k: np.ascontiguousarray(t[k], np.float64 if k == "clique_potentials" else np.int32)
            for k in ["var_dims", "table_nvars", "table_vars", "clique_potentials", "sep_parent", "sep_child"]}
    p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    d = capi.TreeDesc(int(t["n_vars"]), p(keep["var_dims"]), int(t["n_cliques"]), int(t["n_separators"]), p(keep["table_nvars"]),
                      p(keep["table_vars"]), p(keep["clique_potentials"]), p(keep["sep_parent"]), p(keep["sep_child"]),
                      int(t["root"]), int(t["query_var"]))
    return d, keep


def run(t, team, ev_off, ev_vars, ev_vals):
    """Returns (post, labels, info) or raises ValueError(planner message)."""
    L = _lib()
    d, keep = tree_desc(t)
    n = len(ev_off) - 1
    sd = int(np.sum(t["var_dims"]))
    post = np.zeros((n, sd))
    lab = np.zeros(n, np.int32)
    info = np.zeros(16, np.int64)
    eo = np.ascontiguousarray(ev_off, np.int64)
    ev = np.ascontiguousarray(ev_vars, np.int32)
    ew = np.ascontiguousarray(ev_vals, np.int32)
    p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    rc = L.emu_run(C.byref(d), team, n, p(eo), p(ev), p(ew), p(lab), p(post), p(info))
    if rc:
        raise ValueError(L.emu_last_error().decode())
    return post, lab, info


def run_resident(t, warps, ev_off, ev_vars, ev_vals, smem_limit=232448):
    """The resident-tile plan (fastbn_b200/csrc/resident.h) of tree t interpreted on the CPU.
    Returns (post, labels, info) or raises ValueError(message): info = [ok, state_rows, max_block_bytes, smem_bytes, n_blocks,
    n_items, balance * 1000, flat tile_rows, constant bytes]."""
    L = _lib()
    d, keep = tree_desc(t)
    n = len(ev_off) - 1
    sd = int(np.sum(t["var_dims"]))
    post = np.zeros((n, sd))
    lab = np.zeros(n, np.int32)
    info = np.zeros(16, np.int64)
    eo = np.ascontiguousarray(ev_off, np.int64)
    ev = np.ascontiguousarray(ev_vars, np.int32)
    ew = np.ascontiguousarray(ev_vals, np.int32)
    p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    rc = L.emu_run_resident(C.byref(d), warps, C.c_int64(smem_limit), C.c_int64(n), p(eo), p(ev), p(ew), p(lab), p(post), p(info))
    if rc:
        raise ValueError(L.emu_last_error().decode())
    return post, lab, info
