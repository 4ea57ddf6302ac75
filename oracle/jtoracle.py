"""TEST INFRASTRUCTURE ONLY: ctypes driver for the C restatement oracle/jt_oracle.c (the parity checker).
Imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg — never by fastbn_b200/."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_build", "libjt_oracle.so")


def build():
    src = os.path.join(_HERE, "jt_oracle.c")
    if not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "_build/libjt_oracle.so"], stdout=subprocess.DEVNULL)
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(build())
        L.jto_create.restype = C.c_void_p
        L.jto_create.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                 C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.jto_destroy.argtypes = [C.c_void_p]
        L.jto_sum_dims.argtypes = [C.c_void_p]
        L.jto_run_case.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.jto_run_batch.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class Oracle:
    """tree: dict of numpy arrays with the fields of fbn_tree_desc (+ optional child_order)."""

    def __init__(self, tree):
        self.L = lib()
        t = tree
        self._keep = [np.ascontiguousarray(t["var_dims"], np.int32), np.ascontiguousarray(t["table_nvars"], np.int32),
                      np.ascontiguousarray(t["table_vars"], np.int32),
                      np.ascontiguousarray(t["clique_potentials"], np.float64),
                      np.ascontiguousarray(t["sep_parent"], np.int32), np.ascontiguousarray(t["sep_child"], np.int32)]
        co = np.ascontiguousarray(t["child_order"], np.int32) if "child_order" in t and t["child_order"] is not None else None
        self._keep.append(co)
        k = self._keep
        self.h = self.L.jto_create(int(t["n_vars"]), _p(k[0]), int(t["n_cliques"]), int(t["n_separators"]), _p(k[1]),
                                   _p(k[2]), _p(k[3]), _p(k[4]), _p(k[5]), int(t["root"]), int(t["query_var"]), _p(co))
        self.sum_dims = self.L.jto_sum_dims(self.h)

    def close(self):
        if self.h:
            self.L.jto_destroy(self.h)
            self.h = None

    def run_batch(self, ev_off, ev_vars, ev_vals):
        ev_off = np.ascontiguousarray(ev_off, np.int64)
        ev_vars = np.ascontiguousarray(ev_vars, np.int32)
        ev_vals = np.ascontiguousarray(ev_vals, np.int32)
        n = len(ev_off) - 1
        post = np.zeros((n, self.sum_dims), np.float64)
        labels = np.zeros(n, np.int32)
        rc = self.L.jto_run_batch(self.h, n, _p(ev_off), _p(ev_vars), _p(ev_vals), _p(post), _p(labels))
        if rc:
            raise RuntimeError("oracle: variable missing from every clique")
        return post, labels
