"""TEST INFRASTRUCTURE ONLY: ctypes driver for oracle/_ref/libfastbn_ref.so (the unmodified reference,
compiled by oracle/build_ref.sh). Imported by tests/, tools/make_golden.py and bench.py's reference /
cpu_baseline legs — never by fastbn_b200/."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libfastbn_ref.so")


def available():
    return os.path.exists(LIB_PATH)


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(LIB_PATH)
        L.fbnref_create.restype = C.c_void_p
        L.fbnref_create.argtypes = [C.c_char_p, C.c_char_p]
        L.fbnref_destroy.argtypes = [C.c_void_p]
        for f in ("fbnref_num_cases", "fbnref_num_vars", "fbnref_sum_dims", "fbnref_max_level"):
            getattr(L, f).restype = C.c_int
            getattr(L, f).argtypes = [C.c_void_p]
        L.fbnref_tree_sizes.argtypes = [C.c_void_p, C.c_void_p]
        L.fbnref_tree_fill.argtypes = [C.c_void_p] + [C.c_void_p] * 8
        L.fbnref_run_case.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.fbnref_run_custom.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.fbnref_run_batch.restype = C.c_double
        L.fbnref_run_batch.argtypes = [C.c_void_p, C.c_int64] + [C.c_void_p] * 5
        L.fbnref_run_batch_threads.restype = C.c_double
        L.fbnref_run_batch_threads.argtypes = [C.c_void_p, C.c_int64] + [C.c_void_p] * 5 + [C.c_int]
        L.fbnref_predict_timed.restype = C.c_double
        L.fbnref_predict_timed.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.c_void_p]
        L.fbnref_evaluate_accuracy.restype = C.c_double
        L.fbnref_evaluate_accuracy.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.c_int]
        L.fbnref_case_evidence.restype = C.c_int
        L.fbnref_case_evidence.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.fbnref_set_quiet.argtypes = [C.c_int]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class Reference:
    """The reference's Dataset + CustomNetwork + JunctionTree for one (xml, libsvm) pair."""

    def __init__(self, xml_path, libsvm_path):
        self.L = lib()
        self.h = self.L.fbnref_create(xml_path.encode(), libsvm_path.encode())
        self.n_cases = self.L.fbnref_num_cases(self.h)
        self.n_vars = self.L.fbnref_num_vars(self.h)
        self.sum_dims = self.L.fbnref_sum_dims(self.h)

    def close(self):
        if self.h:
            self.L.fbnref_destroy(self.h)
            self.h = None

    def tree(self):
        """Flat tree description read from the live reference objects (dict of numpy arrays)."""
        sz = np.zeros(8, np.int64)
        self.L.fbnref_tree_sizes(self.h, _p(sz))
        nv, nc, ns, tv, ce, root, query, max_level = [int(x) for x in sz]
        var_dims = np.zeros(nv, np.int32)
        table_nvars = np.zeros(nc + ns, np.int32)
        table_vars = np.zeros(tv, np.int32)
        pots = np.zeros(ce, np.float64)
        sep_parent = np.zeros(ns, np.int32)
        sep_child = np.zeros(ns, np.int32)
        clique_level = np.zeros(nc, np.int32)
        child_order = np.zeros(ns, np.int32)
        self.L.fbnref_tree_fill(self.h, _p(var_dims), _p(table_nvars), _p(table_vars), _p(pots), _p(sep_parent),
                                _p(sep_child), _p(clique_level), _p(child_order))
        return dict(n_vars=np.int32(nv), n_cliques=np.int32(nc), n_separators=np.int32(ns), var_dims=var_dims,
                    table_nvars=table_nvars, table_vars=table_vars, clique_potentials=pots, sep_parent=sep_parent,
                    sep_child=sep_child, root=np.int32(root), query_var=np.int32(query),
                    max_level=np.int32(max_level), clique_level=clique_level, child_order=child_order)

    def run_case(self, i):
        post = np.zeros(self.sum_dims, np.float64)
        lab = np.zeros(1, np.int32)
        self.L.fbnref_run_case(self.h, i, _p(post), _p(lab))
        return post, int(lab[0])

    def run_custom(self, vars_, vals):
        vars_ = np.ascontiguousarray(vars_, np.int32)
        vals = np.ascontiguousarray(vals, np.int32)
        post = np.zeros(self.sum_dims, np.float64)
        lab = np.zeros(1, np.int32)
        self.L.fbnref_run_custom(self.h, len(vars_), _p(vars_), _p(vals), _p(post), _p(lab))
        return post, int(lab[0])

    def run_batch(self, ev_off, ev_vars, ev_vals, want_post=True):
        ev_off = np.ascontiguousarray(ev_off, np.int64)
        ev_vars = np.ascontiguousarray(ev_vars, np.int32)
        ev_vals = np.ascontiguousarray(ev_vals, np.int32)
        n = len(ev_off) - 1
        post = np.zeros((n, self.sum_dims), np.float64) if want_post else None
        labels = np.zeros(n, np.int32)
        secs = self.L.fbnref_run_batch(self.h, n, _p(ev_off), _p(ev_vars), _p(ev_vals), _p(post), _p(labels))
        return post, labels, secs

    def run_batch_threads(self, ev_off, ev_vars, ev_vals, threads, want_post=False):
        """Per-case sequence with the reference's intra-case OpenMP parallelism (-t N). Returns (post|None, labels, seconds)."""
        ev_off = np.ascontiguousarray(ev_off, np.int64)
        ev_vars = np.ascontiguousarray(ev_vars, np.int32)
        ev_vals = np.ascontiguousarray(ev_vals, np.int32)
        n = len(ev_off) - 1
        post = np.zeros((n, self.sum_dims), np.float64) if want_post else None
        labels = np.zeros(n, np.int32)
        secs = self.L.fbnref_run_batch_threads(self.h, n, _p(ev_off), _p(ev_vars), _p(ev_vals), _p(post), _p(labels), int(threads))
        return post, labels, secs

    def predict_timed(self, pt_path, threads=1):
        """The stock PredictUseJTInfer(num_threads) loop over the loaded LIBSVM cases (>= 20), timed like the `jt` timer.
        Returns (accuracy, seconds)."""
        secs = np.zeros(1, np.float64)
        acc = self.L.fbnref_predict_timed(self.h, pt_path.encode(), int(threads), _p(secs))
        return float(acc), float(secs[0])

    def evaluate_accuracy(self, pt_path, threads=1, quiet=True):
        return self.L.fbnref_evaluate_accuracy(self.h, pt_path.encode(), threads, 1 if quiet else 0)

    def case_evidence(self, i):
        vars_ = np.zeros(4096, np.int32)
        vals = np.zeros(4096, np.int32)
        truth = np.zeros(1, np.int32)
        k = self.L.fbnref_case_evidence(self.h, i, 4096, _p(vars_), _p(vals), _p(truth))
        return vars_[:k].copy(), vals[:k].copy(), int(truth[0])
