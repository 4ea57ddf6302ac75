"""ctypes binding of include/fastbn_b200_host.h: the C++ host mirror of the reference's front end (XMLBIF network ->
junction tree, LIBSVM test cases, EvaluateAccuracy). Plumbing for tests and bench.py; the tree building is native C++."""
import ctypes as C

import numpy as np

from . import capi

HOST_ABI_SYMBOLS = ["fbn_host_model_create", "fbn_host_model_destroy", "fbn_host_model_tree", "fbn_host_model_info",
                    "fbn_host_model_cases", "fbn_host_model_load_golden", "fbn_host_model_evaluate", "fbn_host_model_write_pt",
                    "fbn_host_format_fixed", "fbn_host_parse_double"]


def _lib():
    L = capi.load_library()
    if not getattr(L, "_host_ready", False):
        L.fbn_host_model_create.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.POINTER(C.c_void_p)]
        L.fbn_host_model_destroy.argtypes = [C.c_void_p]
        L.fbn_host_model_tree.restype = C.POINTER(capi.TreeDesc)
        L.fbn_host_model_tree.argtypes = [C.c_void_p]
        L.fbn_host_model_info.argtypes = [C.c_void_p, C.c_void_p]
        L.fbn_host_model_cases.argtypes = [C.c_void_p] + [C.POINTER(C.c_void_p)] * 4
        L.fbn_host_model_evaluate.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.fbn_host_model_load_golden.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p]
        L.fbn_host_model_write_pt.argtypes = [C.c_void_p, C.c_char_p]
        L._host_ready = True
    return L


def _arr(ptr, n, ctype, dtype):
    if n == 0:
        return np.zeros(0, dtype)
    return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(ctype)), shape=(n,)).astype(dtype, copy=True)


class HostModel:
    """CustomNetwork + Dataset + JunctionTree built by the C++ host mirror."""

    def __init__(self, xmlbif_path, libsvm_path=None, quiet=True):
        self.L = _lib()
        h = C.c_void_p()
        rc = self.L.fbn_host_model_create(xmlbif_path.encode(), (libsvm_path or "").encode(), 1 if quiet else 0, C.byref(h))
        if rc:
            raise capi.FastBNError("fbn_host_model_create failed: %d" % rc)
        self.h = h
        info = np.zeros(8, np.int64)
        self.L.fbn_host_model_info(self.h, info.ctypes.data_as(C.c_void_p))
        (self.n_vars, self.n_cliques, self.n_separators, self._tv, self.clique_entries, self.max_level, self.n_cases,
         self.max_clique) = [int(x) for x in info]

    def close(self):
        if getattr(self, "h", None):
            self.L.fbn_host_model_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def tree(self):
        """dict with the fields of fbn_tree_desc (numpy copies)."""
        d = self.L.fbn_host_model_tree(self.h).contents
        nc, ns = d.n_cliques, d.n_separators
        return dict(n_vars=np.int32(d.n_vars), n_cliques=np.int32(nc), n_separators=np.int32(ns),
                    var_dims=_arr(d.var_dims, d.n_vars, C.c_int32, np.int32),
                    table_nvars=_arr(d.table_nvars, nc + ns, C.c_int32, np.int32),
                    table_vars=_arr(d.table_vars, self._tv, C.c_int32, np.int32),
                    clique_potentials=_arr(d.clique_potentials, self.clique_entries, C.c_double, np.float64),
                    sep_parent=_arr(d.sep_parent, ns, C.c_int32, np.int32), sep_child=_arr(d.sep_child, ns, C.c_int32, np.int32),
                    root=np.int32(d.root), query_var=np.int32(d.query_var), max_level=np.int32(self.max_level))

    def cases(self):
        """(ev_off int64[n+1], ev_vars, ev_vals, truths) of the loaded LIBSVM file."""
        p = [C.c_void_p() for _ in range(4)]
        self.L.fbn_host_model_cases(self.h, *[C.byref(x) for x in p])
        n = self.n_cases
        off = _arr(p[0], n + 1, C.c_int64, np.int64)
        ne = int(off[-1]) if n else 0
        return off, _arr(p[1], ne, C.c_int32, np.int32), _arr(p[2], ne, C.c_int32, np.int32), _arr(p[3], n, C.c_int32, np.int32)

    def golden(self, pt_path):
        """Inference::LoadGroundTruthProbabilityTable: (n_cases, sum_dims) array, -1 in the first slot of evidence nodes."""
        sd = int(np.sum(self.tree()["var_dims"]))
        g = np.zeros((self.n_cases, sd), np.float64)
        rc = self.L.fbn_host_model_load_golden(self.h, pt_path.encode(), g.ctypes.data_as(C.c_void_p))
        if rc:
            raise capi.FastBNError("fbn_host_model_load_golden failed: %d" % rc)
        return g

    def evaluate(self, pt_path, devices=None, want_post=False):
        """JunctionTree::EvaluateAccuracy. Returns dict(accuracy, average_mse, average_hd, jt_seconds, kernel_ms, labels, post)."""
        sd = None
        labels = np.empty(self.n_cases, np.int32)
        post = None
        if want_post:
            sd = int(np.sum(self.tree()["var_dims"]))
            post = np.empty((self.n_cases, sd), np.float64)
        out = np.zeros(5, np.float64)
        dv = np.ascontiguousarray(devices if devices is not None else [0], np.int32)
        rc = self.L.fbn_host_model_evaluate(self.h, pt_path.encode(), dv.ctypes.data_as(C.c_void_p), len(dv),
                                            labels.ctypes.data_as(C.c_void_p),
                                            post.ctypes.data_as(C.c_void_p) if post is not None else None,
                                            out.ctypes.data_as(C.c_void_p))
        if rc:
            raise capi.FastBNError("fbn_host_model_evaluate failed: %d" % rc)
        return dict(accuracy=out[0], average_mse=out[1], average_hd=out[2], jt_seconds=out[3], kernel_ms=out[4], labels=labels, post=post)

    def write_pt(self, path):
        self.L.fbn_host_model_write_pt(self.h, path.encode())
