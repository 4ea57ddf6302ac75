"""ctypes binding of include/fastbn_b200.h (the C ABI of the CUDA library). Plumbing only — every number is
computed by the sm_100a kernels in fastbn_b200/csrc/. Fails loudly when the library is missing."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# FASTBN_B200_LIB selects another build of the same library (kernel tuning experiments, tools/prof_run.py)
LIB_PATH = os.environ.get("FASTBN_B200_LIB") or os.path.join(_HERE, "libfastbn_b200.so")

# every symbol include/fastbn_b200.h declares
ABI_SYMBOLS = ["fbn_last_error", "fbn_device_count", "fbn_device_warmup", "fbn_plan_create", "fbn_plan_destroy", "fbn_plan_get_info",
               "fbn_run", "fbn_run_device", "fbn_evaluate", "fbn_score_device", "fbn_score_device_static",
               "fbn_comm_unique_id", "fbn_comm_create", "fbn_comm_create_all", "fbn_reduce", "fbn_reduce_device", "fbn_comm_info",
               "fbn_comm_destroy"]


MODE_AUTO, MODE_FUSED, MODE_LEVEL = 0, 1, 2
TILE_CASES = 64      # test cases per tile: 32 lanes x kW = 2 cases per lane (fastbn_b200/csrc/plan.h; fbn_plan_info.tile_cases)
COMM_ID_BYTES = 128  # FBN_COMM_ID_BYTES


class FastBNError(RuntimeError):
    pass


class TreeDesc(C.Structure):
    _fields_ = [("n_vars", C.c_int32), ("var_dims", C.c_void_p), ("n_cliques", C.c_int32), ("n_separators", C.c_int32),
                ("table_nvars", C.c_void_p), ("table_vars", C.c_void_p), ("clique_potentials", C.c_void_p),
                ("sep_parent", C.c_void_p), ("sep_child", C.c_void_p), ("root", C.c_int32), ("query_var", C.c_int32)]


class PlanOptions(C.Structure):
    _fields_ = [("device", C.c_int32), ("team_warps", C.c_int32), ("max_resident", C.c_int32), ("mode", C.c_int32), ("cases_hint", C.c_int32), ("reserved", C.c_int32 * 3)]


class PlanInfo(C.Structure):
    _fields_ = [("n_vars", C.c_int32), ("n_cliques", C.c_int32), ("n_separators", C.c_int32), ("sum_dims", C.c_int32),
                ("n_levels", C.c_int32), ("team_warps", C.c_int32), ("resident_tiles", C.c_int32), ("n_phases", C.c_int32),
                ("mode", C.c_int32), ("launches_per_wave", C.c_int32),
                ("clique_entries", C.c_int64), ("separator_entries", C.c_int64), ("parent_entries", C.c_int64),
                ("marginal_entries", C.c_int64), ("scratch_bytes", C.c_int64), ("plan_bytes", C.c_int64),
                ("alg_bytes_per_case", C.c_double), ("kernel_bytes_per_case", C.c_double),
                ("flat_records", C.c_int32), ("kernel_rev", C.c_int32), ("tile_cases", C.c_int32), ("reserved", C.c_int32),
                ("view_entries", C.c_int64), ("gather_rows", C.c_int64)]


class Stats(C.Structure):
    _fields_ = [("kernel_ms", C.c_double), ("total_ms", C.c_double), ("n_launches", C.c_int64),
                ("h2d_bytes", C.c_int64), ("d2h_bytes", C.c_int64)]


_lib = None


def load_library():
    """Loads the in-tree CUDA library; raises FastBNError (no fallback) when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FastBNError("fastbn_b200: %s is missing — run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(there is no CPU fallback)" % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    L.fbn_last_error.restype = C.c_char_p
    L.fbn_device_count.restype = C.c_int
    L.fbn_plan_create.argtypes = [C.POINTER(TreeDesc), C.POINTER(PlanOptions), C.POINTER(C.c_void_p)]
    L.fbn_plan_destroy.argtypes = [C.c_void_p]
    L.fbn_plan_get_info.argtypes = [C.c_void_p, C.POINTER(PlanInfo)]
    L.fbn_run.argtypes = [C.c_void_p, C.c_int64] + [C.c_void_p] * 5 + [C.POINTER(Stats)]
    L.fbn_run_device.argtypes = [C.c_void_p, C.c_int64] + [C.c_void_p] * 6 + [C.POINTER(Stats)]
    L.fbn_evaluate.argtypes = [C.c_void_p, C.c_int64] + [C.c_void_p] * 8 + [C.POINTER(Stats)]
    L.fbn_score_device.argtypes = [C.c_void_p, C.c_int64] + [C.c_void_p] * 6
    L.fbn_score_device_static.argtypes = [C.c_int32, C.c_int64, C.c_int32] + [C.c_void_p] * 7
    L.fbn_comm_unique_id.argtypes = [C.c_void_p]
    L.fbn_comm_create.argtypes = [C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.POINTER(C.c_void_p)]
    L.fbn_comm_create_all.argtypes = [C.c_int32, C.c_void_p, C.POINTER(C.c_void_p)]
    L.fbn_reduce.argtypes = [C.c_void_p, C.c_void_p]
    L.fbn_reduce_device.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    L.fbn_comm_info.argtypes = [C.c_void_p] + [C.POINTER(C.c_int32)] * 3
    L.fbn_comm_destroy.argtypes = [C.c_void_p]
    _lib = L
    return L


def _check(L, rc):
    if rc != 0:
        raise FastBNError("fastbn_b200 error %d: %s" % (rc, L.fbn_last_error().decode()))


def tree_from_npz(path):
    z = np.load(path)
    return {k: z[k] for k in z.files}


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class Plan:
    """Device plan of one junction tree (fbn_plan). `tree` is a dict with the fields of fbn_tree_desc."""

    def __init__(self, tree, device=-1, team_warps=0, max_resident=0, mode=0, cases_hint=0):
        self.L = load_library()
        t = tree
        self._keep = dict(
            var_dims=np.ascontiguousarray(t["var_dims"], np.int32), table_nvars=np.ascontiguousarray(t["table_nvars"], np.int32),
            table_vars=np.ascontiguousarray(t["table_vars"], np.int32),
            clique_potentials=np.ascontiguousarray(t["clique_potentials"], np.float64),
            sep_parent=np.ascontiguousarray(t["sep_parent"], np.int32), sep_child=np.ascontiguousarray(t["sep_child"], np.int32))
        k = self._keep
        d = TreeDesc(int(t["n_vars"]), _ptr(k["var_dims"]), int(t["n_cliques"]), int(t["n_separators"]), _ptr(k["table_nvars"]),
                     _ptr(k["table_vars"]), _ptr(k["clique_potentials"]), _ptr(k["sep_parent"]), _ptr(k["sep_child"]),
                     int(t["root"]), int(t["query_var"]))
        opt = PlanOptions(device, team_warps, max_resident, mode, cases_hint, (C.c_int32 * 3)())
        h = C.c_void_p()
        _check(self.L, self.L.fbn_plan_create(C.byref(d), C.byref(opt), C.byref(h)))
        self.h = h
        info = PlanInfo()
        _check(self.L, self.L.fbn_plan_get_info(self.h, C.byref(info)))
        self.info = info
        self.sum_dims = info.sum_dims

    @property
    def kernel_name(self):
        """Name of the kernel a fused-mode plan launches (level mode launches one kernel per tree level)."""
        if self.info.mode == MODE_LEVEL:
            return "lvl_sum_kernel"
        if self.info.flat_records == 2:
            return "jt_resident_kernel"
        return "jt_flat_kernel" if self.info.flat_records else "jt_fused_kernel<%d>" % self.info.team_warps

    def close(self):
        if getattr(self, "h", None):
            self.L.fbn_plan_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def run(self, ev_off, ev_vars, ev_vals, want_post=True, labels_out=None, post_out=None):
        """Host buffers (numpy, or pinned torch tensors via .numpy()). Returns (labels, posteriors, Stats)."""
        ev_off = np.ascontiguousarray(ev_off, np.int64)
        ev_vars = np.ascontiguousarray(ev_vars, np.int32)
        ev_vals = np.ascontiguousarray(ev_vals, np.int32)
        n = len(ev_off) - 1
        labels = labels_out if labels_out is not None else np.empty(n, np.int32)
        post = post_out if post_out is not None else (np.empty((n, self.sum_dims), np.float64) if want_post else None)
        st = Stats()
        _check(self.L, self.L.fbn_run(self.h, n, _ptr(ev_off), _ptr(ev_vars), _ptr(ev_vals), _ptr(labels), _ptr(post), C.byref(st)))
        return labels, post, st

    def evaluate(self, ev_off, ev_vars, ev_vals, golden, truths, want_post=False):
        """fbn_evaluate: inference + MSE/HD/accuracy vs golden posteriors. Returns (labels, posteriors|None, score dict, Stats)."""
        ev_off = np.ascontiguousarray(ev_off, np.int64)
        ev_vars = np.ascontiguousarray(ev_vars, np.int32)
        ev_vals = np.ascontiguousarray(ev_vals, np.int32)
        golden = np.ascontiguousarray(golden, np.float64)
        truths = np.ascontiguousarray(truths, np.int32)
        n = len(ev_off) - 1
        labels = np.empty(n, np.int32)
        post = np.empty((n, self.sum_dims), np.float64) if want_post else None
        out = np.zeros(4, np.float64)
        st = Stats()
        _check(self.L, self.L.fbn_evaluate(self.h, n, _ptr(ev_off), _ptr(ev_vars), _ptr(ev_vals), _ptr(golden), _ptr(truths),
                                           _ptr(labels), _ptr(post), _ptr(out), C.byref(st)))
        return labels, post, dict(n_correct=int(out[0]), n_total=int(out[1]), sum_mse=float(out[2]), sum_hd=float(out[3])), st

    def run_device(self, n_cases, ev_off_ptr, ev_vars_ptr, ev_vals_ptr, labels_ptr, post_ptr, stream=0, timed=True):
        """Device pointers (ints, e.g. torch.Tensor.data_ptr())."""
        st = Stats()
        _check(self.L, self.L.fbn_run_device(self.h, n_cases, ev_off_ptr, ev_vars_ptr, ev_vals_ptr, labels_ptr, post_ptr,
                                             stream, C.byref(st) if timed else None))
        return st

    def score_device(self, n_cases, post_ptr, labels_ptr, golden_ptr, truths_ptr, stream=0):
        out = np.zeros(4, np.float64)
        _check(self.L, self.L.fbn_score_device(self.h, n_cases, post_ptr, labels_ptr, golden_ptr, truths_ptr, _ptr(out), stream))
        return dict(n_correct=int(out[0]), n_total=int(out[1]), sum_mse=float(out[2]), sum_hd=float(out[3]))


STATIC_SCORE_MAX_CASES = 65536  # FBN_STATIC_SCORE_MAX_CASES


def score_device_static(device, n_cases, n_vars, var_offsets_ptr, post_ptr, labels_ptr, golden_ptr, truths_ptr, out4_ptr, stream=0):
    """fbn_score_device_static: scoring without any allocation (static device scratch); every pointer is a device pointer,
    the four results land in out4_ptr (device) in stream order."""
    L = load_library()
    _check(L, L.fbn_score_device_static(device, n_cases, n_vars, var_offsets_ptr, post_ptr, labels_ptr, golden_ptr, truths_ptr,
                                        out4_ptr, stream))


def comm_unique_id():
    """fbn_comm_unique_id: 128 bytes rank 0 hands to the other ranks."""
    L = load_library()
    buf = (C.c_ubyte * COMM_ID_BYTES)()
    _check(L, L.fbn_comm_unique_id(buf))
    return bytes(buf)


class Comm:
    """fbn_comm: NCCL communicator of the path's one exchange step (sum of the four scores over the ranks)."""

    def __init__(self, device, rank, world, unique_id):
        self.L = load_library()
        h = C.c_void_p()
        idbuf = (C.c_ubyte * COMM_ID_BYTES).from_buffer_copy(unique_id)
        _check(self.L, self.L.fbn_comm_create(device, rank, world, idbuf, C.byref(h)))
        self.h = h
        self.rank, self.world = rank, world

    def reduce(self, v4):
        """In-place sum over all ranks of four doubles (collective). Returns a list of four floats."""
        a = np.ascontiguousarray(v4, np.float64).copy()
        assert a.shape == (4,)
        _check(self.L, self.L.fbn_reduce(self.h, _ptr(a)))
        return [float(x) for x in a]

    def nccl_version(self):
        v = C.c_int32()
        _check(self.L, self.L.fbn_comm_info(self.h, None, None, C.byref(v)))
        return int(v.value)

    def close(self):
        if getattr(self, "h", None):
            self.L.fbn_comm_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
