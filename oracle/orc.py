"""ctypes face of the CPU oracle (oracle/gbt_cv_oracle.c).

TEST INFRASTRUCTURE ONLY.  Importable from tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs; the product package gentun_b200 never imports it.
Parity status: unpinned against real xgboost (see the header of gbt_cv_oracle.c).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liborc.so")
MAX_BIN = 256
MAX_NODES = 2047
GENE_ORDER = ("eta", "min_child_weight", "max_depth", "gamma", "max_delta_step", "subsample",
              "colsample_bytree", "colsample_bylevel", "lambda", "alpha", "scale_pos_weight")
OBJECTIVES = {"reg:linear": 0, "reg:squarederror": 0, "binary:logistic": 1}
METRICS = {"rmse": 0, "logloss": 1}


class Opts(C.Structure):
    _fields_ = [("objective", C.c_int32), ("metric", C.c_int32), ("nfold", C.c_int32),
                ("num_boost_round", C.c_int32), ("early_stopping_rounds", C.c_int32), ("nthread", C.c_int32)]


class Result(C.Structure):
    _fields_ = [("fitness", C.c_double), ("rounds_run", C.c_int32), ("best_iteration", C.c_int32),
                ("stopped_early", C.c_int32)]


NODE_DUMP = np.dtype([("feat", np.int32), ("bin", np.int32), ("value", np.float32), ("gain", np.float32)])


def build(force=False):
    src = os.path.join(_HERE, "gbt_cv_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        L.orc_ctx_create.restype = C.c_void_p
        L.orc_ctx_create.argtypes = [C.c_void_p, C.c_int64, C.c_int32, C.c_int64, C.c_void_p, C.c_int32, C.c_int32]
        L.orc_ctx_destroy.argtypes = [C.c_void_p]
        for name, rt in (("orc_ctx_ntrain", C.c_int64), ("orc_ctx_ntest", C.c_int64),
                         ("orc_ctx_train_rows", C.c_void_p), ("orc_ctx_test_rows", C.c_void_p),
                         ("orc_ctx_cuts", C.c_void_p), ("orc_ctx_ncuts", C.c_void_p),
                         ("orc_ctx_bins_col", C.c_void_p)):
            fn = getattr(L, name)
            fn.restype = rt
            fn.argtypes = [C.c_void_p, C.c_int]
        L.orc_cv.restype = C.c_int
        L.orc_cv.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(Opts), C.POINTER(Result), C.c_void_p, C.c_void_p,
                             C.c_int32, C.c_void_p, C.c_int64]
        L.orc_cross_validate.restype = C.c_int
        L.orc_cross_validate.argtypes = [C.c_void_p, C.c_int64, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p,
                                         C.POINTER(Opts), C.POINTER(Result)]
        L.orc_make_folds.argtypes = [C.c_int64, C.c_int32, C.c_uint32, C.c_void_p, C.c_void_p]
        L.orc_max_threads.restype = C.c_int
        L.orc_threads_in_use.restype = C.c_int
        L.orc_threads_in_use.argtypes = [C.c_int]
        _lib = L
    return _lib


def genes_vector(genes):
    return np.array([float(genes[k]) for k in GENE_ORDER], dtype=np.float64)


def make_opts(objective="reg:linear", eval_metric="rmse", nfold=5, num_boost_round=5000,
              early_stopping_rounds=100, nthread=0, **_ignored):
    return Opts(OBJECTIVES[objective], METRICS[eval_metric], int(nfold), int(num_boost_round),
                int(early_stopping_rounds), int(nthread))


def make_folds(n, nfold, seed=0):
    perm = np.empty(n, dtype=np.int64)
    fb = np.empty(nfold + 1, dtype=np.int64)
    lib().orc_make_folds(n, nfold, seed, perm.ctypes.data, fb.ctypes.data)
    return perm, fb


def _as_f32_matrix(x):
    a = np.ascontiguousarray(np.asarray(x, dtype=np.float32))
    if a.ndim != 2:
        raise ValueError("x must be 2-d")
    return a


class Context:
    """Everything xgb.cv derives from (data, nfold) alone: slices, cuts, bin matrices."""

    def __init__(self, x, y, nfold, nthread=0):
        self.x = _as_f32_matrix(x)
        self.y = np.ascontiguousarray(np.asarray(y, dtype=np.float32).reshape(-1))
        self.n, self.f = self.x.shape
        self.nfold = int(nfold)
        self._h = lib().orc_ctx_create(self.x.ctypes.data, self.n, self.f, self.f, self.y.ctypes.data, self.nfold, nthread)

    def close(self):
        if self._h:
            lib().orc_ctx_destroy(self._h)
            self._h = None

    __del__ = close

    def _arr(self, ptr, shape, dtype):
        n = int(np.prod(shape))
        buf = (C.c_char * (n * np.dtype(dtype).itemsize)).from_address(ptr)
        return np.frombuffer(buf, dtype=dtype).reshape(shape).copy()

    def ntrain(self, k):
        return lib().orc_ctx_ntrain(self._h, k)

    def ntest(self, k):
        return lib().orc_ctx_ntest(self._h, k)

    def train_rows(self, k):
        return self._arr(lib().orc_ctx_train_rows(self._h, k), (self.ntrain(k),), np.int64)

    def test_rows(self, k):
        return self._arr(lib().orc_ctx_test_rows(self._h, k), (self.ntest(k),), np.int64)

    def cuts(self, k):
        nc = self._arr(lib().orc_ctx_ncuts(self._h, k), (self.f,), np.int32)
        cu = self._arr(lib().orc_ctx_cuts(self._h, k), (self.f, MAX_BIN), np.float32)
        return cu, nc

    def train_bins(self, k):
        """[ntrain, f] uint8, rows in DMatrix.slice order."""
        return self._arr(lib().orc_ctx_bins_col(self._h, k), (self.f, self.ntrain(k)), np.uint8).T.copy()

    def all_bins(self, k):
        """[n, f] uint8 in fold order: train slice rows then test slice rows, binned with fold k's cuts."""
        cu, nc = self.cuts(k)
        rows = np.concatenate([self.train_rows(k), self.test_rows(k)])
        out = np.empty((self.n, self.f), dtype=np.uint8)
        for j in range(self.f):
            b = np.searchsorted(cu[j, :nc[j]], self.x[rows, j], side="right")
            out[:, j] = np.minimum(b, nc[j] - 1)
        return out

    def cv(self, genes, opts, trace=False, trees_rounds=0):
        g = genes_vector(genes) if isinstance(genes, dict) else np.asarray(genes, dtype=np.float64)
        res = Result()
        tr = np.full(opts.num_boost_round, np.nan) if trace else None
        trees = np.zeros((self.nfold, trees_rounds, MAX_NODES), dtype=NODE_DUMP) if trees_rounds else None
        stride = max(self.ntrain(k) for k in range(self.nfold))
        rows = np.full((self.nfold, stride), -1, dtype=np.int64) if trees_rounds else None
        rc = lib().orc_cv(self._h, g.ctypes.data, C.byref(opts), C.byref(res),
                          tr.ctypes.data if trace else None, trees.ctypes.data if trees_rounds else None,
                          trees_rounds, rows.ctypes.data if trees_rounds else None, stride)
        if rc != 0:
            raise ValueError("oracle rejected genes (rc=%d)" % rc)
        out = {"fitness": res.fitness, "rounds": res.rounds_run, "best_iteration": res.best_iteration,
               "stopped_early": bool(res.stopped_early)}
        if trace:
            out["trace"] = tr[:res.rounds_run]
        if trees_rounds:
            out["trees"] = trees
            out["final_rows"] = rows
        return out


def cross_validate(x, y, genes, nthread=0, **additional_parameters):
    """The reference's per-fit path (DMatrix + cv rebuilt for every individual)."""
    xa = _as_f32_matrix(x)
    ya = np.ascontiguousarray(np.asarray(y, dtype=np.float32).reshape(-1))
    opts = make_opts(nthread=nthread, **additional_parameters)
    g = genes_vector(genes)
    res = Result()
    rc = lib().orc_cross_validate(xa.ctypes.data, xa.shape[0], xa.shape[1], xa.shape[1], ya.ctypes.data,
                                  g.ctypes.data, C.byref(opts), C.byref(res))
    if rc != 0:
        raise ValueError("oracle rejected genes (rc=%d)" % rc)
    return res.fitness


def max_threads():
    return lib().orc_max_threads()


def host_cores():
    """Cores this process may run on (the affinity mask, not what OMP_NUM_THREADS says)."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def threads_in_use(nthread=0):
    """Threads an OpenMP region of the oracle really runs with when `nthread` is requested."""
    return lib().orc_threads_in_use(int(nthread))
