"""ctypes binding of libgentun_b200.so (C ABI in include/gentun_b200.h).

This is the only route from the Python drop-in to the GPU.  There is no CPU fallback: if the
shared library is missing or no CUDA device is present, construction raises.
"""
import ctypes as C
import os
import threading
import weakref

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GENTUN_B200_LIB", os.path.join(_HERE, "libgentun_b200.so"))   # override: kernel experiments only

GENE_ORDER = ("eta", "min_child_weight", "max_depth", "gamma", "max_delta_step", "subsample",
              "colsample_bytree", "colsample_bylevel", "lambda", "alpha", "scale_pos_weight")
# objective / metric spellings XgboostModel forwards to xgboost (xgboost_models.py:16-21)
OBJECTIVES = {"reg:linear": 0, "reg:squarederror": 0, "binary:logistic": 1}
METRICS = {"rmse": 0, "logloss": 1}
MAX_BIN = 256
MAX_NODES = 2047
NUM_GENES = 11
_DTYPES = {np.float32: 0, np.float64: 1, np.int64: 2}

TREE_NODE = np.dtype([("feat", np.int32), ("bin", np.int32), ("value", np.float32), ("gain", np.float32)])
NODE_ROWS = np.dtype([("begin", np.uint32), ("count", np.uint32), ("buf", np.int32), ("state", np.int32)])


class EngineError(RuntimeError):
    pass


class DataOpts(C.Structure):
    _fields_ = [("device", C.c_int32), ("x_on_device", C.c_int32), ("mem_budget_bytes", C.c_int64),
                ("reserved", C.c_int32 * 4)]


class FitOpts(C.Structure):
    _fields_ = [("objective", C.c_int32), ("metric", C.c_int32), ("nfold", C.c_int32),
                ("num_boost_round", C.c_int32), ("early_stopping_rounds", C.c_int32), ("reserved", C.c_int32 * 3)]


class Dump(C.Structure):
    _fields_ = [("tree_rounds", C.c_int32), ("trees", C.c_void_p), ("trace_means", C.c_void_p),
                ("rows_round", C.c_int32), ("row_sets", C.c_void_p), ("node_rows", C.c_void_p),
                ("row_stride", C.c_int64)]


class Stats(C.Structure):
    _fields_ = [("hist_ms", C.c_double), ("hist_alg_bytes", C.c_double), ("hist_row_features", C.c_double),
                ("hist_launches", C.c_int64), ("kernel_launches", C.c_int64), ("eval_ms", C.c_double),
                ("prepare_ms", C.c_double), ("waves", C.c_int64), ("jobs", C.c_int64)]


_lib = None
_lib_lock = threading.Lock()

EXPORTS = ("gt_engine_create", "gt_engine_create_from", "gt_engine_prepare", "gt_engine_eval", "gt_engine_eval_ex", "gt_engine_destroy",
           "gt_engine_fold_sizes", "gt_engine_get_fold_rows", "gt_engine_get_cuts", "gt_engine_get_bins",
           "gt_engine_get_stats", "gt_make_folds", "gt_last_error", "gt_version",
           "gt_column_cuts_noalloc", "gt_column_cuts_noalloc_max_rows")


def load_library():
    """dlopen the CUDA library; raise loudly when it is not there (no fallback path exists)."""
    global _lib
    with _lib_lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise EngineError(
                "gentun_b200: %s is missing. Build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "or gentun_b200/csrc/build.sh; there is no CPU fallback." % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        vp, i32, i64 = C.c_void_p, C.c_int32, C.c_int64
        L.gt_engine_create.argtypes = [vp, i64, i32, i64, vp, C.POINTER(DataOpts), C.POINTER(vp)]
        L.gt_engine_create_from.argtypes = [vp, i32, i64, i32, i64, i64, vp, i32, C.POINTER(DataOpts), C.POINTER(vp)]
        L.gt_engine_prepare.argtypes = [vp, i32]
        L.gt_engine_eval.argtypes = [vp, i32, vp, C.POINTER(FitOpts), vp, vp]
        L.gt_engine_eval_ex.argtypes = [vp, i32, vp, C.POINTER(FitOpts), vp, vp, C.POINTER(Dump)]
        L.gt_engine_destroy.argtypes = [vp]
        L.gt_engine_destroy.restype = None
        L.gt_engine_fold_sizes.argtypes = [vp, i32, i32, C.POINTER(i64), C.POINTER(i64)]
        L.gt_engine_get_fold_rows.argtypes = [vp, i32, i32, vp, vp]
        L.gt_engine_get_cuts.argtypes = [vp, i32, i32, vp, vp]
        L.gt_engine_get_bins.argtypes = [vp, i32, i32, vp]
        L.gt_engine_get_stats.argtypes = [vp, C.POINTER(Stats)]
        L.gt_make_folds.argtypes = [i64, i32, C.c_uint32, vp, vp]
        L.gt_last_error.restype = C.c_char_p
        L.gt_version.restype = C.c_int
        L.gt_column_cuts_noalloc.argtypes = [vp, i64, vp, vp, vp]
        L.gt_column_cuts_noalloc_max_rows.restype = i64
        _lib = L
        return _lib


def _check(rc):
    if rc != 0:
        msg = load_library().gt_last_error().decode("utf-8", "replace")
        if rc == -1:
            raise ValueError("gentun_b200: " + msg)
        raise EngineError("gentun_b200 (code %d): %s" % (rc, msg))


def make_folds(n, nfold, seed=0):
    """Host-only: the fold permutation xgb.cv draws (np.random.seed(seed); permutation; array_split)."""
    perm = np.empty(n, dtype=np.int64)
    fb = np.empty(nfold + 1, dtype=np.int64)
    _check(load_library().gt_make_folds(n, nfold, seed, perm.ctypes.data, fb.ctypes.data))
    return perm, fb


# what xgboost uses for a parameter the caller leaves out (XgboostModel only dict.update()s the genes it is given into
# its parameter dict, xgboost_models.py:16-22, so a genome that tunes two genes leaves the rest at these)
XGB_DEFAULTS = {"eta": 0.3, "min_child_weight": 1.0, "max_depth": 6, "gamma": 0.0, "max_delta_step": 0.0, "subsample": 1.0,
                "colsample_bytree": 1.0, "colsample_bylevel": 1.0, "lambda": 1.0, "alpha": 0.0, "scale_pos_weight": 1.0}
# parameter names xgboost would honour but this engine does not implement: refusing beats silently ignoring them
_UNSUPPORTED = {"tree_method", "max_bin", "grow_policy", "max_leaves", "colsample_bynode", "num_parallel_tree",
                "monotone_constraints", "interaction_constraints", "base_score", "seed", "sampling_method",
                "learning_rate", "min_split_loss", "reg_alpha", "reg_lambda"}
_PASSIVE = {"booster", "objective", "eval_metric", "silent", "nthread", "n_jobs", "verbosity"}


def complete_genes(genes):
    """The 11 tuned parameters of one individual: absent ones take xgboost's defaults; a key that would change what
    xgboost trains but that this engine cannot honour raises."""
    bad = [k for k in genes if k in _UNSUPPORTED]
    if bad:
        raise NotImplementedError("gentun_b200 does not implement the xgboost parameter(s) %s" % sorted(bad))
    unknown = [k for k in genes if k not in XGB_DEFAULTS and k not in _PASSIVE]
    if unknown:
        raise ValueError("unknown hyperparameter(s) %s" % sorted(unknown))
    out = dict(XGB_DEFAULTS)
    out.update({k: v for k, v in genes.items() if k in XGB_DEFAULTS})
    return out


def genes_matrix(genes_list):
    """[n, 11] float64 in the genome order of XgboostIndividual (absent genes at xgboost's defaults)."""
    out = np.empty((len(genes_list), NUM_GENES), dtype=np.float64)
    for i, g in enumerate(genes_list):
        full = complete_genes(g)
        for j, name in enumerate(GENE_ORDER):
            out[i, j] = float(full[name])
    return out


def fit_opts(booster="gbtree", objective="reg:linear", eval_metric="rmse", nfold=5, num_boost_round=5000,
             early_stopping_rounds=100):
    if booster != "gbtree":
        raise NotImplementedError("gentun_b200 evaluates booster='gbtree' only (got %r)" % (booster,))
    if objective not in OBJECTIVES:
        raise NotImplementedError("objective %r is not implemented (have %s)" % (objective, sorted(OBJECTIVES)))
    if eval_metric not in METRICS:
        raise NotImplementedError("eval_metric %r is not implemented (have %s)" % (eval_metric, sorted(METRICS)))
    esr = 0 if early_stopping_rounds is None else int(early_stopping_rounds)
    return FitOpts(OBJECTIVES[objective], METRICS[eval_metric], int(nfold), int(num_boost_round), esr)


class Engine(object):
    """One dataset resident on one GPU: replaces the per-individual xgb.DMatrix + xgb.cv set-up."""

    def __init__(self, x, y, device=0, mem_budget_bytes=0, collect_timing=False, device_pointers=None, streaming_only=False):
        self._lib = load_library()
        self._h = C.c_void_p()
        opts = DataOpts()
        opts.device = int(device)
        opts.mem_budget_bytes = int(mem_budget_bytes)
        opts.reserved[0] = 1 if collect_timing else 0
        opts.reserved[1] = 1 if streaming_only else 0   # never the shared-memory-resident kernel (regime comparison tests)
        if device_pointers is not None:
            xp, yp, n, f, ldx = device_pointers
            opts.x_on_device = 1
            self.n, self.f = int(n), int(f)
            self._keep = (x, y)  # keep the owning tensors alive
            _check(self._lib.gt_engine_create(C.c_void_p(xp), self.n, self.f, int(ldx), C.c_void_p(yp), C.byref(opts),
                                              C.byref(self._h)))
        else:
            # the arrays go to the library as they lie in memory (a float64 DataFrame is column-major): the cast to
            # float32 that xgb.DMatrix applies and the row-major layout are made on the device
            xa = np.asarray(x)
            if xa.dtype not in (np.float32, np.float64):
                xa = xa.astype(np.float64 if xa.dtype.kind in "iub" and xa.dtype.itemsize > 2 else np.float32)
            if xa.ndim != 2:
                raise ValueError("x must be [n, f] and y [n]")
            if not (xa.flags.c_contiguous or xa.flags.f_contiguous):
                xa = np.ascontiguousarray(xa)
            ya = np.asarray(y).reshape(-1)
            if ya.dtype not in (np.float32, np.float64, np.int64):
                ya = ya.astype(np.float32)
            ya = np.ascontiguousarray(ya)
            if xa.shape[0] != ya.shape[0]:
                raise ValueError("x must be [n, f] and y [n]")
            self.n, self.f = xa.shape
            self._keep = None
            rs, cs = (self.f, 1) if xa.flags.c_contiguous else (1, self.n)
            _check(self._lib.gt_engine_create_from(xa.ctypes.data, _DTYPES[xa.dtype.type], self.n, self.f, rs, cs,
                                                   ya.ctypes.data, _DTYPES[ya.dtype.type], C.byref(opts), C.byref(self._h)))
        self.device = int(device)

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            self._lib.gt_engine_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def prepare(self, nfold):
        _check(self._lib.gt_engine_prepare(self._h, int(nfold)))

    def evaluate(self, genes_list, trace=False, tree_rounds=0, rows_round=-1, **additional_parameters):
        """Fitness of each gene dict (mean n-fold validation metric), in the order given."""
        opts = fit_opts(**additional_parameters)
        g = genes_matrix(genes_list) if not isinstance(genes_list, np.ndarray) else np.ascontiguousarray(genes_list, dtype=np.float64)
        n_ind = g.shape[0]
        fit = np.empty(n_ind, dtype=np.float64)
        rounds = np.zeros(n_ind, dtype=np.int32)
        if not (trace or tree_rounds or rows_round >= 0):
            _check(self._lib.gt_engine_eval(self._h, n_ind, g.ctypes.data, C.byref(opts), fit.ctypes.data, rounds.ctypes.data))
            return fit, rounds
        d = Dump()
        extra = {}
        nfold = opts.nfold
        if trace:
            extra["trace"] = np.full((n_ind, opts.num_boost_round), np.nan)
            d.trace_means = extra["trace"].ctypes.data
        if tree_rounds:
            extra["trees"] = np.zeros((n_ind, nfold, tree_rounds, MAX_NODES), dtype=TREE_NODE)
            d.tree_rounds = tree_rounds
            d.trees = extra["trees"].ctypes.data
        d.rows_round = -1
        if rows_round >= 0:
            stride = (max(self.fold_sizes(nfold, k)[0] for k in range(nfold)) + 31) // 32 * 32
            extra["row_sets"] = np.zeros((n_ind, nfold, 2, stride), dtype=np.uint32)
            extra["node_rows"] = np.zeros((n_ind, nfold, MAX_NODES), dtype=NODE_ROWS)
            d.rows_round = rows_round
            d.row_sets = extra["row_sets"].ctypes.data
            d.node_rows = extra["node_rows"].ctypes.data
            d.row_stride = stride
        _check(self._lib.gt_engine_eval_ex(self._h, n_ind, g.ctypes.data, C.byref(opts), fit.ctypes.data,
                                           rounds.ctypes.data, C.byref(d)))
        return fit, rounds, extra

    def fold_sizes(self, nfold, k):
        a, b = C.c_int64(), C.c_int64()
        _check(self._lib.gt_engine_fold_sizes(self._h, nfold, k, C.byref(a), C.byref(b)))
        return a.value, b.value

    def fold_rows(self, nfold, k):
        ntr, nte = self.fold_sizes(nfold, k)
        tr = np.empty(ntr, dtype=np.int64)
        te = np.empty(nte, dtype=np.int64)
        _check(self._lib.gt_engine_get_fold_rows(self._h, nfold, k, tr.ctypes.data, te.ctypes.data))
        return tr, te

    def cuts(self, nfold, k):
        cu = np.zeros((self.f, MAX_BIN), dtype=np.float32)
        nc = np.zeros(self.f, dtype=np.int32)
        _check(self._lib.gt_engine_get_cuts(self._h, nfold, k, cu.ctypes.data, nc.ctypes.data))
        return cu, nc

    def bins(self, nfold, k):
        out = np.empty((self.n, self.f), dtype=np.uint8)
        _check(self._lib.gt_engine_get_bins(self._h, nfold, k, out.ctypes.data))
        return out

    def stats(self):
        s = Stats()
        _check(self._lib.gt_engine_get_stats(self._h, C.byref(s)))
        return {k: getattr(s, k) for k, _ in Stats._fields_}


# ---- engine cache: the GA re-creates Population objects every generation and every tournament
# (gentun/algorithms.py:38-41, 51-55), so engines live outside the populations, keyed by the identity of the data
# objects.  An entry dies with its data (weakref.finalize), is rebuilt when the data were changed in place (a cheap
# fingerprint is compared on every lookup), and takes its memoised fitness values with it: the memo is keyed by a
# per-engine serial number that is never reused, not by an address.
_engines = {}
_memo = {}
_serial = [0]


def default_device():
    return int(os.environ.get("GENTUN_B200_DEVICE", os.environ.get("LOCAL_RANK", "0")))


def _fingerprint(a):
    """Shape, type, address and a strided sample of the values: enough to notice in-place rescaling or relabelling."""
    v = np.asarray(a)
    flat = v.reshape(-1) if v.flags.c_contiguous else v.T.reshape(-1) if v.flags.f_contiguous else v.ravel()
    step = max(1, flat.shape[0] // 61)
    sample = np.ascontiguousarray(flat[::step][:64])
    return (v.shape, str(v.dtype), v.__array_interface__["data"][0], sample.tobytes(), flat[-1:].tobytes())


def _drop(key):
    hit = _engines.pop(key, None)
    if hit is None:
        return
    eng = hit["engine"]
    for k in [k for k in _memo if k[0] == eng.serial]:
        del _memo[k]
    eng.close()


def engine_for(x, y, device=None):
    dev = default_device() if device is None else int(device)
    if x is None or y is None:
        raise ValueError("x_train / y_train are required to evaluate fitness")
    key = (id(x), id(y), dev)
    fp = (_fingerprint(x), _fingerprint(y))
    hit = _engines.get(key)
    if hit is not None:
        if hit["x"]() is x and hit["y"]() is y and hit["fp"] == fp:
            return hit["engine"]
        _drop(key)   # a recycled id, or data changed in place: the old engine and its memo entries are void
    eng = Engine(x, y, device=dev)
    _serial[0] += 1
    eng.serial = _serial[0]

    def _ref(o):
        try:
            r = weakref.ref(o)
            weakref.finalize(o, _drop, key)   # the data died: free the device copy and the per-job buffers now
            return r
        except TypeError:
            return lambda o=o: o
    _engines[key] = {"x": _ref(x), "y": _ref(y), "fp": fp, "engine": eng}
    return eng


def clear_engines():
    for key in list(_engines):
        _drop(key)
    _memo.clear()


def memo_key(eng, genes, params):
    return (getattr(eng, "serial", -1), tuple(float(genes[k]) for k in GENE_ORDER), tuple(sorted(params.items())))


def evaluate_many(x, y, genes_list, params, use_memo=True):
    """Fitness for many gene dicts sharing `params`; identical (genes, params) are trained once
    (xgb.cv is deterministic for fixed genes)."""
    eng = engine_for(x, y)
    genes_list = [complete_genes(g) for g in genes_list]
    keys = [memo_key(eng, g, params) for g in genes_list]
    todo, todo_keys = [], []
    seen = set()
    for g, k in zip(genes_list, keys):
        if (not use_memo or k not in _memo) and k not in seen:
            todo.append(g)
            todo_keys.append(k)
            seen.add(k)
    fresh = {}
    if todo:
        from . import sharding
        fit = sharding.evaluate_sharded(eng, todo, params)
        for k, v in zip(todo_keys, fit):
            fresh[k] = float(v)
            if use_memo:
                _memo[k] = float(v)
    return [(_memo[k] if (use_memo and k in _memo) else fresh[k]) for k in keys]
