"""ctypes binding of the CPU oracle (oracle/mllib_oracle.cpp).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
``--impl reference`` legs.  The product package never imports this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass, field

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libb2rf_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "mllib_oracle.cpp")
    if force or not os.path.exists(_SO) or (os.path.exists(src) and os.path.getmtime(_SO) < os.path.getmtime(src)):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = C.CDLL(_SO)
        _lib.b2rf_oracle_last_error.restype = C.c_char_p
        for n in ("b2rf_oracle_num_nodes", "b2rf_oracle_num_cats", "b2rf_oracle_num_trace", "b2rf_oracle_updates"):
            getattr(_lib, n).restype = C.c_int64
            getattr(_lib, n).argtypes = [C.c_void_p]
        _lib.b2rf_oracle_num_groups.argtypes = [C.c_void_p]
        _lib.b2rf_oracle_free.argtypes = [C.c_void_p]
    return _lib


class OracleError(RuntimeError):
    pass


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class Params(C.Structure):
    _fields_ = [("maxDepth", C.c_int32), ("maxBins", C.c_int32), ("minInstancesPerNode", C.c_int32), ("numTrees", C.c_int32),
                ("featuresPerNode", C.c_int32), ("maxMemoryInMB", C.c_int32), ("bootstrap", C.c_int32), ("numThreads", C.c_int32),
                ("minInfoGain", C.c_double), ("minWeightFractionPerNode", C.c_double), ("subsamplingRate", C.c_double),
                ("seed", C.c_int64), ("impurity", C.c_int32), ("reserved", C.c_int32)]


IMPURITIES = {"variance": 0, "gini": 1, "entropy": 2}


# ---------------- RNG known-answer hooks ----------------
def java_random(seed):
    a = C.c_int32(); b = C.c_int64()
    lib().b2rf_oracle_java_random(C.c_int64(seed), C.byref(a), C.byref(b))
    return a.value, b.value


def java_random_longs(seed, n):
    out = np.zeros(n, np.int64)
    lib().b2rf_oracle_java_random_longs(C.c_int64(seed), C.c_int(n), _p(out))
    return out


def xorshift_doubles(seed, n):
    out = np.zeros(n)
    lib().b2rf_oracle_xorshift_doubles(C.c_int64(seed), C.c_int(n), _p(out))
    return out


def xorshift_nextint(seed):
    lib().b2rf_oracle_xorshift_nextint.restype = C.c_int32
    return lib().b2rf_oracle_xorshift_nextint(C.c_int64(seed))


def well_doubles(seed, n):
    out = np.zeros(n)
    lib().b2rf_oracle_well_doubles(C.c_int64(seed), C.c_int(n), _p(out))
    return out


def poisson(seed, mean, n):
    out = np.zeros(n, np.int32)
    lib().b2rf_oracle_poisson(C.c_int64(seed), C.c_double(mean), C.c_int(n), _p(out))
    return out


def reservoir(F, k, seed):
    out = np.zeros(min(k, F), np.int32)
    lib().b2rf_oracle_reservoir(C.c_int(F), C.c_int(k), C.c_int64(seed), _p(out))
    return out


def group_order(keys):
    keys = np.ascontiguousarray(keys, np.int32)
    out = np.zeros(len(keys), np.int32)
    lib().b2rf_oracle_group_order(_p(keys), C.c_int(len(keys)), _p(out))
    return out


# ---------------- stage functions ----------------
def find_splits(sampleX, arity, max_possible_bins, weighted_num_samples):
    sampleX = np.ascontiguousarray(sampleX, np.float64)
    n, F = sampleX.shape
    arity = np.ascontiguousarray(arity, np.int32)
    ms = max(max_possible_bins - 1, 1)
    thr = np.zeros((F, ms))
    ns = np.zeros(F, np.int32)
    lib().b2rf_oracle_find_splits(_p(sampleX), C.c_int64(n), C.c_int(F), _p(arity), C.c_int(max_possible_bins),
                                  C.c_double(weighted_num_samples), _p(thr), _p(ns))
    return thr, ns


def sample_keep(seed, rows, fraction):
    lib().b2rf_oracle_sample_keep.restype = C.c_int
    return np.array([lib().b2rf_oracle_sample_keep(C.c_int64(seed), C.c_int64(int(r)), C.c_double(fraction)) for r in rows], bool)


def bin_features(X, thresholds, nsplits, arity):
    X = np.ascontiguousarray(X, np.float64)
    N, F = X.shape
    out = np.zeros((N, F), np.int32)
    thresholds = np.ascontiguousarray(thresholds, np.float64)
    rc = lib().b2rf_oracle_bin(_p(X), C.c_int64(N), C.c_int(F), _p(thresholds), _p(np.ascontiguousarray(nsplits, np.int32)),
                               _p(np.ascontiguousarray(arity, np.int32)), C.c_int(thresholds.shape[1]), _p(out))
    if rc:
        raise OracleError(lib().b2rf_oracle_last_error().decode())
    return out


def bag(seed, part_offsets, T, rate, bootstrap):
    po = np.ascontiguousarray(part_offsets, np.int64)
    N = int(po[-1])
    out = np.zeros((N, T), np.uint8)
    lib().b2rf_oracle_bag(C.c_int64(seed), _p(po), C.c_int(len(po) - 1), C.c_int(T), C.c_double(rate), C.c_int(int(bootstrap)), _p(out))
    return out


def hist(bins, y, counts, tree_ids, slot, feats, off, total_cells):
    bins = np.ascontiguousarray(bins, np.int32); N, F = bins.shape
    y = np.ascontiguousarray(y, np.float64); counts = np.ascontiguousarray(counts, np.uint8)
    tree_ids = np.ascontiguousarray(tree_ids, np.int32); slot = np.ascontiguousarray(slot, np.int32)
    feats = np.ascontiguousarray(feats, np.int32); off = np.ascontiguousarray(off, np.int64)
    out = np.zeros((total_cells, 4))
    lib().b2rf_oracle_hist(_p(bins), C.c_int64(N), C.c_int(F), _p(y), _p(counts), C.c_int(counts.shape[1]), _p(tree_ids),
                           C.c_int(len(tree_ids)), _p(slot), _p(feats), C.c_int(feats.shape[1]), _p(off), _p(out))
    return out


@dataclass
class OracleForest:
    tree_off: np.ndarray
    feature: np.ndarray
    threshold: np.ndarray
    left: np.ndarray
    right: np.ndarray
    prediction: np.ndarray
    impurity: np.ndarray
    gain: np.ndarray
    stats: np.ndarray
    ncat: np.ndarray
    cat_off: np.ndarray
    cat_len: np.ndarray
    cats: np.ndarray
    thresholds: np.ndarray
    nsplits: np.ndarray
    group_sizes: np.ndarray
    trace_i: np.ndarray
    trace_d: np.ndarray
    timers: dict = field(default_factory=dict)
    updates: int = 0
    _h: object = None
    impurity_kind: str = "variance"
    is_single_tree: bool = False

    @property
    def num_trees(self):
        return len(self.tree_off) - 1

    def predict(self, X):
        X = np.ascontiguousarray(X, np.float64)
        if self.impurity_kind != "variance":   # raw2prediction: argmax of rawPrediction (first maximum)
            return np.argmax(self.predict_raw(X), axis=1).astype(np.float64)
        out = np.zeros(X.shape[0])
        lib().b2rf_oracle_predict(C.c_void_p(self._h), _p(X), C.c_int64(X.shape[0]), C.c_int(X.shape[1]), _p(out))
        return out

    def predict_raw(self, X, single_tree=None):
        """Classification rawPrediction [n, 2]: RandomForestClassificationModel.predictRaw (per-tree normalised votes summed in
        tree order) or, for a single DecisionTreeClassificationModel (single_tree=True), the leaf's class counts."""
        X = np.ascontiguousarray(X, np.float64)
        out = np.zeros((X.shape[0], 2))
        mode = 1 if (self.is_single_tree if single_tree is None else single_tree) else 0
        lib().b2rf_oracle_predict_raw(C.c_void_p(self._h), _p(X), C.c_int64(X.shape[0]), C.c_int(X.shape[1]), C.c_int(mode), _p(out))
        return out

    def __del__(self):
        if self._h:
            try:
                lib().b2rf_oracle_free(C.c_void_p(self._h))
            except Exception:
                pass
            self._h = None


def features_per_node(strategy: str, F: int, num_trees: int, regression: bool = True) -> int:
    """DecisionTreeMetadata.buildMetadata featureSubsetStrategy table."""
    import math
    s = strategy
    if s == "auto":
        s = "all" if num_trees == 1 else ("onethird" if regression else "sqrt")
    if s == "all":
        return F
    if s == "sqrt":
        return int(math.ceil(math.sqrt(F)))
    if s == "log2":
        return max(1, int(math.ceil(math.log(F) / math.log(2))))
    if s == "onethird":
        return int(math.ceil(F / 3.0))
    try:
        n = int(s)
        if n > 0:
            return min(n, F)
    except ValueError:
        pass
    try:
        x = float(s)
        if 0.0 < x <= 1.0:
            return int(math.ceil(x * F))
    except ValueError:
        pass
    raise ValueError(f"Supported values: auto, all, onethird, sqrt, log2, (0.0-1.0], [1-n]. Got {strategy!r}")


def fit(X, y, arity, *, maxDepth=5, maxBins=32, minInstancesPerNode=1, numTrees=1, featureSubsetStrategy="auto",
        maxMemoryInMB=256, bootstrap=False, minInfoGain=0.0, minWeightFractionPerNode=0.0, subsamplingRate=1.0, seed=0,
        part_offsets=None, sample_idx=None, fraction=1.0, numThreads=1, impurity="variance", single_tree=None) -> OracleForest:
    """impurity: "variance" (regression) or "gini" / "entropy" (binary classification, labels 0.0 / 1.0; featureSubsetStrategy
    "auto" then resolves to sqrt for numTrees > 1).  single_tree: the model is a DecisionTree*Model (default: numTrees == 1)."""
    X = np.ascontiguousarray(X, np.float64); y = np.ascontiguousarray(y, np.float64)
    N, F = X.shape
    kind = impurity
    arity = np.ascontiguousarray(arity, np.int32)
    if part_offsets is None:
        part_offsets = np.array([0, N], np.int64)
    po = np.ascontiguousarray(part_offsets, np.int64)
    pr = Params(maxDepth, maxBins, minInstancesPerNode, numTrees, features_per_node(featureSubsetStrategy, F, numTrees, regression=(kind == "variance")),
                maxMemoryInMB, int(bootstrap), numThreads, minInfoGain, minWeightFractionPerNode, subsamplingRate, seed, IMPURITIES[kind], 0)
    h = C.c_void_p()
    si = None if sample_idx is None else np.ascontiguousarray(sample_idx, np.int64)
    rc = lib().b2rf_oracle_fit(_p(X), C.c_int64(N), C.c_int(F), _p(y), _p(arity), C.byref(pr), _p(po), C.c_int(len(po) - 1),
                               None if si is None else _p(si), C.c_int64(0 if si is None else len(si)), C.c_double(fraction), C.byref(h))
    if rc:
        raise OracleError(lib().b2rf_oracle_last_error().decode())
    L = lib()
    n = L.b2rf_oracle_num_nodes(h); nc = L.b2rf_oracle_num_cats(h); ng = L.b2rf_oracle_num_groups(h); nt = L.b2rf_oracle_num_trace(h)
    T = numTrees
    tree_off = np.zeros(T + 1, np.int64)
    feature = np.zeros(n, np.int32); threshold = np.zeros(n); left = np.zeros(n, np.int32); right = np.zeros(n, np.int32)
    prediction = np.zeros(n); impurity = np.zeros(n); gain = np.zeros(n); stats = np.zeros((n, 4))
    ncat = np.zeros(n, np.int32); cat_off = np.zeros(n, np.int32); cat_len = np.zeros(n, np.int32); cats = np.zeros(max(nc, 1))
    L.b2rf_oracle_export(h, _p(tree_off), _p(feature), _p(threshold), _p(left), _p(right), _p(prediction), _p(impurity), _p(gain),
                         _p(stats), _p(ncat), _p(cat_off), _p(cat_len), _p(cats))
    mpb = min(maxBins, N)
    thr = np.zeros((F, max(mpb - 1, 1))); ns = np.zeros(F, np.int32)
    L.b2rf_oracle_export_splits(h, _p(thr), _p(ns))
    gs = np.zeros(ng, np.int32); L.b2rf_oracle_export_groups(h, _p(gs))
    ti = np.zeros((nt, 5), np.int32); td = np.zeros((nt, 5)); L.b2rf_oracle_export_trace(h, _p(ti), _p(td))
    tm = np.zeros(5); L.b2rf_oracle_timers(h, _p(tm))
    return OracleForest(tree_off, feature, threshold, left, right, prediction, impurity, gain, stats, ncat, cat_off, cat_len, cats[:nc],
                        thr, ns, gs, ti, td, dict(init=tm[0], findSplits=tm[1], findBestSplits=tm[2], chooseSplits=tm[3], total=tm[4]),
                        int(L.b2rf_oracle_updates(h)), h.value, kind, (numTrees == 1) if single_tree is None else bool(single_tree))
