"""ctypes binding of libb2rf.so — the C-ABI declared in include/b2rf.h.  Fails loudly when the CUDA library is missing:
there is no CPU fallback in this package."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.environ.get("B2RF_LIB") or os.path.join(HERE, "libb2rf.so")   # B2RF_LIB: A/B-test another build of the library

B2RF_OK, ERR_ARG, ERR_MAXBINS, ERR_DATA, ERR_CUDA, ERR_STATE = 0, -1, -2, -3, -4, -5


class B2rfParams(C.Structure):
    _fields_ = [("maxDepth", C.c_int32), ("maxBins", C.c_int32), ("minInstancesPerNode", C.c_int32), ("numTrees", C.c_int32),
                ("featuresPerNode", C.c_int32), ("maxMemoryInMB", C.c_int32), ("bootstrap", C.c_int32), ("histMode", C.c_int32),
                ("minInfoGain", C.c_double), ("minWeightFractionPerNode", C.c_double), ("subsamplingRate", C.c_double),
                ("seed", C.c_int64), ("impurity", C.c_int32), ("numClasses", C.c_int32)]


class B2rfError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(msg)
        self.code = code


class IllegalArgumentException(B2rfError, ValueError):
    """Raised where MLlib raises IllegalArgumentException (pyspark re-raises it under this name)."""


_lib = None
vp, i32, i64, f64, sz = C.c_void_p, C.c_int32, C.c_int64, C.c_double, C.c_size_t

_SIGS = {
    "b2rf_version": (C.c_int, []),
    "b2rf_last_error": (C.c_char_p, []),
    "b2rf_features_per_node": (C.c_int, [C.c_char_p, i32, i32]),
    "b2rf_check_metadata": (C.c_int, [vp, i32, vp, i64]),
    "b2rf_samples_fraction": (f64, [i64, i32]),
    "b2rf_sample_rows": (C.c_int, [vp, i64, i64, i64, i32, i64, f64, vp, i64, vp, vp]),
    "b2rf_sample_indices": (i64, [i64, i64, i64, f64, vp, i64]),
    "b2rf_sample_row_ids": (C.c_int, [i64, i64, i64, f64, vp, i64, vp, vp]),
    "b2rf_find_splits_workspace_bytes": (sz, [i64, i32]),
    "b2rf_find_splits": (C.c_int, [vp, i64, i32, vp, i32, f64, vp, vp, vp, sz, vp]),
    "b2rf_find_splits_noalloc_max_bytes": (sz, []),
    "b2rf_find_splits_noalloc": (C.c_int, [vp, i64, i32, vp, i32, f64, vp, vp, vp]),
    "b2rf_bin": (C.c_int, [vp, i64, i64, i32, vp, vp, vp, i32, vp, i64, i64, vp, vp, i64, i32, vp]),
    "b2rf_bag": (C.c_int, [i64, vp, i32, i32, i32, f64, i32, vp, i64, vp, i64, vp]),
    "b2rf_trainer_create": (C.c_int, [vp, i32, vp, vp, vp, i64, vp]),
    "b2rf_trainer_destroy": (None, [vp]),
    "b2rf_trainer_max_group_cells": (i64, [vp]),
    "b2rf_trainer_code_bytes": (C.c_int, [vp]),
    "b2rf_trainer_bind": (C.c_int, [vp, vp, i64, vp, i64, vp, vp, i64, vp, i64, i64, vp, i64]),
    "b2rf_trainer_begin_group": (C.c_int, [vp, vp, vp]),
    "b2rf_trainer_end_group": (C.c_int, [vp, vp]),
    "b2rf_trainer_stats": (C.c_int, [vp, vp, vp]),
    "b2rf_trainer_forest_sizes": (C.c_int, [vp, vp, vp]),
    "b2rf_trainer_export": (C.c_int, [vp] * 14),
    "b2rf_trainer_set_trace": (C.c_int, [vp, i32]),
    "b2rf_trainer_trace_size": (i64, [vp]),
    "b2rf_trainer_trace": (C.c_int, [vp, vp, vp]),
    "b2rf_k4_launch_counts": (C.c_int, [vp, i32]),
    "b2rf_trainer_set_shard": (C.c_int, [vp, i32, i32]),
    "b2rf_trainer_shard_block_doubles": (i64, [vp]),
    "b2rf_trainer_end_group_shard_launch": (C.c_int, [vp, vp, vp]),
    "b2rf_trainer_end_group_shard_finish": (C.c_int, [vp, vp, vp]),
    "b2rf_trainer_group_info": (C.c_int, [vp] * 6),
    "b2rf_trainers_begin_group_shared": (C.c_int, [vp, i32, vp, vp, vp]),
    "b2rf_trainers_end_group_shared": (C.c_int, [vp, i32, vp]),
    "b2rf_set_option": (C.c_int, [C.c_char_p, i64]),
    "b2rf_hist_build": (C.c_int, [vp, i64, i32, i32, vp, vp, i64, vp, i64, vp, i32, i64, vp, vp, i32, vp, vp]),
    "b2rf_predict": (C.c_int, [vp, i64, i64, i32] + [vp] * 12),
    "b2rf_forest_pack_bytes": (C.c_size_t, [i64]),
    "b2rf_forest_pack": (C.c_int, [i64] + [vp] * 10),
    "b2rf_predict_packed": (C.c_int, [vp, i64, i64, i32] + [vp] * 4 + [i32] + [vp] * 3),
    "b2rf_predict_coded": (C.c_int, [vp, i64, i32, i64, i32] + [vp] * 6),
    "b2rf_predict_raw_mc": (C.c_int, [vp, i64, i64, i32] + [vp] * 6 + [i32] + [vp] * 4 + [i32, vp, vp]),
    "b2rf_trainer_num_classes": (C.c_int, [vp]),
    "b2rf_trainer_export_class_counts": (C.c_int, [vp, vp]),
    "b2rf_predict_raw": (C.c_int, [vp, i64, i64, i32] + [vp] * 10 + [i32, vp, vp]),
}

EXPORTS = tuple(_SIGS)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(SO):
            raise ImportError(f"{SO} is missing: build it with __graft_entry__.build() (nvcc, sm_100a). "
                              "This package has no CPU fallback.")
        L = C.CDLL(SO)
        for name, (res, args) in _SIGS.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc: int) -> int:
    if rc < 0:
        msg = lib().b2rf_last_error().decode("utf-8", "replace")
        if rc in (ERR_MAXBINS, ERR_DATA):
            raise IllegalArgumentException(rc, msg)
        raise B2rfError(rc, msg or f"b2rf error {rc}")
    return rc
