"""b2rf — B200-native MLlib-compatible DecisionTreeRegressor / RandomForestRegressor training.

Hot path only: quantile split finding, binning, Poisson bagging, per-group histogram build, variance
split scan, row routing and forest prediction as hand-written sm_100a CUDA kernels behind the C-ABI in include/b2rf.h.
"""
from .params import TreeParams  # noqa: F401
from .forest import Forest, java_double_str  # noqa: F401
from ._lib import B2rfError, IllegalArgumentException  # noqa: F401
from . import synth  # noqa: F401


def __getattr__(name):
    if name in ("fit_tensors", "fit_many_tensors", "fit_host", "find_splits", "bin_features", "bag", "hist_build", "sample_rows", "default_partitions"):
        from . import trainer
        return getattr(trainer, name)
    raise AttributeError(name)
