"""Tree-training parameters with pyspark.ml's names and defaults (reference call sites
`ML 06 - Decision Trees.py:75,118`, `ML 07 - Random Forests and Hyperparameter Tuning.py:41,74-77`, `ML 08 - Hyperopt.py:71,97`)."""
from __future__ import annotations

from dataclasses import dataclass, replace

from ._lib import B2rfParams, check, lib


def _to_int(v, name):
    """pyspark TypeConverters.toInt: accepts integral floats (`ML 08 - Hyperopt.py:97,120-121` passes hp.quniform floats)."""
    if isinstance(v, bool):
        raise TypeError(f"Could not convert {v!r} to int for Param {name}")
    if isinstance(v, int):
        return v
    if isinstance(v, float) and v.is_integer():
        return int(v)
    try:
        import numpy as np
        if isinstance(v, np.integer):
            return int(v)
        if isinstance(v, np.floating) and float(v).is_integer():
            return int(v)
    except ImportError:
        pass
    raise TypeError(f"Could not convert {v!r} to int for Param {name}")


@dataclass
class TreeParams:
    maxDepth: int = 5
    maxBins: int = 32
    minInstancesPerNode: int = 1
    minWeightFractionPerNode: float = 0.0
    minInfoGain: float = 0.0
    maxMemoryInMB: int = 256
    cacheNodeIds: bool = False        # accepted, ignored (row->node ids are always cached on the device)
    checkpointInterval: int = 10      # accepted, ignored
    impurity: str = "variance"
    subsamplingRate: float = 1.0
    numTrees: int = 1
    featureSubsetStrategy: str = "auto"
    bootstrap: bool = False           # DecisionTreeRegressor: False; RandomForestRegressor default: True
    seed: int = 0
    numClasses: int = 0               # classification only: 0 / 2 = binary; 3..32 = multiclass (per-class histograms)

    def copy(self, **kw) -> "TreeParams":
        return replace(self, **kw)

    IMPURITIES = {"variance": 0, "gini": 1, "entropy": 2}

    @property
    def is_classification(self) -> bool:
        return self.impurity != "variance"

    @property
    def num_classes(self) -> int:
        return 0 if not self.is_classification else max(2, int(self.numClasses))

    def validate(self):
        if self.impurity not in self.IMPURITIES:
            raise ValueError(f"impurity must be 'variance' (regression) or 'gini' / 'entropy' (classification), got {self.impurity!r}")
        if self.numClasses and (not self.is_classification or not (2 <= int(self.numClasses) <= 32)):
            raise ValueError(f"numClasses = {self.numClasses}: tree classifiers here support 2..32 classes (MLlib allows up to 100), and "
                             f"regression takes none")
        if self.maxDepth < 0 or self.maxDepth > 30:
            raise ValueError(f"DecisionTree currently only supports maxDepth <= 30, but was given maxDepth = {self.maxDepth}.")
        if self.maxBins < 2:
            raise ValueError(f"maxBins must be >= 2, got {self.maxBins}")
        if not (0.0 < self.subsamplingRate <= 1.0):
            raise ValueError(f"subsamplingRate must be in (0, 1], got {self.subsamplingRate}")
        if self.minInstancesPerNode < 1:
            raise ValueError("minInstancesPerNode must be >= 1")
        if not (0.0 <= self.minWeightFractionPerNode < 0.5):
            raise ValueError("minWeightFractionPerNode must be in [0, 0.5)")

    def features_per_node(self, F: int) -> int:
        strategy = str(self.featureSubsetStrategy)
        if strategy == "auto" and self.is_classification:   # DecisionTreeMetadata: classification forests default to sqrt
            strategy = "all" if _to_int(self.numTrees, "numTrees") == 1 else "sqrt"
        return check(lib().b2rf_features_per_node(strategy.encode(), int(F), _to_int(self.numTrees, 'numTrees')))

    def to_c(self, F: int) -> B2rfParams:
        self.validate()
        return B2rfParams(_to_int(self.maxDepth, "maxDepth"), _to_int(self.maxBins, "maxBins"),
                          _to_int(self.minInstancesPerNode, "minInstancesPerNode"), _to_int(self.numTrees, "numTrees"),
                          self.features_per_node(F), int(self.maxMemoryInMB), int(bool(self.bootstrap)), 0, float(self.minInfoGain),
                          float(self.minWeightFractionPerNode), float(self.subsamplingRate), int(self.seed), self.IMPURITIES[self.impurity],
                          int(self.numClasses) if int(self.numClasses) >= 3 else 0)
