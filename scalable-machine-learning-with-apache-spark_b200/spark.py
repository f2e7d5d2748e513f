"""Drop-in pyspark.ml Estimator / Model classes for the tree stage of the labs' Pipelines.

    from b2rf.spark import B200DecisionTreeRegressor as DecisionTreeRegressor      # `ML 06 - Decision Trees.py:73-75`
    from b2rf.spark import B200RandomForestRegressor as RandomForestRegressor      # `ML 07 ...:25,41`, `ML 08 - Hyperopt.py:59,71`

They accept the same Params with the same defaults and converters as pyspark 3.1.2 (kwargs constructor, setX/getX,
setParams, copy(extra) keeping the uid, explainParams), read categorical arities from the `ml_attr` metadata that
StringIndexer -> VectorAssembler attach to the features column (`ML 06:47,60-63`), override fitMultiple so a
CrossValidator placed around the tree stage trains every candidate on ONE binned copy of the fold
(`ML 07:143-149`), and append a double `prediction` column in _transform.

pyspark cannot be installed in this image (no JVM, no wheel): when it is missing the same classes are built on
_pyspark_stub.py, which is what tests/test_cpu_spark_adapter.py and the GPU adapter test exercise.
"""
from __future__ import annotations

import threading

import numpy as np

try:  # real Spark
    from pyspark.ml import Estimator, Model
    from pyspark.ml.linalg import DenseVector, SparseVector
    from pyspark.ml.param import Param, Params, TypeConverters
    HAVE_PYSPARK = True
except ImportError:  # this image
    from ._pyspark_stub import DenseVector, Estimator, Model, Param, Params, SparseVector, TypeConverters
    HAVE_PYSPARK = False

from .forest import Forest
from .params import TreeParams


def _java_string_hash(s: str) -> int:
    """pyspark's HasSeed default is hash(type(self).__name__) — process-dependent; Scala's is className.hashCode.
    A stable value is used here so an unseeded fit (`ML 07:41`) is at least reproducible across processes."""
    h = 0
    for ch in s:
        h = (31 * h + ord(ch)) & 0xFFFFFFFF
    return h - (1 << 32) if h >= (1 << 31) else h


class _TreeRegressorParams(Params):
    """Param declarations with pyspark.ml.tree._DecisionTreeParams/_TreeEnsembleParams/_RandomForestParams names and docs."""
    featuresCol = Param(Params, "featuresCol", "features column name.", TypeConverters.toString)
    labelCol = Param(Params, "labelCol", "label column name.", TypeConverters.toString)
    predictionCol = Param(Params, "predictionCol", "prediction column name.", TypeConverters.toString)
    maxDepth = Param(Params, "maxDepth", "Maximum depth of the tree. (>= 0) E.g., depth 0 means 1 leaf node; depth 1 means 1 internal "
                     "node + 2 leaf nodes. Must be in range [0, 30].", TypeConverters.toInt)
    maxBins = Param(Params, "maxBins", "Max number of bins for discretizing continuous features.  Must be >=2 and >= number of categories "
                    "for any categorical feature.", TypeConverters.toInt)
    minInstancesPerNode = Param(Params, "minInstancesPerNode", "Minimum number of instances each child must have after split. If a split "
                                "causes the left or right child to have fewer than minInstancesPerNode, the split will be discarded as "
                                "invalid. Should be >= 1.", TypeConverters.toInt)
    minWeightFractionPerNode = Param(Params, "minWeightFractionPerNode", "Minimum fraction of the weighted sample count that each child must "
                                     "have after split. Should be in interval [0.0, 0.5).", TypeConverters.toFloat)
    minInfoGain = Param(Params, "minInfoGain", "Minimum information gain for a split to be considered at a tree node.", TypeConverters.toFloat)
    maxMemoryInMB = Param(Params, "maxMemoryInMB", "Maximum memory in MB allocated to histogram aggregation. If too small, then 1 node will "
                          "be split per iteration, and its aggregates may exceed this size.", TypeConverters.toInt)
    cacheNodeIds = Param(Params, "cacheNodeIds", "If false, the algorithm will pass trees to executors to match instances with nodes. If "
                         "true, the algorithm will cache node IDs for each instance.", TypeConverters.toBoolean)
    checkpointInterval = Param(Params, "checkpointInterval", "set checkpoint interval (>= 1) or disable checkpoint (-1).", TypeConverters.toInt)
    impurity = Param(Params, "impurity", "Criterion used for information gain calculation (case-insensitive). Supported options: variance "
                     "(regressors); entropy, gini (classifiers)",
                     TypeConverters.toString)
    seed = Param(Params, "seed", "random seed.", TypeConverters.toInt)
    leafCol = Param(Params, "leafCol", "Leaf indices column name. Predicted leaf index of each instance in each tree by preorder.",
                    TypeConverters.toString)
    weightCol = Param(Params, "weightCol", "weight column name. If this is not set or empty, we treat all instance weights as 1.0.",
                      TypeConverters.toString)

    def _base_defaults(self):
        self._setDefault(featuresCol="features", labelCol="label", predictionCol="prediction", maxDepth=5, maxBins=32, minInstancesPerNode=1,
                         minWeightFractionPerNode=0.0, minInfoGain=0.0, maxMemoryInMB=256, cacheNodeIds=False, checkpointInterval=10,
                         impurity="variance", leafCol="", seed=_java_string_hash(type(self).__name__))


def _accessors(cls, names):
    for n in names:
        cap = n[0].upper() + n[1:]

        def getter(self, _n=n):
            return self.getOrDefault(_n)

        def setter(self, value, _n=n):
            return self._set(**{_n: value})
        if not hasattr(cls, "get" + cap):
            setattr(cls, "get" + cap, getter)
        if not hasattr(cls, "set" + cap):
            setattr(cls, "set" + cap, setter)
    return cls


_COMMON = ["featuresCol", "labelCol", "predictionCol", "maxDepth", "maxBins", "minInstancesPerNode", "minWeightFractionPerNode", "minInfoGain",
           "maxMemoryInMB", "cacheNodeIds", "checkpointInterval", "impurity", "seed", "leafCol", "weightCol"]


# ------------------------------------------------------------------------------------------------------------------
# dataset plumbing
# ------------------------------------------------------------------------------------------------------------------
def categorical_arities(dataset, featuresCol: str, num_features: int) -> np.ndarray:
    """MetadataUtils.getCategoricalFeatures: nominal attrs -> len(vals) (or num_vals); binary -> 2; no metadata -> continuous."""
    arity = np.zeros(num_features, np.int32)
    try:
        meta = dataset.schema[featuresCol].metadata
    except Exception:  # noqa: BLE001
        meta = {}
    attrs = (meta or {}).get("ml_attr", {}).get("attrs", {})
    for a in attrs.get("nominal", []):
        k = len(a["vals"]) if "vals" in a else int(a.get("num_vals", 0))
        if 0 <= a["idx"] < num_features:
            arity[a["idx"]] = k
    for a in attrs.get("binary", []):
        if 0 <= a["idx"] < num_features:
            arity[a["idx"]] = 2
    return arity


def _extract(dataset, featuresCol, labelCol=None):
    """(X f64 [n, F] host, y f64 [n] host or None, partition sizes).  Real Spark: ONE job — Arrow toPandas() of (features array,
    label, spark_partition_id()); rows arrive partition by partition, so the partition sizes bagging needs (A.4) come from the same
    evaluation of the plan as the rows themselves.  Stub: LocalFrame."""
    if HAVE_PYSPARK and hasattr(dataset, "sparkSession"):
        from pyspark.ml.functions import vector_to_array
        from pyspark.sql import functions as F
        cols = [vector_to_array(F.col(featuresCol)).alias("f"), F.spark_partition_id().alias("p")]
        if labelCol:
            cols.append(F.col(labelCol).cast("double").alias("y"))
        pdf = dataset.select(*cols).toPandas()
        X = np.ascontiguousarray(np.stack(pdf["f"].to_numpy()), dtype=np.float64) if len(pdf) else np.zeros((0, 0))
        y = pdf["y"].to_numpy(np.float64) if labelCol else None
        pid = pdf["p"].to_numpy(np.int64)
        if len(pid) and (np.diff(pid) < 0).any():
            raise RuntimeError("toPandas() did not return the rows in partition order; cannot reproduce MLlib's per-partition bagging streams")
        sizes = np.bincount(pid, minlength=int(pid.max()) + 1 if len(pid) else 1).tolist()
        return X, y, sizes
    X = np.ascontiguousarray(dataset.col(featuresCol), dtype=np.float64)
    y = np.ascontiguousarray(dataset.col(labelCol), dtype=np.float64) if labelCol else None
    n = X.shape[0]
    P = max(1, int(getattr(dataset, "num_partitions", 1)))
    sizes = [n // P + (1 if i < n % P else 0) for i in range(P)]
    return X, y, sizes


def _score_dataset(dataset, featuresCol, num_features, score_fn):
    """Model._transform: appends the columns `score_fn(X f64 [n, F]) -> {name: 1-D array (double column) | 2-D array (vector column)}`.

    Real Spark: one pass over the data, no driver collect and no positional join — a scalar pandas UDF receives each Arrow batch of
    vector_to_array(features), scores it on the executor's GPU and returns every output column as one struct, which is then
    unpacked (asNondeterministic keeps Catalyst from inlining the UDF once per extracted field).  [Written against the pyspark 3.1
    API; pyspark is unavailable in this image, so this branch is untested here — INTEGRATION.md.]  Stub: LocalFrame columns."""
    if HAVE_PYSPARK and hasattr(dataset, "sparkSession"):
        import pandas as pd
        from pyspark.ml.functions import array_to_vector, vector_to_array
        from pyspark.sql import functions as F
        from pyspark.sql import types as T
        probe = score_fn(np.zeros((0, num_features)))
        fields = [T.StructField(n, T.DoubleType() if np.asarray(v).ndim == 1 else T.ArrayType(T.DoubleType())) for n, v in probe.items()]
        if not fields:
            return dataset

        def score(feats: pd.Series) -> pd.DataFrame:
            X = np.ascontiguousarray(np.stack(feats.to_numpy()), np.float64) if len(feats) else np.zeros((0, num_features))
            cols = score_fn(X)
            return pd.DataFrame({n: (np.asarray(v, np.float64) if np.asarray(v).ndim == 1 else [list(map(float, r)) for r in np.asarray(v)])
                                 for n, v in cols.items()})
        udf = F.pandas_udf(score, returnType=T.StructType(fields)).asNondeterministic()
        out = dataset.withColumn("__b2rf_s", udf(vector_to_array(F.col(featuresCol))))
        for n, v in probe.items():
            c = F.col("__b2rf_s")[n]
            out = out.withColumn(n, c if np.asarray(v).ndim == 1 else array_to_vector(c))
        return out.drop("__b2rf_s")
    X = np.ascontiguousarray(dataset.col(featuresCol), dtype=np.float64)
    for name, v in score_fn(X).items():
        dataset = dataset.withColumn(name, np.asarray(v, np.float64))
    return dataset


class _ForestModelBase(Model, _TreeRegressorParams):
    def __init__(self, forest: Forest = None):
        super().__init__()
        self._base_defaults()
        self._forest = forest

    # Model.copy(extra) must carry the fitted forest: CrossValidator evaluates with model.transform(validation, epm[i])
    def copy(self, extra=None):
        that = super().copy(extra)
        that._forest = self._forest
        return that

    @property
    def numFeatures(self):
        return self._forest.num_features

    @property
    def featureImportances(self):
        imp = self._forest.feature_importances()
        nz = np.nonzero(imp)[0]
        return SparseVector(len(imp), nz, imp[nz])

    @property
    def toDebugString(self):
        self._forest.uid = self.uid
        return self._forest.to_debug_string(ensemble=self._ensemble)

    # ---- MLWritable / MLReadable (MLlib 3.1.2 directory layout, persistence.py) ----
    def write(self):
        from .persistence import ModelWriter
        return ModelWriter(self)

    def save(self, path):
        self.write().save(path)

    @classmethod
    def load(cls, path):
        from .persistence import load_forest
        meta, forest = load_forest(path)
        m = cls(forest)
        m.uid = meta["uid"]
        for p in list(m.params):
            setattr(m, p.name, p._copy_new_parent(m)) if hasattr(p, "_copy_new_parent") else None
        m._paramMap, m._defaultParamMap = {}, {}
        m._base_defaults()
        if getattr(m, "_ensemble", False):
            m._setDefault(numTrees=20, featureSubsetStrategy="auto", subsamplingRate=1.0, bootstrap=True)
        for k, v in meta.get("paramMap", {}).items():
            if m.hasParam(k):
                m._set(**{k: v})
        return m

    @classmethod
    def read(cls):
        class _R:
            @staticmethod
            def load(path):
                return cls.load(path)
        return _R()

    def predict(self, value):
        import torch
        v = np.asarray(value.toArray() if hasattr(value, "toArray") else value, np.float64).reshape(1, -1)
        return float(self._forest.predict(torch.as_tensor(v, device=self._device())).item())

    @staticmethod
    def _device():
        import torch
        if not torch.cuda.is_available():
            raise RuntimeError("b2rf models predict on the GPU: no CUDA device is visible and there is no CPU fallback")
        return torch.device("cuda", torch.cuda.current_device())

    def _leaf_columns(self, X):
        """leafCol (every tree model: DecisionTreeModel.predictLeaf, the leaf's rank among its tree's leaves, a double for a single
        tree and a vector for an ensemble) and varianceCol (DecisionTreeRegressionModel.predictVariance: the leaf's impurity)."""
        import torch
        cols = {}
        leaf_col = self.getOrDefault("leafCol") if self.hasParam("leafCol") and self.isDefined("leafCol") else ""
        var_col = self.getOrDefault("varianceCol") if self.hasParam("varianceCol") and self.isDefined("varianceCol") else ""
        if not leaf_col and not var_col:
            return cols
        fo = self._forest
        if X.shape[0] == 0:
            leaves = np.zeros((0, fo.num_trees), np.int64)
        else:
            leaves = fo.predict_leaves(torch.as_tensor(X, device=self._device())).cpu().numpy().astype(np.int64)
        flat = leaves + fo.tree_off[:-1][None, :]
        if leaf_col:
            ords = fo.leaf_ordinals()[flat].astype(np.float64)
            cols[leaf_col] = ords if self._ensemble else ords[:, 0]
        if var_col:
            cols[var_col] = fo.impurity[flat[:, 0]]
        return cols

    def _score_columns(self, X):
        """Output columns for a batch of feature rows (predictionCol, leafCol, varianceCol)."""
        import torch
        cols = {}
        if self.getOrDefault("predictionCol"):
            cols[self.getOrDefault("predictionCol")] = (self._forest.predict(torch.as_tensor(X, device=self._device())).cpu().numpy()
                                                        if X.shape[0] else np.zeros(0))
        cols.update(self._leaf_columns(X))
        return cols

    def _transform(self, dataset):
        return _score_dataset(dataset, self.getOrDefault("featuresCol"), self._forest.num_features, self._score_columns)


_VARIANCE_COL = Param(Params, "varianceCol", "column name for the biased sample variance of prediction.", TypeConverters.toString)


class B200DecisionTreeRegressionModel(_ForestModelBase):
    _ensemble = False
    varianceCol = _VARIANCE_COL

    def predictVariance(self, value):
        import torch
        v = np.asarray(value.toArray() if hasattr(value, "toArray") else value, np.float64).reshape(1, -1)
        leaf = int(self._forest.predict_leaves(torch.as_tensor(v, device=self._device()))[0, 0].item())
        return float(self._forest.impurity[leaf])

    def predictLeaf(self, value):
        import torch
        v = np.asarray(value.toArray() if hasattr(value, "toArray") else value, np.float64).reshape(1, -1)
        leaf = int(self._forest.predict_leaves(torch.as_tensor(v, device=self._device()))[0, 0].item())
        return float(self._forest.leaf_ordinals()[leaf])

    @property
    def depth(self):
        return self._forest.depth(0)

    @property
    def numNodes(self):
        return self._forest.num_nodes(0)


class B200RandomForestRegressionModel(_ForestModelBase):
    _ensemble = True
    _tree_model_cls = B200DecisionTreeRegressionModel
    numTrees = Param(Params, "numTrees", "Number of trees to train (>= 1).", TypeConverters.toInt)
    featureSubsetStrategy = Param(Params, "featureSubsetStrategy", "The number of features to consider for splits at each tree node.",
                                  TypeConverters.toString)
    subsamplingRate = Param(Params, "subsamplingRate", "Fraction of the training data used for learning each decision tree, in range (0, 1].",
                            TypeConverters.toFloat)
    bootstrap = Param(Params, "bootstrap", "Whether bootstrap samples are used when building trees.", TypeConverters.toBoolean)

    def __init__(self, forest=None):
        super().__init__(forest)
        self._setDefault(numTrees=20, featureSubsetStrategy="auto", subsamplingRate=1.0, bootstrap=True)

    @property
    def getNumTrees(self):
        return self._forest.num_trees

    @property
    def treeWeights(self):
        return [1.0] * self._forest.num_trees

    @property
    def totalNumNodes(self):
        return self._forest.num_nodes()

    @property
    def trees(self):
        out = []
        fo = self._forest
        for t in range(fo.num_trees):
            s = fo.tree_slice(t)
            sub = Forest(fo.num_features, np.array([0, s.stop - s.start], np.int64), fo.feature[s], fo.threshold[s], fo.left[s], fo.right[s],
                         fo.prediction[s], fo.impurity[s], fo.gain[s], fo.stats[s], fo.ncat[s], fo.cat_off[s], fo.cat_len[s], fo.cats,
                         impurity_kind=fo.impurity_kind, counts=None if fo.counts is None else fo.counts[s])
            out.append(self._tree_model_cls(sub))
        return out


class _TreeEstimatorBase(Estimator, _TreeRegressorParams):
    _model_cls = None

    def _tree_params(self) -> TreeParams:
        g = self.getOrDefault
        return TreeParams(maxDepth=g("maxDepth"), maxBins=g("maxBins"), minInstancesPerNode=g("minInstancesPerNode"),
                          minWeightFractionPerNode=g("minWeightFractionPerNode"), minInfoGain=g("minInfoGain"), maxMemoryInMB=g("maxMemoryInMB"),
                          impurity=g("impurity").lower(), seed=g("seed"), **self._ensemble_params())

    def _make_model(self, forest):
        m = self._model_cls(forest)
        m.uid = self.uid                       # pyspark models carry their estimator's uid, so shared Params route to them
        for p in list(m.params):
            setattr(m, p.name, p._copy_new_parent(m)) if hasattr(p, "_copy_new_parent") else None
        m._paramMap = {}
        m._defaultParamMap = {}
        m._base_defaults()
        if getattr(m, "_ensemble", False):
            m._setDefault(numTrees=20, featureSubsetStrategy="auto", subsamplingRate=1.0, bootstrap=True)
        self._copyValues(m)
        return m

    def _fit_many(self, dataset, estimators):
        """Train several parameterisations on one extraction of the dataset (one host->device copy, shared by all)."""
        import torch
        from . import trainer
        est0 = estimators[0]
        for e in estimators:
            if e.isDefined("weightCol") and e.getOrDefault("weightCol"):
                raise NotImplementedError("weightCol is not supported: K4 accumulates integer bag counts (every instance weight 1.0), "
                                          "which is what every tree fit of the reference uses")
        X, y, sizes = _extract(dataset, est0.getOrDefault("featuresCol"), est0.getOrDefault("labelCol"))
        if X.shape[0] == 0:
            raise ValueError("DecisionTree requires size of input RDD > 0, but was given by empty one.")
        arity = categorical_arities(dataset, est0.getOrDefault("featuresCol"), X.shape[1])
        po = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
        dev = _ForestModelBase._device()
        plist = [e._tree_params() for e in estimators]
        if plist[0].is_classification:
            # Classifier.getNumClasses without label metadata: max label + 1 (MLlib stops at 100 classes, this build at 32)
            ok = len(y) > 0 and bool(np.all((y >= 0) & (y == np.floor(y))))
            nc = int(y.max()) + 1 if ok else 0
            if not ok or nc > 32:
                from ._lib import IllegalArgumentException
                raise IllegalArgumentException(-3, "requirement failed: Classifier inferred " + str(nc) + " from label values in column "
                                               + est0.getOrDefault("labelCol") + ", but labels must be the integers 0 .. numClasses - 1 with "
                                               "numClasses <= 32 here (MLlib: <= 100).")
            plist = [p.copy(numClasses=max(nc, 2)) for p in plist]
        if len(estimators) == 1:
            # one fit: the host-buffer entry point streams X through two staging buffers (copy of one row chunk while the previous one is
            # binned), so the f64 feature matrix is never resident on the GPU
            yield estimators[0]._make_model(trainer.fit_host(X, y, arity, plist[0], device=dev, part_offsets=po))
            return
        # several candidates (fitMultiple): one device copy of the fold, binned once, histogram passes shared (fit_many_tensors)
        Xd = torch.as_tensor(X, device=dev)
        yd = torch.as_tensor(y, device=dev)
        forests = trainer.fit_many_tensors(Xd, yd, arity, plist, part_offsets=po)   # shared passes
        for est, forest in zip(estimators, forests):
            yield est._make_model(forest)

    def _fit(self, dataset):
        return next(self._fit_many(dataset, [self]))

    def fitMultiple(self, dataset, paramMaps):
        """Thread-safe iterator of (index, model) (CrossValidator's ThreadPool calls next() concurrently, `ML 07:130,143-144`).
        All candidates share one extraction / device copy of the fold."""
        ests = [self.copy(pm) for pm in paramMaps]
        gen = self._fit_many(dataset, ests)
        lock = threading.Lock()
        counter = iter(range(len(ests)))

        class _It:
            def __iter__(s):
                return s

            def __next__(s):
                with lock:
                    i = next(counter)
                    return i, next(gen)
            next = __next__
        return _It()


class B200DecisionTreeRegressor(_TreeEstimatorBase):
    """pyspark.ml.regression.DecisionTreeRegressor, trained on the GPU (`ML 06 - Decision Trees.py:73-75,118,126`)."""
    _model_cls = B200DecisionTreeRegressionModel
    varianceCol = _VARIANCE_COL

    def __init__(self, **kwargs):
        super().__init__()
        self._base_defaults()
        self.setParams(**kwargs)

    def setParams(self, **kwargs):
        return self._set(**kwargs)

    def _ensemble_params(self):
        return dict(numTrees=1, featureSubsetStrategy="all", subsamplingRate=1.0, bootstrap=False)


class B200RandomForestRegressor(_TreeEstimatorBase):
    """pyspark.ml.regression.RandomForestRegressor, trained on the GPU (`ML 07 ...:41,74-77`, `ML 08 - Hyperopt.py:71,97`)."""
    _model_cls = B200RandomForestRegressionModel
    numTrees = B200RandomForestRegressionModel.numTrees
    featureSubsetStrategy = B200RandomForestRegressionModel.featureSubsetStrategy
    subsamplingRate = B200RandomForestRegressionModel.subsamplingRate
    bootstrap = B200RandomForestRegressionModel.bootstrap

    def __init__(self, **kwargs):
        super().__init__()
        self._base_defaults()
        self._setDefault(numTrees=20, featureSubsetStrategy="auto", subsamplingRate=1.0, bootstrap=True)
        self.setParams(**kwargs)

    def setParams(self, **kwargs):
        return self._set(**kwargs)

    def _ensemble_params(self):
        g = self.getOrDefault
        return dict(numTrees=g("numTrees"), featureSubsetStrategy=g("featureSubsetStrategy"), subsamplingRate=g("subsamplingRate"),
                    bootstrap=g("bootstrap"))


# ------------------------------------------------------------------------------------------------------------------
# Binary classifiers: pyspark.ml.classification.DecisionTreeClassifier / RandomForestClassifier as the labs'
# `Solutions/Labs/ML 07L - Hyperparameter Tuning Lab.py:80-82,105-108,140-155` use them (label priceClass in {0, 1}).
# Binary labels ride on the regression statistics (sum w, sum w*y); three or more classes take the per-class kernels of
# k8_multiclass.cu (numClasses = max label + 1 as Classifier.getNumClasses infers it, at most 32 here).
# ------------------------------------------------------------------------------------------------------------------
class _ClassifierParams:
    rawPredictionCol = Param(Params, "rawPredictionCol", "raw prediction (a.k.a. confidence) column name.", TypeConverters.toString)
    probabilityCol = Param(Params, "probabilityCol", "Column name for predicted class conditional probabilities. Note: Not all models output "
                           "well-calibrated probability estimates! These probabilities should be treated as confidences, not precise "
                           "probabilities.", TypeConverters.toString)
    thresholds = Param(Params, "thresholds", "Thresholds in multi-class classification to adjust the probability of predicting each class. "
                       "Array must have length equal to the number of classes, with values > 0, excepting that at most one value may be 0. "
                       "The class with largest value p/t is predicted, where p is the original probability of that class and t is the "
                       "class's threshold.", TypeConverters.toListFloat)

    def _base_defaults(self):
        super()._base_defaults()
        self._setDefault(impurity="gini", rawPredictionCol="rawPrediction", probabilityCol="probability")


class _ClassifierModelBase(_ClassifierParams, _ForestModelBase):
    @property
    def numClasses(self):
        return self._forest.num_classes

    def _raw(self, X):
        import torch
        return self._forest.predict_raw(torch.as_tensor(X, device=self._device()), single_tree=not self._ensemble).cpu().numpy()

    @staticmethod
    def _normalize(raw):      # ProbabilisticClassificationModel.normalizeToProbabilitiesInPlace
        tot = raw.sum(axis=1, keepdims=True)
        return raw / np.where(tot != 0, tot, 1.0)

    def _predict_from_probability(self, prob):
        """probability2prediction: argmax p (first maximum), or argmax p / t when thresholds are set."""
        if self.isDefined("thresholds") and self.getOrDefault("thresholds") is not None:
            t = np.asarray(self.getOrDefault("thresholds"), np.float64)
            if t.shape != (prob.shape[1],) or (t < 0).any() or (t == 0).sum() > 1:
                raise ValueError(f"thresholds must have length numClasses = {prob.shape[1]}, values >= 0, at most one 0; got {t.tolist()}")
            with np.errstate(divide="ignore", invalid="ignore"):
                scaled = prob / t
            scaled = np.where(np.isnan(scaled), -np.inf, scaled)
            return np.argmax(scaled, axis=1).astype(np.float64)
        return np.argmax(prob, axis=1).astype(np.float64)

    def _one(self, value):
        return np.asarray(value.toArray() if hasattr(value, "toArray") else value, np.float64).reshape(1, -1)

    def predictRaw(self, value):
        return DenseVector(self._raw(self._one(value))[0])

    def predictProbability(self, value):
        return DenseVector(self._normalize(self._raw(self._one(value)))[0])

    def predict(self, value):
        return float(self._predict_from_probability(self._normalize(self._raw(self._one(value))))[0])

    def _score_columns(self, X):
        raw = self._raw(X) if X.shape[0] else np.zeros((0, self.numClasses))
        prob = self._normalize(raw)
        cols = {}
        if self.getOrDefault("rawPredictionCol"):
            cols[self.getOrDefault("rawPredictionCol")] = raw
        if self.getOrDefault("probabilityCol"):
            cols[self.getOrDefault("probabilityCol")] = prob
        if self.getOrDefault("predictionCol"):
            cols[self.getOrDefault("predictionCol")] = self._predict_from_probability(prob)
        cols.update(self._leaf_columns(X))
        return cols


class B200DecisionTreeClassificationModel(_ClassifierModelBase):
    _ensemble = False

    @property
    def depth(self):
        return self._forest.depth(0)

    @property
    def numNodes(self):
        return self._forest.num_nodes(0)


class B200RandomForestClassificationModel(_ClassifierModelBase):
    _ensemble = True
    _tree_model_cls = B200DecisionTreeClassificationModel
    numTrees = B200RandomForestRegressionModel.numTrees
    featureSubsetStrategy = B200RandomForestRegressionModel.featureSubsetStrategy
    subsamplingRate = B200RandomForestRegressionModel.subsamplingRate
    bootstrap = B200RandomForestRegressionModel.bootstrap

    def __init__(self, forest=None):
        super().__init__(forest)
        self._setDefault(numTrees=20, featureSubsetStrategy="auto", subsamplingRate=1.0, bootstrap=True)

    getNumTrees = B200RandomForestRegressionModel.getNumTrees
    treeWeights = B200RandomForestRegressionModel.treeWeights
    totalNumNodes = B200RandomForestRegressionModel.totalNumNodes
    trees = B200RandomForestRegressionModel.trees


class B200DecisionTreeClassifier(_ClassifierParams, _TreeEstimatorBase):
    """pyspark.ml.classification.DecisionTreeClassifier (binary labels), trained on the GPU."""
    _model_cls = B200DecisionTreeClassificationModel

    def __init__(self, **kwargs):
        super().__init__()
        self._base_defaults()
        self.setParams(**kwargs)

    def setParams(self, **kwargs):
        return self._set(**kwargs)

    def _ensemble_params(self):
        return dict(numTrees=1, featureSubsetStrategy="all", subsamplingRate=1.0, bootstrap=False)


class B200RandomForestClassifier(_ClassifierParams, _TreeEstimatorBase):
    """pyspark.ml.classification.RandomForestClassifier (binary labels), trained on the GPU (`Solutions/Labs/ML 07L:80-82`)."""
    _model_cls = B200RandomForestClassificationModel
    numTrees = B200RandomForestRegressionModel.numTrees
    featureSubsetStrategy = B200RandomForestRegressionModel.featureSubsetStrategy
    subsamplingRate = B200RandomForestRegressionModel.subsamplingRate
    bootstrap = B200RandomForestRegressionModel.bootstrap

    def __init__(self, **kwargs):
        super().__init__()
        self._base_defaults()
        self._setDefault(numTrees=20, featureSubsetStrategy="auto", subsamplingRate=1.0, bootstrap=True)
        self.setParams(**kwargs)

    def setParams(self, **kwargs):
        return self._set(**kwargs)

    _ensemble_params = B200RandomForestRegressor._ensemble_params


_CLS = ["rawPredictionCol", "probabilityCol", "thresholds"]
for _c in (B200DecisionTreeClassifier, B200DecisionTreeClassificationModel):
    _accessors(_c, _COMMON + _CLS)
for _c in (B200RandomForestClassifier, B200RandomForestClassificationModel):
    _accessors(_c, _COMMON + _CLS + ["numTrees", "featureSubsetStrategy", "subsamplingRate", "bootstrap"])
for _c in (B200DecisionTreeRegressor, B200DecisionTreeRegressionModel):
    _accessors(_c, _COMMON + ["varianceCol"])
for _c in (B200RandomForestRegressor, B200RandomForestRegressionModel):
    _accessors(_c, _COMMON + ["numTrees", "featureSubsetStrategy", "subsamplingRate", "bootstrap"])
