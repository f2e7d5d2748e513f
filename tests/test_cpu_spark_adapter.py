"""Host-side contract of the pyspark.ml drop-in classes, exercised without a GPU.  pyspark is not
installable in this image, so the classes run on the in-repo stub of the pyspark.ml.param surface."""
import numpy as np
import pytest


@pytest.fixture(scope="module")
def sp(b2rf):
    from b2rf import spark
    return spark


def test_defaults_match_pyspark_3_1(sp):
    dt = sp.B200DecisionTreeRegressor()
    want = dict(maxDepth=5, maxBins=32, minInstancesPerNode=1, minWeightFractionPerNode=0.0, minInfoGain=0.0, maxMemoryInMB=256,
                cacheNodeIds=False, checkpointInterval=10, impurity="variance", featuresCol="features", labelCol="label",
                predictionCol="prediction", leafCol="")
    for k, v in want.items():
        assert dt.getOrDefault(k) == v, k
    rf = sp.B200RandomForestRegressor()
    for k, v in dict(numTrees=20, featureSubsetStrategy="auto", subsamplingRate=1.0, bootstrap=True, **want).items():
        assert rf.getOrDefault(k) == v, k
    assert isinstance(rf.getSeed(), int)


def test_ml06_setter_and_kwargs(sp):
    dt = sp.B200DecisionTreeRegressor(labelCol="price")          # `ML 06 - Decision Trees.py:75`
    assert dt.getLabelCol() == "price" and dt.getMaxBins() == 32
    assert dt.setMaxBins(40) is dt and dt.getMaxBins() == 40     # `ML 06:118`
    rf = sp.B200RandomForestRegressor(labelCol="price", maxBins=40, seed=42)   # `ML 08 - Hyperopt.py:71`
    assert (rf.getMaxBins(), rf.getSeed()) == (40, 42)
    assert "maxDepth: Maximum depth of the tree." in rf.explainParams()        # `ML 07:56`
    assert "current: 40" in rf.explainParam("maxBins")


def test_copy_keeps_uid_and_converts_integral_floats(sp):
    rf = sp.B200RandomForestRegressor(labelCol="price", maxBins=40, seed=42)
    c = rf.copy({rf.maxDepth: 4.0, rf.numTrees: 37.0})           # `ML 08:97,120-121`: hp.quniform yields floats
    assert c.uid == rf.uid and c is not rf
    assert c.getMaxDepth() == 4 and isinstance(c.getMaxDepth(), int) and c.getNumTrees() == 37
    assert rf.getMaxDepth() == 5 and rf.getNumTrees() == 20      # original untouched
    with pytest.raises(TypeError):
        rf.copy({rf.maxDepth: 4.5})
    tp = c._tree_params()
    assert (tp.maxDepth, tp.numTrees, tp.bootstrap, tp.featureSubsetStrategy) == (4, 37, True, "auto")
    dt = sp.B200DecisionTreeRegressor(maxDepth=3)
    tp = dt._tree_params()
    assert (tp.numTrees, tp.bootstrap, tp.featureSubsetStrategy, tp.subsamplingRate) == (1, False, "all", 1.0)


def test_param_grid_style_maps(sp):
    rf = sp.B200RandomForestRegressor()
    grid = [{rf.maxDepth: d, rf.numTrees: t} for d in (2, 5) for t in (5, 10)]      # `ML 07:74-77`
    got = [(rf.copy(pm).getMaxDepth(), rf.copy(pm).getNumTrees()) for pm in grid]
    assert got == [(2, 5), (2, 10), (5, 5), (5, 10)]
    other = sp.B200RandomForestRegressor()
    assert other.copy({rf.maxDepth: 9}).getMaxDepth() == 5       # a Param owned by another uid is not applied


def test_categorical_arities_from_ml_attr(sp, b2rf):
    from b2rf._pyspark_stub import LocalFrame, ml_attr_metadata
    arity = [2, 0, 36, 0]
    lf = LocalFrame({"features": np.zeros((3, 4)), "price": np.zeros(3)}, {"features": ml_attr_metadata(arity)})
    assert sp.categorical_arities(lf, "features", 4).tolist() == arity
    meta = {"ml_attr": {"attrs": {"binary": [{"idx": 1, "name": "b"}], "nominal": [{"idx": 0, "name": "n", "num_vals": 7}]}, "num_attrs": 3}}
    lf = LocalFrame({"features": np.zeros((3, 3))}, {"features": meta})
    assert sp.categorical_arities(lf, "features", 3).tolist() == [7, 2, 0]
    assert sp.categorical_arities(LocalFrame({"features": np.zeros((3, 3))}), "features", 3).tolist() == [0, 0, 0]


def test_model_copy_carries_forest_and_api(sp, b2rf, oracle):
    X = np.array([[1.0], [0.0]]); y = np.array([1.0, 0.0])
    o = oracle.fit(X, y, [0], maxDepth=2)
    fo = b2rf.Forest(1, o.tree_off, o.feature, o.threshold, o.left, o.right, o.prediction, o.impurity, o.gain, o.stats, o.ncat, o.cat_off,
                     o.cat_len, o.cats)
    est = sp.B200DecisionTreeRegressor(maxDepth=2)
    m = est._make_model(fo)
    assert m.uid == est.uid and m.getMaxDepth() == 2
    assert (m.depth, m.numNodes, m.numFeatures) == (1, 3, 1)
    assert list(m.featureImportances) == [1.0] and "SparseVector(1, {0: 1.0})" == repr(m.featureImportances)
    assert m.toDebugString.startswith(f"DecisionTreeRegressionModel: uid={est.uid}, depth=1, numNodes=3, numFeatures=1\n  If (feature 0 <= 0.5)")
    m2 = m.copy({m.predictionCol: "p2"})
    assert m2._forest is m._forest and m2.getPredictionCol() == "p2" and m.getPredictionCol() == "prediction"


def test_persistence_round_trip_mllib_layout(sp, b2rf, oracle, tmp_path):
    """model.write().overwrite().save(path) (`Labs/ML 07L - Hyperparameter Tuning Lab.py:209`) writes MLlib 3.1.2's directory
    layout; load() restores an identical forest, params and uid."""
    import json
    import pyarrow.parquet as pq
    X, y, arity = b2rf.synth.airbnb(1500)
    for T in (1, 4):
        o = oracle.fit(X.numpy(), y.numpy(), arity, maxDepth=5, maxBins=40, numTrees=T, bootstrap=T > 1, seed=3)
        fo = b2rf.Forest(33, o.tree_off, o.feature, o.threshold, o.left, o.right, o.prediction, o.impurity, o.gain, o.stats, o.ncat, o.cat_off,
                         o.cat_len, o.cats)
        est = (sp.B200DecisionTreeRegressor if T == 1 else sp.B200RandomForestRegressor)(maxDepth=5, maxBins=40, labelCol="price")
        m = est._make_model(fo)
        path = str(tmp_path / f"model{T}")
        m.write().overwrite().save(path)
        with pytest.raises(IOError, match="already exists"):
            m.write().save(path)
        meta = json.loads(open(path + "/metadata/part-00000").readline())
        assert meta["class"].endswith("DecisionTreeRegressionModel" if T == 1 else "RandomForestRegressionModel")
        assert meta["uid"] == m.uid and meta["numFeatures"] == 33 and meta["paramMap"]["maxBins"] == 40 and meta["sparkVersion"] == "3.1.2"
        tbl = pq.read_table(path + "/data/part-00000.parquet")
        names = tbl.schema.names
        assert names == (["id", "prediction", "impurity", "impurityStats", "rawCount", "gain", "leftChild", "rightChild", "split"] if T == 1
                         else ["treeID", "nodeData"])
        assert tbl.num_rows == len(o.feature)
        if T > 1:
            assert meta["numTrees"] == T and pq.read_table(path + "/treesMetadata/part-00000.parquet").num_rows == T
        m2 = type(m).load(path)
        f2 = m2._forest
        assert m2.uid == m.uid and m2.getMaxBins() == 40 and m2.getLabelCol() == "price"
        for a in ("tree_off", "feature", "threshold", "left", "right", "prediction", "ncat", "cat_len"):
            assert np.array_equal(getattr(f2, a), getattr(fo, a)), a
        assert np.array_equal(f2.stats[:, :3], fo.stats[:, :3]) and np.array_equal(f2.gain, fo.gain)
        assert [sorted(f2.cats[o_:o_ + n]) for o_, n in zip(f2.cat_off, f2.cat_len)] == [sorted(fo.cats[o_:o_ + n]) for o_, n in zip(fo.cat_off, fo.cat_len)]
        assert m2.toDebugString == m.toDebugString


def _classifier_forest(b2rf, o, F):
    """Product-side Forest from the oracle's classification arrays: the trainer's statistics are (sum w, sum w*y, sum w*y*y, raw)
    with y in {0, 1}; the oracle keeps (count0, count1, 0, raw)."""
    st = np.stack([o.stats[:, 0] + o.stats[:, 1], o.stats[:, 1], o.stats[:, 1], o.stats[:, 3]], axis=1)
    return b2rf.Forest(F, o.tree_off, o.feature, o.threshold, o.left, o.right, o.prediction, o.impurity, o.gain, st, o.ncat, o.cat_off,
                       o.cat_len, o.cats, impurity_kind=o.impurity_kind)


def test_classifier_defaults_and_params(sp):
    """pyspark 3.1.2 DecisionTreeClassifier / RandomForestClassifier defaults (`Solutions/Labs/ML 07L:80-82`)."""
    dt = sp.B200DecisionTreeClassifier()
    assert (dt.getImpurity(), dt.getMaxDepth(), dt.getMaxBins(), dt.getRawPredictionCol(), dt.getProbabilityCol(), dt.getPredictionCol()) == \
        ("gini", 5, 32, "rawPrediction", "probability", "prediction")
    rf = sp.B200RandomForestClassifier(labelCol="priceClass", maxBins=40, seed=42)
    assert (rf.getNumTrees(), rf.getFeatureSubsetStrategy(), rf.getBootstrap(), rf.getImpurity()) == (20, "auto", True, "gini")
    p = rf.copy({rf.maxDepth: 4.0, rf.numTrees: 10, rf.impurity: "Entropy"})._tree_params()
    assert (p.maxDepth, p.numTrees, p.impurity, p.bootstrap) == (4, 10, "entropy", True)
    assert p.features_per_node(33) == 6                               # classification forests: auto -> sqrt
    assert "thresholds:" in rf.explainParams() and "rawPredictionCol:" in rf.explainParams()
    rf.setThresholds([0.3, 0.7])
    assert rf.getThresholds() == [0.3, 0.7]


def test_classifier_model_api_and_persistence(sp, b2rf, oracle, tmp_path):
    import json
    X, y, arity = b2rf.synth.airbnb(1500)
    X, y = X.numpy(), (y.numpy() > np.median(y.numpy())).astype(np.float64)
    for T in (1, 4):
        o = oracle.fit(X, y, arity, maxDepth=4, maxBins=40, numTrees=T, bootstrap=T > 1, seed=3, impurity="gini")
        fo = _classifier_forest(b2rf, o, 33)
        est = (sp.B200DecisionTreeClassifier if T == 1 else sp.B200RandomForestClassifier)(maxDepth=4, maxBins=40, labelCol="priceClass")
        m = est._make_model(fo)
        assert m.numClasses == 2 and m.numFeatures == 33 and abs(sum(m.featureImportances) - 1.0) < 1e-12
        head = "DecisionTreeClassificationModel: uid=" if T == 1 else "RandomForestClassificationModel: uid="
        assert m.toDebugString.startswith(head) and "numClasses=2, numFeatures=33" in m.toDebugString.splitlines()[0]
        if T > 1:
            assert len(m.trees) == T and m.trees[0].numClasses == 2 and m.treeWeights == [1.0] * T
        path = str(tmp_path / f"cls{T}")
        m.write().overwrite().save(path)
        meta = json.loads(open(path + "/metadata/part-00000").readline())
        assert meta["class"] == "org.apache.spark.ml.classification." + ("DecisionTreeClassificationModel" if T == 1 else "RandomForestClassificationModel")
        assert meta["numClasses"] == 2 and meta["paramMap"]["maxBins"] == 40
        m2 = type(m).load(path)
        f2 = m2._forest
        assert f2.is_classification and f2.impurity_kind == "gini" and m2.getLabelCol() == "priceClass"
        assert np.array_equal(f2.class_counts, fo.class_counts) and np.array_equal(f2.class_counts, o.stats[:, :2])
        for a in ("tree_off", "feature", "threshold", "left", "right", "prediction"):
            assert np.array_equal(getattr(f2, a), getattr(fo, a)), a
        assert m2.toDebugString == m.toDebugString


def test_classifier_thresholds_rule(sp, b2rf, oracle):
    """probability2prediction: argmax of p / t; without thresholds the first maximum of p."""
    o = oracle.fit(np.array([[1.0], [0.0]]), np.array([1.0, 0.0]), [0], maxDepth=2, impurity="gini")
    m = sp.B200DecisionTreeClassifier()._make_model(_classifier_forest(b2rf, o, 1))
    prob = np.array([[0.5, 0.5], [0.6, 0.4], [0.2, 0.8]])
    assert m._predict_from_probability(prob).tolist() == [0.0, 0.0, 1.0]
    m.setThresholds([0.8, 0.2])
    assert m._predict_from_probability(prob).tolist() == [1.0, 1.0, 1.0]
    m.setThresholds([0.0, 1.0])
    assert m._predict_from_probability(np.array([[0.1, 0.9], [0.0, 1.0]])).tolist() == [0.0, 1.0]   # 0.1 / 0 = inf wins; 0 / 0 is skipped


def test_persistence_schema_matches_committed_fixture(sp, b2rf, tmp_path):
    """The written model directory (Parquet schemas of data/ and treesMetadata/, metadata keys, class names) against the committed
    fixture tests/golden/persistence_schema.json, which restates MLlib 3.1.2's NodeData / SplitData / EnsembleNodeData layout.
    Spark is unavailable on this box, so this pins the format that only this package has read back so far."""
    import importlib.util
    import json
    import os
    here = os.path.dirname(os.path.abspath(__file__))
    spec = importlib.util.spec_from_file_location("make_persistence_schema", os.path.join(here, "golden", "make_persistence_schema.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    got = mod.describe(str(tmp_path))
    want = json.load(open(os.path.join(here, "golden", "persistence_schema.json")))
    assert got == want
