"""GPU parity of the binary tree classifiers (reference call sites `Solutions/Labs/ML 07L:80-82,105-108,140-155`)
against the oracle's independent class-count restatement, the pyspark doctest values and scikit-learn's exact search."""
import numpy as np
import pytest
import torch

from helpers import assert_classifier_parity

pytestmark = pytest.mark.gpu


def _fit(b2rf, cuda, X, y, arity, part_offsets=None, **kw):
    p = b2rf.TreeParams(**kw)
    return b2rf.trainer.fit_tensors(torch.as_tensor(X, device=cuda), torch.as_tensor(y, device=cuda), arity, p, part_offsets=part_offsets, keep_trace=True)


def test_pyspark_doctest_decision_tree_classifier(b2rf, cuda):
    """pyspark.ml.classification.DecisionTreeClassifier doctest: two rows, maxDepth=2."""
    X = np.array([[1.0], [0.0]]); y = np.array([1.0, 0.0])
    f = _fit(b2rf, cuda, X, y, [0], maxDepth=2, impurity="gini")
    assert f.depth(0) == 1 and f.num_nodes(0) == 3 and f.num_classes == 2
    assert f.feature_importances().tolist() == [1.0]
    T = torch.tensor([[-1.0], [1.0]], dtype=torch.float64, device=cuda)
    assert f.predict(T).tolist() == [0.0, 1.0]
    assert f.predict_raw(T).tolist() == [[1.0, 0.0], [0.0, 1.0]]
    assert f.predict_proba(T).tolist() == [[1.0, 0.0], [0.0, 1.0]]
    assert f.to_debug_string().startswith("DecisionTreeClassificationModel: uid=, depth=1, numNodes=3, numClasses=2, numFeatures=1\n")


def test_pyspark_doctest_random_forest_classifier(b2rf, cuda):
    """pyspark.ml.classification.RandomForestClassifier doctest: numTrees=3, maxDepth=2, seed=42 on the same two rows."""
    X = np.array([[1.0], [0.0]]); y = np.array([1.0, 0.0])
    f = _fit(b2rf, cuda, X, y, [0], part_offsets=[0, 1, 2], maxDepth=2, numTrees=3, bootstrap=True, seed=42, impurity="gini")
    T = torch.tensor([[-1.0], [1.0]], dtype=torch.float64, device=cuda)
    assert f.num_trees == 3
    assert f.predict(T).tolist() == [0.0, 1.0]
    raw = f.predict_raw(T).cpu().numpy()
    assert int(np.argmax(raw[0])) == 0 and int(np.argmax(raw[1])) == 1
    assert f.feature_importances().tolist() == [1.0]


@pytest.mark.parametrize("impurity,root", [("gini", 0.48), ("entropy", 0.9709505944546686)])
def test_categorical_stump_600_400(b2rf, cuda, impurity, root):
    """DecisionTreeSuite 'Binary classification stump with ordered categorical features' data: 600 x (label 1, features (0, 1)),
    400 x (label 0, features (1, 0)), arity 3 each."""
    X = np.array([[0.0, 1.0]] * 600 + [[1.0, 0.0]] * 400); y = np.array([1.0] * 600 + [0.0] * 400)
    f = _fit(b2rf, cuda, X, y, [3, 3], maxDepth=2, maxBins=100, impurity=impurity)
    assert f.feature.tolist() == [0, -1, -1] and f.ncat[0] == 3
    assert f.cats[f.cat_off[0]:f.cat_off[0] + f.cat_len[0]].tolist() == [1.0]        # ordered by the class-1 count: category 1 has none
    assert f.impurity[0] == pytest.approx(root, rel=1e-12) and f.gain[0] == pytest.approx(root, rel=1e-12)
    assert f.class_counts.tolist() == [[400.0, 600.0], [400.0, 0.0], [0.0, 600.0]]
    assert f.prediction.tolist() == [1.0, 0.0, 1.0]


def _airbnb_binary(b2rf, n, seed=0):
    X, y, arity = b2rf.synth.airbnb(n, device="cpu")
    X, y = np.asarray(X), np.asarray(y)
    return X, (y > np.median(y)).astype(np.float64), list(arity)      # ML 07L: priceClass = price above the median


@pytest.mark.parametrize("impurity", ["gini", "entropy"])
def test_decision_tree_classifier_parity(b2rf, oracle, cuda, impurity):
    X, y, arity = _airbnb_binary(b2rf, 5716)
    kw = dict(maxDepth=6, maxBins=40, impurity=impurity)
    g = _fit(b2rf, cuda, X, y, arity, **kw)
    o = oracle.fit(X, y, arity, **kw)
    assert_classifier_parity(g, o, exact=(impurity == "gini"))
    Xd = torch.as_tensor(X, device=cuda)
    assert np.array_equal(g.predict(Xd).cpu().numpy(), o.predict(X))
    assert np.array_equal(g.predict_raw(Xd).cpu().numpy(), o.predict_raw(X))
    assert float((g.predict(Xd).cpu().numpy() == y).mean()) > 0.7


@pytest.mark.parametrize("impurity,strategy,mem", [("gini", "auto", 256), ("entropy", "log2", 256), ("gini", "0.5", 256), ("gini", "all", 1)])
def test_random_forest_classifier_parity(b2rf, oracle, cuda, impurity, strategy, mem):
    """mem = 1: node groups are cut by maxMemoryInMB, where classification counts numClasses (not 3) statistics per bin."""
    X, y, arity = _airbnb_binary(b2rf, 5716)
    po = list(range(0, 5716, 1429)) + [5716]
    kw = dict(maxDepth=7 if mem == 1 else 5, maxBins=40, numTrees=20, bootstrap=True, seed=42, featureSubsetStrategy=strategy, impurity=impurity,
              minInstancesPerNode=2, maxMemoryInMB=mem)
    g = _fit(b2rf, cuda, X, y, arity, part_offsets=po, **kw)
    o = oracle.fit(X, y, arity, part_offsets=np.array(po), **kw)
    assert_classifier_parity(g, o, exact=(impurity == "gini"))
    if mem == 1:
        assert g.counters["groups"] == len(o.group_sizes) > 8
    Xd = torch.as_tensor(X, device=cuda)
    np.testing.assert_allclose(g.predict_raw(Xd).cpu().numpy(), o.predict_raw(X), rtol=1e-15, atol=0)
    assert np.array_equal(g.predict(Xd).cpu().numpy(), o.predict(X))
    p = g.predict_proba(Xd).cpu().numpy()
    np.testing.assert_allclose(p.sum(1), 1.0, rtol=1e-12)


def test_classifier_rejects_other_labels(b2rf, cuda):
    X = np.random.default_rng(0).normal(size=(50, 3)); y = np.arange(50) % 3.0
    with pytest.raises(b2rf.IllegalArgumentException, match="numClasses == 2"):
        _fit(b2rf, cuda, X, y, [0, 0, 0], impurity="gini")


def test_classifier_matches_sklearn_exact_search(b2rf, cuda):
    sk = pytest.importorskip("sklearn.tree")
    rng = np.random.default_rng(5); n = 20000
    X = rng.integers(0, 12, size=(n, 6)).astype(np.float64)
    logit = 1.2 * (X[:, 0] > 4) + 0.8 * (X[:, 1] > 7) - 1.0 * (X[:, 2] <= 2) * (X[:, 3] > 5) + 0.05 * X[:, 4] - 0.6
    y = (rng.random(n) < 1 / (1 + np.exp(-logit))).astype(np.float64)
    for crit in ["gini", "entropy"]:
        g = _fit(b2rf, cuda, X, y, [0] * 6, maxDepth=4, maxBins=32, impurity=crit)
        m = sk.DecisionTreeClassifier(max_depth=4, criterion=crit, random_state=0).fit(X, y)
        assert np.array_equal(g.predict(torch.as_tensor(X, device=cuda)).cpu().numpy(), m.predict(X))


def test_spark_adapter_classifier_flow(b2rf, oracle, cuda, tmp_path):
    """`Solutions/Labs/ML 07L`: RandomForestClassifier(labelCol="priceClass", maxBins=40, seed=42) inside a grid over maxDepth / numTrees,
    evaluated on the transformed test frame; then the best model is saved and loaded back."""
    from b2rf import spark as sp
    from b2rf._pyspark_stub import LocalFrame, ml_attr_metadata
    X, y, arity = _airbnb_binary(b2rf, 7146)
    ntr = 5716
    meta = {"features": ml_attr_metadata(arity)}
    train = LocalFrame({"features": X[:ntr], "priceClass": y[:ntr]}, meta, num_partitions=4)
    test = LocalFrame({"features": X[ntr:], "priceClass": y[ntr:]}, meta)
    po = np.array([0, 1429, 2858, 4287, 5716])
    rf = sp.B200RandomForestClassifier(labelCol="priceClass", maxBins=40, seed=42)
    grid = [{rf.maxDepth: d, rf.numTrees: t} for d in (2, 5) for t in (5, 10)]
    got = dict(rf.fitMultiple(train, grid))
    assert sorted(got) == [0, 1, 2, 3]
    for i, pm in enumerate(grid):
        m = got[i]
        o = oracle.fit(X[:ntr], y[:ntr], arity, maxDepth=pm[rf.maxDepth], maxBins=40, numTrees=pm[rf.numTrees], bootstrap=True, seed=42,
                       part_offsets=po, impurity="gini")
        assert m.getNumTrees == pm[rf.numTrees] and m.totalNumNodes == len(o.feature) and m.numClasses == 2
        out = m.transform(test)
        np.testing.assert_allclose(out.col("rawPrediction"), o.predict_raw(X[ntr:]), rtol=1e-15, atol=0)
        assert np.array_equal(out.col("prediction"), o.predict(X[ntr:]))
        np.testing.assert_allclose(out.col("probability").sum(1), 1.0, rtol=1e-12)
    best = got[3]
    acc = float((best.transform(test).col("prediction") == y[ntr:]).mean())
    assert acc > 0.6                                                       # held-out rows of the synthetic listing data
    v = X[ntr]
    assert best.predict(v) == best.transform(test).col("prediction")[0]
    assert list(best.predictProbability(v)) == list(best.transform(test).col("probability")[0])
    path = str(tmp_path / "rfc")
    best.write().overwrite().save(path)
    again = sp.B200RandomForestClassificationModel.load(path)
    assert np.array_equal(again.transform(test).col("rawPrediction"), best.transform(test).col("rawPrediction"))
    dt = sp.B200DecisionTreeClassifier(labelCol="priceClass", maxBins=40, impurity="entropy").fit(train)
    od = oracle.fit(X[:ntr], y[:ntr], arity, maxDepth=5, maxBins=40, part_offsets=po, impurity="entropy")
    assert dt.numNodes == len(od.feature) and np.array_equal(dt.transform(test).col("rawPrediction"), od.predict_raw(X[ntr:]))
    with pytest.raises(b2rf.IllegalArgumentException, match="labels must be the integers"):
        sp.B200DecisionTreeClassifier(labelCol="priceClass").fit(LocalFrame({"features": X[:30], "priceClass": np.arange(30) * 0.5}))
