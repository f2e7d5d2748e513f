"""GPU parity of multiclass tree classifiers (numClasses > 2: per-class count histograms, impurity-ordered and UNORDERED categorical
splits) against oracle/multiclass.py, the numpy restatement that is itself pinned to the C++ oracle for
numClasses == 2 (tests/test_cpu_oracle.py)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _data(b2rf, n, C):
    X, y, arity = b2rf.synth.airbnb(n, device="cpu")
    X, y = np.asarray(X), np.asarray(y)
    qs = np.quantile(y, np.linspace(0, 1, C + 1)[1:-1])
    return X, np.digitize(y, qs).astype(np.float64), list(arity)


def _parity(g, m):
    assert np.array_equal(g.tree_off, m.tree_off), f"tree sizes differ: {np.diff(g.tree_off)[:8]} vs {np.diff(m.tree_off)[:8]}"
    assert np.array_equal(g.feature, m.feature)
    assert np.array_equal(g.left, m.left) and np.array_equal(g.right, m.right)
    assert np.array_equal(g.threshold, m.threshold)
    assert np.array_equal(g.ncat, m.ncat)
    for i in np.nonzero(g.ncat > 0)[0]:
        assert sorted(g.cats[g.cat_off[i]:g.cat_off[i] + g.cat_len[i]]) == sorted(m.cats[m.cat_off[i]:m.cat_off[i] + m.cat_len[i]]), i
    assert np.array_equal(g.class_counts, m.class_counts)
    assert np.array_equal(g.prediction, m.prediction)
    assert np.array_equal(g.trace_i, m.trace_i)


@pytest.mark.parametrize("C,impurity,kw", [
    (3, "gini", dict(maxDepth=5, maxBins=40)),
    (5, "entropy", dict(maxDepth=4, maxBins=40)),
    (4, "gini", dict(maxDepth=5, maxBins=40, numTrees=8, bootstrap=True, seed=42, minInstancesPerNode=2)),
    (3, "gini", dict(maxDepth=6, maxBins=64, numTrees=5, bootstrap=True, seed=7, featureSubsetStrategy="all", maxMemoryInMB=1)),
    (7, "gini", dict(maxDepth=4, maxBins=40, numTrees=4, bootstrap=True, seed=3, featureSubsetStrategy="log2")),
])
def test_multiclass_parity(b2rf, cuda, C, impurity, kw):
    from oracle import multiclass as M
    X, y, arity = _data(b2rf, 3000, C)
    po = np.array([0, 750, 1500, 2250, 3000])
    m = M.fit(X, y, arity, C, part_offsets=po, impurity=impurity, **kw)
    assert len(m.unordered) >= 4                                  # the small-arity categorical features take the subset-split path
    p = b2rf.TreeParams(impurity=impurity, numClasses=C, **kw)
    g = b2rf.trainer.fit_tensors(torch.as_tensor(X, device=cuda), torch.as_tensor(y, device=cuda), arity, p, part_offsets=po, keep_trace=True)
    assert g.num_classes == C
    _parity(g, m)
    exact = impurity == "gini"
    mi = m.impurity >= 0
    if exact:
        assert np.array_equal(g.impurity[mi], m.impurity[mi]) and np.array_equal(g.gain, m.gain)
    else:
        np.testing.assert_allclose(g.impurity[mi], m.impurity[mi], rtol=1e-12, atol=1e-15)
        np.testing.assert_allclose(g.gain, m.gain, rtol=1e-9, atol=1e-13)
    assert g.counters["groups"] == len(m.group_sizes)
    Xt = X[:300]
    raw = g.predict_raw(torch.as_tensor(Xt, device=cuda)).cpu().numpy()
    np.testing.assert_allclose(raw, m.predict_raw(Xt), rtol=1e-15, atol=0)
    assert np.array_equal(g.predict(torch.as_tensor(Xt, device=cuda)).cpu().numpy(), m.predict(Xt))
    assert (g.ncat > 0).any()


@pytest.mark.parametrize("F,C,kw,staged", [
    (50, 4, dict(maxDepth=5, maxBins=40, numTrees=6, bootstrap=True, seed=11), True),                                      # 64-byte code rows
    (50, 3, dict(maxDepth=4, maxBins=64, numTrees=3, bootstrap=True, seed=5, featureSubsetStrategy="all"), True),          # two feature groups
    (100, 5, dict(maxDepth=4, maxBins=36, numTrees=4, bootstrap=True, seed=2, featureSubsetStrategy="onethird"), True),    # 128-byte code rows
    (70, 3, dict(maxDepth=4, maxBins=48, numTrees=3, bootstrap=True, seed=9), False),                                      # 80-byte rows: gather kernel
])
def test_multiclass_staged_histograms(b2rf, cuda, monkeypatch, F, C, kw, staged):
    """The class-count kernel that stages code rows through shared memory (64- / 128-byte rows, labels in the row lists) and the
    gather kernel behind it produce the oracle's forest; `k4_launch_counts` proves which one ran."""
    from oracle import multiclass as M
    if not staged:
        monkeypatch.setattr(b2rf.trainer, "CODE_ROW_ALIGN", 16)    # code rows padded to 16 instead of 64 / 128
    n = 6000                                                       # the root of every tree spans two 4096-row chunks
    X, y, arity = _data(b2rf, n, C)
    rng = np.random.default_rng(F)
    extra = rng.normal(size=(n, F - X.shape[1]))
    extra[:, ::3] = np.round(extra[:, ::3] * 2)                    # some columns with few distinct values
    X = np.concatenate([X, extra], axis=1); arity = arity + [0] * extra.shape[1]
    y = np.where(extra[:, 1] + extra[:, 2] > 0.8, (y + 1) % C, y)  # the added columns carry signal
    po = np.array([0, 1500, 3000, 4500, 6000])
    m = M.fit(X, y, arity, C, part_offsets=po, impurity="gini", **kw)
    before = b2rf.trainer.k4_launch_counts()
    g = b2rf.trainer.fit_tensors(torch.as_tensor(X, device=cuda), torch.as_tensor(y, device=cuda), arity,
                                 b2rf.TreeParams(impurity="gini", numClasses=C, **kw), part_offsets=po, keep_trace=True)
    after = b2rf.trainer.k4_launch_counts()
    assert after["class_counts"] > before["class_counts"]
    assert (after["class_staged"] > before["class_staged"]) == staged
    if staged:
        assert after["class_staged"] - before["class_staged"] == after["class_counts"] - before["class_counts"]
    _parity(g, m)
    mi = m.impurity >= 0
    assert np.array_equal(g.impurity[mi], m.impurity[mi]) and np.array_equal(g.gain, m.gain)
    assert (g.feature >= 33).any()


def test_multiclass_label_check(b2rf, cuda):
    X = np.random.default_rng(0).normal(size=(60, 3)); y = np.arange(60) % 4.0
    with pytest.raises(b2rf.IllegalArgumentException, match="numClasses == 3"):
        b2rf.trainer.fit_tensors(torch.as_tensor(X, device=cuda), torch.as_tensor(y, device=cuda), [0, 0, 0], b2rf.TreeParams(impurity="gini", numClasses=3))


def test_multiclass_adapter_and_persistence(b2rf, cuda, tmp_path):
    """pyspark-style flow with a three-level label: numClasses is inferred as max label + 1, the model emits rawPrediction /
    probability vectors of that length, honours thresholds, and survives save / load in MLlib's layout (impurityStats = class counts)."""
    import json
    from b2rf import spark as sp
    from b2rf._pyspark_stub import LocalFrame, ml_attr_metadata
    from oracle import multiclass as M
    X, y, arity = _data(b2rf, 3000, 3)
    meta = {"features": ml_attr_metadata(arity)}
    train = LocalFrame({"features": X[:2400], "tier": y[:2400]}, meta, num_partitions=3)
    test = LocalFrame({"features": X[2400:], "tier": y[2400:]}, meta)
    po = np.array([0, 800, 1600, 2400])
    rf = sp.B200RandomForestClassifier(labelCol="tier", maxBins=40, maxDepth=4, numTrees=6, seed=11)
    model = rf.fit(train)
    m = M.fit(X[:2400], y[:2400], arity, 3, part_offsets=po, maxDepth=4, maxBins=40, numTrees=6, bootstrap=True, seed=11, impurity="gini")
    assert model.numClasses == 3 and model.totalNumNodes == len(m.feature)
    out = model.transform(test)
    np.testing.assert_allclose(out.col("rawPrediction"), m.predict_raw(X[2400:]), rtol=1e-15, atol=0)
    assert np.array_equal(out.col("prediction"), m.predict(X[2400:]))
    assert out.col("probability").shape == (600, 3)
    np.testing.assert_allclose(out.col("probability").sum(1), 1.0, rtol=1e-12)
    biased = model.copy({model.thresholds: [1.0, 1.0, 0.01]}).transform(test).col("prediction")
    assert (biased == 2.0).mean() > (out.col("prediction") == 2.0).mean()
    assert len(model.trees) == 6 and model.trees[0].numClasses == 3
    path = str(tmp_path / "rfc3")
    model.write().overwrite().save(path)
    md = json.loads(open(path + "/metadata/part-00000").readline())
    assert md["numClasses"] == 3 and md["class"].endswith("RandomForestClassificationModel")
    again = sp.B200RandomForestClassificationModel.load(path)
    assert again.numClasses == 3 and np.array_equal(again._forest.class_counts, model._forest.class_counts)
    assert np.array_equal(again.transform(test).col("rawPrediction"), out.col("rawPrediction"))
    assert "numClasses=3" in model.toDebugString.splitlines()[0]
    dt = sp.B200DecisionTreeClassifier(labelCol="tier", maxBins=40, maxDepth=3, impurity="entropy").fit(train)
    md_ = M.fit(X[:2400], y[:2400], arity, 3, part_offsets=po, maxDepth=3, maxBins=40, impurity="entropy")
    assert np.array_equal(dt.transform(test).col("rawPrediction"), md_.predict_raw(X[2400:]))
