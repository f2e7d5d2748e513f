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


def test_multiclass_label_check(b2rf, cuda):
    X = np.random.default_rng(0).normal(size=(60, 3)); y = np.arange(60) % 4.0
    with pytest.raises(b2rf.IllegalArgumentException, match="numClasses == 3"):
        b2rf.trainer.fit_tensors(torch.as_tensor(X, device=cuda), torch.as_tensor(y, device=cuda), [0, 0, 0], b2rf.TreeParams(impurity="gini", numClasses=3))
