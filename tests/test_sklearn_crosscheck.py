"""Independent structural cross-check: scikit-learn's exact-split DecisionTreeRegressor shares no code
with the oracle or the CUDA path.  On integer-valued features with fewer distinct values than bins, MLlib's candidate thresholds
are the midpoints between all consecutive distinct values, so the histogram scan evaluates exactly the partitions sklearn's
exhaustive search evaluates; with untied gains both must grow the same tree (same partition of the training rows at every
node).  Thresholds differ by construction (sklearn: midpoint of the values present in the node; MLlib: global midpoints), so
the comparison is made on the training rows: per-row predictions, leaf count, and per-node (feature, sample count)."""
import numpy as np
import pytest

sklearn_tree = pytest.importorskip("sklearn.tree")


def _data(seed, n=3000, F=6, K=12):
    rng = np.random.default_rng(seed)
    X = rng.integers(0, K, size=(n, F)).astype(np.float64)
    y = 3.0 * (X[:, 0] > 4) + 1.5 * (X[:, 1] > 7) - 2.0 * (X[:, 2] <= 2) * (X[:, 3] > 5) + 0.1 * X[:, 4] + rng.normal(0, 1.0, n)
    return X, y


def _sk(X, y, depth):
    return sklearn_tree.DecisionTreeRegressor(max_depth=depth, criterion="squared_error", random_state=0).fit(X, y)


def _node_signature_sklearn(m):
    t = m.tree_
    return sorted((int(t.feature[i]), int(t.n_node_samples[i])) for i in range(t.node_count) if t.children_left[i] != -1)


def _node_signature_ours(f):
    internal = f.left >= 0
    return sorted((int(a), int(b)) for a, b in zip(f.feature[internal], f.stats[internal, 0]))


@pytest.mark.parametrize("seed,depth", [(1, 3), (2, 5), (3, 6)])
def test_oracle_matches_sklearn_exact_search(oracle, seed, depth):
    X, y = _data(seed)
    o = oracle.fit(X, y, [0] * X.shape[1], maxDepth=depth, maxBins=32)
    m = _sk(X, y, depth)
    np.testing.assert_allclose(o.predict(X), m.predict(X), rtol=1e-9, atol=1e-9)
    assert int((o.left < 0).sum()) == m.get_n_leaves()
    assert _node_signature_ours(o) == _node_signature_sklearn(m)


@pytest.mark.gpu
@pytest.mark.parametrize("seed,depth", [(1, 3), (2, 5), (3, 6)])
def test_cuda_path_matches_sklearn_exact_search(b2rf, cuda, seed, depth):
    import torch
    X, y = _data(seed)
    f = b2rf.trainer.fit_tensors(torch.as_tensor(X, device=cuda), torch.as_tensor(y, device=cuda), [0] * X.shape[1],
                                 b2rf.TreeParams(maxDepth=depth, maxBins=32))
    m = _sk(X, y, depth)
    np.testing.assert_allclose(f.predict(torch.as_tensor(X, device=cuda)).cpu().numpy(), m.predict(X), rtol=1e-9, atol=1e-9)
    assert int((f.left < 0).sum()) == m.get_n_leaves()
    assert _node_signature_ours(f) == _node_signature_sklearn(m)


@pytest.mark.parametrize("crit", ["gini", "entropy"])
def test_oracle_classifier_matches_sklearn_exact_search(oracle, crit):
    """Binary classification: same predictions on the training rows as sklearn's exhaustive search (MLlib additionally merges
    sibling leaves that predict the same class, so leaf counts differ)."""
    rng = np.random.default_rng(5); n = 20000
    X = rng.integers(0, 12, size=(n, 6)).astype(np.float64)
    logit = 1.2 * (X[:, 0] > 4) + 0.8 * (X[:, 1] > 7) - 1.0 * (X[:, 2] <= 2) * (X[:, 3] > 5) + 0.05 * X[:, 4] - 0.6
    y = (rng.random(n) < 1 / (1 + np.exp(-logit))).astype(np.float64)
    o = oracle.fit(X, y, [0] * 6, maxDepth=4, maxBins=32, impurity=crit)
    m = sklearn_tree.DecisionTreeClassifier(max_depth=4, criterion=crit, random_state=0).fit(X, y)
    assert np.array_equal(o.predict(X), m.predict(X))
