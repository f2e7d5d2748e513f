"""CPU tests of the oracle (oracle/mllib_oracle.cpp) against every known-answer vector available for this path
(upstream Spark unit-test values and pyspark doctest values) and the committed golden fixtures."""
import json
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def test_java_random(oracle):
    assert oracle.java_random(42) == (-1170105035, -5025562857975149833)


def test_xorshift(oracle):
    d = oracle.xorshift_doubles(42, 2)
    assert d[0] == 0.619189370225301 and d[1] == 0.5096018842446481
    # pyspark docs: rand(seed=42) * 3 on a 2-row frame
    assert d[0] * 3 == 1.8575681106759028 and d[1] * 3 == 1.5288056527339444


def test_scala_map_orders(oracle):
    assert oracle.group_order(np.arange(1, 11)).tolist() == [5, 10, 1, 6, 9, 2, 7, 3, 8, 4]
    assert oracle.group_order([1, 2, 3]).tolist() == [2, 1, 3]
    assert oracle.group_order([0, 1, 2, 3]).tolist() == [2, 1, 3, 0]


def test_well19937c_poisson(oracle):
    np.testing.assert_allclose(oracle.well_doubles(43, 3), [0.105981, 0.663391, 0.315666], atol=1e-6)
    assert oracle.poisson(43, 1.0, 6).tolist() == [0, 1, 1, 1, 1, 0]
    np.testing.assert_allclose(oracle.well_doubles(44, 3), [0.303765, 0.632236, 0.80593], atol=1e-6)
    assert oracle.poisson(44, 1.0, 6).tolist() == [0, 4, 1, 0, 0, 1]


def test_reservoir(oracle):
    seeds = oracle.java_random_longs(42, 3)
    want = [[0, 1, 28, 23, 30, 31, 29, 7, 22, 9, 10], [12, 1, 20, 23, 25, 5, 11, 7, 8, 14, 27], [27, 30, 2, 25, 32, 14, 6, 22, 18, 21, 12]]
    for s, w in zip(seeds, want):
        assert oracle.reservoir(33, 11, int(s)).tolist() == w
    assert oracle.reservoir(5, 5, 1).tolist() == [0, 1, 2, 3, 4]


@pytest.mark.parametrize("vals,nbins,want", [
    ([1, 1, 2, 2, 3, 3, 3, 3], 5, [1.5, 2.5]), ([2] * 10 + [3, 4, 5], 3, [2.5, 3.5]), ([0, 1] + [2] * 10, 2, [1.5]), ([0, 0, 0], 3, []),
    ([0, 1, 0, 0, 1, 0, 1, 1], 3, [0.5]), ([0, 0, 1, 1, 2, 2, 3, 3], 3, [0.5, 2.5])])
def test_find_splits_upstream_kats(oracle, vals, nbins, want):
    """RandomForestSuite 'find splits for a continuous feature' cases (incl. SPARK-16957)."""
    thr, ns = oracle.find_splits(np.array(vals, float).reshape(-1, 1), [0], nbins, float(len(vals)))
    assert thr[0, : ns[0]].tolist() == want


def test_feature_subset_strategy_table(oracle):
    F = 33
    assert oracle.features_per_node("onethird", F, 10) == 11
    assert oracle.features_per_node("sqrt", F, 10) == 6
    assert oracle.features_per_node("log2", F, 10) == 6
    assert oracle.features_per_node("auto", F, 1) == 33 and oracle.features_per_node("auto", F, 2) == 11
    for bad in ["-.1", "0", "0.0", "1.1"]:
        with pytest.raises(ValueError):
            oracle.features_per_node(bad, F, 10)


def test_pyspark_doctests(oracle):
    X = np.array([[1.0], [0.0]]); y = np.array([1.0, 0.0])
    dt = oracle.fit(X, y, [0], maxDepth=2)
    assert len(dt.feature) == 3 and dt.feature.tolist() == [0, -1, -1]
    assert dt.predict(np.array([[-1.0], [1.0]])).tolist() == [0.0, 1.0]
    rf = oracle.fit(X, y, [0], maxDepth=2, numTrees=2, bootstrap=True, seed=42, part_offsets=[0, 1, 2])
    assert rf.predict(np.array([[-1.0], [1.0]])).tolist() == [0.0, 0.5]
    assert np.diff(rf.tree_off).tolist() == [1, 3]          # tree 0 drew an empty bag: single leaf predicting 0.0


def test_categorical_stump_and_importance(oracle):
    X = np.zeros((1000, 2)); y = np.zeros(1000)
    X[:600, 1] = 1; y[:600] = 1; X[600:, 0] = 1
    f = oracle.fit(X, y, [3, 3], maxDepth=2, maxBins=100)
    assert f.feature[0] == 0 and f.cats[: f.cat_len[0]].tolist() == [1.0]
    assert f.prediction[0] == pytest.approx(0.6) and f.impurity[0] == pytest.approx(0.24) and f.gain[0] == pytest.approx(0.24)
    rows = [(0, [1, 0, 0, 0, 1]), (1, [1, 1, 0, 1, 0]), (1, [1, 1, 0, 0, 0]), (0, [1, 0, 0, 0, 0]), (1, [1, 1, 0, 0, 0])]
    f = oracle.fit(np.array([r[1] for r in rows], float), np.array([r[0] for r in rows], float), [0] * 5, maxDepth=3)
    assert f.feature[0] == 1


def test_maxbins_requirement(oracle):
    X = np.zeros((100, 2)); X[:, 1] = np.arange(100) % 40
    with pytest.raises(oracle.OracleError, match=r"maxBins \(= 32\) to be at least as large"):
        oracle.fit(X, np.zeros(100), [0, 40], maxBins=32)
    oracle.fit(X, np.arange(100.0), [0, 40], maxBins=40)


def test_grouping_independent_without_feature_subsampling(oracle, b2rf):
    """Upstream 'Second level node building with vs. without groups': with all features the forest does not depend on maxMemoryInMB."""
    X, y, arity = b2rf.synth.airbnb(1500)
    a = oracle.fit(X.numpy(), y.numpy(), arity, maxDepth=6, maxBins=40, maxMemoryInMB=256)
    b = oracle.fit(X.numpy(), y.numpy(), arity, maxDepth=6, maxBins=40, maxMemoryInMB=0)   # one node per group
    assert len(b.group_sizes) > len(a.group_sizes)
    assert np.array_equal(a.feature, b.feature) and np.array_equal(a.threshold, b.threshold) and np.array_equal(a.prediction, b.prediction)


def test_multithreaded_oracle_matches_single_thread(oracle, b2rf):
    X, y, arity = b2rf.synth.airbnb(2000)
    po = np.arange(0, 2001, 250)
    a = oracle.fit(X.numpy(), y.numpy(), arity, maxDepth=6, maxBins=40, numTrees=8, bootstrap=True, seed=3, part_offsets=po, numThreads=1)
    b = oracle.fit(X.numpy(), y.numpy(), arity, maxDepth=6, maxBins=40, numTrees=8, bootstrap=True, seed=3, part_offsets=po, numThreads=4)
    assert np.array_equal(a.feature, b.feature) and np.array_equal(a.threshold, b.threshold)
    assert np.array_equal(a.stats[:, 0], b.stats[:, 0])


def test_golden_fixtures(oracle, b2rf):
    """tests/golden/*.json were written by tests/golden/make_golden.py from this oracle; they pin it against regressions and
    travel to the GPU box where the CUDA path is checked against the same files."""
    files = sorted(f for f in os.listdir(GOLD) if f.endswith(".json") and f[0] == "c")   # c1_*, c2_*: forest fixtures (the persistence schema fixture has its own test)
    assert files, "no golden fixtures"
    for fn in files:
        gold = json.load(open(os.path.join(GOLD, fn)))
        c = gold["config"]
        X, y, arity = b2rf.synth.airbnb(c["rows"], seed=c["data_seed"])
        f = oracle.fit(X.numpy(), y.numpy(), arity, part_offsets=c["part_offsets"], **c["params"])
        assert f.tree_off.tolist() == gold["tree_off"]
        assert f.feature.tolist() == gold["feature"]
        assert [float(v).hex() for v in f.threshold] == gold["threshold_hex"]
        assert [float(v).hex() for v in f.prediction] == gold["prediction_hex"]
        assert f.stats[:, 0].tolist() == gold["count"]


def test_classifier_known_answers(oracle):
    """Binary classification restatement against the values available without a JVM: the pyspark DecisionTreeClassifier doctest
    (two rows -> depth 1, 3 nodes, rawPrediction [1, 0] / [0, 1]) and the upstream 600/400 categorical stump data (root gini
    1 - 0.6^2 - 0.4^2 = 0.48, categories ordered by their class-1 count so category 1 goes left)."""
    X = np.array([[1.0], [0.0]]); y = np.array([1.0, 0.0])
    f = oracle.fit(X, y, [0], maxDepth=2, impurity="gini")
    assert len(f.feature) == 3 and f.threshold[0] == 0.5
    assert f.predict_raw(np.array([[-1.0], [1.0]])).tolist() == [[1.0, 0.0], [0.0, 1.0]]
    assert f.predict(np.array([[-1.0], [1.0]])).tolist() == [0.0, 1.0]
    Xs = np.array([[0.0, 1.0]] * 600 + [[1.0, 0.0]] * 400); ys = np.array([1.0] * 600 + [0.0] * 400)
    g = oracle.fit(Xs, ys, [3, 3], maxDepth=2, maxBins=100, impurity="gini")
    assert g.feature.tolist() == [0, -1, -1] and g.cats.tolist() == [1.0]
    assert g.impurity[0] == pytest.approx(0.48, rel=1e-15) and g.gain[0] == pytest.approx(0.48, rel=1e-15)
    assert g.stats[:, :2].tolist() == [[400.0, 600.0], [400.0, 0.0], [0.0, 600.0]] and g.prediction.tolist() == [1.0, 0.0, 1.0]
    e = oracle.fit(Xs, ys, [3, 3], maxDepth=2, maxBins=100, impurity="entropy")
    assert e.impurity[0] == pytest.approx(-(0.6 * np.log2(0.6) + 0.4 * np.log2(0.4)), rel=1e-14)
    with pytest.raises(oracle.OracleError):
        oracle.fit(np.zeros((3, 1)), np.array([0.0, 1.0, 2.0]), [0], impurity="gini")


def test_numpy_multiclass_restatement_matches_cpp_oracle_on_binary(oracle, b2rf):
    """oracle/multiclass.py (any numClasses, unordered categorical splits) is an independent restatement of the training loop; for
    numClasses == 2 it must reproduce the C++ oracle bit for bit (trees, class counts, impurities, gains, evaluation order, groups)."""
    from oracle import multiclass as M
    X, y, arity = b2rf.synth.airbnb(3000)
    X, y = X.numpy(), y.numpy()
    yb = (y > np.median(y)).astype(np.float64)
    po = np.array([0, 750, 1500, 2250, 3000])
    for kw in [dict(maxDepth=5, maxBins=40, impurity="gini"),
               dict(maxDepth=4, maxBins=40, numTrees=6, bootstrap=True, seed=42, part_offsets=po, impurity="entropy"),
               dict(maxDepth=6, maxBins=40, numTrees=5, bootstrap=True, seed=7, part_offsets=po, impurity="gini", featureSubsetStrategy="all",
                    maxMemoryInMB=1, minInstancesPerNode=3)]:
        m = M.fit(X, yb, arity, 2, **kw)
        o = oracle.fit(X, yb, arity, **kw)
        for a in ("tree_off", "feature", "threshold", "left", "right", "prediction", "impurity", "gain", "trace_i"):
            assert np.array_equal(getattr(m, a), getattr(o, a)), a
        assert np.array_equal(m.class_counts, o.stats[:, :2]) and list(m.group_sizes) == list(o.group_sizes)
        assert np.array_equal(m.predict_raw(X[:100]), o.predict_raw(X[:100]))


def test_multiclass_metadata_and_subset_splits(b2rf):
    """DecisionTreeMetadata: with maxBins = 32 categorical features of arity <= floor(log2(32/2 + 1) + 1) = 5 are unordered and count
    2^(arity-1) - 1 bins; RandomForest.extractMultiClassCategories(splitIndex + 1, arity) lists the set bits."""
    from oracle import multiclass as M
    assert M.unordered_left_categories(0, 3) == [0] and M.unordered_left_categories(1, 3) == [1] and M.unordered_left_categories(2, 3) == [0, 1]
    rng = np.random.default_rng(1)
    n = 600
    X = np.stack([rng.integers(0, 3, n), rng.integers(0, 5, n), rng.integers(0, 6, n), rng.normal(size=n)], axis=1).astype(np.float64)
    y = ((X[:, 0] == 1) * 1 + (X[:, 1] >= 3) * 1 + (X[:, 3] > 0.5) * 1).astype(np.float64) % 3
    m = M.fit(X, y, [3, 5, 6, 0], 3, maxDepth=3, maxBins=32)
    assert m.unordered == [0, 1]                                   # arity 6 > 5 stays ordered (sorted by category impurity)
    assert (m.predict(X) == y).mean() > 0.6
    # the three-class toy: a pure split on an unordered feature must be found at the root when one exists
    Xp = np.array([[0.0], [1.0], [2.0]] * 30); yp = np.array([0.0, 1.0, 2.0] * 30)
    t = M.fit(Xp, yp, [3], 3, maxDepth=2, maxBins=32)
    assert t.feature[0] == 0 and t.ncat[0] == 3 and len(t.feature) == 5 and (t.predict(Xp) == yp).all()
    assert t.impurity[0] == pytest.approx(1 - 3 * (1 / 3) ** 2)
