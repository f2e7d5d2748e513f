"""CPU tests of the drop-in boundary: libb2rf.so loads without a GPU and exports every symbol include/b2rf.h declares;
host-only entry points (parameter tables, metadata checks, Java-style number formatting) behave like MLlib."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "b2rf.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b2rf_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(b2rf):
    from b2rf._lib import EXPORTS, SO
    assert os.path.exists(SO), "libb2rf.so not built (python __graft_entry__.py)"
    L = C.CDLL(SO)
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/b2rf.h but not exported"
    assert set(EXPORTS) == set(names), f"python binding out of sync with the header: {set(EXPORTS) ^ set(names)}"


def test_version_and_error_string(b2rf):
    from b2rf._lib import lib
    assert lib().b2rf_version() >= 100
    assert isinstance(lib().b2rf_last_error(), bytes)


def test_feature_subset_strategy(b2rf):
    P = b2rf.TreeParams
    assert P(numTrees=1).features_per_node(33) == 33
    assert P(numTrees=100).features_per_node(33) == 11           # auto -> onethird for regression forests
    assert P(featureSubsetStrategy="sqrt").features_per_node(33) == 6
    assert P(featureSubsetStrategy="log2").features_per_node(33) == 6
    assert P(featureSubsetStrategy="0.5").features_per_node(33) == 17
    assert P(featureSubsetStrategy="7").features_per_node(33) == 7
    assert P(featureSubsetStrategy="100").features_per_node(33) == 33
    for bad in ["-.1", "0", "0.0", "1.1", "foo"]:
        with pytest.raises(b2rf.B2rfError, match="Supported values"):
            P(featureSubsetStrategy=bad).features_per_node(33)


def test_check_metadata_matches_mllib_message(b2rf):
    from b2rf._lib import check, lib
    p = b2rf.TreeParams(maxBins=32).to_c(3)
    ar = np.array([0, 36, 2], np.int32)
    with pytest.raises(b2rf.IllegalArgumentException) as e:
        check(lib().b2rf_check_metadata(C.byref(p), 3, ar.ctypes.data_as(C.c_void_p), 5000))
    assert "DecisionTree requires maxBins (= 32) to be at least as large as the number of values in each categorical feature, " \
           "but categorical feature 1 has 36 values." in str(e.value)
    p = b2rf.TreeParams(maxBins=40).to_c(3)
    assert check(lib().b2rf_check_metadata(C.byref(p), 3, ar.ctypes.data_as(C.c_void_p), 5000)) == 40
    assert check(lib().b2rf_check_metadata(C.byref(p), 3, ar.ctypes.data_as(C.c_void_p), 37)) == 37   # maxPossibleBins = min(maxBins, N)
    assert lib().b2rf_samples_fraction(5716, 40) == 1.0
    assert lib().b2rf_samples_fraction(10_000_000, 64) == 10000 / 10_000_000
    assert lib().b2rf_samples_fraction(100_000_000, 256) == 65536 / 100_000_000


def test_params_validation(b2rf):
    with pytest.raises(ValueError):
        b2rf.TreeParams(maxDepth=31).to_c(4)
    with pytest.raises(ValueError):
        b2rf.TreeParams(impurity="mse").to_c(4)
    g = b2rf.TreeParams(impurity="gini", numTrees=10).to_c(16)     # classification forests: featureSubsetStrategy auto -> sqrt
    assert g.impurity == 1 and g.featuresPerNode == 4
    assert b2rf.TreeParams(impurity="entropy").to_c(16).impurity == 2 and b2rf.TreeParams(impurity="entropy").to_c(16).featuresPerNode == 16
    with pytest.raises(ValueError):
        b2rf.TreeParams(subsamplingRate=0.0).to_c(4)
    c = b2rf.TreeParams(maxDepth=4.0, numTrees=37.0).to_c(4)      # `ML 08 - Hyperopt.py:97`: integral floats from hp.quniform
    assert c.maxDepth == 4 and c.numTrees == 37


def test_no_cpu_fallback_without_gpu(b2rf):
    import torch
    from b2rf import trainer
    with pytest.raises(TypeError, match="CUDA tensor"):
        trainer.fit_tensors(torch.zeros((4, 2), dtype=torch.float64), torch.zeros(4, dtype=torch.float64), [0, 0], b2rf.TreeParams())


def test_java_double_str(b2rf):
    f = b2rf.java_double_str
    assert [f(v) for v in (0.5, 1e-4, 123456789.0, 100.0, 0.001, 1.5e10, -37.76123, 1234567.0, 0.0, 1e7, 9999999.0, 2.5e-5)] == \
        ["0.5", "1.0E-4", "1.23456789E8", "100.0", "0.001", "1.5E10", "-37.76123", "1234567.0", "0.0", "1.0E7", "9999999.0", "2.5E-5"]


def test_forest_model_api_on_oracle_arrays(b2rf, oracle):
    """featureImportances / toDebugString / depth on a forest (arrays from the oracle: no GPU needed)."""
    X = np.array([[1.0], [0.0]]); y = np.array([1.0, 0.0])
    o = oracle.fit(X, y, [0], maxDepth=2)
    fo = b2rf.Forest(1, o.tree_off, o.feature, o.threshold, o.left, o.right, o.prediction, o.impurity, o.gain, o.stats, o.ncat, o.cat_off,
                     o.cat_len, o.cats, uid="DecisionTreeRegressor_abc")
    assert fo.depth(0) == 1 and fo.num_nodes(0) == 3
    assert fo.feature_importances().tolist() == [1.0]
    assert fo.to_debug_string() == ("DecisionTreeRegressionModel: uid=DecisionTreeRegressor_abc, depth=1, numNodes=3, numFeatures=1\n"
                                    "  If (feature 0 <= 0.5)\n   Predict: 0.0\n  Else (feature 0 > 0.5)\n   Predict: 1.0\n")
    rows = [(0, [1, 0, 0, 0, 1]), (1, [1, 1, 0, 1, 0]), (1, [1, 1, 0, 0, 0]), (0, [1, 0, 0, 0, 0]), (1, [1, 1, 0, 0, 0])]
    o = oracle.fit(np.array([r[1] for r in rows], float), np.array([r[0] for r in rows], float), [0] * 5, maxDepth=3)
    fo = b2rf.Forest(5, o.tree_off, o.feature, o.threshold, o.left, o.right, o.prediction, o.impurity, o.gain, o.stats, o.ncat, o.cat_off,
                     o.cat_len, o.cats)
    imp = fo.feature_importances()
    assert imp.sum() == pytest.approx(1.0) and int(np.argmax(imp)) == 1
