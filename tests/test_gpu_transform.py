"""K7 in its packed form and the `_transform` outputs around it: predictions, leaf indices (leafCol),
leaf variances (varianceCol), checked against a plain numpy traversal of the oracle's trees."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _walk(o, X):
    """numpy traversal of an OracleForest: tree-local leaf index [n, T]."""
    T = len(o.tree_off) - 1
    out = np.zeros((X.shape[0], T), np.int64)
    for i in range(X.shape[0]):
        for t in range(T):
            b = int(o.tree_off[t]); n = 0
            while o.feature[b + n] >= 0:
                x = X[i, o.feature[b + n]]
                if o.ncat[b + n] == 0:
                    left = x <= o.threshold[b + n]
                else:
                    left = x in o.cats[o.cat_off[b + n]:o.cat_off[b + n] + o.cat_len[b + n]]
                n = int(o.left[b + n] if left else o.right[b + n])
            out[i, t] = n
    return out


def _forest(b2rf, o, F):
    return b2rf.Forest(F, o.tree_off, o.feature, o.threshold, o.left, o.right, o.prediction, o.impurity, o.gain, o.stats, o.ncat, o.cat_off,
                       o.cat_len, o.cats)


def test_packed_predict_and_leaves_match_numpy_walk(b2rf, oracle, cuda):
    X, y, arity = b2rf.synth.airbnb(3000)
    X, y = X.numpy(), y.numpy()
    for T in (1, 7):      # 7: a tail group shorter than the trees walked at once
        o = oracle.fit(X[:2000], y[:2000], arity, maxDepth=6, maxBins=40, numTrees=T, bootstrap=T > 1, seed=5)
        fo = _forest(b2rf, o, 33)
        assert fo.is_preorder()
        Xt = torch.as_tensor(X[2000:], device=cuda)
        leaves = fo.predict_leaves(Xt).cpu().numpy()
        want = _walk(o, X[2000:])
        assert np.array_equal(leaves, want)
        np.testing.assert_array_equal(fo.predict(Xt).cpu().numpy(), o.predict(X[2000:]))     # same tree-order sum, same division


def test_plain_kernel_still_serves_non_preorder_forests(b2rf, oracle, cuda):
    """A forest whose node numbering is not pre-order (children swapped in storage) goes through b2rf_predict."""
    X, y, arity = b2rf.synth.airbnb(1500)
    X, y = X.numpy(), y.numpy()
    o = oracle.fit(X, y, arity, maxDepth=4, maxBins=40)
    n = len(o.feature)
    perm = np.arange(n); perm[1:] = perm[1:][::-1]          # new position of old node k is perm[k]; the root stays first
    inv = np.argsort(perm)
    def mv(a): return np.asarray(a)[inv]
    remap = lambda c: np.where(c >= 0, perm[np.maximum(c, 0)], c).astype(np.int32)
    fo = b2rf.Forest(33, o.tree_off, mv(o.feature), mv(o.threshold), remap(mv(o.left)), remap(mv(o.right)), mv(o.prediction), mv(o.impurity),
                     mv(o.gain), mv(o.stats), mv(o.ncat), mv(o.cat_off), mv(o.cat_len), o.cats)
    assert not fo.is_preorder()
    np.testing.assert_array_equal(fo.predict(torch.as_tensor(X, device=cuda)).cpu().numpy(), o.predict(X))


def test_leaf_and_variance_columns(b2rf, oracle, cuda):
    from b2rf import spark as sp
    from b2rf._pyspark_stub import LocalFrame, ml_attr_metadata
    X, y, arity = b2rf.synth.airbnb(3000)
    X, y = X.numpy(), y.numpy()
    meta = {"features": ml_attr_metadata(arity)}
    train = LocalFrame({"features": X[:2000], "price": y[:2000]}, meta)
    test = LocalFrame({"features": X[2000:], "price": y[2000:]}, meta)
    dt = sp.B200DecisionTreeRegressor(labelCol="price", maxBins=40, maxDepth=4, leafCol="leaf", varianceCol="var").fit(train)
    o = oracle.fit(X[:2000], y[:2000], arity, maxDepth=4, maxBins=40)
    want = _walk(o, X[2000:])[:, 0]
    leaf_rank = np.cumsum(o.feature < 0) - 1                 # pre-order: rank among leaves == leaves seen so far
    out = dt.transform(test)
    assert np.array_equal(out.col("leaf"), leaf_rank[want].astype(np.float64))
    np.testing.assert_allclose(out.col("var"), o.impurity[want], rtol=1e-9)
    assert dt.predictLeaf(X[2000]) == out.col("leaf")[0] and dt.predictVariance(X[2000]) == out.col("var")[0]
    rf = sp.B200RandomForestRegressor(labelCol="price", maxBins=40, maxDepth=3, numTrees=5, seed=7, leafCol="leaves").fit(train)
    orf = oracle.fit(X[:2000], y[:2000], arity, maxDepth=3, maxBins=40, numTrees=5, bootstrap=True, seed=7)
    w = _walk(orf, X[2000:])
    ranks = np.zeros_like(w, dtype=np.float64)
    for t in range(5):
        s = slice(int(orf.tree_off[t]), int(orf.tree_off[t + 1]))
        ranks[:, t] = (np.cumsum(orf.feature[s] < 0) - 1)[w[:, t]]
    got = rf.transform(test).col("leaves")
    assert got.shape == (1000, 5) and np.array_equal(got, ranks)
    with pytest.raises(NotImplementedError, match="weightCol"):
        sp.B200DecisionTreeRegressor(labelCol="price", weightCol="w").fit(train)


def test_coded_traversal_equals_f64_traversal(b2rf, cuda):
    """K7 on bin codes (all-continuous forests, large batches) reaches the same leaves as the f64 traversal: identical leaf indices and
    bit-identical predictions, for uint8 codes and, with more than 255 distinct thresholds on a feature, uint16 codes."""
    X, y, arity = b2rf.synth.c3(60000, device=cuda, F=16)
    for maxBins, depth, T, want_dt in [(64, 8, 9, torch.uint8), (1000, 12, 3, torch.int16)]:
        p = b2rf.TreeParams(maxDepth=depth, maxBins=maxBins, numTrees=T, bootstrap=True, seed=5)
        f = b2rf.trainer.fit_tensors(X[:30000], y[:30000], arity, p)
        cd = f._coded(cuda)
        assert cd is not None and cd["code_dt"] == want_dt
        a, la = f._predict_packed(X, True, True)
        b, lb = f._predict_coded(X, True, True, chunk_rows=25000)        # three chunks, the last one ragged
        assert torch.equal(la, lb)
        assert torch.equal(a, b)
        assert torch.equal(f.predict(X), a)                                # the public entry takes the coded path for this batch size
        edge = X[:5000].clone()
        thr = torch.as_tensor(f.threshold[f.feature >= 0][:16], device=cuda)
        edge[:16, int(f.feature[0])] = thr[0]                              # values exactly on a threshold go left in both forms
        assert torch.equal(f._predict_packed(edge, True, False)[0], f._predict_coded(edge, True, False)[0])


def test_coded_traversal_not_used_with_categorical_splits(b2rf, oracle, cuda):
    X, y, arity = b2rf.synth.airbnb(6000)
    o = oracle.fit(X.numpy()[:3000], y.numpy()[:3000], arity, maxDepth=5, maxBins=40)
    fo = _forest(b2rf, o, 33)
    assert (fo.ncat > 0).any() and fo._coded(cuda) is None
    np.testing.assert_array_equal(fo.predict(X.to(cuda)).cpu().numpy(), o.predict(X.numpy()))


def test_nan_features_take_the_last_bin_and_go_right(b2rf, oracle, cuda):
    """TreePoint.findBin uses java.util.Arrays.binarySearch, which orders NaN above every double: a NaN feature lands in the last bin;
    at predict time `NaN <= threshold` is false, so both K7 forms send it right."""
    X = np.array([[0.0], [1.0], [2.0], [3.0], [np.nan]]); arity = [0]
    thr, ns = oracle.find_splits(X[:4], arity, 32, 4.0)
    want = np.asarray(oracle.bin_features(X, thr, ns, arity)).reshape(-1)
    assert want[-1] == ns[0]
    tr = b2rf.trainer
    got = tr.bin_features(torch.as_tensor(X, device=cuda), torch.as_tensor(thr, device=cuda), torch.as_tensor(ns, device=cuda),
                          torch.as_tensor(np.asarray(arity, np.int32), device=cuda))
    assert got[0, :5].cpu().numpy().tolist() == want.tolist()
    Xc, yc, ar = b2rf.synth.c3(20000, device=cuda, F=8)
    f = tr.fit_tensors(Xc, yc, ar, b2rf.TreeParams(maxDepth=5, maxBins=32, numTrees=3, bootstrap=True, seed=1))
    Xn = Xc.clone(); Xn[::7, 3] = float("nan"); Xn[::11, 0] = float("nan")
    assert torch.equal(f._predict_packed(Xn, True, False)[0], f._predict_coded(Xn, True, False)[0])


def test_packed_categorical_bit_masks(b2rf, oracle, cuda):
    """Categorical splits in the packed kernel are one bit test; values that are not a category (non-integral, negative, beyond the
    arity) go right, as `categories.contains(x)` does."""
    X, y, arity = b2rf.synth.airbnb(4000)
    X, y = X.numpy(), y.numpy()
    o = oracle.fit(X[:3000], y[:3000], arity, maxDepth=7, maxBins=40, numTrees=5, bootstrap=True, seed=2)
    fo = _forest(b2rf, o, 33)
    assert (fo.ncat > 0).sum() > 10
    Xt = X[3000:].copy()
    cat_cols = [j for j, a in enumerate(arity) if a > 1]
    Xt[::5, cat_cols[0]] = 2.5
    Xt[::7, cat_cols[1]] = -1.0
    Xt[::9, cat_cols[2]] = 1000.0
    np.testing.assert_array_equal(fo.predict(torch.as_tensor(Xt, device=cuda)).cpu().numpy(), o.predict(Xt))
    assert np.array_equal(fo.predict_leaves(torch.as_tensor(Xt, device=cuda)).cpu().numpy(), _walk(o, Xt))
