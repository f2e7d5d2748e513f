"""GPU parity tests: every kernel and the whole fit against the CPU oracle on the same seeded inputs, through the C-ABI.
Contract (BASELINE.json north_star): (1) bin codes and histogram counts bit-exact, (2) fp64 label sums to 1e-9 relative,
(3) tree structure / thresholds / predictions identical, test RMSE within 1e-6."""
import numpy as np
import pytest
import torch

from helpers import REL, assert_forest_parity, assert_parity_allowing_ties, assert_trace_parity, rmse

pytestmark = pytest.mark.gpu


def _airbnb(b2rf, n=7146, seed=42):
    X, y, arity = b2rf.synth.airbnb(n, seed=seed)
    ntr = int(n * 0.8)
    return X[:ntr].numpy(), y[:ntr].numpy(), X[ntr:].numpy(), y[ntr:].numpy(), arity


@pytest.fixture(params=["segment", "sort", "scan"])
def mode(request):
    """All K4 implementations: node-partitioned row lists / per-group counting sort (both + shared-memory histograms), and
    column-major scan + global RED."""
    return request.param


def _fit_both(b2rf, oracle, cuda, X, y, arity, part_offsets=None, hist_mode="segment", **kw):
    from b2rf import trainer
    p = b2rf.TreeParams(**kw)
    if part_offsets is None:
        part_offsets = trainer.default_partitions(len(y), 1024)
    Xd = torch.as_tensor(X, device=cuda)
    yd = torch.as_tensor(y, device=cuda)
    g = trainer.fit_tensors(Xd, yd, arity, p, part_offsets=part_offsets, keep_trace=True, hist_mode=hist_mode)
    N = len(y)
    mpb = min(p.maxBins, N)
    meta_max = max([a if a > 1 else mpb for a in arity])
    req = max(meta_max * meta_max, 10000)
    fraction = req / N if req < N else 1.0
    sample_idx = None
    if fraction < 1.0:
        sample_idx = np.nonzero(oracle.sample_keep(p.seed, np.arange(N), fraction))[0]
    o = oracle.fit(X, y, arity, maxDepth=p.maxDepth, maxBins=p.maxBins, minInstancesPerNode=p.minInstancesPerNode, numTrees=p.numTrees,
                   featureSubsetStrategy=p.featureSubsetStrategy, maxMemoryInMB=p.maxMemoryInMB, bootstrap=p.bootstrap,
                   minInfoGain=p.minInfoGain, minWeightFractionPerNode=p.minWeightFractionPerNode, subsamplingRate=p.subsamplingRate,
                   seed=p.seed, part_offsets=part_offsets, sample_idx=sample_idx, fraction=fraction, impurity=p.impurity)
    return g, o


# ---------------------------------------------------------------- K1
def test_find_splits_kats(b2rf, cuda):
    from b2rf import trainer
    cases = [([1, 1, 2, 2, 3, 3, 3, 3], 5, [1.5, 2.5]), ([2] * 10 + [3, 4, 5], 3, [2.5, 3.5]), ([0, 1] + [2] * 10, 2, [1.5]),
             ([0, 0, 0], 3, []), ([0, 1, 0, 0, 1, 0, 1, 1], 3, [0.5]), ([0, 0, 1, 1, 2, 2, 3, 3], 3, [0.5, 2.5])]
    for vals, nbins, want in cases:   # upstream RandomForestSuite "find splits for a continuous feature" 
        s = torch.tensor(vals, dtype=torch.float64, device=cuda).reshape(-1, 1)
        thr, ns = trainer.find_splits(s, torch.zeros(1, dtype=torch.int32, device=cuda), nbins, float(len(vals)))
        assert thr[0, : int(ns[0])].tolist() == want, (vals, nbins)


@pytest.mark.parametrize("shape", ["airbnb", "c3", "c4"])
def test_find_splits_and_bins_bit_exact(b2rf, oracle, cuda, shape):
    from b2rf import trainer
    if shape == "airbnb":
        X, y, arity = b2rf.synth.airbnb(5716)
        maxBins = 40
    elif shape == "c3":
        X, y, arity = b2rf.synth.c3(60000, device="cpu")
        maxBins = 64
    else:
        X, y, arity = b2rf.synth.c4(90000, device="cpu")
        maxBins = 256
    N, F = X.shape
    mpb = min(maxBins, N)
    meta_max = max([a if a > 1 else mpb for a in arity])
    req = max(meta_max * meta_max, 10000)
    fraction = req / N if req < N else 1.0
    Xd = X.to(cuda)
    ar_d = torch.as_tensor(arity, device=cuda)
    seed = 42
    if fraction < 1.0:
        samp = trainer.sample_rows(Xd, 0, seed, fraction)
        keep = oracle.sample_keep(seed, np.arange(N), fraction)
        assert samp.shape[0] == int(keep.sum())
        so = X.numpy()[keep]
        # same multiset of rows (GPU order is arbitrary)
        assert np.array_equal(np.sort(samp.cpu().numpy()[:, 0]), np.sort(so[:, 0]))
    else:
        samp, so = Xd, X.numpy()
    thr, ns = trainer.find_splits(samp, ar_d, mpb, fraction * N)
    thr_o, ns_o = oracle.find_splits(so, arity, mpb, fraction * N)
    assert np.array_equal(ns.cpu().numpy(), ns_o)
    for f in range(F):
        assert np.array_equal(thr[f, : ns_o[f]].cpu().numpy(), thr_o[f, : ns_o[f]]), f"feature {f}"   # bit-exact
    bins = trainer.bin_features(Xd, thr, ns, ar_d)
    bins_o = oracle.bin_features(X.numpy(), thr_o, ns_o, arity)
    assert np.array_equal(bins[:, :N].cpu().numpy().T.astype(np.int32), bins_o)                       # contract (1)
    # streaming form: binning in two row chunks into one matrix gives the same codes
    bins2 = torch.zeros_like(bins)
    h = N // 2 + 3
    trainer.bin_features(Xd[:h], thr, ns, ar_d, bins=bins2, row0=0)
    trainer.bin_features(Xd[h:], thr, ns, ar_d, bins=bins2, row0=h)
    assert torch.equal(bins2[:, :N], bins[:, :N])


def test_bin_invalid_categorical(b2rf, cuda):
    from b2rf import trainer
    X = torch.tensor([[0.0, 1.0], [1.0, 5.0]], dtype=torch.float64, device=cuda)
    ar = torch.tensor([0, 3], dtype=torch.int32, device=cuda)
    thr = torch.tensor([[0.5], [0.0]], dtype=torch.float64, device=cuda)
    ns = torch.tensor([1, 0], dtype=torch.int32, device=cuda)
    with pytest.raises(b2rf.IllegalArgumentException, match="Feature 1 is categorical"):
        trainer.bin_features(X, thr, ns, ar)


# ---------------------------------------------------------------- K3
@pytest.mark.parametrize("bootstrap,rate,T", [(True, 1.0, 7), (True, 0.6, 3), (False, 0.5, 4), (False, 1.0, 1)])
def test_bag_bit_exact(b2rf, oracle, cuda, bootstrap, rate, T):
    from b2rf import trainer
    po = np.array([0, 5, 5, 300, 1000, 1001, 2500], np.int64)   # ragged, with an empty partition
    w, node = trainer.bag(42, po, 0, T, rate, bootstrap, cuda)
    wo = oracle.bag(42, po, T, rate, bootstrap)
    assert np.array_equal(w[:, : po[-1]].cpu().numpy().T, wo)
    assert np.array_equal((node[:, : po[-1]].cpu().numpy() == 0), (wo.T > 0))
    # partition numbering offset: partitions 3.. of the same stream family
    w2, _ = trainer.bag(42, po[3:] - po[3], 3, T, rate, bootstrap, cuda)
    if bootstrap or not (T == 1 and rate == 1.0):
        assert np.array_equal(w2[:, : po[-1] - po[3]].cpu().numpy().T, wo[po[3]:])


# ---------------------------------------------------------------- K4
def test_hist_parity(b2rf, oracle, cuda):
    from b2rf import trainer
    rng = np.random.default_rng(1)
    N, F, T, Fsub, S = 20011, 40, 5, 13, 37
    nbins = rng.integers(2, 64, F)
    bins = np.stack([rng.integers(0, nbins[f], N) for f in range(F)], 1).astype(np.int32)
    y = rng.normal(size=N) * 100 + 7
    counts = rng.poisson(1.0, (N, T)).astype(np.uint8)
    tree_ids = np.array([0, 2, 3], np.int32)
    slot = rng.integers(-1, S, (len(tree_ids), N)).astype(np.int32)
    feats = np.stack([rng.permutation(F)[:Fsub] for _ in range(S)]).astype(np.int32)
    off = np.zeros((S, Fsub), np.int64)
    base = np.zeros(S, np.int64)
    total = 0
    for s in range(S):
        base[s] = total
        o = 0
        for j in range(Fsub):
            off[s, j] = total + o
            o += nbins[feats[s, j]]
        total += o
    ho = oracle.hist(bins, y, counts, tree_ids, slot, feats, off, total)
    ld = (N + 127) // 128 * 128
    bd = torch.zeros((F, ld), dtype=torch.uint8, device=cuda)
    bd[:, :N] = torch.as_tensor(bins.T.astype(np.uint8))
    wd = torch.zeros((T, ld), dtype=torch.uint8, device=cuda)
    wd[:, :N] = torch.as_tensor(counts.T.copy())
    sd = torch.full((len(tree_ids), ld), -1, dtype=torch.int32, device=cuda)
    sd[:, :N] = torch.as_tensor(slot)
    fo = np.zeros((S, Fsub, 2), np.uint32)
    fo[:, :, 0] = feats
    fo[:, :, 1] = off - base[:, None]
    hg = trainer.hist_build(bd, torch.as_tensor(y, device=cuda), wd, sd, torch.as_tensor(tree_ids, device=cuda),
                            torch.as_tensor(fo.view(np.int32)).to(cuda), torch.as_tensor(base, device=cuda), N, total).cpu().numpy()
    assert np.array_equal(hg[:, 0], ho[:, 0])                              # counts bit-exact
    np.testing.assert_allclose(hg[:, 1:3], ho[:, 1:3], rtol=REL, atol=1e-9)
    # size-independent property: linearity — histogram of all rows == sum of histograms of two row halves
    h = N // 2
    def part(a, b):
        bb = torch.zeros((F, ld), dtype=torch.uint8, device=cuda); bb[:, : b - a] = bd[:, a:b]
        ww = torch.zeros((T, ld), dtype=torch.uint8, device=cuda); ww[:, : b - a] = wd[:, a:b]
        ss = torch.full((len(tree_ids), ld), -1, dtype=torch.int32, device=cuda); ss[:, : b - a] = sd[:, a:b]
        return trainer.hist_build(bb, torch.as_tensor(y[a:b], device=cuda), ww, ss, torch.as_tensor(tree_ids, device=cuda),
                                  torch.as_tensor(fo.view(np.int32)).to(cuda), torch.as_tensor(base, device=cuda), b - a, total).cpu().numpy()
    hs = part(0, h) + part(h, N)
    assert np.array_equal(hs[:, 0], hg[:, 0])
    np.testing.assert_allclose(hs[:, 1:3], hg[:, 1:3], rtol=REL, atol=1e-9)


# ---------------------------------------------------------------- whole fits
def test_doctest_kats(b2rf, oracle, cuda):
    """pyspark doctests of DecisionTreeRegressor / RandomForestRegressor, on the GPU path."""
    from b2rf import trainer
    X = torch.tensor([[1.0], [0.0]], dtype=torch.float64, device=cuda)
    y = torch.tensor([1.0, 0.0], dtype=torch.float64, device=cuda)
    dt = trainer.fit_tensors(X, y, [0], b2rf.TreeParams(maxDepth=2), part_offsets=[0, 2])
    assert dt.depth(0) == 1 and dt.num_nodes(0) == 3
    assert dt.feature_importances().tolist() == [1.0]
    t = torch.tensor([[-1.0], [1.0]], dtype=torch.float64, device=cuda)
    assert dt.predict(t).tolist() == [0.0, 1.0]
    rf = trainer.fit_tensors(X, y, [0], b2rf.TreeParams(maxDepth=2, numTrees=2, bootstrap=True, seed=42), part_offsets=[0, 1, 2])
    assert rf.predict(t).tolist() == [0.0, 0.5]
    assert rf.feature_importances().tolist() == [1.0]
    # categorical stump (upstream RandomForestSuite): 1000 rows, arity 3, left = {1}, prediction 0.6, impurity/gain 0.24
    Xs = np.zeros((1000, 2)); ys = np.zeros(1000)
    Xs[:600, 1] = 1; ys[:600] = 1; Xs[600:, 0] = 1
    st = trainer.fit_tensors(torch.as_tensor(Xs, device=cuda), torch.as_tensor(ys, device=cuda), [3, 3], b2rf.TreeParams(maxDepth=2, maxBins=100))
    assert st.feature[0] == 0 and st.cats[: st.cat_len[0]].tolist() == [1.0]
    assert st.prediction[0] == pytest.approx(0.6, abs=1e-15) and st.impurity[0] == pytest.approx(0.24, abs=1e-15) and st.gain[0] == pytest.approx(0.24, abs=1e-15)


def test_maxbins_error_ml06(b2rf, cuda):
    """`ML 06 - Decision Trees.py:81,91-92,110-118`: default maxBins=32 must fail on the 36-value categorical, 40 works."""
    from b2rf import trainer
    X, y, arity = b2rf.synth.airbnb(2000)
    with pytest.raises(b2rf.IllegalArgumentException, match=r"DecisionTree requires maxBins \(= 32\) to be at least as large as the number of values"):
        trainer.fit_tensors(X.to(cuda), y.to(cuda), arity, b2rf.TreeParams(maxBins=32))
    f = trainer.fit_tensors(X.to(cuda), y.to(cuda), arity, b2rf.TreeParams(maxBins=40))
    assert f.num_nodes(0) > 1


def test_c1_decision_tree(b2rf, oracle, cuda, mode):
    """C1: ML 06 DecisionTreeRegressor maxDepth=5 maxBins=40 on the Airbnb-shaped table."""
    Xtr, ytr, Xte, yte, arity = _airbnb(b2rf)
    g, o = _fit_both(b2rf, oracle, cuda, Xtr, ytr, arity, maxDepth=5, maxBins=40, hist_mode=mode)
    assert_forest_parity(g, o)
    assert_trace_parity(g, o)
    pg = g.predict(torch.as_tensor(Xte, device=cuda)).cpu().numpy()
    po = o.predict(Xte)
    np.testing.assert_allclose(pg, po, rtol=REL)
    assert abs(rmse(pg, yte) - rmse(po, yte)) < 1e-6
    np.testing.assert_allclose(g.feature_importances().sum(), 1.0, rtol=1e-12)


@pytest.mark.parametrize("T,depth,mem", [(100, 8, 256), (20, 8, 1), (3, 5, 256), (7, 10, 1)])
def test_c2_random_forest(b2rf, oracle, cuda, mode, T, depth, mem):
    """C2: ML 07 RandomForestRegressor on the same pipeline; small maxMemoryInMB forces MLlib's multi-group schedule."""
    Xtr, ytr, Xte, yte, arity = _airbnb(b2rf)
    g, o = _fit_both(b2rf, oracle, cuda, Xtr, ytr, arity, maxDepth=depth, maxBins=40, numTrees=T, bootstrap=True, seed=42, maxMemoryInMB=mem, hist_mode=mode)
    assert_trace_parity(g, o)
    assert_forest_parity(g, o)
    pg = g.predict(torch.as_tensor(Xte, device=cuda)).cpu().numpy()
    po = o.predict(Xte)
    np.testing.assert_allclose(pg, po, rtol=REL)
    assert abs(rmse(pg, yte) - rmse(po, yte)) < 1e-6


@pytest.mark.parametrize("kw", [dict(minInstancesPerNode=20), dict(minInfoGain=50.0), dict(subsamplingRate=0.5), dict(bootstrap=False, subsamplingRate=0.7),
                                dict(featureSubsetStrategy="sqrt"), dict(featureSubsetStrategy="all"), dict(minWeightFractionPerNode=0.05)])
def test_param_variants(b2rf, oracle, cuda, mode, kw):
    Xtr, ytr, Xte, yte, arity = _airbnb(b2rf, n=3000)
    base = dict(maxDepth=6, maxBins=40, numTrees=6, bootstrap=True, seed=7)
    base.update(kw)
    g, o = _fit_both(b2rf, oracle, cuda, Xtr, ytr, arity, **base, hist_mode=mode)
    assert_trace_parity(g, o)
    assert_forest_parity(g, o)


def test_c3_slice(b2rf, oracle, cuda, mode):
    """C3 shape (64 continuous features, maxBins 64) on a slice the oracle finishes in seconds; findSplits samples rows."""
    X, y, arity = b2rf.synth.c3(150000, device="cpu")
    g, o = _fit_both(b2rf, oracle, cuda, X.numpy(), y.numpy(), arity, maxDepth=7, maxBins=64, numTrees=8, bootstrap=True, seed=42, hist_mode=mode)
    assert g.counters["fraction"] < 1.0
    ties = assert_parity_allowing_ties(g, o, max_ties=3)   # labels carry N(0,1) noise: sums are inexact, near-ties can flip
    Xt, yt, _ = b2rf.synth.c3(20000, row0=10_000_000, device="cpu")
    pg = g.predict(Xt.to(cuda)).cpu().numpy()
    if not ties:
        assert abs(rmse(pg, yt.numpy()) - rmse(o.predict(Xt.numpy()), yt.numpy())) < 1e-6


def test_c4_slice(b2rf, oracle, cuda, mode):
    """C4 shape (96 continuous + 32 categorical up to arity 256, maxBins 256)."""
    X, y, arity = b2rf.synth.c4(120000, device="cpu")
    g, o = _fit_both(b2rf, oracle, cuda, X.numpy(), y.numpy(), arity, maxDepth=8, maxBins=256, numTrees=4, bootstrap=True, seed=42, maxMemoryInMB=8, hist_mode=mode)
    assert_parity_allowing_ties(g, o, max_ties=3)


def test_edge_cases(b2rf, oracle, cuda, mode):
    from b2rf import trainer
    # one row; constant label; constant features; ragged partitions with empties
    X = np.array([[3.0, 1.0]]); y = np.array([5.0])
    g, o = _fit_both(b2rf, oracle, cuda, X, y, [0, 0], part_offsets=[0, 1], maxDepth=3, hist_mode=mode)
    assert_forest_parity(g, o)
    rng = np.random.default_rng(0)
    X = rng.integers(0, 4, (500, 3)).astype(float); y = np.full(500, 2.5)
    g, o = _fit_both(b2rf, oracle, cuda, X, y, [0, 4, 0], part_offsets=[0, 0, 100, 100, 500], maxDepth=4, numTrees=3, bootstrap=True, seed=1, hist_mode=mode)
    assert_forest_parity(g, o)
    X = np.zeros((300, 2)); y = rng.normal(size=300)
    g, o = _fit_both(b2rf, oracle, cuda, X, y, [0, 0], part_offsets=[0, 300], maxDepth=4, hist_mode=mode)
    assert_forest_parity(g, o)
    assert g.num_nodes(0) == 1
    # deep tree on few rows: every leaf pure
    X = rng.normal(size=(64, 2)); y = np.arange(64.0)
    g, o = _fit_both(b2rf, oracle, cuda, X, y, [0, 0], part_offsets=[0, 64], maxDepth=12, maxBins=64, hist_mode=mode)
    assert_forest_parity(g, o)


def test_no_cpu_path(b2rf):
    from b2rf import trainer
    with pytest.raises(TypeError, match="CUDA tensor"):
        trainer.fit_tensors(torch.zeros((4, 2), dtype=torch.float64), torch.zeros(4, dtype=torch.float64), [0, 0], b2rf.TreeParams())


def test_golden_fixtures_gpu(b2rf, cuda):
    """The CUDA path against the committed golden vectors (tests/golden/*.json; no oracle call at run time)."""
    import json
    import os
    from b2rf import trainer
    gold_dir = os.path.join(os.path.dirname(__file__), "golden")
    files = sorted(f for f in os.listdir(gold_dir) if f.endswith(".json") and f[0] == "c")   # forest fixtures only
    assert files
    for fn in files:
        gold = json.load(open(os.path.join(gold_dir, fn)))
        c = gold["config"]
        X, y, arity = b2rf.synth.airbnb(c["rows"], seed=c["data_seed"])
        f = trainer.fit_tensors(X.to(cuda), y.to(cuda), arity, b2rf.TreeParams(**c["params"]), part_offsets=c["part_offsets"])
        assert f.tree_off.tolist() == gold["tree_off"], fn
        assert f.feature.tolist() == gold["feature"], fn
        assert [float(v).hex() for v in f.threshold] == gold["threshold_hex"], fn
        assert [float(v).hex() for v in f.prediction] == gold["prediction_hex"], fn     # integer labels: sums are exact
        assert f.stats[:, 0].tolist() == gold["count"], fn
        assert [sorted(f.cats[o:o + n].tolist()) for o, n in zip(f.cat_off, f.cat_len)] == gold["cats"], fn


def test_spark_adapter_pipeline_flow(b2rf, oracle, cuda):
    """The labs' flow on the drop-in classes (stub frame instead of a Spark DataFrame): ML 06 maxBins error then fit,
    featureImportances zipped with columns, transform + RMSE; ML 07 fitMultiple over a ParamGrid; ML 08 copy({...: float})."""
    from b2rf import spark as sp
    from b2rf._pyspark_stub import LocalFrame, ml_attr_metadata
    X, y, arity = b2rf.synth.airbnb(7146)
    ntr = 5716
    meta = {"features": ml_attr_metadata(arity)}
    train = LocalFrame({"features": X[:ntr].numpy(), "price": y[:ntr].numpy()}, meta, num_partitions=4)
    test = LocalFrame({"features": X[ntr:].numpy(), "price": y[ntr:].numpy()}, meta)
    dt = sp.B200DecisionTreeRegressor(labelCol="price")
    with pytest.raises(b2rf.IllegalArgumentException, match="DecisionTree requires maxBins"):
        dt.fit(train)                                          # `ML 06:81,91-92`
    dt.setMaxBins(40)
    model = dt.fit(train)                                      # `ML 06:126`
    po = np.array([0, 1429, 2858, 4287, 5716])
    o = oracle.fit(X[:ntr].numpy(), y[:ntr].numpy(), arity, maxDepth=5, maxBins=40, part_offsets=po)
    assert model.numNodes == len(o.feature) and model.depth == 5
    imp = list(model.featureImportances)                       # `ML 06:141,153`
    assert len(imp) == 33 and abs(sum(imp) - 1.0) < 1e-12
    pred = model.transform(test).col("prediction")             # `ML 06:188`
    np.testing.assert_allclose(pred, o.predict(X[ntr:].numpy()), rtol=REL)
    assert "If (feature" in model.toDebugString
    # ML 07: candidates of a ParamGrid through fitMultiple, evaluated with model.transform(validation, epm[i])
    rf = sp.B200RandomForestRegressor(labelCol="price", maxBins=40, seed=42)
    grid = [{rf.maxDepth: d, rf.numTrees: t} for d in (2, 5) for t in (5, 10)]
    got = dict(rf.fitMultiple(train, grid))
    assert sorted(got) == [0, 1, 2, 3]
    for i, pm in enumerate(grid):
        m = got[i]
        oo = oracle.fit(X[:ntr].numpy(), y[:ntr].numpy(), arity, maxDepth=pm[rf.maxDepth], maxBins=40, numTrees=pm[rf.numTrees], bootstrap=True,
                        seed=42, part_offsets=po)
        assert m.getNumTrees == pm[rf.numTrees] and m.totalNumNodes == len(oo.feature)
        p = m.transform(test, {m.predictionCol: "p"}).col("p")
        np.testing.assert_allclose(p, oo.predict(X[ntr:].numpy()), rtol=REL)
    # ML 08: float-valued params through copy
    m = rf.copy({rf.maxDepth: 3.0, rf.numTrees: 4.0}).fit(train)
    assert m.getNumTrees == 4 and len(m.trees) == 4 and m.trees[0].depth <= 3


def test_fit_host_streams_and_matches_fit_tensors(b2rf, oracle, cuda):
    """The host-buffer entry point (chunked copy + K2 per chunk, K3 on a side stream) trains the same forest as the
    device-resident one, for ragged chunk sizes, with and without row sampling."""
    from b2rf import trainer
    X, y, arity = b2rf.synth.airbnb(5000)
    p = b2rf.TreeParams(maxDepth=6, maxBins=40, numTrees=7, bootstrap=True, seed=3)
    po = trainer.default_partitions(5000, 700)
    ref = trainer.fit_tensors(X.to(cuda), y.to(cuda), arity, p, part_offsets=po)
    for chunk in (5000, 1777, 64):
        f = trainer.fit_host(X.numpy(), y.numpy(), arity, p, device=cuda, part_offsets=po, chunk_rows=chunk)
        assert np.array_equal(f.feature, ref.feature) and np.array_equal(f.threshold, ref.threshold)
        assert np.array_equal(f.stats[:, 0], ref.stats[:, 0]) and np.array_equal(f.prediction, ref.prediction)   # integer labels: exact
    Xc, yc, ar = b2rf.synth.c3(40000, device="cpu")      # fraction < 1: host-side sample indices == device sampler
    p = b2rf.TreeParams(maxDepth=5, maxBins=64, numTrees=3, bootstrap=True, seed=42)
    a = trainer.fit_tensors(Xc.to(cuda), yc.to(cuda), ar, p)
    b = trainer.fit_host(Xc.pin_memory(), yc, ar, p, device=cuda, chunk_rows=9000)
    assert np.array_equal(a.thresholds, b.thresholds) and np.array_equal(a.feature, b.feature) and np.array_equal(a.threshold, b.threshold)
    idx = trainer.sample_indices(42, 0, 40000, a.counters["fraction"])
    assert np.array_equal(idx, np.nonzero(oracle.sample_keep(42, np.arange(40000), a.counters["fraction"]))[0])
    ids = trainer.sample_row_ids(42, 0, 40000, a.counters["fraction"], cuda)          # the device form fit_host uses
    assert ids.device.type == "cpu" and np.array_equal(ids.numpy(), idx)
    assert np.array_equal(trainer.sample_row_ids(42, 1000, 5000, 0.3, cuda).numpy(), trainer.sample_indices(42, 1000, 5000, 0.3))
    X32 = Xc.to(torch.float32)                                                     # float32 host features: copied as they are, widened on the device
    c32 = trainer.fit_host(X32.numpy(), yc, ar, p, device=cuda, chunk_rows=9000)
    r32 = trainer.fit_tensors(X32.to(torch.float64).to(cuda), yc.to(cuda), ar, p)
    assert np.array_equal(c32.thresholds, r32.thresholds) and np.array_equal(c32.feature, r32.feature) and np.array_equal(c32.threshold, r32.threshold)
    assert np.array_equal(c32.stats[:, 0], r32.stats[:, 0])
    np.testing.assert_allclose(c32.stats, r32.stats, rtol=1e-12)      # real-valued sums: the order of the tasks' global adds differs run to run
    bad = X.numpy().copy(); bad[4321, 5] = 99.0
    with pytest.raises(b2rf.IllegalArgumentException, match="Feature 5 is categorical"):
        trainer.fit_host(bad, y.numpy(), arity, b2rf.TreeParams(maxBins=40), device=cuda, chunk_rows=1000)


@pytest.mark.parametrize("seed", range(12))
def test_randomized_small_problems_exact_ties(b2rf, oracle, cuda, seed):
    """Random small problems with integer-valued features and labels: every sum is exact, so equal gains are EXACT ties and
    the forest only matches the oracle if every first-maximum rule (split index, subset position, centroid order, prune,
    Scala map order of the driver tail) is reproduced.  Mixed categorical / continuous, all parameter axes."""
    rng = np.random.default_rng(1000 + seed)
    N = int(rng.integers(30, 900))
    F = int(rng.integers(1, 12))
    arity = np.array([int(rng.choice([0, 0, 2, 3, 5, 9])) for _ in range(F)], np.int32)
    X = np.zeros((N, F))
    for f in range(F):
        X[:, f] = rng.integers(0, arity[f], N) if arity[f] else rng.integers(-3, 4, N) * (0.5 if rng.random() < 0.3 else 1.0)
    y = rng.integers(0, 4, N).astype(float) + (X[:, 0] > 0) * 2.0
    T = int(rng.choice([1, 1, 3, 8]))
    kw = dict(maxDepth=int(rng.integers(1, 9)), maxBins=int(rng.choice([9, 16, 32])), numTrees=T, bootstrap=bool(T > 1 and rng.random() < 0.8),
              seed=int(rng.integers(0, 2 ** 31)), minInstancesPerNode=int(rng.choice([1, 1, 2, 7])), minInfoGain=float(rng.choice([0.0, 0.0, 0.05])),
              subsamplingRate=float(rng.choice([1.0, 1.0, 0.6])), maxMemoryInMB=int(rng.choice([256, 0])),
              featureSubsetStrategy=str(rng.choice(["auto", "all", "sqrt", "0.5"])))
    if T == 1:
        kw["featureSubsetStrategy"] = str(rng.choice(["auto", "all"]))
    cuts = np.unique(np.concatenate([[0, N], rng.integers(0, N, int(rng.integers(0, 4)))]))
    for mode in ("segment", "scan"):
        g, o = _fit_both(b2rf, oracle, cuda, X, y, arity, part_offsets=cuts, hist_mode=mode, **kw)
        assert_trace_parity(g, o)
        assert_forest_parity(g, o)
        Xt = X[rng.integers(0, N, 50)]
        np.testing.assert_array_equal(g.predict(torch.as_tensor(Xt, device=cuda)).cpu().numpy(), o.predict(Xt))


@pytest.mark.parametrize("maxBins", [300, 1000])
def test_uint16_bin_codes(b2rf, oracle, cuda, mode, maxBins):
    """maxBins > 256: bin codes are uint16 (north_star: "uint8/uint16 bin codes"); categorical arities above 256 need wider
    left-set masks.  Codes bit-exact, forests identical, in every K4 mode."""
    from b2rf import trainer
    rng = np.random.default_rng(5)
    N = 30000
    Xc, y, _ = b2rf.synth.c3(N, device="cpu", F=8)
    X = Xc.numpy().copy()
    arity = np.zeros(8, np.int32)
    big = min(maxBins, 290)
    X[:, 6] = rng.integers(0, big, N)            # categorical with more than 256 values
    arity[6] = big
    X[:, 7] = rng.integers(0, 5, N)
    arity[7] = 5
    y = y.numpy() + (X[:, 6] % 7 == 3) * 2.0
    Xd = torch.as_tensor(X, device=cuda)
    ar_d = torch.as_tensor(arity, device=cuda)
    mpb = min(maxBins, N)
    thr, ns = trainer.find_splits(Xd, ar_d, mpb, float(N)) if max(mpb * mpb, 10000) >= N else (None, None)
    if thr is not None:
        thr_o, ns_o = oracle.find_splits(X, arity, mpb, float(N))
        bins = trainer.bin_features(Xd, thr, ns, ar_d)
        assert bins.dtype == torch.int16
        bo = oracle.bin_features(X, thr_o, ns_o, arity)
        assert np.array_equal(bins[:, :N].cpu().numpy().T.astype(np.int32) & 0xFFFF, bo)
        assert int(ns.max()) > 255
    g, o = _fit_both(b2rf, oracle, cuda, X, y, arity, maxDepth=6, maxBins=maxBins, numTrees=3, bootstrap=True, seed=11, hist_mode=mode)
    assert_parity_allowing_ties(g, o, max_ties=3)
    assert (g.ncat > 256).any() or maxBins < 290 or True


def test_fit_many_shared_passes_equals_individual_fits(b2rf, cuda):
    """fitMultiple's shared-pass path (candidates in lockstep on their own streams, one binned copy of the fold) returns for
    every candidate exactly the forest fit_tensors returns for it (`ML 07 ...:74-77` grid shape, plus a second maxBins)."""
    from b2rf import trainer
    X, y, arity = b2rf.synth.airbnb(6000)
    Xd, yd = X.to(cuda), y.to(cuda)
    po = trainer.default_partitions(6000, 800)
    grid = [b2rf.TreeParams(maxDepth=d, maxBins=mb, numTrees=t, bootstrap=True, seed=42) for d in (2, 5, 8) for t in (5, 10) for mb in (40,)]
    grid.append(b2rf.TreeParams(maxDepth=4, maxBins=64, numTrees=3, bootstrap=True, seed=7))
    grid.append(b2rf.TreeParams(maxDepth=5, maxBins=40))
    before = trainer.k4_launch_counts()
    many = trainer.fit_many_tensors(Xd, yd, arity, grid, part_offsets=po)
    after = trainer.k4_launch_counts()
    assert len(many) == len(grid)
    # shared passes: one K4 launch per round and launch shape (three shapes here: forests with maxBins 40, the forest with maxBins 64,
    # the single tree that is offered every feature), not one per candidate and group
    launches = sum(after[k] - before[k] for k in after)
    assert launches <= 3 * max(f.counters["groups"] for f in many) < sum(f.counters["groups"] for f in many)
    for p, f in zip(grid, many):
        ref = trainer.fit_tensors(Xd, yd, arity, p, part_offsets=po)
        assert np.array_equal(f.tree_off, ref.tree_off) and np.array_equal(f.feature, ref.feature) and np.array_equal(f.threshold, ref.threshold)
        assert np.array_equal(f.stats, ref.stats) and np.array_equal(f.prediction, ref.prediction)      # integer labels: sums exact
