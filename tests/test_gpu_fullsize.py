"""Full-size checks (BASELINE config 3 shape: 10M rows x 64 features on one GPU) through properties that do not need the
oracle, which would take minutes at this size: sortedness / idempotence of the binning, conservation of counts and sums
through every split of every tree (a checksum of checksums), bagging statistics, and agreement between the default
(segment) and scan-mode trainers."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

N, F = 10_000_000, 64


@pytest.fixture(scope="module")
def data(cuda):
    import b2rf
    X, y, arity = b2rf.synth.c3(N, device=cuda)
    return X, y, arity


def test_binning_sorted_and_idempotent_full_size(b2rf, cuda, data):
    from b2rf import trainer
    X, y, arity = data
    ar_d = torch.as_tensor(arity, device=cuda)
    frac = 10000 / N
    samp = trainer.sample_rows(X, 0, 42, frac)
    assert abs(samp.shape[0] - 10000) < 600                                 # Bernoulli(1e-3) over 1e7 rows: sd = 100
    thr, ns = trainer.find_splits(samp, ar_d, 64, frac * N)
    assert int(ns.max()) <= 63 and int(ns.min()) >= 1
    for f in range(F):
        t = thr[f, : int(ns[f])]
        assert bool((t[1:] > t[:-1]).all()), f"thresholds of feature {f} not strictly increasing"
    bins = trainer.bin_features(X, thr, ns, ar_d)
    # sortedness: the code is a non-decreasing function of the value (check along the value-sorted order of 3 features)
    for f in (0, 40, 63):
        order = torch.argsort(X[:, f])
        b = bins[f, :N][order].to(torch.int16)
        assert bool((b[1:] >= b[:-1]).all())
        # definition: code = #thresholds strictly below x
        ref = torch.searchsorted(thr[f, : int(ns[f])].contiguous(), X[:, f].contiguous(), right=False)
        assert torch.equal(ref.to(torch.uint8), bins[f, :N])
    # idempotence / streaming: binning again in uneven chunks gives the same matrix
    bins2 = torch.zeros_like(bins)
    for r0, r1 in ((0, 3_333_333), (3_333_333, 9_999_999), (9_999_999, N)):
        trainer.bin_features(X[r0:r1], thr, ns, ar_d, bins=bins2, row0=r0)
    assert torch.equal(bins2[:, :N], bins[:, :N])


def test_fit_conservation_laws_full_size(b2rf, cuda, data):
    from b2rf import trainer
    X, y, arity = data
    T = 8
    p = b2rf.TreeParams(maxDepth=6, maxBins=64, numTrees=T, bootstrap=True, seed=42)
    po = trainer.default_partitions(N, 4096)
    f = trainer.fit_tensors(X, y, arity, p, part_offsets=po)
    w, _ = trainer.bag(42, po, 0, T, 1.0, True, cuda, want_node=False)
    wsum = w[:, :N].to(torch.int64).sum(1).cpu().numpy()
    assert abs(wsum.mean() / N - 1.0) < 5e-3                                # Poisson(1) mean
    wy = (w[:, :N].to(torch.float64) * y[None, :]).sum(1).cpu().numpy()
    for t in range(T):
        s = f.tree_slice(t)
        feat, left, right, st = f.feature[s], f.left[s], f.right[s], f.stats[s]
        assert st[0, 0] == wsum[t]                                           # root count == total bag weight of the tree (exact)
        np.testing.assert_allclose(st[0, 1], wy[t], rtol=1e-9)              # root label sum
        internal = np.nonzero(feat >= 0)[0]
        assert len(internal) > 20
        # every split conserves count exactly and the two sums to 1e-9 (children's stats come from the parent's histogram)
        assert np.array_equal(st[internal, 0], st[left[internal], 0] + st[right[internal], 0])
        np.testing.assert_allclose(st[internal, 1], st[left[internal], 1] + st[right[internal], 1], rtol=1e-9, atol=1e-6)
        np.testing.assert_allclose(st[internal, 2], st[left[internal], 2] + st[right[internal], 2], rtol=1e-9, atol=1e-6)
        leaves = feat < 0
        assert st[leaves, 0].sum() == wsum[t]                                # checksum of checksums: leaves partition the bag
        np.testing.assert_allclose(f.prediction[s][leaves], st[leaves, 1] / st[leaves, 0], rtol=1e-12)
        assert (f.gain[s][internal] > 0).all()
    # the scan-mode trainer (column scan + global RED, different code path end to end) grows the same forest
    g = trainer.fit_tensors(X, y, arity, p, part_offsets=po, hist_mode="scan")
    same = np.array_equal(f.tree_off, g.tree_off) and np.array_equal(f.feature, g.feature) and np.array_equal(f.threshold, g.threshold)
    if same:
        assert np.array_equal(f.stats[:, 0], g.stats[:, 0])
        np.testing.assert_allclose(f.prediction, g.prediction, rtol=1e-9)
    else:   # only a documented gain tie may separate them: identical counts at the first differing node's parent level
        i = int(np.nonzero(f.feature[: min(len(f.feature), len(g.feature))] != g.feature[: min(len(f.feature), len(g.feature))])[0][0])
        assert abs(f.gain[i] - g.gain[i]) <= 1e-12 * max(1.0, abs(g.gain[i])), "forests differ beyond a gain tie"
    # prediction on 1M training rows is the mean of the trees' leaf means: bounded by the label range, finite
    pred = f.predict(X[:1_000_000])
    assert bool(torch.isfinite(pred).all()) and float(pred.min()) >= float(y.min()) and float(pred.max()) <= float(y.max())
