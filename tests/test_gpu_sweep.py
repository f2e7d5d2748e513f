"""Seeded sweep of fit shapes against the oracle: feature counts on both sides of the 16 / 32 / 64 / 128 boundaries (feature-group width,
code-row width, register vs TMA staging), bins, label kinds (real / integer / 0-1), depth, memory budget (group cuts), partition layout
and featureSubsetStrategy drawn from a seeded generator.  Every case must equal the oracle's forest up to documented gain ties; cases
without ties must match bit for bit in structure and to 1e-9 in sums (helpers.assert_parity_allowing_ties).
B2RF_SWEEP_CASES sets the number of cases (default 24; profiles/r02_parity_sweep.md records a 600-case run)."""
import os

import numpy as np
import pytest
import torch

from helpers import assert_parity_allowing_ties, assert_classifier_parity
from test_gpu_parity import _fit_both

pytestmark = pytest.mark.gpu
N_CASES = int(os.environ.get("B2RF_SWEEP_CASES", "24"))


def _case(i):
    rng = np.random.default_rng(1000 + i)
    F = int(rng.choice([3, 8, 9, 15, 16, 17, 24, 31, 32, 33, 48, 63, 64, 65, 96, 127, 128, 129, 140]))
    n = int(rng.choice([600, 2500, 5000, 9000, 14000]))
    bins = int(rng.choice([8, 16, 32, 64, 100, 256]))
    depth = int(rng.integers(2, 9))
    T = int(rng.choice([1, 1, 2, 3, 5]))
    strat = str(rng.choice(["all", "auto", "sqrt", "onethird", "log2", "0.4"])) if T > 1 else "all"
    labels = str(rng.choice(["real", "integer", "binary", "tiny", "shifted"]))
    impurity = str(rng.choice(["gini", "entropy"])) if labels == "binary" and rng.random() < 0.6 else "variance"
    mem = int(rng.choice([256, 256, 1]))
    parts = int(rng.choice([1, 3, 7]))
    mi = int(rng.choice([1, 1, 5]))
    return dict(F=F, n=n, bins=bins, depth=depth, T=T, strat=strat, labels=labels, impurity=impurity, mem=mem, parts=parts, mi=mi, seed=int(rng.integers(1, 10_000)))


@pytest.mark.parametrize("i", range(N_CASES))
def test_sweep_case(b2rf, oracle, cuda, i):
    c = _case(i)
    X, y, arity = b2rf.synth.c3(c["n"], device="cpu", F=c["F"], seed=c["seed"])
    rng = np.random.default_rng(c["seed"])
    if c["labels"] == "integer":
        y = torch.round(y * 4.0)
    elif c["labels"] == "binary":
        y = (y > y.median()).to(torch.float64)
    elif c["labels"] == "tiny":
        y = y * 1e-140
    elif c["labels"] == "shifted":
        y = y + 20.0       # a mean away from zero.  Larger shifts make the variance formula itself ill-conditioned: with +300 three of 400
                           # cases differ from the oracle by near-ties of 5e-12 relative gain (beyond the helpers' 1e-12 tie criterion), with
                           # +1e6 impurities differ by 1e-4 — MLlib's own fp64 sums move that much between partition layouts
    # a few categorical columns (small arities) when the bin budget allows
    Xn = X.numpy().copy()
    arity = list(arity)
    for f in rng.choice(c["F"], size=min(3, c["F"] // 3), replace=False):
        k = int(rng.integers(2, min(c["bins"], 12) + 1))
        Xn[:, f] = rng.integers(0, k, size=c["n"]).astype(np.float64)
        arity[int(f)] = k
    cuts = np.sort(rng.choice(np.arange(1, c["n"]), size=c["parts"] - 1, replace=False)) if c["parts"] > 1 else []
    po = np.concatenate([[0], cuts, [c["n"]]]).astype(np.int64)
    kw = dict(maxDepth=c["depth"], maxBins=c["bins"], numTrees=c["T"], bootstrap=c["T"] > 1, seed=c["seed"], featureSubsetStrategy=c["strat"],
              maxMemoryInMB=c["mem"], minInstancesPerNode=c["mi"], impurity=c["impurity"])
    g, o = _fit_both(b2rf, oracle, cuda, Xn, y.numpy(), arity, part_offsets=po, **kw)
    print("case", i, c, "nodes", int(g.num_nodes()))
    if c["impurity"] == "variance":
        sub = c["T"] > 1 and c["strat"] != "all"
        # many features over few rows make identical partitions (genuine gain ties) common: each tie is verified to be one by the helper
        assert_parity_allowing_ties(g, o, max_ties=6 + c["F"] // 4, subsampling=sub, min_fraction=0.0)
    else:
        assert_classifier_parity(g, o, exact=(c["impurity"] == "gini"))


N_MC = int(os.environ.get("B2RF_SWEEP_MC_CASES", "6"))


@pytest.mark.parametrize("i", range(N_MC))
def test_sweep_multiclass_case(b2rf, cuda, i):
    """The same for multiclass classification (class-count histograms: staged kernel for 64- / 128-byte code rows, gather kernel
    otherwise; unordered and impurity-ordered categorical splits) against oracle/multiclass.py."""
    from oracle import multiclass as M
    from test_gpu_multiclass import _parity
    rng = np.random.default_rng(5000 + i)
    F = int(rng.choice([5, 9, 16, 17, 33, 50, 64, 70, 100, 130]))
    n = int(rng.choice([800, 3000, 6000]))
    C = int(rng.choice([3, 4, 5, 7]))
    bins = int(rng.choice([16, 32, 64]))
    T = int(rng.choice([1, 3, 4]))
    kw = dict(maxDepth=int(rng.integers(2, 6)), maxBins=bins, numTrees=T, bootstrap=T > 1, seed=int(rng.integers(1, 999)),
              featureSubsetStrategy=str(rng.choice(["all", "auto", "onethird"])) if T > 1 else "all", maxMemoryInMB=int(rng.choice([256, 1])))
    impurity = str(rng.choice(["gini", "gini", "entropy"]))
    X, y, arity = b2rf.synth.c3(n, device="cpu", F=F, seed=int(rng.integers(1, 9999)))
    Xn = X.numpy().copy()
    arity = list(arity)
    for f in rng.choice(F, size=min(3, F // 3), replace=False):
        k = int(rng.integers(2, min(bins, 9) + 1))          # small arities: unordered splits for some, ordered for others
        Xn[:, f] = rng.integers(0, k, size=n).astype(np.float64)
        arity[int(f)] = k
    yn = y.numpy()
    lab = np.digitize(yn, np.quantile(yn, np.linspace(0, 1, C + 1)[1:-1])).astype(np.float64)
    po = np.array([0, n // 3, n // 2, n], np.int64)
    m = M.fit(Xn, lab, arity, C, part_offsets=po, impurity=impurity, **kw)
    g = b2rf.trainer.fit_tensors(torch.as_tensor(Xn, device=cuda), torch.as_tensor(lab, device=cuda), arity,
                                 b2rf.TreeParams(impurity=impurity, numClasses=C, **kw), part_offsets=po, keep_trace=True)
    print("multiclass case", i, dict(F=F, n=n, C=C, bins=bins, impurity=impurity, **kw), "nodes", int(g.num_nodes()))
    _parity(g, m)
    mi = m.impurity >= 0
    if impurity == "gini":
        assert np.array_equal(g.impurity[mi], m.impurity[mi]) and np.array_equal(g.gain, m.gain)
    else:
        np.testing.assert_allclose(g.impurity[mi], m.impurity[mi], rtol=1e-12, atol=1e-15)
        np.testing.assert_allclose(g.gain, m.gain, rtol=1e-9, atol=1e-13)
