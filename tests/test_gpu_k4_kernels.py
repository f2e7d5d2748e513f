"""K4 (DTStatsAggregator.update) kernel-by-kernel parity: every histogram implementation of the default segment mode is forced
in turn and compared with the CPU oracle at the CONFIGS' depth (BASELINE.json configs[2] / [3]: depth 10 / 12), on row counts
the oracle finishes in seconds, so the deep levels run the small-node paths (fewer updates than cells -> direct global atomics,
single-row chunks) as well as the shared-memory ones.  `b2rf_k4_launch_counts` proves which kernel ran.

Also contract (1) "histogram counts bit-exact" cell by cell on the DEFAULT path: the reduced histogram of every node group of a
fit (b2rf_trainer_group_info + the caller's d_hist) against the oracle's DTStatsAggregator restatement."""
import numpy as np
import pytest
import torch

from helpers import REL, assert_forest_parity, assert_parity_allowing_ties, assert_trace_parity
from test_gpu_parity import _fit_both

pytestmark = pytest.mark.gpu

KERNELS = ["limb", "fp64"]


@pytest.fixture(params=KERNELS)
def kernel(request, b2rf):
    """limb: native 32-bit shared-memory adds on fixed-point limbs (default).  fp64: the compare-and-swap kernels
    (bucket_hist32_kernel for <= 32 features per node, uint8 codes; the general kernel otherwise)."""
    from b2rf import trainer
    trainer.set_option("k4_fp64", 1 if request.param == "fp64" else 0)
    yield request.param
    trainer.set_option("k4_fp64", 0)


def _ran(b2rf, before, kernel, general=False):
    from b2rf import trainer
    after = trainer.k4_launch_counts()
    d = {k: after[k] - before[k] for k in after}
    want = "limb" if kernel == "limb" else ("fp64_general" if general else "fp64_32")
    assert d[want] > 0, f"expected the {want} kernel to run, launches: {d}"
    assert sum(v for k, v in d.items() if k != want) == 0, f"unexpected kernels ran: {d}"


@pytest.mark.parametrize("mem", [256, 1])
def test_c3_config_depth(b2rf, oracle, cuda, kernel, mem):
    """C3's shape and depth: F = 64 -> 22 features per node, maxBins 64, depth 10; 40k rows x 5 trees, so level-9 nodes hold a few
    dozen rows (direct branch) while the top levels fill whole 4096-row chunks.  maxMemoryInMB = 1 cuts MLlib's node groups."""
    from b2rf import trainer
    X, y, arity = b2rf.synth.c3(40000, device="cpu")
    before = trainer.k4_launch_counts()
    g, o = _fit_both(b2rf, oracle, cuda, X.numpy(), y.numpy(), arity, maxDepth=10, maxBins=64, numTrees=5, bootstrap=True, seed=42, maxMemoryInMB=mem)
    _ran(b2rf, before, kernel)
    assert g.depth(0) == 10
    assert_parity_allowing_ties(g, o, max_ties=4)


def test_c3_config_depth_all_features(b2rf, oracle, cuda, kernel):
    """The same depth with every feature offered at every node (two 32-feature groups per slot): no feature-subset draws, so a gain
    tie — two splits that cut a small node's rows identically — only changes its own subtree and everything else must match."""
    from b2rf import trainer
    X, y, arity = b2rf.synth.c3(30000, device="cpu")
    before = trainer.k4_launch_counts()
    g, o = _fit_both(b2rf, oracle, cuda, X.numpy(), y.numpy(), arity, maxDepth=10, maxBins=64, numTrees=2, bootstrap=True, seed=3,
                     featureSubsetStrategy="all")
    _ran(b2rf, before, kernel, general=True)
    assert_parity_allowing_ties(g, o, max_ties=12, subsampling=False)


@pytest.mark.parametrize("fsub", ["16", "17", "32", "8", "5"])
def test_feature_width_edges(b2rf, oracle, cuda, kernel, fsub):
    """Lane == feature layouts at their edges: 17 and 32 features per node (one row per warp instruction), 16 / 8 / 5 (two / four
    rows per instruction in the group forms)."""
    from b2rf import trainer
    X, y, arity = b2rf.synth.c3(30000, device="cpu")
    before = trainer.k4_launch_counts()
    g, o = _fit_both(b2rf, oracle, cuda, X.numpy(), y.numpy(), arity, maxDepth=10, maxBins=64, numTrees=4, bootstrap=True, seed=7, featureSubsetStrategy=fsub)
    _ran(b2rf, before, kernel, general=int(fsub) <= 16)
    assert_parity_allowing_ties(g, o, max_ties=4)


def test_c4_config_depth(b2rf, oracle, cuda, kernel):
    """C4's shape and depth: 96 continuous + 32 categorical features (arity up to 256), maxBins 256, 43 features per node, depth 12:
    feature groups of 16 (plane-split lane mapping: one row per warp instruction, even / odd planes on the two half-warps), one 768-thread
    CTA per SM."""
    from b2rf import trainer
    X, y, arity = b2rf.synth.c4(60000, device="cpu")
    before = trainer.k4_launch_counts()
    g, o = _fit_both(b2rf, oracle, cuda, X.numpy(), y.numpy(), arity, maxDepth=12, maxBins=256, numTrees=3, bootstrap=True, seed=42)
    _ran(b2rf, before, kernel, general=True)
    assert g.depth(0) == 12
    assert_parity_allowing_ties(g, o, max_ties=4)


@pytest.mark.parametrize("labels", ["negative", "wide", "tiny", "huge", "integer", "mixed_huge"])
def test_label_scales(b2rf, oracle, cuda, kernel, labels):
    """The limb kernel scales every task by the largest label magnitude of its own rows.  Labels of both signs, a dynamic range of
    many decades, magnitudes far from 1, integers (sums exact: forests bit-identical), and magnitudes whose square leaves the fp64
    range (those tasks take the fp64 path) must all keep contract (2): sums within 1e-9 relative of the oracle's."""
    rng = np.random.default_rng(3)
    N = 20000
    X, _, arity = b2rf.synth.c3(N, device="cpu", F=12)
    X = X.numpy()
    step = (X[:, 0] > 0.3) * 1.0 + (X[:, 3] > -0.5) * 2.0 + (X[:, 7] > 1.0) * 4.0
    if labels == "negative":
        y = -50.0 + step * 13.0 + rng.normal(size=N)
    elif labels == "wide":
        y = np.exp(rng.normal(size=N) * 6.0) * (1.0 + step)          # ~ 20 decades between the smallest and the largest label
    elif labels == "tiny":
        y = (step + rng.normal(size=N) * 0.1) * 1e-200
    elif labels == "huge":
        y = (step + rng.normal(size=N) * 0.1) * 1e140                # y*y ~ 1e280: still finite
    elif labels == "mixed_huge":
        y = step + rng.normal(size=N) * 0.1
        y[::97] *= 1e160                                               # y*y overflows in some tasks only
    else:
        y = np.round(step * 3 + rng.integers(0, 5, N)).astype(np.float64)
    kw = dict(maxDepth=8, maxBins=32, numTrees=3, bootstrap=True, seed=5)
    g, o = _fit_both(b2rf, oracle, cuda, X, y, arity, **kw)
    if labels == "integer":
        assert_trace_parity(g, o)
        assert_forest_parity(g, o)
        assert np.array_equal(g.stats[:, 1:3], o.stats[:, 1:3])        # exact sums
    elif labels == "mixed_huge":
        # sum w*y*y overflows to +inf on both sides; structure is then decided by NaN gains identically or not at all: compare what is finite
        assert np.array_equal(g.stats[:, 0][: 1], o.stats[:, 0][: 1])
        np.testing.assert_allclose(g.stats[0, 1], o.stats[0, 1], rtol=REL)
    else:
        assert_parity_allowing_ties(g, o, max_ties=4)


def test_group_histograms_cell_by_cell(b2rf, oracle, cuda, kernel):
    """Contract (1) on the default path: every cell of every node group's histogram (segment mode, shared-memory kernels) equals
    DTStatsAggregator's: counts bit-exact, sums to 1e-9 relative.  The slot table comes from b2rf_trainer_group_info, row -> node
    membership from the oracle-side replay of the fit's own decision trace (all-continuous data: row goes left iff bin <= split)."""
    from b2rf import trainer
    N, T, depth = 30000, 4, 9
    X, y, arity = b2rf.synth.c3(N, device="cpu")
    X, y = X.numpy(), y.numpy()
    po = trainer.default_partitions(N, 1024)
    groups = []
    p = b2rf.TreeParams(maxDepth=depth, maxBins=64, numTrees=T, bootstrap=True, seed=42, maxMemoryInMB=4)
    g = trainer.fit_tensors(torch.as_tensor(X, device=cuda), torch.as_tensor(y, device=cuda), arity, p, part_offsets=po, keep_trace=True,
                            hist_hook=lambda info, h: groups.append((info, h.cpu().numpy().copy())))
    assert len(groups) == g.counters["groups"] and len(groups) > 10
    # oracle-side inputs: same sample -> same thresholds -> same bin codes; same bag counts
    mpb = 64
    fraction = max(mpb * mpb, 10000) / N
    keep = oracle.sample_keep(42, np.arange(N), fraction)
    thr_o, ns_o = oracle.find_splits(X[keep], arity, mpb, fraction * N)
    assert np.array_equal(thr_o, g.thresholds)
    bins = oracle.bin_features(X, thr_o, ns_o, arity)
    counts = oracle.bag(42, po, T, 1.0, True)
    # heap node id of every row per tree, level by level, replaying the trace; rows stop at nodes that were not split
    split = [dict() for _ in range(T)]
    for r, d in zip(g.trace_i, g.trace_d):
        if d[0] > 0 and (int(r[2]).bit_length() - 1) < depth:        # gain > 0 and not at maxDepth: the node was split
            split[int(r[1])][int(r[2])] = (int(r[3]), int(r[4]))
    node = np.ones((T, N), np.int64)
    per_level = [node.copy()]
    for _ in range(depth):
        nxt = node.copy()
        for t in range(T):
            for nid, (f, s) in split[t].items():
                m = node[t] == nid
                if m.any():
                    nxt[t, m] = np.where(bins[m, f] <= s, 2 * nid, 2 * nid + 1)
        node = nxt
        per_level.append(node.copy())
    checked = 0
    for info, hg in groups:
        S = len(info["tree"])
        tree_ids = np.unique(info["tree"]).astype(np.int32)
        slot = np.full((len(tree_ids), N), -1, np.int32)
        for s in range(S):
            t, nid = int(info["tree"][s]), int(info["node_id"][s])
            lvl = nid.bit_length() - 1
            slot[np.searchsorted(tree_ids, t), per_level[lvl][t] == nid] = s
        feats = info["slot_fo"][:, :, 0].astype(np.int32)
        off = info["slot_fo"][:, :, 1].astype(np.int64) + info["slot_base"][:, None]
        ho = oracle.hist(bins, y, counts, tree_ids, slot, feats, off, hg.shape[0])
        assert np.array_equal(hg[:, 0], ho[:, 0]), "histogram counts differ"            # contract (1)
        np.testing.assert_allclose(hg[:, 1:3], ho[:, 1:3], rtol=REL, atol=1e-9)          # contract (2)
        checked += hg.shape[0]
    assert checked > 100000


def test_labels_gathered_when_not_carried(b2rf, oracle, cuda):
    """carry_labels = 0 (what a device short of memory falls back to): the row lists hold no labels, K4 gathers them, and the limb
    kernel does not apply; the forest is the same."""
    from b2rf import trainer
    X, y, arity = b2rf.synth.c3(25000, device="cpu")
    trainer.set_option("carry_labels", 0)
    try:
        before = trainer.k4_launch_counts()
        g, o = _fit_both(b2rf, oracle, cuda, X.numpy(), y.numpy(), arity, maxDepth=8, maxBins=64, numTrees=3, bootstrap=True, seed=1)
        after = trainer.k4_launch_counts()
        assert after["limb"] == before["limb"] and after["fp64_32"] > before["fp64_32"]
    finally:
        trainer.set_option("carry_labels", 1)
    assert_parity_allowing_ties(g, o, max_ties=4)


def test_constant_nominal_column(b2rf, oracle, cuda):
    """A nominal feature with ONE category (a constant StringIndexer output).  MLlib 3.1.2 only overwrites numBins for categorical
    features with more than one category but keeps the feature in featureArity: it stays categorical (no continuous splits are
    sampled, values must be 0), keeps maxPossibleBins bins in the group-memory accounting, and never yields a valid split.  Trainer
    and oracle must agree on that, including under a memory limit that cuts groups (bins of the feature count towards the limit)."""
    rng = np.random.default_rng(11)
    N = 4000
    X, y, arity = b2rf.synth.airbnb(N)
    X = X.numpy().copy()
    arity = arity.copy()
    X[:, 2] = 0.0
    arity[2] = 1
    for mem in (256, 1):
        g, o = _fit_both(b2rf, oracle, cuda, X, y.numpy(), arity, maxDepth=6, maxBins=40, numTrees=4, bootstrap=True, seed=9, maxMemoryInMB=mem)
        assert_trace_parity(g, o)
        assert_forest_parity(g, o)
        assert not (g.feature == 2).any()
    bad = X.copy(); bad[17, 2] = 1.0
    from b2rf import trainer
    with pytest.raises(b2rf.IllegalArgumentException, match="Feature 2 is categorical"):
        trainer.fit_tensors(torch.as_tensor(bad, device=cuda), y.to(cuda), arity, b2rf.TreeParams(maxBins=40))


@pytest.mark.parametrize("tma", [0, 1])
def test_integer_labels_64_byte_rows(b2rf, oracle, cuda, tma):
    """Integer labels on 64-feature rows: the narrow (7-plane) limb form, with register-staged code rows (the default for 64- and
    128-byte rows) and with TMA bulk copies.  Sums are exact, so the forest must equal the oracle's bit for bit (all features offered:
    no subset draws, a tie stays local)."""
    from b2rf import trainer
    X, y, arity = b2rf.synth.c3(30000, device="cpu")
    yi = torch.round(y * 3.0)
    trainer.set_option("k4_tma", tma)
    try:
        before = trainer.k4_launch_counts()
        g, o = _fit_both(b2rf, oracle, cuda, X.numpy(), yi.numpy(), arity, maxDepth=9, maxBins=64, numTrees=3, bootstrap=True, seed=4)
        assert trainer.k4_launch_counts()["limb"] > before["limb"]
    finally:
        trainer.set_option("k4_tma", 0)
    k = first_tie = None
    from helpers import first_tie_in_order
    k = first_tie_in_order(g, o)
    upto = len(o.trace_i) if k is None else k
    assert upto > 0.5 * len(o.trace_i)
    assert np.array_equal(g.trace_d[:upto, 2:5], o.trace_d[:upto, 2:5])      # counts AND both sums bit-identical
    if k is None:
        assert_forest_parity(g, o)
        assert np.array_equal(g.stats[:, 1:3], o.stats[:, 1:3])


def test_tma_staging_real_labels(b2rf, oracle, cuda):
    """The TMA-staged form of the wide limb kernel (one bulk async copy per code row, mbarrier ring) on 64-byte rows."""
    from b2rf import trainer
    X, y, arity = b2rf.synth.c3(30000, device="cpu")
    trainer.set_option("k4_tma", 1)
    try:
        g, o = _fit_both(b2rf, oracle, cuda, X.numpy(), y.numpy(), arity, maxDepth=9, maxBins=64, numTrees=3, bootstrap=True, seed=4)
    finally:
        trainer.set_option("k4_tma", 0)
    assert_parity_allowing_ties(g, o, max_ties=4)


@pytest.mark.parametrize("F", [12, 40, 70])
def test_odd_code_row_widths(b2rf, oracle, cuda, monkeypatch, F):
    """Code rows that are neither 64 nor 128 bytes (matrices the caller pre-binned, or more than 128 features): the limb kernel
    stages them with one bulk async copy per row at a run-time width.  Forced here by padding rows to 16 bytes."""
    from b2rf import trainer
    monkeypatch.setattr(trainer, "CODE_ROW_ALIGN", 16)
    assert trainer.code_row_width(F) == (F + 15) // 16 * 16
    X, y, arity = b2rf.synth.c3(20000, device="cpu", F=F)
    before = trainer.k4_launch_counts()
    g, o = _fit_both(b2rf, oracle, cuda, X.numpy(), y.numpy(), arity, maxDepth=8, maxBins=64, numTrees=3, bootstrap=True, seed=5)
    assert trainer.k4_launch_counts()["limb"] > before["limb"]
    assert_parity_allowing_ties(g, o, max_ties=4)


def test_more_than_128_features(b2rf, oracle, cuda):
    """150 features: 160-byte code rows (default padding), TMA staging at a run-time width."""
    from b2rf import trainer
    assert trainer.code_row_width(150) == 160 and trainer.code_row_width(64) == 64 and trainer.code_row_width(65) == 128
    X, y, arity = b2rf.synth.c3(12000, device="cpu", F=150)
    before = trainer.k4_launch_counts()
    g, o = _fit_both(b2rf, oracle, cuda, X.numpy(), y.numpy(), arity, maxDepth=6, maxBins=32, numTrees=2, bootstrap=True, seed=1)
    assert trainer.k4_launch_counts()["limb"] > before["limb"]
    assert_parity_allowing_ties(g, o, max_ties=4)


@pytest.mark.parametrize("impurity,tma,F", [("gini", 0, 64), ("variance", 0, 64), ("gini", 1, 64), ("entropy", 0, 100), ("gini", 0, 12)])
def test_binary_labels_two_plane_form(b2rf, oracle, cuda, impurity, tma, F):
    """Labels that are all 0 or 1 (binary classification, or a 0 / 1 regression target): the limb kernel keeps two planes per cell
    (count and sum of w*y; w*y*y == w*y) instead of seven.  All sums are small integers, so counts and sums must be bit-identical to
    the oracle's, and the forest must be the one the general integer-label form builds."""
    from b2rf import trainer
    from helpers import first_tie_in_order
    X, y, arity = b2rf.synth.c3(30000, device="cpu", F=F)
    yb = (y > y.median()).to(torch.float64)
    kw = dict(maxDepth=8, maxBins=64, numTrees=3, bootstrap=True, seed=6, impurity=impurity, featureSubsetStrategy="all")
    trainer.set_option("k4_tma", tma)
    try:
        before = trainer.k4_launch_counts()
        g, o = _fit_both(b2rf, oracle, cuda, X.numpy(), yb.numpy(), arity, **kw)
        assert trainer.k4_launch_counts()["limb"] > before["limb"]
        trainer.set_option("k4_binary", 0)
        g7 = b2rf.trainer.fit_tensors(X.to(cuda), yb.to(cuda), arity, b2rf.TreeParams(**kw), part_offsets=trainer.default_partitions(30000, 1024), keep_trace=True)
    finally:
        trainer.set_option("k4_tma", 0)
        trainer.set_option("k4_binary", 1)
    if impurity == "variance":
        k = first_tie_in_order(g, o)
        upto = len(o.trace_i) if k is None else k
        assert upto > 0.5 * len(o.trace_i)
        assert np.array_equal(g.trace_d[:upto, 2:5], o.trace_d[:upto, 2:5])      # counts AND both sums bit-identical
        if k is None:
            assert_forest_parity(g, o)
    else:   # class counts are exact integers on both sides and Gini / entropy ties resolve identically: full parity (the oracle's trace
            # holds (count of class 0, count of class 1), the device's (count, sum of w*y, sum of w*y*y))
        from helpers import assert_classifier_parity
        assert_classifier_parity(g, o, exact=(impurity == "gini"))
        assert np.array_equal(g.trace_d[:, 2] - g.trace_d[:, 3], o.trace_d[:, 2]) and np.array_equal(g.trace_d[:, 3], o.trace_d[:, 3])
        assert np.array_equal(g.trace_d[:, 3], g.trace_d[:, 4])
    # exact integer sums in both forms: identical node statistics as long as the two device fits take the same path
    m = min(len(g.trace_i), len(g7.trace_i))
    same = np.all(g.trace_i[:m] == g7.trace_i[:m], axis=1)
    first_diff = m if same.all() else int(np.argmin(same))
    assert first_diff > 0.5 * m
    assert np.array_equal(g.trace_d[:first_diff, 2:5], g7.trace_d[:first_diff, 2:5])
