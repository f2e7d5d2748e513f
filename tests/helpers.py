"""Comparison helpers shared by the parity tests: product Forest (GPU) vs OracleForest (CPU restatement)."""
import numpy as np

REL = 1e-9   # north_star contract (2): fp64 label sums agree to 1e-9 relative


def assert_forest_parity(g, o, check_stats=True):
    """g: b2rf Forest, o: oracle OracleForest.  Structure, thresholds and category sets must be identical."""
    assert np.array_equal(g.tree_off, o.tree_off), f"tree sizes differ: {np.diff(g.tree_off)[:10]} vs {np.diff(o.tree_off)[:10]}"
    assert np.array_equal(g.feature, o.feature), _first_diff("feature", g, o, g.feature, o.feature)
    assert np.array_equal(g.left, o.left) and np.array_equal(g.right, o.right)
    assert np.array_equal(g.threshold, o.threshold), _first_diff("threshold", g, o, g.threshold, o.threshold)   # bit-exact
    assert np.array_equal(g.ncat, o.ncat)
    for i in np.nonzero(g.ncat > 0)[0]:
        gs = sorted(g.cats[g.cat_off[i]:g.cat_off[i] + g.cat_len[i]])
        os_ = sorted(o.cats[o.cat_off[i]:o.cat_off[i] + o.cat_len[i]])
        assert gs == os_, f"node {i}: left categories differ {gs} vs {os_}"
    if check_stats:
        assert np.array_equal(g.stats[:, 0], o.stats[:, 0]), "node counts differ"          # contract (1): counts bit-exact
        np.testing.assert_allclose(g.stats[:, 1:3], o.stats[:, 1:3], rtol=REL, atol=0)      # contract (2)
        np.testing.assert_allclose(g.prediction, o.prediction, rtol=REL, atol=1e-300)
        m = o.impurity > -0.5          # -1 marks a leaf whose statistics are invalid; a valid impurity can be a rounding-level negative
        assert np.array_equal(m, g.impurity > -0.5)
        np.testing.assert_allclose(g.impurity[m], o.impurity[m], rtol=1e-6, atol=1e-9)
        np.testing.assert_allclose(g.gain, o.gain, rtol=1e-6, atol=1e-9)


def _first_diff(what, g, o, a, b):
    i = int(np.nonzero(a != b)[0][0])
    t = int(np.searchsorted(g.tree_off, i, side="right") - 1)
    return (f"{what} differs first at flat node {i} (tree {t}, local {i - int(g.tree_off[t])}): gpu={a[i]} oracle={b[i]}; "
            f"gpu gain={g.gain[i]!r} oracle gain={o.gain[i]!r}")


def assert_trace_parity(g, o):
    """Same nodes evaluated in the same order with the same decisions (group schedule + RNG order, Appendix A.6)."""
    assert g.trace_i.shape == o.trace_i.shape, f"trace lengths differ {g.trace_i.shape} vs {o.trace_i.shape}"
    assert np.array_equal(g.trace_i, o.trace_i), f"first differing trace row: {np.nonzero((g.trace_i != o.trace_i).any(1))[0][:1]}"
    np.testing.assert_allclose(g.trace_d[:, 2], o.trace_d[:, 2], rtol=0, atol=0)          # counts exact
    np.testing.assert_allclose(g.trace_d[:, 3:5], o.trace_d[:, 3:5], rtol=REL, atol=0)


def rmse(pred, y):
    return float(np.sqrt(np.mean((np.asarray(pred) - np.asarray(y)) ** 2)))


# ---------------------------------------------------------------------------------------------------------------
# Tie-aware comparison.  fp64 sums are order-dependent (MLlib itself is not bit-reproducible across cluster shapes), so
# two candidate splits whose gains agree to ~1 ulp can be ranked differently: the contract allows "documented gain
# ties".  A tie is accepted only when both sides evaluated the SAME node and their best gains agree to 1e-12 relative;
# everything outside the subtrees under tie nodes must still be identical.
# ---------------------------------------------------------------------------------------------------------------
def find_ties(g, o, rel=1e-12):
    gi = {(int(r[1]), int(r[2])): k for k, r in enumerate(g.trace_i)}
    oi = {(int(r[1]), int(r[2])): k for k, r in enumerate(o.trace_i)}
    ties = []

    def under_tie(key):
        t, nid = key
        return any(t == tt and _is_descendant(nid, tn) for tt, tn in ties)

    for key in sorted(set(gi) | set(oi), key=lambda k: (k[0], k[1])):   # ancestors (smaller heap ids) first within a tree
        if under_tie(key):
            continue
        assert key in gi and key in oi, f"node {key} evaluated by only one side"
        a, b = g.trace_i[gi[key]], o.trace_i[oi[key]]
        da, db = g.trace_d[gi[key]], o.trace_d[oi[key]]
        assert da[2] == db[2], f"node {key}: counts differ {da[2]} vs {db[2]}"
        np.testing.assert_allclose(da[3:5], db[3:5], rtol=REL, atol=0)
        if a[3] != b[3] or a[4] != b[4]:
            assert abs(da[0] - db[0]) <= rel * max(1.0, abs(db[0])), f"node {key}: different split and gains differ: {da[0]!r} vs {db[0]!r}"
            ties.append(key)
    return ties


def _is_descendant(nid, root):
    while nid > root:
        nid >>= 1
    return nid == root and True


def assert_forest_parity_with_ties(g, o, ties):
    assert g.num_trees == o.num_trees
    for t in range(g.num_trees):
        gb, ob = int(g.tree_off[t]), int(o.tree_off[t])
        stack = [(0, 0, 1)]
        while stack:
            ig, io, hid = stack.pop()
            if (t, hid) in ties:
                continue
            a, b = gb + ig, ob + io
            assert g.feature[a] == o.feature[b], f"tree {t} node {hid}: feature {g.feature[a]} vs {o.feature[b]}"
            assert g.threshold[a] == o.threshold[b]
            assert g.stats[a, 0] == o.stats[b, 0]
            np.testing.assert_allclose(g.stats[a, 1:3], o.stats[b, 1:3], rtol=REL, atol=0)
            np.testing.assert_allclose(g.prediction[a], o.prediction[b], rtol=REL, atol=1e-300)
            if g.ncat[a] > 0:
                assert sorted(g.cats[g.cat_off[a]:g.cat_off[a] + g.cat_len[a]]) == sorted(o.cats[o.cat_off[b]:o.cat_off[b] + o.cat_len[b]])
            if g.feature[a] >= 0:
                stack.append((int(g.left[a]), int(o.left[b]), 2 * hid))
                stack.append((int(g.right[a]), int(o.right[b]), 2 * hid + 1))


def first_tie_in_order(g, o, rel=1e-12):
    """Walks both decision traces in processing order (pop order of MLlib's node stack).  Returns None when they are identical, else
    the index of the first differing row after checking that it is a genuine gain tie: the same node, the same count and sums, best
    gains equal to `rel` — two candidate splits that separate the node's rows identically have the same gain in exact arithmetic,
    and which one wins is decided by the last bits of fp64 sums (in MLlib itself: by partition layout).  With feature subsampling
    every node popped after such a tie draws its subset from a master stream that the tie's subtree may have advanced differently,
    so nothing after the tie can be compared node by node."""
    n = min(len(g.trace_i), len(o.trace_i))
    neq = np.nonzero((g.trace_i[:n] != o.trace_i[:n]).any(1))[0]
    k = int(neq[0]) if len(neq) else (n if len(g.trace_i) != len(o.trace_i) else None)
    upto = n if k is None else k
    np.testing.assert_array_equal(g.trace_d[:upto, 2], o.trace_d[:upto, 2])                      # counts exact
    np.testing.assert_allclose(g.trace_d[:upto, 3:5], o.trace_d[:upto, 3:5], rtol=REL, atol=0)   # sums
    if k is None:
        return None
    assert k < n, "one side evaluated more nodes without any differing decision before"
    a, b, da, db = g.trace_i[k], o.trace_i[k], g.trace_d[k], o.trace_d[k]
    assert tuple(a[:3]) == tuple(b[:3]), f"trace row {k}: different nodes were evaluated: {a[:3]} vs {b[:3]}"
    assert da[2] == db[2], f"trace row {k}: counts differ {da[2]} vs {db[2]}"
    np.testing.assert_allclose(da[3:5], db[3:5], rtol=REL, atol=0)
    assert abs(da[0] - db[0]) <= rel * max(1.0, abs(db[0])), f"trace row {k} node {a[:3]}: different split and gains differ: {da[0]!r} vs {db[0]!r}"
    return k


def assert_parity_allowing_ties(g, o, max_ties, subsampling=None, min_fraction=0.0):
    """Forest parity up to documented gain ties.  Without feature subsampling (one tree, or featureSubsetStrategy=all) a tie only
    changes its own subtree: everything outside the subtrees of at most `max_ties` tie nodes must match.  With subsampling the
    comparison is exact up to the first genuine tie in processing order (see first_tie_in_order); `min_fraction` is the share of the
    decision trace that must have been compared by then."""
    if subsampling is None:
        subsampling = g.num_trees > 1 and getattr(g, "subsampling", True)
    if subsampling:
        k = first_tie_in_order(g, o)
        if k is None:
            assert_trace_parity(g, o)
            assert_forest_parity(g, o)
            return []
        frac = k / max(1, len(o.trace_i))
        print(f"documented gain tie at decision {k} of {len(o.trace_i)} (group, tree, node) = {tuple(int(v) for v in o.trace_i[k][:3])}; "
              f"{100 * frac:.1f}% of the decisions compared exactly before it")
        assert frac >= min_fraction, f"first gain tie after only {100 * frac:.1f}% of the decisions"
        return [(int(o.trace_i[k][1]), int(o.trace_i[k][2]))]
    ties = find_ties(g, o)
    assert len(ties) <= max_ties, f"too many gain ties: {ties}"
    if not ties:
        assert_trace_parity(g, o)
        assert_forest_parity(g, o)
    else:
        print(f"documented gain ties (tree, heap node id): {ties}")
        assert_forest_parity_with_ties(g, o, set(ties))
    return ties


# ---------------------------------------------------------------------------------------------------------------
# Binary classification (gini / entropy).  The product keeps (sum w, sum w*y) with y in {0, 1}; the oracle accumulates
# genuine per-class counts.  Both are exact integers, so everything except Entropy's log must agree bit for bit.
# ---------------------------------------------------------------------------------------------------------------
def assert_classifier_parity(g, o, exact=True):
    assert np.array_equal(g.tree_off, o.tree_off), f"tree sizes differ: {np.diff(g.tree_off)[:10]} vs {np.diff(o.tree_off)[:10]}"
    assert np.array_equal(g.feature, o.feature), _first_diff("feature", g, o, g.feature, o.feature)
    assert np.array_equal(g.left, o.left) and np.array_equal(g.right, o.right)
    assert np.array_equal(g.threshold, o.threshold), _first_diff("threshold", g, o, g.threshold, o.threshold)
    assert np.array_equal(g.ncat, o.ncat)
    for i in np.nonzero(g.ncat > 0)[0]:
        gs = sorted(g.cats[g.cat_off[i]:g.cat_off[i] + g.cat_len[i]])
        os_ = sorted(o.cats[o.cat_off[i]:o.cat_off[i] + o.cat_len[i]])
        assert gs == os_, f"node {i}: left categories differ {gs} vs {os_}"
    assert np.array_equal(g.class_counts, o.stats[:, :2]), "class counts differ"
    assert np.array_equal(g.prediction, o.prediction), "class predictions differ"
    m = o.impurity >= 0
    assert np.array_equal(m, g.impurity >= 0)
    if exact:      # gini: divisions, products and differences of exact integers, -fmad=false on the device
        assert np.array_equal(g.impurity[m], o.impurity[m]) and np.array_equal(g.gain, o.gain)
    else:          # entropy: device log vs libm log, each within an ulp or so
        np.testing.assert_allclose(g.impurity[m], o.impurity[m], rtol=1e-12, atol=1e-15)
        np.testing.assert_allclose(g.gain, o.gain, rtol=1e-9, atol=1e-13)
    if getattr(g, "trace_i", None) is not None and len(g.trace_i):
        assert np.array_equal(g.trace_i, o.trace_i), "nodes evaluated in a different order or with different decisions"
