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
        m = o.impurity >= 0
        assert np.array_equal(m, g.impurity >= 0)
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
