"""TEST INFRASTRUCTURE ONLY (imported by tests/): numpy restatement of MLlib 3.1.2 tree-ensemble CLASSIFICATION training for any
numClasses >= 2, including the unordered-categorical path that only multiclass problems reach.  Small inputs only (Python loops
over nodes; N <= 10,000 so findSplits sees every row).

It is a second, independent restatement of the training loop (RandomForest.run / selectNodesToSplit / findBestSplits /
binsToBestSplit / LearningNode.toNode) built on the primitives the C++ oracle already pins with known answers
(findSplits, findBin, the bagging streams, java.util.Random, reservoir sampling, Scala map order).  For numClasses == 2 it must
agree with the C++ oracle's binary path, which is how this file itself is checked (tests/test_cpu_oracle.py).

Upstream anchors (org.apache.spark.ml.tree.impl / mllib.tree.impurity; no sources in this image):
DecisionTreeMetadata.buildMetadata (unordered features, numBins), RandomForest.binsToBestSplit (three feature kinds),
Gini.calculate / Entropy.calculate, ImpurityCalculator.predict (indexOfLargestArrayElement), RandomForest.extractMultiClassCategories.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np

from . import oracle as O

DBL_MAX = 1.7976931348623157e308
EPS = 2.0 ** -52


def impurity_of(kind: str, counts) -> float:
    total = 0.0
    for c in counts:
        total += float(c)
    if total == 0:
        return 0.0
    if kind == "gini":
        imp = 1.0
        for c in counts:
            f = float(c) / total
            imp -= f * f
        return imp
    imp = 0.0
    for c in counts:
        if c != 0:
            f = float(c) / total
            imp -= f * (math.log(f) / math.log(2.0))
    return imp


def predict_of(counts) -> float:
    return float(int(np.argmax(counts))) if float(np.sum(counts)) != 0 else 0.0     # first maximum


def unordered_left_categories(split_index: int, arity: int):
    """RandomForest.extractMultiClassCategories(splitIndex + 1, arity): categories whose bit is set."""
    v = split_index + 1
    return [j for j in range(arity) if (v >> j) & 1]


@dataclass
class _Node:
    id: int = 1
    left: int = -1
    right: int = -1
    is_leaf: bool = False
    feature: int = -1
    split_bin: int = -1
    left_cats: list = field(default_factory=list)
    threshold: float = 0.0
    counts: np.ndarray = None     # impurityStats
    impurity: float = 0.0
    gain: float = float("nan")
    valid: bool = True
    has_stats: bool = False


@dataclass
class McForest:
    num_classes: int
    tree_off: np.ndarray
    feature: np.ndarray
    threshold: np.ndarray
    left: np.ndarray
    right: np.ndarray
    prediction: np.ndarray
    impurity: np.ndarray
    gain: np.ndarray
    class_counts: np.ndarray
    ncat: np.ndarray
    cat_off: np.ndarray
    cat_len: np.ndarray
    cats: np.ndarray
    trace_i: np.ndarray
    group_sizes: list
    unordered: list

    def _leaf(self, x, t):
        b = int(self.tree_off[t]); n = 0
        while self.feature[b + n] >= 0:
            v = x[self.feature[b + n]]
            if self.ncat[b + n] == 0:
                goes_left = v <= self.threshold[b + n]
            else:
                goes_left = v in self.cats[self.cat_off[b + n]:self.cat_off[b + n] + self.cat_len[b + n]]
            n = int(self.left[b + n] if goes_left else self.right[b + n])
        return b + n

    def predict_raw(self, X, single_tree=None):
        T = len(self.tree_off) - 1
        single = (T == 1) if single_tree is None else single_tree
        out = np.zeros((X.shape[0], self.num_classes))
        for i in range(X.shape[0]):
            if single:
                out[i] = self.class_counts[self._leaf(X[i], 0)]
                continue
            for t in range(T):
                c = self.class_counts[self._leaf(X[i], t)]
                total = 0.0
                for v in c:
                    total += float(v)
                if total != 0:
                    for k in range(self.num_classes):
                        out[i, k] += float(c[k]) / total
        return out

    def predict(self, X):
        return np.argmax(self.predict_raw(X), axis=1).astype(np.float64)


def fit(X, y, arity, num_classes, *, maxDepth=5, maxBins=32, minInstancesPerNode=1, numTrees=1, featureSubsetStrategy="auto",
        maxMemoryInMB=256, bootstrap=False, minInfoGain=0.0, minWeightFractionPerNode=0.0, subsamplingRate=1.0, seed=0,
        part_offsets=None, impurity="gini") -> McForest:
    X = np.ascontiguousarray(X, np.float64); y = np.ascontiguousarray(y, np.float64)
    N, F = X.shape
    C, T = int(num_classes), int(numTrees)
    assert N <= 10000, "the numpy restatement has no findSplits row sampling"
    lab = y.astype(np.int64)
    assert np.array_equal(lab.astype(np.float64), y) and lab.min() >= 0 and lab.max() < C
    arity = np.asarray(arity, np.int32)
    po = np.array([0, N], np.int64) if part_offsets is None else np.asarray(part_offsets, np.int64)

    # ---- DecisionTreeMetadata.buildMetadata ----
    mpb = min(maxBins, N)
    unordered = set()
    num_bins = np.full(F, mpb, np.int64)
    if C > 2:
        max_cat_unordered = int(math.floor(math.log(mpb // 2 + 1) / math.log(2.0) + 1))
    for f in range(F):
        if arity[f] > 1:
            if C > 2 and arity[f] <= max_cat_unordered:
                unordered.add(f)
                num_bins[f] = (1 << (int(arity[f]) - 1)) - 1
            else:
                if arity[f] > mpb:
                    raise O.OracleError(f"DecisionTree requires maxBins (= {mpb}) to be at least as large as the number of values in each "
                                        f"categorical feature, but categorical feature {f} has {arity[f]} values.")
                num_bins[f] = arity[f]
    thr, ns = O.find_splits(X, arity, mpb, float(N))
    for f in range(F):
        if arity[f] == 0:
            num_bins[f] = ns[f] + 1
    num_splits = np.array([num_bins[f] if f in unordered else num_bins[f] - 1 for f in range(F)])
    phys_bins = np.array([arity[f] if arity[f] > 1 else num_bins[f] for f in range(F)])      # cells the aggregation really needs
    bins = np.asarray(O.bin_features(X, thr, ns, arity)).reshape(N, F).astype(np.int64)
    counts = np.asarray(O.bag(seed, po, T, subsamplingRate, bootstrap)).reshape(N, T).astype(np.float64)
    fsub = O.features_per_node(featureSubsetStrategy, F, T, regression=False)
    subsampling = fsub != F
    min_weight = minWeightFractionPerNode * float(N)
    longs = iter(O.java_random_longs(seed, 200000)) if subsampling else None

    trees = [[_Node()] for _ in range(T)]
    assign = np.zeros((T, N), np.int64)               # index into trees[t] of the node every row currently sits in
    stack = [(t, 0) for t in range(T)]                 # the last element is the head of MLlib's nodeStack
    max_mem = maxMemoryInMB * 1024 * 1024
    trace, group_sizes, group_no = [], [], 0

    def impurity_stats(running, L, R):
        """calculateImpurityStats; running = (counts, impurity) of the parent or None."""
        if running is None:
            parent = L + R
            imp = impurity_of(impurity, parent)
        else:
            parent, imp = running
        lc, rc = float(np.sum(L)), float(np.sum(R))
        res = dict(parent=parent, impurity=imp, left=L, right=R)
        if lc < minInstancesPerNode or rc < minInstancesPerNode or lc < min_weight or rc < min_weight:
            res.update(gain=-DBL_MAX, valid=False)
            return res
        li, ri = impurity_of(impurity, L), impurity_of(impurity, R)
        total = lc + rc
        gain = imp - (lc / total) * li - (rc / total) * ri
        if gain < minInfoGain:
            res.update(gain=-DBL_MAX, valid=False)
            return res
        res.update(gain=gain, valid=True, li=li, ri=ri)
        return res

    while stack:
        # ---- selectNodesToSplit ----
        group, tree_order, mem = [], [], 0
        while stack:
            t, ni = stack[-1]
            if subsampling:
                feats = list(O.reservoir(F, fsub, int(next(longs))))       # drawn before the memory check
            else:
                feats = list(range(F))
            node_mem = C * int(sum(num_bins[f] for f in feats)) * 8
            if mem + node_mem <= max_mem or mem == 0:
                stack.pop()
                if t not in tree_order:
                    tree_order.append(t)
                group.append((t, ni, feats))
                mem += node_mem
            else:
                break
        group_sizes.append(len(group))
        # ---- findBestSplits: aggregate + binsToBestSplit per node ----
        bests = []
        for t, ni, feats in group:
            node = trees[t][ni]
            rows = np.nonzero((assign[t] == ni) & (counts[:, t] > 0))[0]
            w = counts[rows, t]
            totals = np.zeros(C); np.add.at(totals, lab[rows], w)            # parentStats of the aggregator
            running = (node.counts, node.impurity) if node.id > 1 else None
            best = None
            for f in feats:
                nsp = int(num_splits[f])
                if nsp == 0:
                    continue
                h = np.zeros((int(phys_bins[f]), C))
                np.add.at(h, (bins[rows, f], lab[rows]), w)
                cand = []                                                    # (split index, stats, left categories)
                if arity[f] == 0:
                    cum = np.cumsum(h, axis=0)
                    for s in range(nsp):
                        running_stats = impurity_stats(running, cum[s].copy(), cum[-1] - cum[s])
                        running = (running_stats["parent"], running_stats["impurity"])
                        cand.append((s, running_stats, None))
                elif f in unordered:
                    for s in range(nsp):
                        cats = unordered_left_categories(s, int(arity[f]))
                        L = h[cats].sum(axis=0)
                        running_stats = impurity_stats(running, L, totals - L)
                        running = (running_stats["parent"], running_stats["impurity"])
                        cand.append((s, running_stats, cats))
                else:
                    def centroid(c):
                        if float(np.sum(h[c])) == 0:
                            return DBL_MAX
                        return impurity_of(impurity, h[c]) if C > 2 else float(h[c][1])
                    order = sorted(range(int(arity[f])), key=centroid)      # stable, like List.sortBy
                    cum = np.cumsum(h[order], axis=0)
                    for s in range(nsp):
                        running_stats = impurity_stats(running, cum[s].copy(), cum[-1] - cum[s])
                        running = (running_stats["parent"], running_stats["impurity"])
                        cand.append((s, running_stats, order[:s + 1]))
                fb = None
                for c in cand:                                               # maxBy: first maximum
                    if fb is None or c[1]["gain"] > fb[1]["gain"]:
                        fb = c
                if best is None or fb[1]["gain"] > best[2]["gain"]:
                    best = (f, fb[0], fb[1], fb[2])
            if best is None:      # no feature of the subset has a split: invalid stats on the node totals
                parent = running[0] if running is not None else totals
                best = (feats[0] if feats else 0, 0, dict(parent=parent, impurity=impurity_of(impurity, parent), gain=-DBL_MAX, valid=False,
                                                          left=None, right=None), None)
            bests.append(best)
        # ---- driver tail: trees in Scala map order, nodes in pop order ----
        for t in [int(v) for v in O.group_order(np.array(tree_order, np.int32))]:
            for (tt, ni, feats), (f, s, st, cats) in zip(group, bests):
                if tt != t:
                    continue
                node = trees[t][ni]
                lvl = node.id.bit_length() - 1
                node.is_leaf = (st["gain"] <= 0) or (lvl == maxDepth)
                node.counts, node.impurity, node.gain, node.valid, node.has_stats = st["parent"], st["impurity"], st["gain"], st["valid"], True
                trace.append((group_no, t, node.id, f, s))
                if node.is_leaf:
                    continue
                node.feature = f
                if arity[f] == 0:
                    node.split_bin, node.threshold = s, float(thr[f, s])
                else:
                    node.left_cats = list(cats)
                child_leaf = (lvl + 1) == maxDepth
                li, ri = st["li"], st["ri"]
                L = _Node(id=2 * node.id, is_leaf=child_leaf or abs(li) < EPS, counts=st["left"], impurity=li, has_stats=True)
                R = _Node(id=2 * node.id + 1, is_leaf=child_leaf or abs(ri) < EPS, counts=st["right"], impurity=ri, has_stats=True)
                node.left, node.right = len(trees[t]), len(trees[t]) + 1
                trees[t] += [L, R]
                here = assign[t] == ni
                if arity[f] == 0:
                    goes_left = bins[:, f] <= s
                else:
                    goes_left = np.isin(bins[:, f], node.left_cats)
                assign[t][here & goes_left] = node.left
                assign[t][here & ~goes_left] = node.right
                if not L.is_leaf:
                    stack.append((t, node.left))
                if not R.is_leaf:
                    stack.append((t, node.right))
        group_no += 1

    # ---- LearningNode.toNode(prune = true), pre-order ----
    out = dict(feature=[], threshold=[], left=[], right=[], prediction=[], impurity=[], gain=[], counts=[], ncat=[], cat_off=[], cat_len=[])
    cats_pool, tree_off = [], [0]

    def flatten(t, ni, base):
        n = trees[t][ni]
        me = len(out["feature"]) - base
        for k, v in (("feature", -1), ("threshold", 0.0), ("left", -1), ("right", -1), ("prediction", predict_of(n.counts)),
                     ("impurity", n.impurity if n.valid else -1.0), ("gain", 0.0), ("counts", np.asarray(n.counts, np.float64)),
                     ("ncat", 0), ("cat_off", len(cats_pool)), ("cat_len", 0)):
            out[k].append(v)
        at = base + me
        if n.left >= 0:
            mark = len(cats_pool)
            l = flatten(t, n.left, base)
            r = flatten(t, n.right, base)
            if out["feature"][base + l] < 0 and out["feature"][base + r] < 0 and out["prediction"][base + l] == out["prediction"][base + r]:
                pred = out["prediction"][base + l]                          # new LeafNode(l.prediction, stats.impurity, stats.impurityCalculator)
                for k in out:
                    del out[k][at + 1:]
                del cats_pool[mark:]
                out["impurity"][at], out["prediction"][at] = n.impurity, pred
                return me
            out["feature"][at], out["threshold"][at], out["left"][at], out["right"][at] = n.feature, n.threshold, l, r
            out["impurity"][at], out["gain"][at] = n.impurity, n.gain
            if arity[n.feature] > 0:
                out["ncat"][at], out["cat_off"][at], out["cat_len"][at] = int(arity[n.feature]), len(cats_pool), len(n.left_cats)
                cats_pool.extend(float(c) for c in sorted(n.left_cats))
        return me

    for t in range(T):
        flatten(t, 0, len(out["feature"]))
        tree_off.append(len(out["feature"]))
    return McForest(C, np.array(tree_off, np.int64), np.array(out["feature"], np.int32), np.array(out["threshold"]), np.array(out["left"], np.int32),
                    np.array(out["right"], np.int32), np.array(out["prediction"]), np.array(out["impurity"]), np.array(out["gain"]),
                    np.array(out["counts"]).reshape(-1, C), np.array(out["ncat"], np.int32), np.array(out["cat_off"], np.int32),
                    np.array(out["cat_len"], np.int32), np.array(cats_pool, np.float64), np.array(trace, np.int32).reshape(-1, 5), group_sizes,
                    sorted(unordered))
