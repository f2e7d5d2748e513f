"""Fitted model arrays (LearningNode.toNode(prune=true) output, pre-order per tree) and the model-side API of
pyspark's DecisionTreeRegressionModel / RandomForestRegressionModel (`ML 06 - Decision Trees.py:137,141,153,188`,
`ML 07 ...:162`): predict, featureImportances, toDebugString, depth, numNodes."""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import numpy as np


def java_double_str(x: float) -> str:
    """java.lang.Double.toString formatting (used by MLlib's toDebugString)."""
    x = float(x)
    if x != x:
        return "NaN"
    if x in (float("inf"), float("-inf")):
        return "Infinity" if x > 0 else "-Infinity"
    if x == 0:
        return "-0.0" if str(x).startswith("-") else "0.0"
    ax = abs(x)
    r = repr(ax)
    if "e" in r or "E" in r:
        mant, exp = r.lower().split("e")
        digits = mant.replace(".", "")
        e10 = int(exp) + (mant.index(".") if "." in mant else len(mant)) - 1
    else:
        ip, _, fp = r.partition(".")
        digits = (ip + fp).lstrip("0")
        e10 = len(ip) - 1 if ip.strip("0") else -(len(fp) - len(fp.lstrip("0")) + 1)
    digits = digits.rstrip("0") or "0"
    sign = "-" if x < 0 else ""
    if 1e-3 <= ax < 1e7:
        if e10 >= 0:
            ip = digits[: e10 + 1].ljust(e10 + 1, "0")
            fp = digits[e10 + 1:] or "0"
        else:
            ip = "0"
            fp = "0" * (-e10 - 1) + digits
        return f"{sign}{ip}.{fp}"
    return f"{sign}{digits[0]}.{digits[1:] or '0'}E{e10}"


@dataclass
class Forest:
    num_features: int
    tree_off: np.ndarray      # int64 [T+1]
    feature: np.ndarray       # int32 [n]; -1 = leaf
    threshold: np.ndarray     # f64 [n]
    left: np.ndarray          # int32 [n] tree-local pre-order index
    right: np.ndarray
    prediction: np.ndarray    # f64 [n]
    impurity: np.ndarray
    gain: np.ndarray
    stats: np.ndarray         # f64 [n,4] weight, sum, sumsq, rawCount
    ncat: np.ndarray          # int32 [n] numCategories (0 = continuous split)
    cat_off: np.ndarray
    cat_len: np.ndarray
    cats: np.ndarray          # f64 left-category lists
    thresholds: np.ndarray = None   # [F, maxSplits] split candidates (findSplits output)
    nsplits: np.ndarray = None
    timers: dict = field(default_factory=dict)
    counters: dict = field(default_factory=dict)
    trace_i: np.ndarray = None
    trace_d: np.ndarray = None
    uid: str = ""
    impurity_kind: str = "variance"   # "variance": regression; "gini" / "entropy": classification
    counts: np.ndarray = None         # multiclass only: f64 [n, numClasses] class counts per node (binary models derive them from stats)
    _dev: dict = field(default_factory=dict, repr=False)

    # ---- structure ----
    @property
    def num_trees(self) -> int:
        return len(self.tree_off) - 1

    @property
    def is_classification(self) -> bool:
        return self.impurity_kind != "variance"

    @property
    def num_classes(self) -> int:
        if not self.is_classification:
            return 0
        return 2 if self.counts is None else int(self.counts.shape[1])

    @property
    def class_counts(self) -> np.ndarray:
        """impurityStats of a classification node: [n, 2] class counts.  The trainer accumulates (sum w, sum w*y) with y in {0, 1},
        so count(class 1) = stats[:, 1] and count(class 0) = stats[:, 0] - stats[:, 1], both exact integers."""
        if not self.is_classification:
            raise ValueError("class_counts: this is a regression model")
        if self.counts is not None:
            return self.counts
        return np.stack([self.stats[:, 0] - self.stats[:, 1], self.stats[:, 1]], axis=1)

    def tree_slice(self, t: int) -> slice:
        return slice(int(self.tree_off[t]), int(self.tree_off[t + 1]))

    def num_nodes(self, t: int | None = None) -> int:
        if t is None:
            return int(self.tree_off[-1])
        return int(self.tree_off[t + 1] - self.tree_off[t])

    def depth(self, t: int = 0) -> int:
        base = int(self.tree_off[t])

        def d(i):
            if self.feature[base + i] < 0:
                return 0
            return 1 + max(d(int(self.left[base + i])), d(int(self.right[base + i])))
        return d(0)

    # ---- TreeEnsembleModel.featureImportances (perTreeNormalization = true) ----
    def feature_importances(self) -> np.ndarray:
        total = np.zeros(self.num_features)
        for t in range(self.num_trees):
            s = self.tree_slice(t)
            f = self.feature[s]
            m = f >= 0
            imp = np.zeros(self.num_features)
            np.add.at(imp, f[m], self.gain[s][m] * self.stats[s][m, 0])
            norm = imp.sum()
            if norm != 0:
                total += imp / norm
        tot = total.sum()
        return total / tot if tot != 0 else total

    # ---- toDebugString ----
    def _subtree(self, base: int, i: int, indent: int, out: list):
        p = " " * indent
        if self.feature[base + i] < 0:
            out.append(f"{p}Predict: {java_double_str(self.prediction[base + i])}\n")
            return
        f = int(self.feature[base + i])
        if self.ncat[base + i] == 0:
            l = f"feature {f} <= {java_double_str(self.threshold[base + i])}"
            r = f"feature {f} > {java_double_str(self.threshold[base + i])}"
        else:
            o, n = int(self.cat_off[base + i]), int(self.cat_len[base + i])
            cs = "{" + ",".join(java_double_str(c) for c in sorted(self.cats[o:o + n])) + "}"
            l, r = f"feature {f} in {cs}", f"feature {f} not in {cs}"
        out.append(f"{p}If ({l})\n")
        self._subtree(base, int(self.left[base + i]), indent + 1, out)
        out.append(f"{p}Else ({r})\n")
        self._subtree(base, int(self.right[base + i]), indent + 1, out)

    def to_debug_string(self, ensemble: bool | None = None) -> str:
        ensemble = self.num_trees > 1 if ensemble is None else ensemble
        out = []
        kind = "Classification" if self.is_classification else "Regression"
        ncls = f"numClasses={self.num_classes}, " if self.is_classification else ""
        if not ensemble:
            out.append(f"DecisionTree{kind}Model: uid={self.uid}, depth={self.depth(0)}, numNodes={self.num_nodes(0)}, "
                       f"{ncls}numFeatures={self.num_features}\n")
            self._subtree(0, 0, 2, out)
        else:
            out.append(f"RandomForest{kind}Model: uid={self.uid}, numTrees={self.num_trees}, {ncls}numFeatures={self.num_features}\n")
            for t in range(self.num_trees):
                out.append(f"  Tree {t} (weight 1.0):\n")
                self._subtree(int(self.tree_off[t]), 0, 4, out)
        return "".join(out)

    # ---- K7 predict on the device ----
    def _device_arrays(self, device):
        import torch
        key = str(device)
        if key not in self._dev:
            def T(a, dt):
                return torch.as_tensor(np.ascontiguousarray(a), dtype=dt).to(device)
            cats = self.cats if len(self.cats) else np.zeros(1)
            self._dev[key] = dict(tree_off=T(self.tree_off, torch.int64), feature=T(self.feature, torch.int32),
                                  threshold=T(self.threshold, torch.float64), left=T(self.left, torch.int32), right=T(self.right, torch.int32),
                                  prediction=T(self.prediction, torch.float64), ncat=T(self.ncat, torch.int32),
                                  cat_off=T(self.cat_off, torch.int32), cat_len=T(self.cat_len, torch.int32), cats=T(cats, torch.float64),
                                  stats=T(self.stats, torch.float64))
        return self._dev[key]

    # ---- packed nodes (b2rf_forest_pack): one 16-byte record per node, built once per device ----
    def is_preorder(self) -> bool:
        """True when every internal node's left child is the next node of its tree (MLlib NodeData numbering; always true for
        forests exported by the trainer)."""
        idx = np.arange(len(self.feature)) - np.repeat(self.tree_off[:-1], np.diff(self.tree_off))
        internal = self.feature >= 0
        return bool(np.all(self.left[internal] == idx[internal] + 1))

    def _packed(self, device):
        import torch
        from ._lib import check, lib
        key = "packed:" + str(device)
        if key not in self._dev:
            d = self._device_arrays(device)
            n = len(self.feature)
            buf = torch.empty(max(int(lib().b2rf_forest_pack_bytes(n)), 16), dtype=torch.uint8, device=device)
            # left-category sets of the categorical nodes as bit masks (one row per categorical node)
            catn = np.nonzero((self.feature >= 0) & (self.ncat > 0))[0]
            mw = max(1, (int(self.ncat.max()) + 31) // 32) if len(catn) else 1
            masks = np.zeros((max(len(catn), 1), mw), np.uint32)
            midx = np.full(n, -1, np.int32)
            midx[catn] = np.arange(len(catn), dtype=np.int32)
            if len(catn):
                lens = self.cat_len[catn].astype(np.int64)
                rows = np.repeat(np.arange(len(catn)), lens)
                pos = np.repeat(self.cat_off[catn].astype(np.int64), lens) + (np.arange(lens.sum()) - np.repeat(np.cumsum(lens) - lens, lens))
                c = self.cats[pos].astype(np.int64)
                np.bitwise_or.at(masks, (rows, c >> 5), (np.uint32(1) << (c & 31).astype(np.uint32)))
            masks_d = torch.as_tensor(masks.view(np.int32)).to(device)
            midx_d = torch.as_tensor(midx).to(device)
            with torch.cuda.device(device):
                st = torch.cuda.current_stream().cuda_stream
                check(lib().b2rf_forest_pack(n, d["feature"].data_ptr(), d["threshold"].data_ptr(), d["right"].data_ptr(), d["prediction"].data_ptr(),
                                             d["ncat"].data_ptr(), d["cat_off"].data_ptr(), d["cat_len"].data_ptr(), midx_d.data_ptr(), buf.data_ptr(),
                                             C.c_void_p(st)))
                torch.cuda.current_stream().synchronize()      # midx_d is only needed by the pack kernel
            self._dev[key] = buf
            self._dev["masks:" + str(device)] = (masks_d, mw)
        return self._dev[key]

    def _predict_packed(self, X, want_pred: bool, want_leaves: bool):
        import torch
        from ._lib import check, lib
        d = self._device_arrays(X.device)
        pk = self._packed(X.device)
        out = torch.empty(X.shape[0], dtype=torch.float64, device=X.device) if want_pred else None
        leaves = torch.empty((X.shape[0], self.num_trees), dtype=torch.int32, device=X.device) if want_leaves else None
        with torch.cuda.device(X.device):
            st = torch.cuda.current_stream().cuda_stream
            masks_d, mw = self._dev["masks:" + str(X.device)]
            check(lib().b2rf_predict_packed(X.data_ptr(), X.stride(0), X.shape[0], self.num_trees, d["tree_off"].data_ptr(), pk.data_ptr(),
                                            d["cats"].data_ptr(), masks_d.data_ptr(), mw, 0 if out is None else out.data_ptr(),
                                            0 if leaves is None else leaves.data_ptr(), C.c_void_p(st)))
        return out, leaves

    # ---- K7 on bin codes (all-continuous forests) ----
    def _coded(self, device):
        """Per-feature tables of the forest's own split thresholds, the nodes' split codes, and the 8-byte node records of
        b2rf_predict_coded; None when the forest has categorical splits, is not in pre-order or is too small to bother."""
        import torch
        key = "coded:" + str(device)
        if key in self._dev:
            return self._dev[key]
        res = None
        internal = self.feature >= 0
        if self.is_preorder() and not (self.ncat[internal] > 0).any() and internal.any() and self.num_features < 0xffff:
            F = self.num_features
            feats, thr = self.feature[internal], self.threshold[internal]
            tabs = [np.unique(thr[feats == f]) for f in range(F)]
            L = max(1, max(len(t) for t in tabs))
            if L <= 65534:
                table = np.full((F, L), np.inf)
                for f, t in enumerate(tabs):
                    table[f, :len(t)] = t
                ns = np.array([len(t) for t in tabs], np.int32)
                code = np.zeros(len(self.feature), np.int64)
                for f in range(F):
                    m = internal & (self.feature == f)
                    if m.any():
                        code[m] = np.searchsorted(tabs[f], self.threshold[m])
                fc = np.where(internal, self.feature.astype(np.int64) | (code << 16), 0xffff).astype(np.uint32)
                rec = np.stack([fc.view(np.int32), np.where(internal, self.right, -1).astype(np.int32)], axis=1)
                res = dict(table=torch.as_tensor(table).to(device), ns=torch.as_tensor(ns).to(device),
                           arity=torch.zeros(F, dtype=torch.int32, device=device), nodes=torch.as_tensor(np.ascontiguousarray(rec)).to(device),
                           code_dt=torch.uint8 if L <= 255 else torch.int16, L=L)
        self._dev[key] = res
        return res

    def _predict_coded(self, X, want_pred: bool, want_leaves: bool, chunk_rows: int = 1 << 22):
        """Bins X against the forest's threshold tables (K2) chunk by chunk, then walks the 8-byte nodes (K7 on codes)."""
        import torch
        from ._lib import check, lib
        from .trainer import bin_features, _pad
        cd = self._coded(X.device)
        d = self._device_arrays(X.device)
        n, F = X.shape
        out = torch.empty(n, dtype=torch.float64, device=X.device) if want_pred else None
        leaves = torch.empty((n, self.num_trees), dtype=torch.int32, device=X.device) if want_leaves else None
        rows = min(n, chunk_rows)
        codes = torch.empty((rows, _pad(F, 16)), dtype=cd["code_dt"], device=X.device)
        with torch.cuda.device(X.device):
            st = torch.cuda.current_stream().cuda_stream
            for r0 in range(0, n, rows):
                m = min(rows, n - r0)
                check(lib().b2rf_bin(X[r0:r0 + m].data_ptr(), X.stride(0), m, F, cd["table"].data_ptr(), cd["ns"].data_ptr(), cd["arity"].data_ptr(),
                                     cd["L"], 0, 0, 0, 0, codes.data_ptr(), codes.stride(0), codes.element_size(), C.c_void_p(st)))
                check(lib().b2rf_predict_coded(codes.data_ptr(), codes.stride(0), codes.element_size(), m, self.num_trees, d["tree_off"].data_ptr(),
                                               cd["nodes"].data_ptr(), d["prediction"].data_ptr(), 0 if out is None else out[r0:].data_ptr(),
                                               0 if leaves is None else leaves[r0:].data_ptr(), C.c_void_p(st)))
        return out, leaves

    def predict_leaves(self, X):
        """Tree-local node index of the leaf every row reaches in every tree: torch.int32 [n, T] (K7, packed form)."""
        import torch
        if not (isinstance(X, torch.Tensor) and X.is_cuda):
            raise TypeError("Forest.predict_leaves needs a CUDA tensor: this package has no CPU path")
        if not self.is_preorder():
            raise ValueError("predict_leaves needs pre-order node numbering (MLlib's NodeData order)")
        return self._predict_packed(X.to(torch.float64).contiguous(), False, True)[1]

    def leaf_ordinals(self) -> np.ndarray:
        """DecisionTreeModel.leafIndices: for every node, its rank among the leaves of its tree in left-to-right order (-1 for internal
        nodes).  With pre-order numbering that is the count of leaves before it."""
        leaf = self.feature < 0
        before = np.cumsum(leaf) - leaf
        start = np.repeat(before[self.tree_off[:-1]], np.diff(self.tree_off))
        return np.where(leaf, before - start, -1).astype(np.int64)

    def predict(self, X):
        """X: torch.float64 CUDA tensor [n, F] (row-major).  Returns torch.float64 [n] on the same device."""
        import torch
        from ._lib import check, lib
        if not (isinstance(X, torch.Tensor) and X.is_cuda):
            raise TypeError("Forest.predict needs a CUDA tensor: this package has no CPU path")
        if self.is_classification:   # raw2prediction without thresholds: argmax of rawPrediction, first maximum
            raw = self.predict_raw(X)
            best, idx = raw[:, 0].clone(), torch.zeros(raw.shape[0], dtype=torch.float64, device=raw.device)
            for k in range(1, raw.shape[1]):
                m = raw[:, k] > best
                idx[m] = float(k)
                best = torch.where(m, raw[:, k], best)
            return idx
        X = X.to(torch.float64).contiguous()
        if X.shape[0] >= 4096 and self._coded(X.device) is not None:     # large batches of an all-continuous forest: walk bin codes
            return self._predict_coded(X, True, False)[0]
        if self.is_preorder():
            return self._predict_packed(X, True, False)[0]
        d = self._device_arrays(X.device)
        out = torch.empty(X.shape[0], dtype=torch.float64, device=X.device)
        with torch.cuda.device(X.device):
            st = torch.cuda.current_stream().cuda_stream
            check(lib().b2rf_predict(X.data_ptr(), X.stride(0), X.shape[0], self.num_trees, d["tree_off"].data_ptr(), d["feature"].data_ptr(),
                                     d["threshold"].data_ptr(), d["left"].data_ptr(), d["right"].data_ptr(), d["prediction"].data_ptr(),
                                     d["ncat"].data_ptr(), d["cat_off"].data_ptr(), d["cat_len"].data_ptr(), d["cats"].data_ptr(),
                                     out.data_ptr(), C.c_void_p(st)))
        return out

    def predict_raw(self, X, single_tree: bool | None = None):
        """Classification rawPrediction, torch.float64 [n, 2] on X's device (K7, b2rf_predict_raw): for an ensemble the per-tree
        normalised class counts summed in tree order (RandomForestClassificationModel.predictRaw); for a single
        DecisionTreeClassificationModel (single_tree, default numTrees == 1) the leaf's class counts."""
        import torch
        from ._lib import check, lib
        if not self.is_classification:
            raise ValueError("predict_raw: this is a regression model")
        if not (isinstance(X, torch.Tensor) and X.is_cuda):
            raise TypeError("Forest.predict_raw needs a CUDA tensor: this package has no CPU path")
        X = X.to(torch.float64).contiguous()
        d = self._device_arrays(X.device)
        mode = 1 if ((self.num_trees == 1) if single_tree is None else single_tree) else 0
        if self.counts is not None:     # multiclass (K7c)
            key = "counts:" + str(X.device)
            if key not in self._dev:
                self._dev[key] = torch.as_tensor(np.ascontiguousarray(self.counts), dtype=torch.float64).to(X.device)
            C_ = self.num_classes
            out = torch.empty((X.shape[0], C_), dtype=torch.float64, device=X.device)
            with torch.cuda.device(X.device):
                st = torch.cuda.current_stream().cuda_stream
                check(lib().b2rf_predict_raw_mc(X.data_ptr(), X.stride(0), X.shape[0], self.num_trees, d["tree_off"].data_ptr(), d["feature"].data_ptr(),
                                                d["threshold"].data_ptr(), d["left"].data_ptr(), d["right"].data_ptr(), self._dev[key].data_ptr(), C_,
                                                d["ncat"].data_ptr(), d["cat_off"].data_ptr(), d["cat_len"].data_ptr(), d["cats"].data_ptr(),
                                                mode, out.data_ptr(), C.c_void_p(st)))
            return out
        out = torch.empty((X.shape[0], 2), dtype=torch.float64, device=X.device)
        with torch.cuda.device(X.device):
            st = torch.cuda.current_stream().cuda_stream
            check(lib().b2rf_predict_raw(X.data_ptr(), X.stride(0), X.shape[0], self.num_trees, d["tree_off"].data_ptr(), d["feature"].data_ptr(),
                                         d["threshold"].data_ptr(), d["left"].data_ptr(), d["right"].data_ptr(), d["stats"].data_ptr(),
                                         d["ncat"].data_ptr(), d["cat_off"].data_ptr(), d["cat_len"].data_ptr(), d["cats"].data_ptr(),
                                         mode, out.data_ptr(), C.c_void_p(st)))
        return out

    def predict_proba(self, X, single_tree: bool | None = None):
        """ProbabilisticClassificationModel.normalizeToProbabilitiesInPlace on rawPrediction (rows whose raw sum is 0 stay 0)."""
        raw = self.predict_raw(X, single_tree)
        tot = raw.sum(dim=1, keepdim=True)
        return raw / tot.where(tot != 0, tot.new_ones(()))
