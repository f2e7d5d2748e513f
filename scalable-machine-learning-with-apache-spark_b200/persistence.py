"""MLlib-compatible model persistence: `model.write().overwrite().save(path)` / `Model.load(path)`
(`Labs/ML 07L - Hyperparameter Tuning Lab.py:209` saves the fitted pipeline; `ML 08 - Hyperopt.py:170` logs it to MLflow).

Layout written (Spark 3.1.2 `DecisionTreeModelReadWrite` / `EnsembleModelReadWrite`):
    <path>/metadata/part-00000          one JSON line: class, timestamp, sparkVersion, uid, paramMap, defaultParamMap,
                                        numFeatures [, numTrees]
    <path>/data/part-00000.parquet      NodeData rows (trees: one row per node; forests: (treeID, nodeData))
    <path>/treesMetadata/part-00000.parquet   forests only: (treeID, metadata JSON of the tree, weights)
NodeData = (id, prediction, impurity, impurityStats[], rawCount, gain, leftChild, rightChild,
            split(featureIndex, leftCategoriesOrThreshold[], numCategories)), ids in pre-order.
Written blind against the 3.1.2 format (no Spark here to read it back); the round trip below is what is tested.
"""
from __future__ import annotations

import json
import os
import shutil
import time

import numpy as np

from .forest import Forest

SPARK_VERSION = "3.1.2"
_DT_CLASS = "org.apache.spark.ml.regression.DecisionTreeRegressionModel"
_RF_CLASS = "org.apache.spark.ml.regression.RandomForestRegressionModel"
_DTC_CLASS = "org.apache.spark.ml.classification.DecisionTreeClassificationModel"
_RFC_CLASS = "org.apache.spark.ml.classification.RandomForestClassificationModel"


def _node_rows(fo: Forest, t: int):
    s = fo.tree_slice(t)
    base = s.start
    rows = []
    for i in range(s.stop - s.start):
        a = base + i
        leaf = fo.feature[a] < 0
        if leaf:
            split = {"featureIndex": -1, "leftCategoriesOrThreshold": [], "numCategories": -1}
        elif fo.ncat[a] > 0:
            o, n = int(fo.cat_off[a]), int(fo.cat_len[a])
            split = {"featureIndex": int(fo.feature[a]), "leftCategoriesOrThreshold": [float(c) for c in fo.cats[o:o + n]],
                     "numCategories": int(fo.ncat[a])}
        else:
            split = {"featureIndex": int(fo.feature[a]), "leftCategoriesOrThreshold": [float(fo.threshold[a])], "numCategories": -1}
        # impurityStats: regression (weight, sum, sum of squares); classification the class counts
        ist = [float(v) for v in fo.class_counts[a]] if fo.is_classification else [float(v) for v in fo.stats[a, :3]]
        rows.append({"id": i, "prediction": float(fo.prediction[a]), "impurity": float(fo.impurity[a]),
                     "impurityStats": ist, "rawCount": int(fo.stats[a, 3]),
                     "gain": -1.0 if leaf else float(fo.gain[a]), "leftChild": -1 if leaf else int(fo.left[a]),
                     "rightChild": -1 if leaf else int(fo.right[a]), "split": split})
    return rows


def _pa():
    import pyarrow as pa
    import pyarrow.parquet as pq
    split_t = pa.struct([("featureIndex", pa.int32()), ("leftCategoriesOrThreshold", pa.list_(pa.float64())), ("numCategories", pa.int32())])
    node_t = pa.struct([("id", pa.int32()), ("prediction", pa.float64()), ("impurity", pa.float64()), ("impurityStats", pa.list_(pa.float64())),
                        ("rawCount", pa.int64()), ("gain", pa.float64()), ("leftChild", pa.int32()), ("rightChild", pa.int32()), ("split", split_t)])
    return pa, pq, node_t


def _param_maps(model):
    pm = {p.name: v for p, v in model._paramMap.items()}
    dm = {p.name: v for p, v in model._defaultParamMap.items()}
    return pm, dm


def save_model(model, path: str, overwrite: bool = False):
    pa, pq, node_t = _pa()
    if os.path.exists(path):
        if not overwrite:
            raise IOError(f"Path {path} already exists. To overwrite it, please use write.overwrite().save(path) for Scala and use "
                          f"write().overwrite().save(path) for Java and Python.")
        shutil.rmtree(path)
    fo: Forest = model._forest
    ensemble = bool(getattr(model, "_ensemble", fo.num_trees > 1))
    os.makedirs(os.path.join(path, "metadata"))
    os.makedirs(os.path.join(path, "data"))
    pm, dm = _param_maps(model)
    cls_dt, cls_rf = (_DTC_CLASS, _RFC_CLASS) if fo.is_classification else (_DT_CLASS, _RF_CLASS)
    meta = {"class": cls_rf if ensemble else cls_dt, "timestamp": int(time.time() * 1000), "sparkVersion": SPARK_VERSION,
            "uid": model.uid, "paramMap": pm, "defaultParamMap": dm, "numFeatures": int(fo.num_features)}
    if fo.is_classification:
        meta["numClasses"] = int(fo.num_classes)
    if ensemble:
        meta["numTrees"] = int(fo.num_trees)
    with open(os.path.join(path, "metadata", "part-00000"), "w") as f:
        f.write(json.dumps(meta) + "\n")
    open(os.path.join(path, "metadata", "_SUCCESS"), "w").close()
    if not ensemble:
        rows = _node_rows(fo, 0)
        tbl = pa.Table.from_pylist(rows, schema=pa.schema(list(node_t)))
    else:
        rows = [{"treeID": t, "nodeData": r} for t in range(fo.num_trees) for r in _node_rows(fo, t)]
        tbl = pa.Table.from_pylist(rows, schema=pa.schema([("treeID", pa.int32()), ("nodeData", node_t)]))
        os.makedirs(os.path.join(path, "treesMetadata"))
        tm = [{"treeID": t, "metadata": json.dumps({"class": cls_dt, "timestamp": meta["timestamp"], "sparkVersion": SPARK_VERSION,
                                                    "uid": f"{'dtc' if fo.is_classification else 'dtr'}_{model.uid}_{t}", "paramMap": {}, "defaultParamMap": dm,
                                                    "numFeatures": int(fo.num_features)}), "weights": 1.0} for t in range(fo.num_trees)]
        pq.write_table(pa.Table.from_pylist(tm, schema=pa.schema([("treeID", pa.int32()), ("metadata", pa.string()), ("weights", pa.float64())])),
                       os.path.join(path, "treesMetadata", "part-00000.parquet"))
        open(os.path.join(path, "treesMetadata", "_SUCCESS"), "w").close()
    pq.write_table(tbl, os.path.join(path, "data", "part-00000.parquet"))
    open(os.path.join(path, "data", "_SUCCESS"), "w").close()


def load_forest(path: str):
    """Returns (metadata dict, Forest) from a directory written by save_model (or by Spark 3.x for the same model classes)."""
    _, pq, _ = _pa()
    meta = json.loads(open(os.path.join(path, "metadata", "part-00000")).readline())
    import glob
    files = sorted(glob.glob(os.path.join(path, "data", "*.parquet")))
    rows = []
    for fn in files:
        rows += pq.read_table(fn).to_pylist()
    ensemble = meta["class"].endswith(("RandomForestRegressionModel", "RandomForestClassificationModel"))
    classification = meta["class"].endswith("ClassificationModel")
    num_classes = int(meta.get("numClasses", 2)) if classification else 0
    if classification and not (2 <= num_classes <= 32):
        raise ValueError(f"{path}: numClasses = {num_classes}; classification models with 2..32 classes can be loaded")
    counts = []
    if ensemble:
        T = int(meta["numTrees"])
        per = [[] for _ in range(T)]
        for r in rows:
            per[r["treeID"]].append(r["nodeData"])
    else:
        per = [rows]
    tree_off = [0]
    cols = {k: [] for k in ("feature", "threshold", "left", "right", "prediction", "impurity", "gain", "stats", "ncat", "cat_off", "cat_len")}
    cats = []
    for nodes in per:
        nodes = sorted(nodes, key=lambda r: r["id"])
        for r in nodes:
            sp = r["split"]
            leaf = r["leftChild"] < 0
            cat = (not leaf) and sp["numCategories"] >= 0
            cols["feature"].append(-1 if leaf else sp["featureIndex"])
            cols["threshold"].append(0.0 if (leaf or cat) else sp["leftCategoriesOrThreshold"][0])
            cols["left"].append(r["leftChild"]); cols["right"].append(r["rightChild"])
            cols["prediction"].append(r["prediction"]); cols["impurity"].append(r["impurity"]); cols["gain"].append(0.0 if leaf else r["gain"])
            st = list(r["impurityStats"]) + [0.0] * 3
            if classification and num_classes > 2:
                cc = [float(v) for v in r["impurityStats"]][:num_classes]
                counts.append(cc + [0.0] * (num_classes - len(cc)))
                cols["stats"].append([float(sum(counts[-1])), 0.0, 0.0, float(r["rawCount"])])
            elif classification:   # class counts (c0, c1) -> the trainer's (sum w, sum w*y, sum w*y*y) with y in {0, 1}
                cols["stats"].append([st[0] + st[1], st[1], st[1], float(r["rawCount"])])
            else:
                cols["stats"].append([st[0], st[1], st[2], float(r["rawCount"])])
            cols["ncat"].append(sp["numCategories"] if cat else 0)
            cols["cat_off"].append(len(cats))
            cols["cat_len"].append(len(sp["leftCategoriesOrThreshold"]) if cat else 0)
            if cat:
                cats += list(sp["leftCategoriesOrThreshold"])
        tree_off.append(len(cols["feature"]))
    fo = Forest(int(meta["numFeatures"]), np.array(tree_off, np.int64), np.array(cols["feature"], np.int32), np.array(cols["threshold"]),
                np.array(cols["left"], np.int32), np.array(cols["right"], np.int32), np.array(cols["prediction"]), np.array(cols["impurity"]),
                np.array(cols["gain"]), np.array(cols["stats"]).reshape(-1, 4), np.array(cols["ncat"], np.int32),
                np.array(cols["cat_off"], np.int32), np.array(cols["cat_len"], np.int32), np.array(cats, np.float64), uid=meta["uid"])
    if classification:
        imp = meta.get("paramMap", {}).get("impurity", meta.get("defaultParamMap", {}).get("impurity", "gini"))
        fo.impurity_kind = str(imp).lower()
        if num_classes > 2:
            fo.counts = np.array(counts, np.float64).reshape(-1, num_classes)
    return meta, fo


class ModelWriter:
    """pyspark.ml.util.MLWriter look-alike: model.write().overwrite().save(path)."""

    def __init__(self, model):
        self._model, self._overwrite = model, False

    def overwrite(self):
        self._overwrite = True
        return self

    def save(self, path):
        save_model(self._model, path, self._overwrite)
