"""Writes tests/golden/persistence_schema.json: the Parquet schemas and metadata keys of the MLlib 3.1.2 model directory layout as
this package writes it (persistence.py).  The fixture pins the written format: Spark itself is unavailable on this box, so nothing
but this package has read these files back; a reader that disagrees with Spark's NodeData / EnsembleNodeData case classes would
show up as a diff against the schema strings below, which restate
  org.apache.spark.ml.tree.DecisionTreeModelReadWrite.NodeData / SplitData and EnsembleModelReadWrite.EnsembleNodeData (3.1.2).
Run from the repo root:  python tests/golden/make_persistence_schema.py"""
import json
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def describe(tmp):
    import numpy as np
    import pyarrow.parquet as pq
    import b2rf
    from b2rf import spark as sp
    out = {}
    # a hand-made two-tree forest: root split on a continuous feature, one categorical split
    tree_off = np.array([0, 3, 6], np.int64)
    feature = np.array([0, -1, -1, 1, -1, -1], np.int32)
    threshold = np.array([0.5, 0, 0, 0, 0, 0], np.float64)
    left = np.array([1, -1, -1, 1, -1, -1], np.int32); right = np.array([2, -1, -1, 2, -1, -1], np.int32)
    pred = np.array([1.0, 0.5, 2.0, 1.0, 0.0, 3.0]); imp = np.array([1.0, 0, 0, 1.0, 0, 0]); gain = np.array([0.5, 0, 0, 0.25, 0, 0])
    stats = np.array([[4, 4, 8, 4], [2, 1, 1, 2], [2, 3, 7, 2], [4, 4, 8, 4], [1, 0, 0, 1], [3, 4, 8, 3]], np.float64)
    ncat = np.array([0, 0, 0, 3, 0, 0], np.int32); cat_off = np.array([0, 0, 0, 0, 0, 0], np.int32); cat_len = np.array([0, 0, 0, 2, 0, 0], np.int32)
    cats = np.array([0.0, 2.0])
    for kind, est, T in (("DecisionTreeRegressionModel", sp.B200DecisionTreeRegressor, 1), ("RandomForestRegressionModel", sp.B200RandomForestRegressor, 2)):
        sl = slice(0, 3) if T == 1 else slice(0, 6)
        fo = b2rf.Forest(2, tree_off[: T + 1], feature[sl], threshold[sl], left[sl], right[sl], pred[sl], imp[sl], gain[sl], stats[sl], ncat[sl], cat_off[sl],
                         cat_len[sl], cats)
        m = est(maxDepth=1)._make_model(fo)
        path = os.path.join(tmp, kind)
        m.write().overwrite().save(path)
        meta = json.loads(open(os.path.join(path, "metadata", "part-00000")).readline())
        rec = {"directories": sorted(d for d in os.listdir(path)), "metadata_keys": sorted(meta), "class": meta["class"], "sparkVersion": meta["sparkVersion"],
               "data_schema": str(pq.read_schema(os.path.join(path, "data", "part-00000.parquet"))).splitlines()}
        if T > 1:
            rec["treesMetadata_schema"] = str(pq.read_schema(os.path.join(path, "treesMetadata", "part-00000.parquet"))).splitlines()
        out[kind] = rec
    return out


if __name__ == "__main__":
    with tempfile.TemporaryDirectory() as tmp:
        d = describe(tmp)
    fn = os.path.join(os.path.dirname(os.path.abspath(__file__)), "persistence_schema.json")
    json.dump(d, open(fn, "w"), indent=1)
    print(json.dumps(d, indent=1))
