"""Writes the golden fixtures from the CPU oracle (run from the repo root: python tests/golden/make_golden.py).
The reference repo stores no outputs for this path and MLlib cannot run here, so the vectors pin the oracle (which itself
reproduces every upstream known answer in SURVEY.md Appendix B); labels are integer-valued, so every sum is exact and the
fixtures are independent of summation order."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import b2rf  # noqa: E402
from oracle import oracle as O  # noqa: E402

CASES = {
    "c1_dt_depth5_bins40": dict(rows=5716, data_seed=42, part_offsets=[0, 1500, 1500, 4000, 5716],
                                params=dict(maxDepth=5, maxBins=40)),
    "c2_rf_t10_depth6_bins40": dict(rows=3000, data_seed=7, part_offsets=[0, 700, 1400, 2100, 3000],
                                    params=dict(maxDepth=6, maxBins=40, numTrees=10, bootstrap=True, seed=42)),
    "c2_rf_t6_depth8_groups": dict(rows=2500, data_seed=11, part_offsets=[0, 2500],
                                   params=dict(maxDepth=8, maxBins=40, numTrees=6, bootstrap=True, seed=5, maxMemoryInMB=1,
                                               featureSubsetStrategy="sqrt", subsamplingRate=0.8)),
}

for name, c in CASES.items():
    X, y, arity = b2rf.synth.airbnb(c["rows"], seed=c["data_seed"])
    f = O.fit(X.numpy(), y.numpy(), arity, part_offsets=c["part_offsets"], **c["params"])
    out = dict(config=c, tree_off=f.tree_off.tolist(), feature=f.feature.tolist(), threshold_hex=[float(v).hex() for v in f.threshold],
               prediction_hex=[float(v).hex() for v in f.prediction], count=f.stats[:, 0].tolist(),
               cats=[sorted(f.cats[o:o + n].tolist()) for o, n in zip(f.cat_off, f.cat_len)])
    json.dump(out, open(os.path.join(os.path.dirname(os.path.abspath(__file__)), name + ".json"), "w"))
    print(name, len(f.feature), "nodes")
