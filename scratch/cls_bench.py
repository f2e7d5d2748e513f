"""Classifier fit throughput on the C3 shape (10M rows x 64 features, 100 trees, depth 10, maxBins 64), labels = the C3 label cut
into 2 / 4 classes at its quantiles.  featureSubsetStrategy auto -> sqrt (8 of 64 features per node)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import b2rf

dev = torch.device("cuda:0")
n, F = 10_000_000, 64
X, y, arity = b2rf.synth.c3(n, device=dev, F=F)
po = list(range(0, n + 1, 4096)) + ([n] if n % 4096 else [])
samp = y[:: 1000].cpu().numpy()
for C_, strategy in [(0, "auto"), (2, "auto"), (4, "auto"), (2, "onethird"), (4, "onethird")]:
    if C_ == 0:
        lab, p = y, b2rf.TreeParams(maxDepth=10, maxBins=64, numTrees=100, bootstrap=True, seed=42)
        name = "regression (variance, onethird = 22 features/node)"
    else:
        cuts = torch.as_tensor(np.quantile(samp, np.linspace(0, 1, C_ + 1)[1:-1]), device=dev)
        lab = torch.bucketize(y, cuts).to(torch.float64)
        p = b2rf.TreeParams(maxDepth=10, maxBins=64, numTrees=100, bootstrap=True, seed=42, impurity="gini", numClasses=C_, featureSubsetStrategy=strategy)
        name = f"gini, {C_} classes, featureSubsetStrategy={strategy} ({p.features_per_node(F)} features/node)"
    ts = []
    for rep in range(3):
        torch.cuda.synchronize(); t = time.perf_counter()
        f = b2rf.trainer.fit_tensors(X, lab, arity, p, part_offsets=po)
        torch.cuda.synchronize(); ts.append(time.perf_counter() - t)
    best = min(ts[1:])
    print(f"{name}: {best*1e3:.1f} ms per fit, {n*100/best/1e9:.2f} G rows*trees/s, K4 {f.timers['findBestSplits']*1e3:.1f} ms, nodes {f.num_nodes()}", flush=True)
