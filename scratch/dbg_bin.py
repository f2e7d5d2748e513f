import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, b2rf
from b2rf import trainer
dev = torch.device("cuda:0")
for F, n in [(64, 30000), (33, 3000), (64, 3000), (64, 5000)]:
    X, y, arity = b2rf.synth.c3(n, device="cpu", F=F)
    yb = (y > y.median()).to(torch.float64)
    for strat in ["all", "sqrt"]:
        for binf in (1, 0):
            trainer.set_option("k4_binary", binf)
            p = b2rf.TreeParams(maxDepth=3, maxBins=64, numTrees=1, bootstrap=False, seed=6, impurity="gini", featureSubsetStrategy=strat)
            g = trainer.fit_tensors(X.to(dev), yb.to(dev), arity, p, part_offsets=trainer.default_partitions(n, 1024), keep_trace=True)
            print(F, n, strat, "binary" if binf else "narrow", g.trace_d[:3, 2:5].tolist(), "label sum", float(yb.sum()))
trainer.set_option("k4_binary", 1)
