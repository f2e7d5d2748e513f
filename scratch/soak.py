"""Repeated fits of different shapes in one process: device memory in use (cudaMemGetInfo) must not creep."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, b2rf
from b2rf import trainer
dev = torch.device("cuda:0")
def used():
    torch.cuda.synchronize(); free, total = torch.cuda.mem_get_info(dev); return (total - free) / 2**30
shapes = [(2_000_000, 64, 50, 10, "auto", "variance"), (1_000_000, 33, 20, 8, "sqrt", "gini"), (500_000, 128, 10, 6, "onethird", "variance"),
          (3_000_000, 12, 30, 9, "all", "variance")]
log = []
for rep in range(12):
    for n, F, T, d, strat, imp in shapes:
        X, y, ar = b2rf.synth.c3(n, device=dev, F=F, seed=rep + 1)
        if imp == "gini":
            y = (y > y.median()).double()
        f = trainer.fit_tensors(X, y, ar, b2rf.TreeParams(maxDepth=d, maxBins=64, numTrees=T, bootstrap=True, seed=rep, impurity=imp, featureSubsetStrategy=strat))
        del X, y, f
    torch.cuda.empty_cache()
    log.append(used())
    print(f"round {rep}: device memory in use after the round {log[-1]:.2f} GiB", flush=True)
assert max(log[3:]) - min(log[3:]) < 0.5, log
print("soak ok")
