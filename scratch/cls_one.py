"""One 4-class RandomForestClassifier-shaped fit on the C3 matrix (10M x 64, 100 trees, depth 10, 8 features per node) — the ncu target for K4c."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, b2rf
dev = torch.device("cuda:0")
n, F = 10_000_000, 64
X, y, arity = b2rf.synth.c3(n, device=dev, F=F)
po = list(range(0, n + 1, 4096)) + ([n] if n % 4096 else [])
cuts = torch.as_tensor(np.quantile(y[::1000].cpu().numpy(), [0.25, 0.5, 0.75]), device=dev)
lab = torch.bucketize(y, cuts).to(torch.float64)
p = b2rf.TreeParams(maxDepth=10, maxBins=64, numTrees=100, bootstrap=True, seed=42, impurity="gini", numClasses=4)
f = b2rf.trainer.fit_tensors(X, lab, arity, p, part_offsets=po)
torch.cuda.synchronize()
print("K4c", f.timers["findBestSplits"], "nodes", f.num_nodes())
