import sys, os, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b2rf
from b2rf import trainer, synth
dev = torch.device('cuda:0')
N = int(sys.argv[1]) if len(sys.argv) > 1 else 300_000
X, y, arity = synth.c3(N, device=dev)
grid = [b2rf.TreeParams(maxDepth=d, maxBins=64, numTrees=t, bootstrap=True, seed=42) for d in (2, 5, 8) for t in (5, 10, 50)]
po = trainer.default_partitions(N, 4096)
many = trainer.fit_many_tensors(X, y, arity, grid, part_offsets=po)
torch.cuda.synchronize()
print("fit_many ok", [f.num_nodes() for f in many])
for p, f in zip(grid, many):
    ref = trainer.fit_tensors(X, y, arity, p, part_offsets=po)
    same = np.array_equal(f.feature, ref.feature) and np.array_equal(f.threshold, ref.threshold) and np.array_equal(f.stats[:, 0], ref.stats[:, 0])
    print(p.maxDepth, p.numTrees, "same structure:", same, "max rel stat diff", float(np.max(np.abs(f.stats[:, 1:3] - ref.stats[:, 1:3]) / np.maximum(1e-300, np.abs(ref.stats[:, 1:3])))) if same else "")
