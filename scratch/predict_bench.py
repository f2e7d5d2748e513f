"""K7 throughput: fit the C3 forest shape on 2M rows, predict 10M rows."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import b2rf

dev = torch.device("cuda:0")
n_fit, n_pred, F = 2_000_000, 10_000_000, 64
X, y, arity = b2rf.synth.c3(n_pred, device=dev, F=F)
p = b2rf.TreeParams(maxDepth=10, maxBins=64, numTrees=100, bootstrap=True, seed=42)
f = b2rf.trainer.fit_tensors(X[:n_fit], y[:n_fit], arity, p, part_offsets=list(range(0, n_fit + 1, 4096)) + ([n_fit] if n_fit % 4096 else []))
print("forest nodes", f.num_nodes(), flush=True)
for rep in range(3):
    torch.cuda.synchronize(); t = time.perf_counter()
    out = f.predict(X)
    torch.cuda.synchronize(); dt = time.perf_counter() - t
    print(f"predict {n_pred} rows x {f.num_trees} trees: {dt*1e3:.1f} ms  {n_pred*f.num_trees/dt/1e9:.2f} G row-trees/s", flush=True)
print("rmse", float(((out - y) ** 2).mean().sqrt()))
