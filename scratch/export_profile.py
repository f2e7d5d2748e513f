import os, sys, time, cProfile, pstats
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, b2rf
from b2rf import trainer
dev = torch.device("cuda:0")
n = 2_000_000
X, y, ar = b2rf.synth.c3(n, device=dev, F=64)
p = b2rf.TreeParams(maxDepth=10, maxBins=64, numTrees=100, bootstrap=True, seed=42)
po = list(range(0, n + 1, 4096)) + ([n] if n % 4096 else [])
trainer.fit_tensors(X, y, ar, p, part_offsets=po)
orig = trainer._export
def timed(*a, **k):
    pr = cProfile.Profile(); pr.enable(); t = time.perf_counter()
    r = orig(*a, **k)
    dt = time.perf_counter() - t; pr.disable()
    print(f"_export {dt*1e3:.1f} ms, nodes {r.num_nodes()}")
    pstats.Stats(pr).sort_stats("cumulative").print_stats(12)
    return r
trainer._export = timed
trainer.fit_tensors(X, y, ar, p, part_offsets=po)
