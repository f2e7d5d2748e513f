import sys, os, time, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b2rf
from b2rf import trainer, synth
dev = torch.device('cuda:0')
N = 6_670_000
X, y, arity = synth.c3(N, device=dev)
grid = [b2rf.TreeParams(maxDepth=d, maxBins=64, numTrees=t, bootstrap=True, seed=42) for d in (2, 5, 8) for t in (5, 10, 50)]
po = trainer.default_partitions(N, 4096)
for it in range(3):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    many = trainer.fit_many_tensors(X, y, arity, grid, part_offsets=po)
    torch.cuda.synchronize(); t1 = time.perf_counter()
    seq = [trainer.fit_tensors(X, y, arity, p, part_offsets=po) for p in grid]
    torch.cuda.synchronize(); t2 = time.perf_counter()
    print(f"iter {it}: shared {t1 - t0:.3f} s   sequential {t2 - t1:.3f} s", flush=True)
