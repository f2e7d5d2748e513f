import sys, os, time, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b2rf
from b2rf import trainer
dev = torch.device("cuda:0")
n, F = 10_000_000, 64
X, y, arity = b2rf.synth.c3(n, device=dev)
p = b2rf.TreeParams(maxDepth=10, maxBins=64, numTrees=100, bootstrap=True, seed=42)
po = trainer.default_partitions(n, 4096)
for it in range(6):
    torch.cuda.synchronize(); t1 = time.perf_counter()
    g = trainer.fit_tensors(X, y, arity, p, part_offsets=po)
    torch.cuda.synchronize(); t2 = time.perf_counter()
    print(f"iter {it}: fit_tensors {t2 - t1:.4f} s", flush=True)
Xh = torch.empty((n, F), dtype=torch.float64, pin_memory=True); Xh.copy_(X)
yh = torch.empty(n, dtype=torch.float64, pin_memory=True); yh.copy_(y)
del X
torch.cuda.empty_cache()
for it in range(14):
    torch.cuda.synchronize(); t1 = time.perf_counter()
    g = trainer.fit_host(Xh, yh, arity, p, device=dev, part_offsets=po)
    torch.cuda.synchronize(); t2 = time.perf_counter()
    tm = g.timers
    print(f"iter {it}: fit_host {t2 - t1:.4f} s  init {tm['init']:.3f} fs {tm['findSplits']:.3f} loop {tm['loop']:.3f}", flush=True)
