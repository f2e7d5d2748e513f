"""Fit latency on the labs' own shapes (C1: ML 06 DT depth 5; C2: ML 07 RF 100 trees depth 8) — GPU vs the CPU oracle."""
import sys, time, numpy as np, torch
sys.path.insert(0, '.')
import b2rf
from b2rf import trainer
from oracle import oracle as O
dev = torch.device('cuda:0')
X, y, arity = b2rf.synth.airbnb(7146)
X, y = X[:5716], y[:5716]
Xd, yd = X.to(dev), y.to(dev)
po = trainer.default_partitions(5716, 1429)
for name, kw in [("C1 DT depth5", dict(maxDepth=5, maxBins=40)), ("C2 RF T=100 depth8", dict(maxDepth=8, maxBins=40, numTrees=100, bootstrap=True, seed=42)),
                 ("ML07 grid RF T=10 depth5", dict(maxDepth=5, maxBins=40, numTrees=10, bootstrap=True, seed=42))]:
    p = b2rf.TreeParams(**kw)
    for _ in range(3): f = trainer.fit_tensors(Xd, yd, arity, p, part_offsets=po)
    torch.cuda.synchronize(); t = time.perf_counter()
    for _ in range(10): f = trainer.fit_tensors(Xd, yd, arity, p, part_offsets=po)
    torch.cuda.synchronize(); g = (time.perf_counter() - t) / 10
    t = time.perf_counter(); o = O.fit(X.numpy(), y.numpy(), arity, part_offsets=po, numThreads=16, **kw); c = time.perf_counter() - t
    print(f"{name}: gpu {g*1e3:.1f} ms (groups {f.counters['groups']}, launches {f.counters['kernel_launches']}, timers {dict((k, round(v,4)) for k,v in f.timers.items())}); oracle 16 threads {c*1e3:.1f} ms")
