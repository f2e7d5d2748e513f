"""BASELINE config 5: ML 07-style CrossValidator, 3 folds x ParamGrid maxDepth{2,5,8} x numTrees{5,10,50} on 10M-row synthetic (C3 data).
Per fold: train the 9 candidates on the fold's training rows (2/3 of the data), predict the validation rows, RMSE; sequential fits
vs the shared-pass fit_many_tensors."""
import sys, time, numpy as np, torch
sys.path.insert(0, '.')
import b2rf
from b2rf import trainer, synth
dev = torch.device('cuda:0')
N = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
X, y, arity = synth.c3(N, device=dev)
idx = torch.arange(N, dtype=torch.int64, device=dev)
fold = torch.clamp((synth.uniform(42, 9999, idx) * 3).to(torch.int64), max=2)
grid = [b2rf.TreeParams(maxDepth=d, maxBins=64, numTrees=t, bootstrap=True, seed=42) for d in (2, 5, 8) for t in (5, 10, 50)]
def run(shared):
    torch.cuda.synchronize(); t0 = time.perf_counter(); rmses = np.zeros(len(grid))
    for k in range(3):
        tr = fold != k
        Xt, yt, Xv, yv = X[tr], y[tr], X[~tr], y[~tr]
        po = trainer.default_partitions(Xt.shape[0], 4096)
        forests = trainer.fit_many_tensors(Xt, yt, arity, grid, part_offsets=po) if shared else [trainer.fit_tensors(Xt, yt, arity, p, part_offsets=po) for p in grid]
        for i, f in enumerate(forests):
            rmses[i] += float(torch.sqrt(torch.mean((f.predict(Xv) - yv) ** 2)).item()) / 3
    torch.cuda.synchronize()
    return time.perf_counter() - t0, rmses
for shared in (False, True, False, True):
    dt, r = run(shared)
    print(f"{'shared-pass fit_many' if shared else 'sequential fits    '}: {dt:.3f} s for 3 folds x 9 candidates (195 trees per fold) on {N} rows; best avg RMSE {r.min():.4f} (candidate {int(r.argmin())})", flush=True)
