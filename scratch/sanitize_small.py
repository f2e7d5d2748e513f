"""Small end-to-end run for compute-sanitizer (memcheck / racecheck): every kernel, every K4 mode, uint8 and uint16 codes."""
import sys, numpy as np, torch
sys.path.insert(0, '.')
import b2rf
from b2rf import trainer
dev = torch.device('cuda:0')
X, y, arity = b2rf.synth.airbnb(2500)
Xd, yd = X.to(dev), y.to(dev)
po = trainer.default_partitions(2500, 600)
ref = None
for mode in ("segment", "sort", "scan"):
    f = trainer.fit_tensors(Xd, yd, arity, b2rf.TreeParams(maxDepth=5, maxBins=40, numTrees=3, bootstrap=True, seed=1), part_offsets=po, hist_mode=mode)
    p = f.predict(Xd[:100]).cpu().numpy()
    if ref is None: ref = f
    assert np.array_equal(f.feature, ref.feature), mode
    print(mode, f.num_nodes(), flush=True)
Xc, yc, ar = b2rf.synth.c3(20000, device="cpu", F=6)
for mode in ("segment", "scan"):
    f = trainer.fit_tensors(Xc.to(dev), yc.to(dev), ar, b2rf.TreeParams(maxDepth=4, maxBins=300, numTrees=2, bootstrap=True, seed=2), hist_mode=mode)
    print("u16", mode, f.num_nodes(), flush=True)
f = trainer.fit_host(X.numpy(), y.numpy(), arity, b2rf.TreeParams(maxDepth=4, maxBins=40, numTrees=2, bootstrap=False, subsamplingRate=0.7, seed=3), device=dev, chunk_rows=700)
m = trainer.fit_many_tensors(Xd, yd, arity, [b2rf.TreeParams(maxDepth=3, maxBins=40, numTrees=2, bootstrap=True, seed=4), b2rf.TreeParams(maxDepth=2, maxBins=40)])
print("ok", f.num_nodes(), [q.num_nodes() for q in m])
# round 2 kernels: limb forms (wide / narrow / two-plane) at feature groups of 32 / 16 / 8 with register and TMA staging, the staged and the
# gather class-count kernels, the device sampler ids, the sharded-K5 pack / unpack path emulated on one GPU
import ctypes as C
for F, strat in [(64, "all"), (64, "onethird"), (40, "onethird"), (12, "all"), (12, "sqrt"), (140, "onethird"), (100, "all")]:
    Xs, ys, ars = b2rf.synth.c3(6000, device="cpu", F=F)
    for lab, imp in [(ys, "variance"), (torch.round(ys * 3), "variance"), ((ys > ys.median()).double(), "gini")]:
        for tma in (0, 1):
            trainer.set_option("k4_tma", tma)
            f = trainer.fit_tensors(Xs.to(dev), lab.to(dev), ars, b2rf.TreeParams(maxDepth=5, maxBins=64, numTrees=2, bootstrap=True, seed=5, impurity=imp,
                                                                                   featureSubsetStrategy=strat), part_offsets=trainer.default_partitions(6000, 2048))
    print("limb", F, strat, f.num_nodes(), flush=True)
trainer.set_option("k4_tma", 0)
for F in (50, 100, 70 + 66):
    Xs, ys, ars = b2rf.synth.c3(6000, device="cpu", F=F)
    lab = torch.bucketize(ys, torch.quantile(ys, torch.tensor([0.25, 0.5, 0.75], dtype=torch.float64))).double()
    f = trainer.fit_tensors(Xs.to(dev), lab.to(dev), ars, b2rf.TreeParams(maxDepth=5, maxBins=32, numTrees=3, bootstrap=True, seed=6, impurity="gini", numClasses=4))
    print("multiclass", F, f.num_nodes(), flush=True)
ids = trainer.sample_row_ids(42, 0, 50000, 0.2, dev)
print("sampler ids", len(ids), flush=True)
print("round 2 kernels ok")
