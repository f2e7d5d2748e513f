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
