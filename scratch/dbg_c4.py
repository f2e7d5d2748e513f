import sys, numpy as np, torch
sys.path.insert(0,'.'); sys.path.insert(0,'tests')
import b2rf
from oracle import oracle
from test_gpu_parity import _fit_both
cuda=torch.device('cuda:0')
X, y, arity = b2rf.synth.c4(120000, device="cpu")
g, o = _fit_both(b2rf, oracle, cuda, X.numpy(), y.numpy(), arity, maxDepth=8, maxBins=256, numTrees=4, bootstrap=True, seed=42, maxMemoryInMB=8)
d = np.nonzero((g.trace_i != o.trace_i).any(1))[0]
print("n diff rows", len(d), "of", len(g.trace_i))
for i in d[:5]:
    print(i, "gpu", g.trace_i[i], [repr(float(v)) for v in g.trace_d[i]])
    print(i, "ora", o.trace_i[i], [repr(float(v)) for v in o.trace_d[i]])
    f=g.trace_i[i][3]; print(" arity gpu feat", arity[g.trace_i[i][3]], "ora feat", arity[o.trace_i[i][3]])
