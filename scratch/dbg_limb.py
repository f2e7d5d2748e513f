"""Debug: compare the reduced group histograms of the limb kernel and the fp64 kernel on the same fit, cell by cell."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import b2rf
from b2rf import trainer

dev = torch.device("cuda:0")
N, T, depth, mem = 40000, 5, 10, int(os.environ.get("MEM", "1"))
X, y, arity = b2rf.synth.c3(N, device="cpu")
Xd, yd = X.to(dev), y.to(dev)
po = trainer.default_partitions(N, 1024)
p = b2rf.TreeParams(maxDepth=depth, maxBins=64, numTrees=T, bootstrap=True, seed=42, maxMemoryInMB=mem)
out = {}
for name, flag in (("fp64", 1), ("limb", 0)):
    trainer.set_option("k4_fp64", flag)
    groups = []
    f = trainer.fit_tensors(Xd, yd, arity, p, part_offsets=po, keep_trace=True, hist_hook=lambda info, h: groups.append((info, h.cpu().numpy().copy())))
    out[name] = (f, groups)
    print(name, "groups", len(groups), "nodes", f.num_nodes(), trainer.k4_launch_counts())
ga, gb = out["fp64"][1], out["limb"][1]
for gi, ((ia, ha), (ib, hb)) in enumerate(zip(ga, gb)):
    if ha.shape != hb.shape or not np.array_equal(ia["node_id"], ib["node_id"]):
        print("group", gi, "structure differs (earlier divergence)"); break
    bad = np.nonzero((ha[:, 0] != hb[:, 0]) | (np.abs(ha[:, 1] - hb[:, 1]) > 1e-9 * np.maximum(1, np.abs(ha[:, 1]))) |
                     (np.abs(ha[:, 2] - hb[:, 2]) > 1e-9 * np.maximum(1, np.abs(ha[:, 2]))))[0]
    if len(bad):
        print("group", gi, "slots", len(ia["tree"]), "cells", ha.shape[0], "bad cells", len(bad))
        base = ia["slot_base"]
        for c in bad[:12]:
            s = int(np.searchsorted(base, c, side="right") - 1)
            fo = ia["slot_fo"][s]
            j = int(np.searchsorted(fo[:, 1], c - base[s], side="right") - 1)
            print(f"  cell {c}: slot {s} (tree {ia['tree'][s]} node {ia['node_id'][s]}) j={j} feat={fo[j,0]} bin={c - base[s] - fo[j,1]}  fp64={ha[c,:3]} limb={hb[c,:3]}")
        # per-slot totals
        sl = sorted(set(int(np.searchsorted(base, c, side='right') - 1) for c in bad))
        print("  bad slots:", sl[:20], "of", len(ia["tree"]))
        for s in sl[:5]:
            fo = ia["slot_fo"][s]
            lo = base[s] + fo[0, 1]; hi = base[s] + (fo[1, 1] if len(fo) > 1 else 0)
            print(f"   slot {s}: feature0 totals fp64 count {ha[lo:hi,0].sum()} limb {hb[lo:hi,0].sum()}")
        break
else:
    print("all groups identical")

# ---- first differing decision ----
fa, fb = out["fp64"][0], out["limb"][0]
m = min(len(fa.trace_i), len(fb.trace_i))
dif = np.nonzero((fa.trace_i[:m] != fb.trace_i[:m]).any(1))[0]
if len(dif):
    k = int(dif[0])
    print("first differing decision at trace row", k)
    print(" fp64:", fa.trace_i[k], fa.trace_d[k])
    print(" limb:", fb.trace_i[k], fb.trace_d[k])
    gi = int(fa.trace_i[k][0]); tr = int(fa.trace_i[k][1]); nid = int(fa.trace_i[k][2])
    (ia, ha), (ib, hb) = ga[gi], gb[gi]
    s = [i for i in range(len(ia["tree"])) if ia["tree"][i] == tr and ia["node_id"][i] == nid][0]
    fo = ia["slot_fo"][s]; base = ia["slot_base"][s]
    nxt = ia["slot_base"][s + 1] if s + 1 < len(ia["tree"]) else ha.shape[0]
    A, B = ha[base:nxt], hb[base:nxt]
    print(" slot", s, "cells", nxt - base, "nan fp64", np.isnan(A).sum(), "nan limb", np.isnan(B).sum())
    d = np.nonzero((A[:, :3] != B[:, :3]).any(1))[0]
    print(" cells not bit-identical:", len(d), "of", len(A))
    for c in d[:10]:
        j = int(np.searchsorted(fo[:, 1], c, side="right") - 1)
        print(f"   cell {c} feat {fo[j,0]} bin {c - fo[j,1]}: fp64 {A[c,:3]!r} limb {B[c,:3]!r}")
    for nm, f_ in (("fp64", fa), ("limb", fb)):
        r = f_.trace_i[k]
        j = [q for q in range(len(fo)) if fo[q, 0] == r[3]]
        print(nm, "chose feature", r[3], "split", r[4], "in subset" if j else "NOT IN SUBSET")
