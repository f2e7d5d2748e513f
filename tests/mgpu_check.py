"""Multi-GPU parity check, run under torchrun on N GPUs (not collected by pytest):
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tests/mgpu_check.py
Rows are sharded contiguously over the ranks; the sharded fit must equal the CPU oracle's fit of the whole dataset."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def run_cases(dev, rank, world, log=print) -> str:
    """Runs the sharded fits on an initialised process group and checks them on rank 0 against the CPU oracle's fit of the whole
    dataset.  Returns "ok" or the first failure (the same string on every rank).  Used by main() below and by bench.py, which
    runs it before its timed region whenever WORLD_SIZE > 1 and reports the result as "mgpu_parity"."""
    import b2rf
    from b2rf import trainer
    from helpers import assert_forest_parity, assert_parity_allowing_ties, assert_trace_parity
    from oracle import oracle as O
    ok = True
    first = ""
    for name, gen, n, kw in [("airbnb-rf", b2rf.synth.airbnb, 6001, dict(maxDepth=7, maxBins=40, numTrees=12, bootstrap=True, seed=42)),
                             ("airbnb-dt", b2rf.synth.airbnb, 5716, dict(maxDepth=5, maxBins=40)),
                             ("c3-slice", lambda m, row0=0: b2rf.synth.c3(m, row0=row0, device="cpu"), 90000,
                              dict(maxDepth=6, maxBins=64, numTrees=6, bootstrap=True, seed=42))]:
        cuts = [n * r // world for r in range(world + 1)]
        cuts[1] += 7 if world > 1 else 0                         # ragged shards
        lo, hi = cuts[rank], cuts[rank + 1]
        X, y, arity = gen(hi - lo, row0=lo)
        # partitions: 512 rows each inside a shard (a partition never spans ranks); global partition list = concatenation
        po_local = np.minimum(np.arange(0, hi - lo + 512, 512), hi - lo).astype(np.int64)
        p = b2rf.TreeParams(**kw)
        g = trainer.fit_tensors(X.to(dev), y.to(dev), arity, p, part_offsets=po_local, keep_trace=True)
        if rank == 0:
            Xa, ya, _ = gen(n, row0=0)
            po = [0]
            for r in range(world):
                m = cuts[r + 1] - cuts[r]
                loc = np.minimum(np.arange(0, m + 512, 512), m)[1:]
                po += [cuts[r] + int(v) for v in loc]
            mpb = min(p.maxBins, n)
            meta_max = max([a if a > 1 else mpb for a in arity])
            req = max(meta_max * meta_max, 10000)
            fraction = req / n if req < n else 1.0
            sidx = np.nonzero(O.sample_keep(p.seed, np.arange(n), fraction))[0] if fraction < 1 else None
            o = O.fit(Xa.numpy(), ya.numpy(), arity, part_offsets=po, sample_idx=sidx, fraction=fraction,
                      **{k: v for k, v in kw.items()})
            try:
                assert_parity_allowing_ties(g, o, max_ties=2)
                log(f"[mgpu] {name}: world={world} OK ({g.num_nodes()} nodes, groups={g.counters['groups']})")
            except AssertionError as e:
                ok = False
                first = first or f"{name}: {str(e)[:200]}"
                log(f"[mgpu] {name}: FAILED {e}")
    # ---- classifiers: binary (class counts ride on the regression statistics) and three classes (per-class histograms, all-reduced
    # as cells of 4 * ceil(C / 4) doubles); integer statistics, so the sharded fit must equal the restatement exactly ----
    from helpers import assert_classifier_parity
    from oracle import multiclass as M
    n = 6001
    cuts = [n * r // world for r in range(world + 1)]
    cuts[1] += 7 if world > 1 else 0
    lo, hi = cuts[rank], cuts[rank + 1]
    Xa, ya, arity = b2rf.synth.airbnb(n, row0=0)
    Xa, ya = Xa.numpy(), ya.numpy()
    po = [0]
    for r in range(world):
        m_ = cuts[r + 1] - cuts[r]
        po += [cuts[r] + int(v) for v in np.minimum(np.arange(0, m_ + 512, 512), m_)[1:]]
    po_local = np.minimum(np.arange(0, hi - lo + 512, 512), hi - lo).astype(np.int64)
    for name, C_ in [("airbnb-rf-binary", 2), ("airbnb-rf-3class", 3)]:
        qs = np.quantile(ya, np.linspace(0, 1, C_ + 1)[1:-1])
        lab = np.digitize(ya, qs).astype(np.float64)
        kw = dict(maxDepth=5, maxBins=40, numTrees=8, bootstrap=True, seed=42)
        p = b2rf.TreeParams(impurity="gini", numClasses=C_, **kw)
        g = trainer.fit_tensors(torch.as_tensor(Xa[lo:hi], device=dev), torch.as_tensor(lab[lo:hi], device=dev), arity, p, part_offsets=po_local,
                                keep_trace=True)
        if rank == 0:
            try:
                if C_ == 2:
                    assert_classifier_parity(g, O.fit(Xa, lab, arity, part_offsets=po, impurity="gini", **kw))
                else:
                    m = M.fit(Xa, lab, arity, C_, part_offsets=np.array(po), impurity="gini", **kw)
                    for a in ("tree_off", "feature", "threshold", "left", "right", "prediction", "impurity", "gain", "class_counts", "trace_i"):
                        assert np.array_equal(getattr(g, a), getattr(m, a)), a
                log(f"[mgpu] {name}: world={world} OK ({g.num_nodes()} nodes, groups={g.counters['groups']})")
            except AssertionError as e:
                ok = False
                first = first or f"{name}: {str(e)[:200]}"
                log(f"[mgpu] {name}: FAILED {e}")
    # ---- fitMultiple shared passes on sharded rows: every candidate's forest equals its own sharded fit_tensors fit (the merged K4
    # launch serves all candidates; each candidate's histogram is all-reduced separately) ----
    Xl = torch.as_tensor(Xa[lo:hi], device=dev)
    yl = torch.as_tensor(ya[lo:hi], device=dev)
    grid = [b2rf.TreeParams(maxDepth=d, maxBins=40, numTrees=t, bootstrap=True, seed=42) for d, t in ((2, 5), (5, 3), (6, 8))]
    many = trainer.fit_many_tensors(Xl, yl, arity, grid, part_offsets=po_local)
    same = True
    for p_, f_ in zip(grid, many):
        r_ = trainer.fit_tensors(Xl, yl, arity, p_, part_offsets=po_local)
        same = same and np.array_equal(f_.feature, r_.feature) and np.array_equal(f_.threshold, r_.threshold) and np.array_equal(f_.stats[:, 0], r_.stats[:, 0]) \
            and np.array_equal(f_.prediction, r_.prediction)
    if rank == 0:
        if same:
            log(f"[mgpu] fit_many shared passes: world={world} OK ({[f_.num_nodes() for f_ in many]} nodes)")
        else:
            ok = False
            first = first or "fit_many shared passes: a candidate differs from its fit_tensors fit"
            log("[mgpu] fit_many shared passes: FAILED")
    msg = [("ok" if ok else "FAILED " + first) if rank == 0 else ""]
    dist.broadcast_object_list(msg, src=0)
    dist.barrier()
    return msg[0]


def main():
    rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    dist.init_process_group("nccl", device_id=dev)
    res = run_cases(dev, rank, world, log=lambda m: print(m, flush=True))
    if rank == 0:
        print(f"[mgpu] result: {res}", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if res == "ok" else 1)


if __name__ == "__main__":
    main()
