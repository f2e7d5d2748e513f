"""world_size-2 gloo tests (CPU) of the multi-rank host logic: global row/partition numbering, the ragged all-gather
of the findSplits row sample, and the property the sharded fit relies on — per-rank histograms of row shards, summed by
all_reduce, equal the histogram of all rows (DTStatsAggregator.merge)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import b2rf
        from b2rf import trainer
        from oracle import oracle as O
        N, T = 3001, 3
        X, y, arity = b2rf.synth.airbnb(N)
        cut = [0, 1200, N]
        lo, hi = cut[rank], cut[rank + 1]
        # the same rows generated from the global row index on each rank
        Xs, ys, _ = b2rf.synth.airbnb(hi - lo, row0=lo)
        assert torch.equal(Xs, X[lo:hi]) and torch.equal(ys, y[lo:hi])
        po_local = np.array([0, 500, hi - lo], np.int64)
        Ng, row0, part0 = trainer.global_numbering(hi - lo, len(po_local) - 1, dist, None, torch.device("cpu"))
        assert (Ng, row0, part0) == (N, lo, 2 * rank)
        # ragged sample gather, rank order
        samp = trainer.gather_samples(Xs[: 3 + 4 * rank], dist, None)
        assert samp.shape[0] == 3 + 7 and torch.equal(samp[:3], X[:3]) and torch.equal(samp[3:], X[1200:1207])
        # bagging streams are per GLOBAL partition: shard-local draws equal the slice of the global draw
        po_global = np.array([0, 500, 1200, 1700, N], np.int64)
        w_all = O.bag(42, po_global, T, 1.0, True)
        w_loc = O.bag(42 + part0, po_local, T, 1.0, True)   # seed + (part0 + p) + 1 == (seed + part0) + p + 1
        assert np.array_equal(w_loc, w_all[lo:hi])
        # histogram linearity under all_reduce
        thr, ns = O.find_splits(X.numpy(), arity, 40, float(N))
        bins = O.bin_features(X.numpy(), thr, ns, arity)
        nb = np.where(arity > 1, arity, ns + 1)
        S, F = 4, X.shape[1]
        feats = np.tile(np.arange(F, dtype=np.int32), (S, 1))
        off = np.zeros((S, F), np.int64)
        tot = 0
        for s in range(S):
            for j in range(F):
                off[s, j] = tot
                tot += nb[j]
        rng = np.random.default_rng(0)
        slot = rng.integers(-1, S, (T, N)).astype(np.int32)
        full = O.hist(bins, y.numpy(), w_all, np.arange(T, dtype=np.int32), slot, feats, off, tot)
        part = O.hist(bins[lo:hi], y.numpy()[lo:hi], w_all[lo:hi], np.arange(T, dtype=np.int32), np.ascontiguousarray(slot[:, lo:hi]), feats, off, tot)
        red = torch.as_tensor(part.copy())
        dist.all_reduce(red)
        assert np.array_equal(red.numpy()[:, 0], full[:, 0])                         # counts exact
        np.testing.assert_allclose(red.numpy()[:, 1:3], full[:, 1:3], rtol=1e-9, atol=1e-9)
        # data-dependent errors are agreed on across ranks before anyone raises: a bad label on rank 1 only must raise on BOTH ranks
        # (a rank that raised alone would leave the others waiting in the next collective), and good labels raise on neither
        p = b2rf.TreeParams(impurity="gini", numTrees=3)
        cpu = torch.device("cpu")
        trainer._check_labels(torch.tensor([0.0, 1.0, 1.0]), p, dist, None, cpu)
        bad = torch.tensor([0.0, 1.0, 2.5]) if rank == 1 else torch.tensor([0.0, 1.0, 0.0])
        try:
            trainer._check_labels(bad, p, dist, None, cpu)
            raise AssertionError("expected IllegalArgumentException on every rank")
        except b2rf.IllegalArgumentException as e:
            assert "Classification labels should be in" in str(e)
        assert trainer._agree_max(7 if rank == 0 else 3, dist, None, cpu) == 7
        q.put((rank, "ok"))
    except Exception as e:  # noqa: BLE001
        import traceback
        q.put((rank, traceback.format_exc()))
    finally:
        dist.destroy_process_group()


def test_two_rank_host_logic_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in res:
        assert msg == "ok", f"rank {rank}: {msg}"
