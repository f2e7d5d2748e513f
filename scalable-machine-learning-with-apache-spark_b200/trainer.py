"""Tensor-level entry point of the hot path: ``fit_tensors`` = RandomForest.run of Spark MLlib 3.1.2 on device tensors.

It is what `DecisionTreeRegressor.fit` / `RandomForestRegressor.fit` (`ML 06 - Decision Trees.py:126`,
`ML 07 - Random Forests and Hyperparameter Tuning.py:116,149`) reach after the Spark adapter (spark.py) has turned the
`features` column into a tensor.  PyTorch here is plumbing only (device memory, streams, torch.distributed); every
pass over rows is a hand-written kernel behind the C-ABI in include/b2rf.h.
"""
from __future__ import annotations

import ctypes as C
import os
import time

import numpy as np
import torch

from ._lib import check, lib
from .forest import Forest
from .params import TreeParams


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


# Row width of the row-major code matrix.  "auto": 64 codes for up to 64 features and 128 for up to 128 — the widths the histogram
# kernels stage through registers with compile-time offsets (k4_limb.cu, k8_multiclass.cu), and rows that start on a 64- / 128-byte
# boundary so a gather touches the fewest 32-byte sectors; wider matrices are padded to 16.  An integer pads to that multiple.
CODE_ROW_ALIGN = "auto"


def code_row_width(F: int) -> int:
    if CODE_ROW_ALIGN == "auto":
        return 64 if F <= 64 else 128 if F <= 128 else _pad(F, 16)
    return _pad(F, int(CODE_ROW_ALIGN))


def _pad(n: int, m: int = 128) -> int:
    return max(m, (n + m - 1) // m * m)


def _require_cuda(t: torch.Tensor, name: str):
    if not (isinstance(t, torch.Tensor) and t.is_cuda):
        raise TypeError(f"{name} must be a CUDA tensor: b2rf has no CPU fallback")


def _dist_info(group):
    import torch.distributed as dist
    if group is None and not (dist.is_available() and dist.is_initialized()):
        return None, 0, 1
    return dist, dist.get_rank(group), dist.get_world_size(group)


# ------------------------------------------------------------------------------------------------
# stage wrappers (K1, K2, K3, K4) — thin; used by fit_tensors and by the parity tests
# ------------------------------------------------------------------------------------------------
def find_splits(sample: torch.Tensor, arity: torch.Tensor, max_possible_bins: int, weighted_num_samples: float):
    """K1.  sample: f64 [n_s, F] CUDA; arity: int32 [F] CUDA.  Returns (thresholds f64 [F, maxSplits], nsplits int32 [F])."""
    _require_cuda(sample, "sample")
    sample = sample.contiguous()
    n_s, F = sample.shape
    ms = max(max_possible_bins - 1, 1)
    thr = torch.zeros((F, ms), dtype=torch.float64, device=sample.device)
    ns = torch.zeros(F, dtype=torch.int32, device=sample.device)
    if n_s == 0:
        return thr, ns
    wsb = lib().b2rf_find_splits_workspace_bytes(n_s, F)
    ws = torch.empty(wsb, dtype=torch.uint8, device=sample.device)
    check(lib().b2rf_find_splits(sample.data_ptr(), n_s, F, arity.data_ptr(), max_possible_bins, float(weighted_num_samples),
                                 thr.data_ptr(), ns.data_ptr(), ws.data_ptr(), wsb, _stream()))
    return thr, ns


def sample_rows(X: torch.Tensor, row0_global: int, seed: int, fraction: float, dist=None, group=None) -> torch.Tensor:
    """Row sample for findSplits when fraction < 1 (hash of the global row index)."""
    _require_cuda(X, "X")
    n, F = X.shape
    cap = int(n * fraction + 6.0 * (n * fraction * (1 - fraction)) ** 0.5 + 64)
    out = torch.empty((cap, F), dtype=torch.float64, device=X.device)
    cnt = torch.zeros(1, dtype=torch.int64, device=X.device)
    check(lib().b2rf_sample_rows(X.data_ptr(), X.stride(0), n, row0_global, F, seed, float(fraction), out.data_ptr(), cap, cnt.data_ptr(),
                                 _stream()))
    k = int(cnt.item())
    if _agree_max(int(k > cap), dist, group, X.device):
        raise RuntimeError("b2rf sample_rows: capacity exceeded")
    return out[:k]


_bin_err = {}


def _agree_max(code: int, dist, group, device) -> int:
    """Largest `code` over the ranks of a sharded fit.  Data-dependent checks (labels, categorical values, sample capacity) sit
    between collectives: if only the rank that holds the bad shard raised, the others would wait in NCCL forever, so every rank
    learns the worst code first and all of them raise the same exception (MLlib: one IllegalArgumentException for the job)."""
    if dist is None or dist.get_world_size(group) == 1:
        return int(code)
    t = torch.tensor([int(code)], dtype=torch.int64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return int(t.item())


def _raise_bin_errors(arity, dist=None, group=None):
    """Raises MLlib's invalid-categorical-value error if any K2 launch on this device — or, in a sharded fit, on any rank —
    flagged one (deferred check)."""
    err = _bin_err.get(arity.device)
    e = 0
    if err is not None:
        e = int(err.item())
        err.zero_()
    e = _agree_max(e, dist, group, arity.device)
    if e:
        from ._lib import ERR_DATA, IllegalArgumentException
        f = e - 1
        raise IllegalArgumentException(ERR_DATA, f"DecisionTree given invalid data: Feature {f} is categorical with values in "
                                       f"{{0,...,{int(arity[f].item()) - 1}}}, but a data point gives it an out-of-range value.")


def bin_features(X: torch.Tensor, thresholds: torch.Tensor, nsplits: torch.Tensor, arity: torch.Tensor, bins: torch.Tensor = None,
                 row0: int = 0, bins_rm: torch.Tensor = None, check_errors: bool = True) -> torch.Tensor:
    """K2.  X f64 [n, F] row-major CUDA -> column-major uint8 bins [F, ldb] (rows row0..row0+n); optionally also the
    row-major twin bins_rm [rows, ldr] that the bucket-mode histogram kernel gathers from."""
    _require_cuda(X, "X")
    n, F = X.shape
    if X.stride(1) != 1:
        X = X.contiguous()
    code_dt = torch.uint8 if thresholds.shape[1] <= 255 else torch.int16      # uint16 codes are stored in int16 tensors
    if bins is None:
        bins = torch.zeros((F, _pad(row0 + n)), dtype=code_dt, device=X.device)
    if bins.dtype != code_dt or (bins_rm is not None and bins_rm.dtype != code_dt):
        raise TypeError(f"bin matrices must be {code_dt} for maxBins = {thresholds.shape[1] + 1}")
    err = _bin_err.get(X.device)
    if err is None:
        err = _bin_err[X.device] = torch.zeros(1, dtype=torch.int32, device=X.device)
    check(lib().b2rf_bin(X.data_ptr(), X.stride(0), n, F, thresholds.data_ptr(), nsplits.data_ptr(), arity.data_ptr(), thresholds.shape[1],
                         bins.data_ptr(), bins.stride(0), row0, err.data_ptr(), 0 if bins_rm is None else bins_rm.data_ptr(),
                         0 if bins_rm is None else bins_rm.stride(0), bins.element_size(), _stream()))
    if check_errors:
        _raise_bin_errors(arity)
    return bins


def bag(seed: int, part_offsets, part0: int, T: int, rate: float, bootstrap: bool, device, ld: int = None, want_node: bool = True):
    """K3.  Returns (w uint8 [T, ld], node int32 [T, ld] or None).  The per-row node cache is only needed by the "sort" and
    "scan" histogram modes; segment mode keeps per-tree row lists instead."""
    po = np.ascontiguousarray(part_offsets, np.int64)
    n = int(po[-1])
    ld = ld or _pad(n)
    w = torch.zeros((T, ld), dtype=torch.uint8, device=device)
    node = torch.full((T, ld), -1, dtype=torch.int32, device=device) if want_node else None
    with torch.cuda.device(device):
        check(lib().b2rf_bag(int(seed), po.ctypes.data_as(C.c_void_p), len(po) - 1, int(part0), T, float(rate), int(bool(bootstrap)),
                             w.data_ptr(), w.stride(0), 0 if node is None else node.data_ptr(), ld if node is None else node.stride(0), _stream()))
    return w, node


def hist_build(bins, y, w, slot_of_row, tree_ids, slot_fo, slot_base, n_rows, total_cells):
    """K4 stand-alone.  Returns hist f64 [total_cells, 4]."""
    hist = torch.zeros((total_cells, 4), dtype=torch.float64, device=bins.device)
    check(lib().b2rf_hist_build(bins.data_ptr(), bins.stride(0), bins.element_size(), bins.shape[0], y.data_ptr(), w.data_ptr(), w.stride(0), slot_of_row.data_ptr(),
                                slot_of_row.stride(0), tree_ids.data_ptr(), tree_ids.numel(), n_rows, slot_fo.data_ptr(), slot_base.data_ptr(),
                                slot_fo.shape[1], hist.data_ptr(), _stream()))
    return hist


# ------------------------------------------------------------------------------------------------
# multi-rank plumbing (device-agnostic: exercised on CPU with gloo in tests/test_cpu_distributed.py)
# ------------------------------------------------------------------------------------------------
def global_numbering(n_local: int, n_parts_local: int, dist, group, device, part0=None):
    """Rows and emulated partitions are numbered rank-major.  Returns (N_global, row0_global, part0_global)."""
    if dist is None or dist.get_world_size(group) == 1:
        return n_local, 0, (part0 or 0)
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    info = torch.tensor([n_local, n_parts_local], dtype=torch.int64, device=device)
    allinfo = [torch.zeros_like(info) for _ in range(world)]
    dist.all_gather(allinfo, info, group=group)
    counts = [int(a[0].item()) for a in allinfo]
    pcounts = [int(a[1].item()) for a in allinfo]
    return sum(counts), sum(counts[:rank]), (sum(pcounts[:rank]) if part0 is None else part0)


def gather_samples(samp: torch.Tensor, dist, group) -> torch.Tensor:
    """findSplits needs the whole (small) row sample on every rank: all-gather of ragged [n_r, F] blocks, rank order."""
    if dist is None or dist.get_world_size(group) == 1:
        return samp
    world = dist.get_world_size(group)
    cnt = torch.tensor([samp.shape[0]], dtype=torch.int64, device=samp.device)
    cnts = [torch.zeros_like(cnt) for _ in range(world)]
    dist.all_gather(cnts, cnt, group=group)
    cnts = [int(c.item()) for c in cnts]
    buf = torch.zeros((max(max(cnts), 1), samp.shape[1]), dtype=samp.dtype, device=samp.device)
    buf[: samp.shape[0]] = samp
    bufs = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(bufs, buf, group=group)
    return torch.cat([b[:c] for b, c in zip(bufs, cnts)], 0)


# ------------------------------------------------------------------------------------------------
# the fit
# ------------------------------------------------------------------------------------------------
def default_partitions(n_rows: int, rows_per_partition: int = 4096):
    """Emulated Spark partition layout used when the caller gives none: contiguous chunks of rows_per_partition rows."""
    P = max(1, (n_rows + rows_per_partition - 1) // rows_per_partition)
    return np.minimum(np.arange(P + 1, dtype=np.int64) * rows_per_partition, n_rows)


_HIST_MODES = {"segment": 0, "bucket": 0, "sort": 1, "scan": 2}
MAX_LIST_ROWS = 1 << 27     # segment / sort mode pack (row, bag weight) into 32 bits: 27-bit row ids per GPU


def _want_node(hist_mode, bins_rm, n_local) -> bool:
    """The per-(tree, row) node cache is what the scan mode routes rows with.  Segment / sort mode keep packed row lists instead,
    which hold 27-bit row ids: a shard of 2^27 (134M) rows or more on one GPU runs the scan mode (slower; shard the rows over
    more GPUs to stay on the list-based kernels) and so needs the cache."""
    return hist_mode not in ("segment", "bucket") or bins_rm is None or n_local >= MAX_LIST_ROWS


def _setup(n_local, F, arity, params, part_offsets, part0, group, dev, hist_mode):
    """Metadata checks and global numbering shared by the device- and host-input entry points (rows a1, e)."""
    L = lib()
    arity_np = np.ascontiguousarray(arity, np.int32)
    if arity_np.shape != (F,):
        raise ValueError("arity must have one entry per feature")
    dist, rank, world = _dist_info(group)
    if part_offsets is None:
        part_offsets = default_partitions(n_local)
    po = np.ascontiguousarray(part_offsets, np.int64)
    if po[0] != 0 or po[-1] != n_local:
        raise ValueError("part_offsets must start at 0 and end at the local row count")
    N, row0, part0 = global_numbering(n_local, len(po) - 1, dist, group, dev, part0)
    cp = params.to_c(F)
    cp.histMode = _HIST_MODES[hist_mode]
    meta_max_bins = check(L.b2rf_check_metadata(C.byref(cp), F, arity_np.ctypes.data_as(C.c_void_p), N))
    fraction = L.b2rf_samples_fraction(N, meta_max_bins)
    return dict(arity_np=arity_np, dist=dist, world=world, po=po, N=N, row0=row0, part0=part0, cp=cp, mpb=min(params.maxBins, N),
                fraction=fraction, arity_d=torch.as_tensor(arity_np, device=dev))


def _check_labels(y, params, dist=None, group=None, device=None):
    """Classification labels must be the integers 0 .. numClasses - 1 (MLlib: 'Classification labels should be in [0 to ...]').
    In a sharded fit every rank raises when any rank holds a bad label."""
    if not params.is_classification:
        return
    C_ = params.num_classes
    bad = bool(y.numel()) and not bool(((y >= 0) & (y < C_) & (y == y.floor())).all())
    if _agree_max(int(bad), dist, group, device if device is not None else y.device):
        from ._lib import IllegalArgumentException
        raise IllegalArgumentException(-3, f"requirement failed: Classification labels should be in [0 to {C_ - 1}] (numClasses == {C_}; pass "
                                           f"TreeParams(numClasses=...) for more than two classes). Found invalid labels.")


def bind_host_thread_to_gpu_node(device):
    """Deployment helper for `fit_host` with several GPUs per host: restrict the calling thread (and the threads it starts later) to
    the CPUs next to the GPU, so that pinned host buffers allocated afterwards are placed on that NUMA node and the host-to-device
    copies of the ranks do not all cross the inter-socket link.  The CPU set comes from sysfs (the PCI device's numa_node) or, where a
    container hides that, from NVML's CPU affinity of the device.  Returns a short description of what was applied, or None when
    neither source says anything (single-socket hosts) — then nothing is changed."""
    import os
    pr = torch.cuda.get_device_properties(device)
    allowed = os.sched_getaffinity(0)
    cpus, how = set(), None
    try:
        addr = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        with open(f"/sys/bus/pci/devices/{addr}/numa_node") as f:
            node = int(f.read().strip())
        if node >= 0:
            with open(f"/sys/devices/system/node/node{node}/cpulist") as f:
                for part in f.read().strip().split(","):
                    lo, _, hi = part.partition("-")
                    cpus.update(range(int(lo), int(hi or lo) + 1))
            how = f"sysfs numa node {node}"
    except (OSError, ValueError, AttributeError):
        pass
    if not cpus:
        try:
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByPciBusId(f"{pr.pci_domain_id:08x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0".encode())
            ncpu = max(os.cpu_count() or 1, max(allowed) + 1)
            words = pynvml.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
            cpus = {i for i in range(ncpu) if (int(words[i // 64]) >> (i % 64)) & 1}
            how = "nvml cpu affinity"
        except Exception:   # noqa: BLE001 — NVML absent or too old: leave the placement to the default policy
            return None
    cpus &= allowed
    if not cpus or cpus == allowed:
        return None          # nothing to narrow
    os.sched_setaffinity(0, cpus)
    return f"{how}: {len(cpus)} of {len(allowed)} cpus"


def k4_launch_counts() -> dict:
    """Launches of each K4 implementation by this process so far (b2rf_k4_launch_counts)."""
    c = np.zeros(6, np.int64)
    check(lib().b2rf_k4_launch_counts(c.ctypes.data_as(C.c_void_p), 6))
    return dict(limb=int(c[0]), fp64_32=int(c[1]), fp64_general=int(c[2]), class_counts=int(c[3]), scan=int(c[4]), class_staged=int(c[5]))


def set_option(name: str, value: int):
    """Process-wide switch of the library (b2rf_set_option): "k4_fp64", "k4_tma", "k4_binary", "carry_labels"."""
    check(lib().b2rf_set_option(name.encode(), int(value)))


def _group_info(h, Fsub):
    L = lib()
    n = C.c_int64()
    check(L.b2rf_trainer_group_info(h, C.byref(n), None, None, None, None))
    n = n.value
    tree, nid = np.zeros(n, np.int32), np.zeros(n, np.int32)
    fo, base = np.zeros((n, Fsub, 2), np.uint32), np.zeros(n, np.int64)
    P = lambda a: a.ctypes.data_as(C.c_void_p)
    check(L.b2rf_trainer_group_info(h, None, P(tree), P(nid), P(fo), P(base)))
    return dict(tree=tree, node_id=nid, slot_fo=fo, slot_base=base)


def _train_loop(S, params, F, n_local, y, bins, bins_rm, w, node, thr_h, ns_h, group, keep_trace, hist_hook=None):
    """RandomForest.run's while (nodeStack.nonEmpty) loop: one begin_group / [all_reduce] / end_group per node group."""
    L = lib()
    dev = y.device
    h = C.c_void_p()
    dbg = os.environ.get("B2RF_TIMING") == "1"
    tc = time.perf_counter()
    check(L.b2rf_trainer_create(C.byref(S["cp"]), F, S["arity_np"].ctypes.data_as(C.c_void_p), ns_h.ctypes.data_as(C.c_void_p),
                                thr_h.ctypes.data_as(C.c_void_p), S["N"], C.byref(h)))
    try:
        # rows sharded over ranks: reduce-scatter the group histogram by node, K5 on the own node slots, all-gather the split records
        # (segment mode; B2RF_SHARD_K5=0 or a histogram hook keep the all-reduce + redundant K5 of every rank)
        if keep_trace:
            check(L.b2rf_trainer_set_trace(h, 1))
        world, dist = S["world"], S["dist"]
        shard = world > 1 and node is None and hist_hook is None and os.environ.get("B2RF_SHARD_K5", "1") != "0"
        if shard:
            check(L.b2rf_trainer_set_shard(h, dist.get_rank(group), world))
        cells = int(L.b2rf_trainer_max_group_cells(h))
        hist = torch.empty(cells * 4, dtype=torch.float64, device=dev)
        free_b, total_b = torch.cuda.mem_get_info(dev)
        if free_b < 0.5 * total_b:          # a large fit: hand torch's cached blocks back so that the trainer's row and label lists fit
            torch.cuda.empty_cache()
        xin = xall = None
        ld = w.stride(0)
        check(L.b2rf_trainer_bind(h, bins.data_ptr(), bins.stride(0), 0 if bins_rm is None else bins_rm.data_ptr(),
                                  0 if bins_rm is None else bins_rm.stride(0), y.data_ptr(), w.data_ptr(), w.stride(0),
                                  0 if node is None else node.data_ptr(), ld if node is None else node.stride(0), n_local,
                                  hist.data_ptr(), cells))
        ncell = C.c_int64()
        st = _stream()
        if dbg:
            torch.cuda.synchronize(dev); tb = time.perf_counter()
        ar_events = []
        while check(L.b2rf_trainer_begin_group(h, C.byref(ncell), st)) == 1:
            blk = int(L.b2rf_trainer_shard_block_doubles(h)) if shard else 0
            if blk > 0:
                ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True),
                      torch.cuda.Event(enable_timing=True))
                chunk = ncell.value * 4 // world                             # doubles per rank: its node slots' cells (padded)
                rank = dist.get_rank(group)
                ev[0].record()
                dist.reduce_scatter_tensor(hist[rank * chunk: (rank + 1) * chunk], hist[: world * chunk], group=group)   # merge by node, row a7
                ev[1].record()
                if xin is None or xin.numel() < blk:
                    xin = torch.empty(blk + blk // 2, dtype=torch.float64, device=dev)
                    xall = torch.empty(world * xin.numel(), dtype=torch.float64, device=dev)
                check(L.b2rf_trainer_end_group_shard_launch(h, xin.data_ptr(), st))
                ev[2].record()
                dist.all_gather_into_tensor(xall[: world * blk], xin[:blk], group=group)                                   # the nodes' best splits
                ev[3].record()
                ar_events.append((ev[0], ev[1])); ar_events.append((ev[2], ev[3]))
                check(L.b2rf_trainer_end_group_shard_finish(h, xall.data_ptr(), st))
                continue
            if S["world"] > 1:
                ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
                ev[0].record()
                S["dist"].all_reduce(hist[: ncell.value * 4], group=group)   # DTStatsAggregator.merge (reduceByKey), row a7
                ev[1].record()
                ar_events.append(ev)
            if hist_hook is not None:   # debug / parity tests: the reduced group histogram next to the group's slot table
                hist_hook(_group_info(h, min(S["cp"].featuresPerNode, F) if S["cp"].featuresPerNode > 0 else F), hist[: ncell.value * 4].view(-1, 4))
            check(L.b2rf_trainer_end_group(h, st))
        torch.cuda.synchronize(dev)
        tl = time.perf_counter()
        forest = _export(h, F, keep_trace)
        te = time.perf_counter()
    finally:
        L.b2rf_trainer_destroy(h)
    if dbg:
        print(f"[b2rf timing] create+bind {tb - tc:.4f} groups {tl - tb:.4f} export {te - tl:.4f} destroy {time.perf_counter() - te:.4f}", flush=True)
    forest.thresholds, forest.nsplits = thr_h, ns_h
    forest.impurity_kind = params.impurity
    forest.timers["allreduce"] = sum(a.elapsed_time(b) for a, b in ar_events) * 1e-3     # device time of the histogram reductions
    forest.counters.update(rows=S["N"], rows_local=n_local, partitions=len(S["po"]) - 1, world=S["world"], fraction=S["fraction"])
    return forest


def fit_tensors(X: torch.Tensor, y: torch.Tensor, arity, params: TreeParams, part_offsets=None, part0: int = None, group=None,
                keep_trace: bool = False, prebinned=None, hist_mode: str = "segment", hist_hook=None) -> Forest:
    """Train a forest on this rank's rows.

    X: f64 [n_local, F] CUDA, row-major (the VectorAssembler `features` column); y: f64 [n_local];
    arity[F]: 0 = continuous, k > 0 = categorical with k values (from the column's ml_attr metadata);
    part_offsets: local row offsets of the emulated Spark partitions (bagging streams are seeded per partition, A.4);
    part0: global index of this rank's first partition (default: partitions are numbered rank-major).
    With torch.distributed initialised, rows are sharded over ranks and only per-group histograms are all-reduced.
    hist_mode: "segment" (default: per-tree row lists kept partitioned by node, shared-memory histograms), "sort" (rows
    bucketed by a per-group counting sort, same histogram kernel) or "scan" (column-major scan + global RED).
    """
    _require_cuda(X, "X")
    _require_cuda(y, "y")
    dev = X.device
    t_start = time.perf_counter()
    with torch.cuda.device(dev):
        y = y.to(torch.float64).contiguous()
        n_local, F = X.shape
        S = _setup(n_local, F, arity, params, part_offsets, part0, group, dev, hist_mode)
        _check_labels(y, params, S["dist"], group, dev)
        T = params.numTrees
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()

        # ---- findSplits (K1) ----
        fraction = S["fraction"]
        samp = sample_rows(X, S["row0"], params.seed, fraction, S["dist"], group) if fraction < 1.0 else X
        samp = gather_samples(samp, S["dist"], group)
        thr, ns = find_splits(samp, S["arity_d"], S["mpb"], fraction * float(S["N"]))
        thr_h = thr.cpu().numpy()
        ns_h = ns.cpu().numpy()
        torch.cuda.synchronize(dev)
        t1 = time.perf_counter()

        # ---- TreePoint (K2) + BaggedPoint (K3) ----
        # (K3 on a side stream next to K1 / K2 was measured: no gain on one GPU — its one wave of long-lived warps leaves the 7 ms of
        # K1 + K2 nothing to hide behind — and a loss on several, where the collectives of findSplits queue behind it)
        ld = _pad(n_local)
        dbg = os.environ.get("B2RF_TIMING") == "1"
        if prebinned is not None:
            bins, bins_rm = prebinned
        else:
            code_dt = torch.uint8 if S["mpb"] <= 256 else torch.int16
            bins = torch.empty((F, ld), dtype=code_dt, device=dev)
            bins_rm = torch.empty((max(n_local, 1), code_row_width(F)), dtype=code_dt, device=dev) if hist_mode != "scan" else None
            if dbg:
                torch.cuda.synchronize(dev); ta = time.perf_counter()
            bin_features(X, thr, ns, S["arity_d"], bins=bins, row0=0, bins_rm=bins_rm, check_errors=False)
            _raise_bin_errors(S["arity_d"], S["dist"], group)
            if dbg:
                torch.cuda.synchronize(dev); tb = time.perf_counter()
        w, node = bag(params.seed, S["po"], S["part0"], T, params.subsamplingRate, params.bootstrap, dev, ld=ld,
                      want_node=_want_node(hist_mode, bins_rm, n_local))
        torch.cuda.synchronize(dev)
        t2 = time.perf_counter()
        if dbg and prebinned is None:
            print(f"[b2rf timing] setup {t0 - t_start:.4f} findSplits {t1 - t0:.4f} alloc {ta - t1:.4f} bin {tb - ta:.4f} bag(+alloc) {t2 - tb:.4f}", flush=True)

        forest = _train_loop(S, params, F, n_local, y, bins, bins_rm, w, node, thr_h, ns_h, group, keep_trace, hist_hook)
        t3 = time.perf_counter()
        forest.timers.update(init=(t0 - t_start) + (t2 - t1), findSplits=t1 - t0, total=t3 - t_start, loop=t3 - t2)
        # stage kernels outside the trainer: sample_rows (if sampling), find_splits, bin, bag
        forest.counters["kernel_launches"] += (1 if fraction < 1.0 else 0) + 3 - (1 if prebinned is not None else 0)
        return forest


def fit_many_tensors(X: torch.Tensor, y: torch.Tensor, arity, params_list, part_offsets=None, part0: int = None, group=None,
                     hist_mode: str = "segment"):
    """Train several parameterisations on the same rows in shared passes (`Estimator.fitMultiple`, CrossValidator's
    candidates of one fold — `ML 07 - Random Forests and Hyperparameter Tuning.py:74-77,143-149`; BASELINE config 5).

    What is shared: the findSplits sample, thresholds and both bin matrices (once per distinct (maxBins, seed): findSplits
    depends on nothing else), the label vector, the device copy of X — and the histogram passes: the candidates' node-group
    loops run in lockstep, and each round's K4 is ONE launch whose task table spans the node slots of every candidate still
    growing (b2rf_trainers_begin_group_shared), so a shallow candidate's few nodes ride with the deep one's.  Bag weights are drawn
    once per distinct (seed, numTrees, subsamplingRate, bootstrap) — MLlib draws them row-major over numTrees, so candidates with
    different numTrees differ, candidates that differ only in depth / bins / feature subsets do not.  Not shared: the row lists.
    Each candidate's forest is identical to what fit_tensors returns for it."""
    _require_cuda(X, "X")
    _require_cuda(y, "y")
    dev = X.device
    L = lib()
    params_list = list(params_list)
    with torch.cuda.device(dev):
        y = y.to(torch.float64).contiguous()
        n_local, F = X.shape
        ld = _pad(n_local)
        shared = {}
        bags = {}
        cands = []
        main = torch.cuda.current_stream()
        for params in params_list:
            S = _setup(n_local, F, arity, params, part_offsets, part0, group, dev, hist_mode)
            _check_labels(y, params, S["dist"], group, dev)
            key = (params.maxBins, params.seed)
            if key not in shared:
                fraction = S["fraction"]
                samp = sample_rows(X, S["row0"], params.seed, fraction, S["dist"], group) if fraction < 1.0 else X
                samp = gather_samples(samp, S["dist"], group)
                thr, ns = find_splits(samp, S["arity_d"], S["mpb"], fraction * float(S["N"]))
                code_dt = torch.uint8 if S["mpb"] <= 256 else torch.int16
                bins = torch.empty((F, ld), dtype=code_dt, device=dev)
                bins_rm = torch.empty((max(n_local, 1), code_row_width(F)), dtype=code_dt, device=dev) if hist_mode != "scan" else None
                bin_features(X, thr, ns, S["arity_d"], bins=bins, row0=0, bins_rm=bins_rm, check_errors=False)
                _raise_bin_errors(S["arity_d"], S["dist"], group)
                shared[key] = (thr.cpu().numpy(), ns.cpu().numpy(), bins, bins_rm)
            thr_h, ns_h, bins, bins_rm = shared[key]
            # bag counts depend on (seed, numTrees, rate, bootstrap) and the partition layout only: candidates that differ in depth,
            # bins or feature subsets share one weight matrix (read-only after bind; the node cache of the scan modes is per candidate)
            want_node = _want_node(hist_mode, bins_rm, n_local)
            bkey = (params.seed, params.numTrees, float(params.subsamplingRate), bool(params.bootstrap))
            if want_node or bkey not in bags:
                w, node = bag(params.seed, S["po"], S["part0"], params.numTrees, params.subsamplingRate, params.bootstrap, dev, ld=ld,
                              want_node=want_node)
                if not want_node:
                    bags[bkey] = w
            else:
                w, node = bags[bkey], None
            cands.append(dict(S=S, params=params, thr_h=thr_h, ns_h=ns_h, bins=bins, bins_rm=bins_rm, w=w, node=node, h=C.c_void_p()))
        torch.cuda.synchronize(dev)
        try:
            for c in cands:
                S = c["S"]
                check(L.b2rf_trainer_create(C.byref(S["cp"]), F, S["arity_np"].ctypes.data_as(C.c_void_p), c["ns_h"].ctypes.data_as(C.c_void_p),
                                            c["thr_h"].ctypes.data_as(C.c_void_p), S["N"], C.byref(c["h"])))
                cells = int(L.b2rf_trainer_max_group_cells(c["h"]))
                c["hist"] = torch.empty(cells * 4, dtype=torch.float64, device=dev)
                w, node, bins, bins_rm = c["w"], c["node"], c["bins"], c["bins_rm"]
                check(L.b2rf_trainer_bind(c["h"], bins.data_ptr(), bins.stride(0), 0 if bins_rm is None else bins_rm.data_ptr(),
                                          0 if bins_rm is None else bins_rm.stride(0), y.data_ptr(), w.data_ptr(), w.stride(0),
                                          0 if node is None else node.data_ptr(), ld if node is None else node.stride(0), n_local,
                                          c["hist"].data_ptr(), cells))
            torch.cuda.synchronize(dev)
            active = list(cands)
            st = _stream()
            dbg = os.environ.get("B2RF_TIMING") == "1"
            t_begin = t_end = 0.0
            rounds = 0
            while active:
                # one round = one node group of every candidate still growing: host scheduling for all of them, ONE K4 launch over
                # the union of their node slots, [all-reduce per candidate], all K5 launches, one synchronisation, the driver tails
                n = len(active)
                hs = (C.c_void_p * n)(*[c["h"].value for c in active])
                ncells = (C.c_int64 * n)()
                started = (C.c_int32 * n)()
                ta = time.perf_counter()
                check(L.b2rf_trainers_begin_group_shared(hs, n, ncells, started, st))
                tb = time.perf_counter()
                for i, c in enumerate(active):
                    if started[i] and c["S"]["world"] > 1:
                        c["S"]["dist"].all_reduce(c["hist"][: ncells[i] * 4], group=group)
                check(L.b2rf_trainers_end_group_shared(hs, n, st))
                t_begin += tb - ta
                t_end += time.perf_counter() - tb
                rounds += 1
                active = [c for i, c in enumerate(active) if started[i]]
            if dbg:
                print(f"[b2rf timing] fit_many: {rounds} rounds, begin (host scheduling + launches) {t_begin:.4f} s, end (K5, wait, tails, K6) {t_end:.4f} s", flush=True)
            torch.cuda.synchronize(dev)
            forests = []
            for c in cands:
                fo = _export(c["h"], F, False)
                fo.thresholds, fo.nsplits = c["thr_h"], c["ns_h"]
                fo.impurity_kind = c["params"].impurity
                S = c["S"]
                fo.counters.update(rows=S["N"], rows_local=n_local, partitions=len(S["po"]) - 1, world=S["world"], fraction=S["fraction"])
                forests.append(fo)
        finally:
            for c in cands:
                if c["h"]:
                    L.b2rf_trainer_destroy(c["h"])
        main.synchronize()
        return forests


def sample_row_ids(seed: int, row0_global: int, n_rows: int, fraction: float, device) -> torch.Tensor:
    """Local ids (ascending, int64, on the host) of the rows the findSplits sampler keeps, decided on the device without touching X:
    what fit_host uses to gather the sample from host memory (10M hashes: microseconds on the GPU, tens of milliseconds on the host)."""
    if fraction >= 1.0:
        return torch.arange(n_rows, dtype=torch.int64)
    cap = int(n_rows * fraction + 6.0 * (n_rows * fraction * (1 - fraction)) ** 0.5 + 64)
    ids = torch.empty(cap, dtype=torch.int64, device=device)
    cnt = torch.zeros(1, dtype=torch.int64, device=device)
    with torch.cuda.device(device):
        check(lib().b2rf_sample_row_ids(int(n_rows), int(row0_global), int(seed), float(fraction), ids.data_ptr(), cap, cnt.data_ptr(), _stream()))
    k = int(cnt.item())
    if k > cap:
        raise RuntimeError("b2rf_sample_row_ids: capacity exceeded")
    return torch.sort(ids[:k]).values.cpu()


def sample_indices(seed: int, row0_global: int, n_rows: int, fraction: float) -> np.ndarray:
    """Local indices of the rows the findSplits sampler keeps (host twin of sample_rows: same hash of the global row index)."""
    if fraction >= 1.0:
        return np.arange(n_rows, dtype=np.int64)
    cap = int(n_rows * fraction + 6.0 * (n_rows * fraction * (1 - fraction)) ** 0.5 + 64)
    out = np.empty(cap, np.int64)
    k = lib().b2rf_sample_indices(int(seed), int(row0_global), int(n_rows), float(fraction), out.ctypes.data_as(C.c_void_p), cap)
    if k < 0 or k > cap:
        raise RuntimeError("b2rf_sample_indices: capacity exceeded")
    return out[:k]


def fit_host(X_host, y_host, arity, params: TreeParams, device="cuda:0", part_offsets=None, part0: int = None, group=None,
             keep_trace: bool = False, hist_mode: str = "segment", chunk_rows: int = 1 << 20) -> Forest:
    """Public host-buffer entry point (what the Spark adapter's single-estimator fit calls): X_host f64 or f32 [n, F] / y_host f64 [n] are
    HOST arrays (numpy or CPU torch tensors; pinned memory makes the copies asynchronous).  float32 features are copied as float32 and
    widened on the device, which trains exactly the forest of the widened matrix.

    The raw feature matrix never has to be resident on the GPU: the findSplits row sample is gathered on the host (5 MB),
    then X streams through up to four staging buffers in `chunk_rows` pieces — copies on one stream (the first ones start before
    the thresholds exist), K2 binning of the arrived pieces on another — while K3 bagging, which does not depend on X, runs
    concurrently on a third stream."""
    dev = torch.device(device)
    Xh = torch.as_tensor(X_host)
    if Xh.dtype != torch.float32:          # float32 features travel as they are (half the PCIe bytes) and are widened on the device: exact
        Xh = Xh.to(torch.float64)
    yh = torch.as_tensor(y_host, dtype=torch.float64)
    if Xh.is_cuda:
        return fit_tensors(Xh.to(torch.float64), yh.to(Xh.device), arity, params, part_offsets=part_offsets, part0=part0, group=group,
                           keep_trace=keep_trace, hist_mode=hist_mode)
    if not Xh.is_contiguous():
        Xh = Xh.contiguous()
    t_start = time.perf_counter()
    with torch.cuda.device(dev):
        n_local, F = Xh.shape
        S = _setup(n_local, F, arity, params, part_offsets, part0, group, dev, hist_mode)
        _check_labels(yh, params, S["dist"], group, dev)
        T = params.numTrees
        main = torch.cuda.current_stream()
        copy_s, bag_s = torch.cuda.Stream(), torch.cuda.Stream()
        yd = torch.empty(n_local, dtype=torch.float64, device=dev)
        yd.copy_(yh, non_blocking=True)
        ld = _pad(n_local)
        # K3 first: it only needs the partition layout
        bag_s.wait_stream(main)
        with torch.cuda.stream(bag_s):
            w, node = bag(params.seed, S["po"], S["part0"], T, params.subsamplingRate, params.bootstrap, dev, ld=ld,
                          want_node=_want_node(hist_mode, True, n_local))
        t0 = time.perf_counter()
        # the first pieces of X start moving before the thresholds exist: `n_stage` staging buffers are filled while K1 runs and its
        # results travel back (K2 of a piece waits for K1 on the main stream anyway)
        chunk_rows = max(1, min(chunk_rows, n_local))
        n_chunks = (n_local + chunk_rows - 1) // chunk_rows
        n_stage = max(2, min(4, n_chunks))
        stage = [torch.empty((chunk_rows, F), dtype=Xh.dtype, device=dev) for _ in range(n_stage)]
        wide = torch.empty((chunk_rows, F), dtype=torch.float64, device=dev) if Xh.dtype != torch.float64 else None
        copied = [torch.cuda.Event() for _ in range(n_stage)]
        binned = [torch.cuda.Event() for _ in range(n_stage)]
        copy_s.wait_stream(main)

        def issue_copy(k):
            b = k % n_stage
            r0 = k * chunk_rows
            r1 = min(n_local, r0 + chunk_rows)
            with torch.cuda.stream(copy_s):
                if k >= n_stage:
                    copy_s.wait_event(binned[b])
                stage[b][: r1 - r0].copy_(Xh[r0:r1], non_blocking=True)
                copied[b].record(copy_s)

        # K1 on the host-gathered sample; its (small) copy is queued ahead of the big pieces, which then stream while K1 runs
        fraction = S["fraction"]
        samp = (Xh[sample_row_ids(params.seed, S["row0"], n_local, fraction, dev)].to(dev) if fraction < 1.0 else Xh.to(dev)).to(torch.float64)
        for k in range(min(n_stage, n_chunks)):
            issue_copy(k)
        samp = gather_samples(samp, S["dist"], group)
        thr, ns = find_splits(samp, S["arity_d"], S["mpb"], fraction * float(S["N"]))
        thr_h = thr.cpu().numpy()
        ns_h = ns.cpu().numpy()
        t1 = time.perf_counter()
        # K2 streamed over row chunks
        code_dt = torch.uint8 if S["mpb"] <= 256 else torch.int16
        bins = torch.empty((F, ld), dtype=code_dt, device=dev)
        bins_rm = torch.empty((max(n_local, 1), code_row_width(F)), dtype=code_dt, device=dev) if hist_mode != "scan" else None
        for k in range(n_chunks):
            b = k % n_stage
            r0 = k * chunk_rows
            r1 = min(n_local, r0 + chunk_rows)
            main.wait_event(copied[b])
            piece = stage[b][: r1 - r0]
            if wide is not None:
                wide[: r1 - r0].copy_(piece)
                piece = wide[: r1 - r0]
            bin_features(piece, thr, ns, S["arity_d"], bins=bins, row0=r0, bins_rm=bins_rm, check_errors=False)
            binned[b].record(main)
            if k + n_stage < n_chunks:
                issue_copy(k + n_stage)
        main.wait_stream(bag_s)
        _raise_bin_errors(S["arity_d"], S["dist"], group)
        torch.cuda.synchronize(dev)
        t2 = time.perf_counter()
        if os.environ.get("B2RF_TIMING") == "1":
            print(f"[b2rf timing] fit_host: setup+bag launch {t0 - t_start:.4f} sample+findSplits {t1 - t0:.4f} stream+bin+bag wait {t2 - t1:.4f} "
                  f"({Xh.nbytes / max(t2 - t1, 1e-9) / 1e9:.1f} GB/s host->device)", flush=True)
        forest = _train_loop(S, params, F, n_local, yd, bins, bins_rm, w, node, thr_h, ns_h, group, keep_trace)
        t3 = time.perf_counter()
        forest.timers.update(init=(t0 - t_start) + (t2 - t1), findSplits=t1 - t0, total=t3 - t_start, loop=t3 - t2)
        forest.counters["kernel_launches"] += 2 + n_chunks
        return forest


def _export(h, F: int, keep_trace: bool) -> Forest:
    L = lib()
    nn, nc = C.c_int64(), C.c_int64()
    check(L.b2rf_trainer_forest_sizes(h, C.byref(nn), C.byref(nc)))
    n, ncat = nn.value, nc.value
    sec = np.zeros(5)
    cnt = np.zeros(9, np.int64)
    L.b2rf_trainer_stats(h, sec.ctypes.data_as(C.c_void_p), cnt.ctypes.data_as(C.c_void_p))
    T = None
    # number of trees = len(tree_off) - 1; query through a first export of tree_off sized generously
    def Z(shape, dt):      # every element is written by the export calls below
        return np.empty(shape, dt)
    # the trainer knows T; tree_off has T+1 entries <= n+1
    tree_off = Z(n + 2, np.int64)
    feature, left, right = Z(n, np.int32), Z(n, np.int32), Z(n, np.int32)
    threshold, prediction, impurity, gain = Z(n, np.float64), Z(n, np.float64), Z(n, np.float64), Z(n, np.float64)
    stats = Z((n, 4), np.float64)
    ncat_a, cat_off, cat_len = Z(n, np.int32), Z(n, np.int32), Z(n, np.int32)
    cats = Z(max(ncat, 1), np.float64)
    P = lambda a: a.ctypes.data_as(C.c_void_p)
    tree_off[:] = -1
    check(L.b2rf_trainer_export(h, P(tree_off), P(feature), P(threshold), P(left), P(right), P(prediction), P(impurity), P(gain), P(stats),
                                P(ncat_a), P(cat_off), P(cat_len), P(cats)))
    T = int(np.argmax(tree_off == n)) if n > 0 else 0
    # tree_off is strictly increasing (every tree has >= 1 node) and ends at n
    T = int(np.nonzero(tree_off == n)[0][0])
    tree_off = tree_off[: T + 1].copy()
    fo = Forest(F, tree_off, feature, threshold, left, right, prediction, impurity, gain, stats, ncat_a, cat_off, cat_len, cats[:ncat])
    ncls = int(L.b2rf_trainer_num_classes(h))
    if ncls > 2:      # multiclass: the per-class counts are the node statistics (binary models derive theirs from stats)
        cc = Z((max(n, 1), ncls), np.float64)
        check(L.b2rf_trainer_export_class_counts(h, P(cc)))
        fo.counts = cc[:n]
    fo.timers = dict(findBestSplits=float(sec[2]), chooseSplits=float(sec[3]))
    fo.counters = dict(groups=int(cnt[0]), hist_launches=int(cnt[1]), rows_trees_scanned=int(cnt[2]), learning_nodes=int(cnt[3]),
                       kernel_launches=int(cnt[4]), hist_alg_bytes=int(cnt[5]), k4_rows_visited=int(cnt[6]),
                       bin_updates=int(cnt[6]) * int(cnt[7]), weighted_bin_updates=int(cnt[8]) * int(cnt[7]))
    if keep_trace:
        nt = int(L.b2rf_trainer_trace_size(h))
        ti, td = Z((nt, 5), np.int32), Z((nt, 5), np.float64)
        check(L.b2rf_trainer_trace(h, P(ti), P(td)))
        fo.trace_i, fo.trace_d = ti, td
    return fo
