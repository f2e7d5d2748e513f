#!/usr/bin/env python
"""bench.py — RF fit throughput (rows*trees/s) of the b2rf hot path on N B200s of one node.

A "step" is one complete RandomForestRegressor fit (findSplits -> binning -> bagging -> the whole node-group loop) of
BASELINE.json configs[2]: synthetic 10M rows x 64 features, numTrees=100, maxDepth=10, maxBins=64, on each GPU's shard
(weak scaling: every rank holds 10M rows; the global dataset is N*10M rows, histograms are all-reduced per node group).

  python bench.py [--gpus N] [--steps K] [--warmup W]            # this repo's CUDA path
  python bench.py --impl reference [...]                         # the CPU restatement of MLlib on the host cores

Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement" for the meaning of every key.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "rf_fit_rows_trees_per_s"
UNIT = "rows*trees/s"
WORKLOADS = {
    # name: (rows per GPU, features, numTrees, maxDepth, maxBins, generator)
    "C3": dict(rows=10_000_000, F=64, numTrees=100, maxDepth=10, maxBins=64, gen="c3"),
    "C3-small": dict(rows=1_000_000, F=64, numTrees=20, maxDepth=8, maxBins=64, gen="c3"),
    "C4-shard": dict(rows=12_500_000, F=128, numTrees=100, maxDepth=12, maxBins=256, gen="c4"),
    "C4-cont": dict(rows=12_500_000, F=128, numTrees=50, maxDepth=10, maxBins=256, gen="c3"),   # C4's shape without categorical features
}
ROWS_PER_PARTITION = 4096   # emulated Spark partition size (bagging streams are per partition; stated in config)
K4_TRAFFIC_FILE = os.path.join(ROOT, "profiles", "k4_traffic.json")   # written from the ncu --set full capture of the dominant kernel


def k4_traffic(workload):
    """DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) of one launch of the dominant kernel from the committed ncu capture,
    with that launch's algorithmic bytes; None when no capture of this workload's kernel is on file."""
    try:
        return json.load(open(K4_TRAFFIC_FILE)).get(workload)
    except Exception:
        return None


def make_config(wl_name, wl, n_local, world, extra=None):
    cfg = {"workload": f"{wl_name}: synthetic {n_local} rows/GPU x {wl['F']} features, RandomForestRegressor numTrees={wl['numTrees']} "
                       f"maxDepth={wl['maxDepth']} maxBins={wl['maxBins']} featureSubsetStrategy=auto seed=42",
           "rows_global": n_local * world, "rows_per_gpu": n_local, "rows_per_partition": ROWS_PER_PARTITION,
           "parallelism": f"rows sharded over {world} GPU(s); per node group: histogram reduce-scatter by node, best splits on the owning rank, all-gather of the split records (NCCL)" if world > 1 else "1 GPU",
           "l2": "inputs (bin matrix 0.64 GB + per-tree arrays 5 GB) exceed the 126 MB L2; no explicit flush"}
    if extra:
        cfg.update(extra)
    return cfg


class ClockSampler(threading.Thread):
    """Samples SM clocks and clock-event (throttle) reasons while the timed region runs.

    The values are the ones `nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,clocks_event_reasons.*` prints, read through NVML
    in-process (nvidia_ml_py): spawning nvidia-smi five times a second from inside the timed region stalls the CUDA driver for tens of
    milliseconds per call and showed up as occasional 0.1-0.3 s outliers in the short host-synchronous phases of a step.  If NVML
    cannot be loaded the sampler falls back to the nvidia-smi subprocess at a lower rate."""

    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index = index
        self.rows = []          # [sm_mhz, sm_max_mhz, hw_slowdown, hw_thermal_slowdown, sw_thermal_slowdown, sw_power_cap] (booleans)
        self.stop_flag = threading.Event()
        self.source = "nvml"
        self._nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = index
            if vis:
                ids = [v.strip() for v in vis.split(",") if v.strip()]
                if index < len(ids) and ids[index].isdigit():
                    phys = int(ids[index])
            self._h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self._nvml = pynvml
            self.index = phys
        except Exception:
            self.source = "nvidia-smi"

    def _sample_nvml(self):
        n = self._nvml
        sm = n.nvmlDeviceGetClockInfo(self._h, n.NVML_CLOCK_SM)
        mx = n.nvmlDeviceGetMaxClockInfo(self._h, n.NVML_CLOCK_SM)
        try:
            r = n.nvmlDeviceGetCurrentClocksEventReasons(self._h)
        except Exception:
            r = n.nvmlDeviceGetCurrentClocksThrottleReasons(self._h)
        bits = [n.nvmlClocksEventReasonHwSlowdown, n.nvmlClocksEventReasonHwThermalSlowdown, n.nvmlClocksEventReasonSwThermalSlowdown,
                n.nvmlClocksEventReasonSwPowerCap]
        self.rows.append([int(sm), int(mx)] + [bool(r & b) for b in bits])

    def _sample_smi(self):
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        out = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                             capture_output=True, text=True, timeout=5).stdout.strip()
        c = [v.strip() for v in out.split(",")]
        if len(c) >= 6 and c[0].isdigit():
            self.rows.append([int(c[0]), int(c[1]) if c[1].isdigit() else 0] + [v.lower().startswith("active") for v in c[2:6]])

    def run(self):
        while not self.stop_flag.is_set():
            try:
                if self._nvml is not None:
                    self._sample_nvml()
                else:
                    self._sample_smi()
            except Exception:
                pass
            self.stop_flag.wait(0.1 if self._nvml is not None else 1.0)

    def summary(self):
        sm = sorted(r[0] for r in self.rows)
        mx = [r[1] for r in self.rows if r[1]]
        reasons = sorted({self.NAMES[i] for r in self.rows for i in range(4) if r[2 + i]})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(sm),
                "source": self.source}


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def cpu_oracle_run(wl, sample_rows, threads, seed=42):
    """Times the CPU restatement of MLlib (oracle/, OpenMP) on a bounded row sample of the same workload."""
    import numpy as np
    import b2rf
    from oracle import oracle as O
    X, y, arity = getattr(b2rf.synth, wl["gen"])(sample_rows, device="cpu", F=wl["F"])
    X = X.numpy(); y = y.numpy()
    po = np.minimum(np.arange(0, sample_rows + ROWS_PER_PARTITION, ROWS_PER_PARTITION), sample_rows)
    mpb = min(wl["maxBins"], sample_rows)
    req = max(mpb * mpb, 10000)
    fraction = req / sample_rows if req < sample_rows else 1.0
    sidx = None
    if fraction < 1.0:
        sidx = np.nonzero(O.sample_keep(seed, np.arange(sample_rows), fraction))[0]
    t0 = time.perf_counter()
    f = O.fit(X, y, arity, maxDepth=wl["maxDepth"], maxBins=wl["maxBins"], numTrees=wl["numTrees"], bootstrap=True, seed=seed,
              part_offsets=po, sample_idx=sidx, fraction=fraction, numThreads=threads)
    dt = time.perf_counter() - t0
    return dt, f


def run_reference(args, wl, wl_name):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    sample = args.cpu_sample_rows
    times = []
    for i in range(args.warmup + args.steps):
        dt, f = cpu_oracle_run(wl, sample, threads)
        if i >= args.warmup:
            times.append(dt)
    t = sum(times) / len(times)
    value = sample * wl["numTrees"] / t
    world = int(os.environ.get("WORLD_SIZE", str(args.gpus)))
    desc = (f"each step trains {sample} rows of the same generator (a bounded sample of the {wl['rows']} rows/GPU workload) with the full "
            f"numTrees={wl['numTrees']} maxDepth={wl['maxDepth']} maxBins={wl['maxBins']}; rows*trees/s of that fit")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": make_config(wl_name, wl, wl["rows"], world),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": desc, "sample_rows": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": "MLlib itself (JVM) is unavailable on this box; this is the CPU restatement of MLlib 3.1.2 in oracle/ (OpenMP, all host threads)"}
    print(json.dumps(line), flush=True)


def run_c4(args, b2rf, trainer, torch, dist, dev, rank, world, timed):
    """C4 sub-record of an N > 1 run: the 100M-row dataset of BASELINE.json configs[3] sharded over the N GPUs."""
    wl = WORKLOADS["C4-shard"]
    n_local = args.c4_rows // world
    try:
        X, y, arity = getattr(b2rf.synth, wl["gen"])(n_local, row0=rank * n_local, seed=42, device=dev, F=wl["F"])
        params = b2rf.TreeParams(maxDepth=wl["maxDepth"], maxBins=wl["maxBins"], numTrees=wl["numTrees"], bootstrap=True, seed=42,
                                 featureSubsetStrategy="auto")
        po = trainer.default_partitions(n_local, ROWS_PER_PARTITION)
        step = lambda: trainer.fit_tensors(X, y, arity, params, part_offsets=po)
        step()
        ms, outs = timed(step, 1)
        f = outs[-1]
        rec = {"workload": f"C4: synthetic {n_local * world} rows x {wl['F']} features (96 continuous + 32 categorical), numTrees={wl['numTrees']} "
                           f"maxDepth={wl['maxDepth']} maxBins={wl['maxBins']}, rows sharded over {world} GPUs ({n_local} each)",
               "scaling": "strong", "ms_per_fit": ms, "value": n_local * world * wl["numTrees"] / (ms * 1e-3), "unit": UNIT, "steps": 1, "warmup": 1,
               "phases_s": {k: round(f.timers[k], 4) for k in ("init", "findSplits", "findBestSplits", "chooseSplits", "total")},
               "allreduce_s": round(f.timers.get("allreduce", 0.0), 4), "node_groups": f.counters["groups"], "forest_nodes": int(f.num_nodes()),
               "kernel_seconds_per_fit": f.timers["findBestSplits"]}
        del X, y
        torch.cuda.empty_cache()
        return rec
    except Exception as e:   # noqa: BLE001 — e.g. out of memory at small N: report it, keep the headline line
        return {"unavailable": f"{type(e).__name__}: {str(e)[:300]}"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b2rf", choices=["b2rf", "reference"])
    ap.add_argument("--workload", default="C3", choices=sorted(WORKLOADS))
    ap.add_argument("--cpu-sample-rows", type=int, default=100_000)
    ap.add_argument("--sklearn-sample-rows", type=int, default=100_000, help="rows for the scikit-learn RF baseline (0 = skip)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--rows-per-gpu", type=int, default=0, help="override the workload's rows per GPU (e.g. strong scaling of C4: 100M / N)")
    ap.add_argument("--no-mgpu-parity", action="store_true", help="N > 1: skip the sharded-fit parity cases run before the timed region")
    ap.add_argument("--no-c4", action="store_true", help="N > 1: skip the C4 strong-scaling sub-record (100M x 128 rows sharded over the N GPUs)")
    ap.add_argument("--c4-rows", type=int, default=100_000_000, help="global row count of the C4 sub-record")
    args = ap.parse_args()
    wl = dict(WORKLOADS[args.workload])
    if args.rows_per_gpu > 0:
        wl["rows"] = args.rows_per_gpu
    if args.impl == "reference":
        return run_reference(args, wl, args.workload)
    if args.warmup < 3:
        print(f"[bench] note: warmup {args.warmup} < 3 — numbers from this run are not valid bench values", file=sys.stderr)

    import numpy as np
    import torch
    import b2rf
    from b2rf import trainer

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: b2rf has no CPU fallback (use --impl reference for the CPU oracle)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    numa_node = trainer.bind_host_thread_to_gpu_node(dev) if world > 1 else None   # pinned e2e buffers on the GPU's own NUMA node
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    # ---- N > 1: sharded fits against the CPU oracle's whole-dataset fits, before anything is timed (tests/mgpu_check.py) ----
    mgpu_parity = None
    if world > 1 and not args.no_mgpu_parity:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        try:
            import mgpu_check
            mgpu_parity = mgpu_check.run_cases(dev, rank, world, log=lambda m: print(m, file=sys.stderr, flush=True) if rank == 0 else None)
        except Exception as e:   # noqa: BLE001 — a broken check must show up in the line, not kill the bench
            mgpu_parity = f"FAILED {type(e).__name__}: {e}"
    n_local = wl["rows"]
    F = wl["F"]
    X, y, arity = getattr(b2rf.synth, wl["gen"])(n_local, row0=rank * n_local, seed=42, device=dev, F=F)
    params = b2rf.TreeParams(maxDepth=wl["maxDepth"], maxBins=wl["maxBins"], numTrees=wl["numTrees"], bootstrap=True, seed=42,
                             featureSubsetStrategy="auto")
    po = trainer.default_partitions(n_local, ROWS_PER_PARTITION)
    N_global = n_local * world

    def barrier():
        torch.cuda.synchronize(dev)
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        outs = [fn() for _ in range(steps)]
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), outs

    # ---- device-resident arm -------------------------------------------------------------------------------
    step = lambda: trainer.fit_tensors(X, y, arity, params, part_offsets=po)
    for _ in range(args.warmup):
        step()
    sampler = ClockSampler(local_rank)
    sampler.start()
    ms, forests = timed(step, args.steps)
    sampler.stop_flag.set()
    sampler.join(timeout=2)
    ms_per_step = ms / args.steps
    value = N_global * wl["numTrees"] / (ms_per_step * 1e-3)
    f0 = forests[-1]
    hist_s = sum(f.timers["findBestSplits"] for f in forests) / len(forests)           # K4 device time per fit (CUDA events)
    hist_bytes = f0.counters["hist_alg_bytes"]
    hist_launches = f0.counters["hist_launches"]
    peak, peak_src = measured_peak_gbs()
    achieved = hist_bytes / hist_s / 1e9 if hist_s > 0 else 0.0
    # bin updates of one fit: every in-bag row of every evaluated node x features per node (from the forest's own node counts)
    fsub = params.features_per_node(F)
    updates_per_fit = float(f0.counters.get("weighted_bin_updates", 0))
    phases = {k: round(sum(f.timers[k] for f in forests) / len(forests), 4) for k in ("init", "findSplits", "findBestSplits", "chooseSplits", "total")}
    phases["allreduce"] = round(sum(f.timers.get("allreduce", 0.0) for f in forests) / len(forests), 4)   # device time of the histogram merges (N > 1)
    phys_updates = float(f0.counters.get("bin_updates", 0))
    gpu_launches = int(sum(f.counters["kernel_launches"] for f in forests))
    cfg_extra = {"node_groups": f0.counters["groups"], "forest_nodes": int(f0.num_nodes())}
    traffic = k4_traffic(args.workload)
    kc = trainer.k4_launch_counts()
    k4_kernel = ("limb_hist_kernel (K4: per-node shared-memory histograms of fixed-point limbs over row segments, native 32-bit adds)"
                 if kc["limb"] else "bucket_hist32_kernel (K4: per-node shared-memory fp64 histograms over row segments)")

    # ---- end-to-end arm: host buffers in, forest arrays out, through the public API --------------------------
    e2e = None
    if not args.no_e2e:
        Xh = torch.empty((n_local, F), dtype=torch.float64, pin_memory=True)
        Xh.copy_(X)
        yh = torch.empty(n_local, dtype=torch.float64, pin_memory=True)
        yh.copy_(y)
        del X
        torch.cuda.empty_cache()
        estep = lambda: trainer.fit_host(Xh, yh, arity, params, device=dev, part_offsets=po)
        for _ in range(min(3, max(2, args.warmup))):   # warm-up: allocator pools (device blocks were just released) and pinned staging
            estep()
        n_e2e = 5 if args.steps >= 3 else max(1, args.steps)   # its own step count (reported in e2e.steps): five fits average out a slow first copy
        ems, eouts = timed(estep, n_e2e)
        ems /= n_e2e
        fe = eouts[-1]
        d2h = sum(a.nbytes for a in (fe.tree_off, fe.feature, fe.threshold, fe.left, fe.right, fe.prediction, fe.impurity, fe.gain, fe.stats,
                                      fe.ncat, fe.cat_off, fe.cat_len, fe.cats))
        e2e = {"value": N_global * wl["numTrees"] / (ems * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(Xh.nbytes + yh.nbytes),
               "d2h_bytes_per_step": int(d2h), "ms_per_step": ems, "steps": n_e2e}

    # ---- N > 1: BASELINE.json configs[3] (C4: 100M x 128 mixed, maxBins 256, depth 12, 100 trees), rows sharded over the N GPUs
    # (strong scaling of one dataset): one warm-up + one timed fit, device-timed, max over ranks ----------------------------------
    c4 = None
    if world > 1 and not args.no_c4:
        X = y = Xh = yh = forests = eouts = fe = estep = step = f0 = None   # release the C3 tensors (device and pinned host)
        import gc
        gc.collect()
        torch.cuda.empty_cache()
        c4 = run_c4(args, b2rf, trainer, torch, dist, dev, rank, world, timed)

    # ---- CPU baseline (rank 0, N=1 only): the oracle on a bounded row sample ---------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        dt, fo = cpu_oracle_run(wl, args.cpu_sample_rows, threads)
        cpu = {"value": args.cpu_sample_rows * wl["numTrees"] / dt, "unit": UNIT, "cores": threads, "kind": "port",
               "sample": f"{args.cpu_sample_rows} rows of the same generator, full numTrees/maxDepth/maxBins; {dt:.1f} s; "
                         f"{fo.updates / dt / 1e6:.0f} M bin-updates/s",
               "seconds": dt}

    # ---- second CPU baseline: scikit-learn's RandomForestRegressor, the reference's own single-node RF
    # (ML 12 trains exactly T=100, depth 10), all host cores, on a bounded row sample; a different algorithm (exact splits,
    # bootstrap by index), so it is context, not a parity target
    skl = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline and args.sklearn_sample_rows > 0:
        try:
            from sklearn.ensemble import RandomForestRegressor
            ns_ = min(args.sklearn_sample_rows, n_local)
            Xs_, ys_, _ = getattr(b2rf.synth, wl["gen"])(ns_, device="cpu", F=wl["F"])
            Xs_, ys_ = np.asarray(Xs_), np.asarray(ys_)
            t_sk = time.perf_counter()
            RandomForestRegressor(n_estimators=wl["numTrees"], max_depth=wl["maxDepth"], max_features=1 / 3, n_jobs=-1, random_state=42).fit(Xs_, ys_)
            t_sk = time.perf_counter() - t_sk
            skl = {"value": ns_ * wl["numTrees"] / t_sk, "unit": UNIT, "cores": os.cpu_count() or 1, "kind": "scikit-learn RandomForestRegressor (exact splits)",
                   "sample": f"{ns_} rows of the same generator, n_estimators={wl['numTrees']} max_depth={wl['maxDepth']} max_features=1/3 n_jobs=-1; {t_sk:.1f} s"}
        except Exception as e:   # a missing or broken sklearn must not cost the bench line
            skl = {"unavailable": f"{type(e).__name__}: {e}"}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": make_config(args.workload, wl, n_local, world, cfg_extra),
            "phases_s": phases,
            "roofline": {"bound": "hbm", "kernel": k4_kernel,
                         "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         # dram__bytes_read + dram__bytes_write of ONE launch of that kernel from the committed ncu --set full capture
                         # (profiles/k4_traffic.json names the capture), next to the same launch's algorithmic bytes
                         "traffic": (traffic or {}).get("dram_bytes_per_launch"),
                         "traffic_launch_algorithmic_bytes": (traffic or {}).get("algorithmic_bytes_per_launch"),
                         "traffic_source": (traffic or {}).get("source"),
                         "peak_source": peak_src, "algorithmic_bytes_per_fit": hist_bytes, "launches_per_fit": hist_launches,
                         "kernel_seconds_per_fit": hist_s, "kernel_share_of_step": hist_s / (ms_per_step * 1e-3),
                         # physical: (row, feature) histogram updates actually executed (rows visited by the K4 launches x features per node);
                         # weighted: the same with every row counted bag-weight times (what MLlib's per-instance loop would execute)
                         "physical_bin_updates_per_s": phys_updates / hist_s if hist_s > 0 and phys_updates else None,
                         "weighted_bin_updates_per_s": updates_per_fit / hist_s if hist_s > 0 else None,
                         "note": "K4 is bound by shared-memory atomic throughput, not HBM: 22 bin updates (9 native 32-bit adds each; 7 for integer labels) per 37 "
                                 "algorithmic bytes; see DESIGN.md section 2 and profiles/"},
            "e2e": e2e, "cpu_baseline": cpu, "cpu_baseline_sklearn": skl, "mgpu_parity": mgpu_parity, "c4": c4,
            "gpu_launches": gpu_launches, "host_numa_node_rank0": numa_node,
            "clocks": sampler.summary(),
        }
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
