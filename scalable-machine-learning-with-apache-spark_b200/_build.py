"""In-tree build of libb2rf.so (hand-written sm_100a CUDA + the C++ host driver) with nvcc: one object per source file,
compiled in parallel and only when stale, then one link."""
from __future__ import annotations

import os
import subprocess
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
SO = os.path.join(HERE, "libb2rf.so")
SOURCES = ["b2rf_core.cu", "k1_find_splits.cu", "k2_bin.cu", "k3_bag.cu", "k4_hist.cu", "k4_limb.cu", "k5_split.cu", "k6_partition.cu", "k7_predict.cu",
           "k8_multiclass.cu", "b2rf_trainer.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo", "-fmad=false", "-Xcompiler", "-fPIC"]


def _nvcc() -> str:
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    return nvcc if os.path.exists(nvcc) else "nvcc"


def _headers_mtime() -> float:
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))] + [os.path.join(HERE, "..", "include", "b2rf.h")]
    return max(os.path.getmtime(d) for d in deps if os.path.exists(d))


def _stale() -> bool:
    if not os.path.exists(SO):
        return True
    t = os.path.getmtime(SO)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", "b2rf.h")]
    return any(os.path.exists(d) and os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not _stale():
        return SO
    os.makedirs(OBJ, exist_ok=True)
    nvcc = _nvcc()
    hdr_t = _headers_mtime()
    extra = ["-Xptxas", "-v"] if verbose else []

    def compile_one(src):
        s = os.path.join(CSRC, src)
        o = os.path.join(OBJ, src.replace(".cu", ".o"))
        if not force and os.path.exists(o) and os.path.getmtime(o) >= max(os.path.getmtime(s), hdr_t):
            return o
        subprocess.check_call([nvcc] + NVCC_FLAGS + extra + ["-c", "-o", o, s])
        return o

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    subprocess.check_call([nvcc, "--shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", SO] + objs)
    return SO


if __name__ == "__main__":
    print(build(force=True, verbose=True))
