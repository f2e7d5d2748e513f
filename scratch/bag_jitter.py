"""Where does the occasional +25 ms of the bagging stage come from?  Times the allocation/memset and the K3 kernel separately."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, ctypes as C
import b2rf
from b2rf import trainer
from b2rf._lib import lib, check
dev = torch.device("cuda:0")
n, T = 10_000_000, 100
po = np.ascontiguousarray(list(range(0, n + 1, 4096)) + ([n] if n % 4096 else []), np.int64)
ld = trainer._pad(n)
X, y, ar = b2rf.synth.c3(n, device=dev, F=64)
p = b2rf.TreeParams(maxDepth=10, maxBins=64, numTrees=T, bootstrap=True, seed=42)
for rep in range(60):
    if rep % 10 == 0:
        f = trainer.fit_tensors(X, y, ar, p, part_offsets=po)      # a full fit in between, as in the bench
    torch.cuda.synchronize(); t0 = time.perf_counter()
    w = torch.zeros((T, ld), dtype=torch.uint8, device=dev)
    torch.cuda.synchronize(); t1 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    check(lib().b2rf_bag(42, po.ctypes.data_as(C.c_void_p), len(po) - 1, 0, T, 1.0, 1, w.data_ptr(), w.stride(0), 0, ld, trainer._stream()))
    e1.record()
    t2 = time.perf_counter()
    torch.cuda.synchronize(); t3 = time.perf_counter()
    if (t3 - t1) > 0.09 or rep % 10 == 0: print(f"rep {rep}: zeros {1e3*(t1-t0):.2f} ms, bag call returns after {1e3*(t2-t1):.2f} ms, done after {1e3*(t3-t1):.2f} ms, device {e0.elapsed_time(e1):.2f} ms", flush=True)
    del w
