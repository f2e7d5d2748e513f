import sys, time, numpy as np, torch
sys.path.insert(0, '.')
import b2rf
from b2rf import trainer
dev = torch.device('cuda:0')
N, T = 10_000_000, 100
po = trainer.default_partitions(N, 4096)
def sync(): torch.cuda.synchronize()
for it in range(4):
    sync(); t = time.perf_counter()
    w, node = trainer.bag(42, po, 0, T, 1.0, True, dev, want_node=False)
    sync(); tb = time.perf_counter() - t
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    w2 = torch.zeros_like(w)
    sync(); e0.record()
    from b2rf._lib import lib, check
    import ctypes as C
    check(lib().b2rf_bag(42, po.ctypes.data_as(C.c_void_p), len(po) - 1, 0, T, 1.0, 1, w2.data_ptr(), w2.stride(0), 0, w2.stride(0), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    e1.record(); sync()
    print(f"iter {it}: trainer.bag wall {tb*1e3:.1f} ms; b2rf_bag device {e0.elapsed_time(e1):.1f} ms; equal={bool(torch.equal(w, w2))}", flush=True)
    del w, node, w2
