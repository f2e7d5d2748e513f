"""The allocation-free split finder, b2rf_find_splits_noalloc, through the C-ABI.

Who needs it: a host that embeds this library but owns every byte of GPU memory through its own pool (a Spark executor
plugin is the typical case) or that cannot afford the device-wide stall of cudaMalloc/cudaFree.  It can pass input and
output buffers from its pool, yet b2rf_find_splits also wants a sort workspace; this entry point takes that workspace from
an 8 MiB array compiled into the library instead.  The test checks that its answers are identical (bit for bit, as for
every split-threshold test here) to the allocating entry point's, on two calls in a row that leave different leftovers
in the shared array, and that oversized requests are refused with a readable message.

Collected last on purpose (file name): a failure here must not hide any other test's result.
"""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _noalloc(lib, sample, arity_d, max_possible_bins, wns):
    n_s, F = sample.shape
    ms = max(max_possible_bins - 1, 1)
    thr = torch.zeros((F, ms), dtype=torch.float64, device=sample.device)      # outputs still come from the caller
    ns = torch.zeros(F, dtype=torch.int32, device=sample.device)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.b2rf_find_splits_noalloc(sample.data_ptr(), n_s, F, arity_d.data_ptr(), max_possible_bins, float(wns), thr.data_ptr(),
                                      ns.data_ptr(), st)
    return rc, thr, ns


def test_noalloc_matches_allocating_entry_point_twice(b2rf, cuda):
    from b2rf import trainer
    from b2rf._lib import lib
    L = lib()
    assert L.b2rf_find_splits_noalloc_max_bytes() == 8 << 20
    # call 1: C3-shaped sample (10k rows x 64 continuous features, maxBins 64) fills the whole built-in workspace
    X1, _, ar1 = b2rf.synth.c3(10000, device="cpu")
    # call 2: a smaller, mixed categorical/continuous sample leaves call 1's data in most of the array
    X2, _, ar2 = b2rf.synth.airbnb(3000)
    for X, arity, mpb in ((X1, ar1, 64), (X2, ar2, 40)):
        s = X.to(cuda).contiguous()
        a = torch.as_tensor(arity, device=cuda)
        rc, thr, ns = _noalloc(L, s, a, mpb, float(s.shape[0]))
        assert rc == 0, L.b2rf_last_error().decode()
        thr_ref, ns_ref = trainer.find_splits(s, a, mpb, float(s.shape[0]))     # the existing entry point (caller workspace)
        torch.cuda.synchronize()
        assert torch.equal(ns, ns_ref)
        assert torch.equal(thr, thr_ref)          # thresholds are exact midpoints: same tolerance as the other K1 tests (none)
        assert int(ns.sum()) > 0


def test_noalloc_rejects_oversized_requests(b2rf, cuda):
    from b2rf._lib import ERR_ARG, lib
    L = lib()
    s = torch.zeros((20000, 64), dtype=torch.float64, device=cuda)              # next_pow2(20000) * 64 * 8 = 16 MiB > 8 MiB
    a = torch.zeros(64, dtype=torch.int32, device=cuda)
    rc, _, _ = _noalloc(L, s, a, 64, 20000.0)
    assert rc == ERR_ARG
    msg = L.b2rf_last_error().decode()
    assert "built-in one holds 8388608 bytes" in msg and "b2rf_find_splits" in msg
