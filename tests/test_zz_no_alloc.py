"""mk_select_flagged_static (the allocation-free variant of mk_select_flagged:
static __device__ scratch inside the library) against mk_select_flagged, through
the C-ABI, on two calls in a row.  Collected last on purpose
(file name)."""
import ctypes as C

import numpy as np
import pytest
import torch

from mink_octtree_stablediffusion_b200 import _lib


def _select(lib, keep, static):
    n = keep.numel()
    kept = torch.full((max(n, 1),), -5, dtype=torch.int32, device="cuda")
    cnt = torch.full((1,), -1, dtype=torch.int64, device="cuda")
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    kp = C.c_void_p(keep.data_ptr()) if n else None
    if static:
        rc = lib.mk_select_flagged_static(kp, n, C.c_void_p(kept.data_ptr()), C.c_void_p(cnt.data_ptr()), stream)
    else:
        ws = torch.empty(int(lib.mk_select_workspace_bytes(n)), dtype=torch.uint8, device="cuda")
        rc = lib.mk_select_flagged(kp, n, C.c_void_p(kept.data_ptr()), C.c_void_p(ws.data_ptr()),
                                   C.c_void_p(cnt.data_ptr()), stream)
    _lib.check(rc)
    k = int(cnt.item())
    return k, kept[:k].cpu().numpy()


@pytest.mark.gpu
def test_static_select_matches_workspace_select_twice(lib):
    rng = np.random.default_rng(7)
    # two calls in a row on different sizes / densities: the second must not see the first's scratch
    for n, p in ((1_000_003, 0.37), (70_001, 0.9), (5, 1.0), (0, 0.5), (2_500_000, 0.02)):
        keep_np = (rng.random(n) < p).astype(np.uint8)
        keep = torch.from_numpy(keep_np).cuda()
        k_ref, rows_ref = _select(lib, keep, static=False)
        for _ in range(2):
            k, rows = _select(lib, keep, static=True)
            assert k == k_ref == int(keep_np.sum())
            assert np.array_equal(rows, rows_ref)            # integer work: bit-exact
            assert np.array_equal(rows, np.nonzero(keep_np)[0].astype(np.int32))


@pytest.mark.gpu
def test_static_select_allocates_nothing(lib):
    n = 300_000
    keep = (torch.rand(n, device="cuda") < 0.5).to(torch.uint8)
    kept = torch.empty(n, dtype=torch.int32, device="cuda")
    cnt = torch.empty(1, dtype=torch.int64, device="cuda")
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    lib.mk_select_flagged_static(C.c_void_p(keep.data_ptr()), n, C.c_void_p(kept.data_ptr()),
                                 C.c_void_p(cnt.data_ptr()), stream)          # set-up call (symbol address)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    stats0 = torch.cuda.memory_stats().get("allocation.all.allocated", 0)
    for _ in range(3):
        _lib.check(lib.mk_select_flagged_static(C.c_void_p(keep.data_ptr()), n, C.c_void_p(kept.data_ptr()),
                                                C.c_void_p(cnt.data_ptr()), stream))
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert torch.cuda.memory_stats().get("allocation.all.allocated", 0) == stats0   # no torch allocation
    assert free1 >= free0                                                            # no driver allocation
    assert int(cnt.item()) == int(keep.sum().item())


def test_static_select_rejects_oversize(lib):
    # argument validation happens before any CUDA call: runs without a GPU
    limit = int(lib.mk_select_static_max_rows())
    assert limit == 8192 * 2048
    rc = lib.mk_select_flagged_static(None, limit + 1, None, None, None)
    assert rc == 1
    assert b"exceeds the static scratch limit" in lib.mk_last_error()
