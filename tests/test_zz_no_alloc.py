"""GPU test (-m gpu) of pm_knn_static, the entry point that allocates nothing (include/pmhot.h: no counterpart in
the reference). Its scratch is a file-scope device array of the
library, shared by every call; the test calls it through the C ABI twice in a row and compares each result with
pm_knn's — bit-exact as (dist² bits, id) sets per query, the bar every kNN test here uses — so that a call that
relied on the array being fresh, or left something behind, shows. Collected last on
purpose."""
import ctypes

import numpy as np
import pytest

from helpers import entries, unique_cloud

pytestmark = pytest.mark.gpu


def _rows(ids, d2, r2):
    """per query: the sorted (dist² bits << 32 | id) entries + the returned radius bits"""
    return entries(ids.cpu().numpy(), d2.cpu().numpy()), r2.cpu().numpy().view(np.uint32)


def _call_static(pkg, ctx, m, q, k, cutoff):
    import torch

    nq = q.shape[0]
    # poisoned outputs: every slot must be written by the call
    ids = torch.full((nq, k), 0x5A5A5A5A, dtype=torch.int32, device="cuda")
    d2 = torch.full((nq, k), float("nan"), dtype=torch.float32, device="cuda")
    r2 = torch.full((nq,), float("nan"), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    rc = pkg.lib().pm_knn_static(ctx.handle, m.handle, q.data_ptr(), nq, k, cutoff, ids.data_ptr(), d2.data_ptr(), r2.data_ptr())
    assert rc == 0, pkg.lib().pm_last_error().decode()
    return ids, d2, r2


@pytest.mark.parametrize("n,scale,k,cutoff", [
    (300000, 2.0**-8, 64, 100.0),   # uniform-like cloud: k_knn_fast + hand-overs (the C2 shape, smaller)
    (60000, 2.0**-8, 50, 1.5),      # tight cutoff: unfilled slots (-1 / cutoff²)
    (2000, 2.0**-8, 16, 100.0),     # map without a grid: the warp kd-tree kernel alone
])
def test_knn_static_equals_knn_twice_in_a_row(pkg, ctx, n, scale, k, cutoff):
    import torch

    a = unique_cloud(n, 4242 + n, scale)
    m = pkg.PhotonMap.build(ctx, a.copy())
    rng = np.random.default_rng(n)
    ext = n * scale
    batches = [
        (rng.random((20000, 3), dtype=np.float32) * np.float32(1.1 * ext) - np.float32(0.05 * ext)),
        (rng.random((7777, 3), dtype=np.float32) * np.float32(ext)),   # second call: other size, other queries
    ]
    for qh in batches:
        q = torch.from_numpy(qh).cuda()
        want = _rows(*m.knn(q, k, cutoff))
        got = _rows(*_call_static(pkg, ctx, m, q, k, cutoff))
        assert np.array_equal(got[0], want[0])
        assert np.array_equal(got[1], want[1])
    # and the first batch again, after the array has been used for something else
    q = torch.from_numpy(batches[0]).cuda()
    want = _rows(*m.knn(q, k, cutoff))
    got = _rows(*_call_static(pkg, ctx, m, q, k, cutoff))
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
    # the call allocates nothing: the device's free memory is the same before and after (outputs made beforehand)
    ids = torch.empty((q.shape[0], k), dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    assert pkg.lib().pm_knn_static(ctx.handle, m.handle, q.data_ptr(), q.shape[0], k, cutoff, ids.data_ptr(), None, None) == 0
    assert torch.cuda.mem_get_info()[0] == free0
    m.close()


def test_knn_static_at_its_maximum_and_beyond(pkg, ctx):
    import torch

    n, k = 200000, 8
    a = unique_cloud(n, 77)
    m = pkg.PhotonMap.build(ctx, a.copy())
    nq = pkg.KNN_STATIC_MAX_QUERIES
    rng = np.random.default_rng(5)
    q = torch.from_numpy(rng.random((nq + 1, 3), dtype=np.float32) * np.float32(n * 2.0**-8)).cuda()
    want = _rows(*m.knn(q[:nq], k, 100.0))
    got = _rows(*_call_static(pkg, ctx, m, q[:nq], k, 100.0))
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
    # one query more than the resident array is sized for: a clear error, nothing written
    ids = torch.zeros((nq + 1, k), dtype=torch.int32, device="cuda")
    rc = pkg.lib().pm_knn_static(ctx.handle, m.handle, q.data_ptr(), nq + 1, k, 100.0, ids.data_ptr(), None, None)
    assert rc == -5  # PM_ERR_CAPACITY
    assert b"PM_KNN_STATIC_MAX_QUERIES" in pkg.lib().pm_last_error()
    assert int(ids.abs().sum()) == 0
    m.close()


def test_knn_static_rejects_host_buffers_and_handles_empty_maps(pkg, ctx):
    import torch

    a = unique_cloud(5000, 3)
    m = pkg.PhotonMap.build(ctx, a.copy())
    qh = np.random.default_rng(1).random((100, 3), dtype=np.float32)
    ids_h = np.zeros((100, 8), np.int32)
    q = torch.from_numpy(qh).cuda()
    ids = torch.zeros((100, 8), dtype=torch.int32, device="cuda")
    L = pkg.lib()
    assert L.pm_knn_static(ctx.handle, m.handle, qh.ctypes.data, 100, 8, 100.0, ids.data_ptr(), None, None) == -1
    assert L.pm_knn_static(ctx.handle, m.handle, q.data_ptr(), 100, 8, 100.0, ids_h.ctypes.data, None, None) == -1
    assert b"device memory" in L.pm_last_error()
    assert L.pm_knn_static(ctx.handle, m.handle, q.data_ptr(), 100, 7, 100.0, ids.data_ptr(), None, None) == -1
    assert L.pm_knn_static(ctx.handle, m.handle, q.data_ptr(), 0, 8, 100.0, None, None, None) == 0
    # no map: every slot -1 / cutoff², as pm_knn answers
    d2 = torch.zeros((100, 8), dtype=torch.float32, device="cuda")
    r2 = torch.zeros(100, dtype=torch.float32, device="cuda")
    assert L.pm_knn_static(ctx.handle, None, q.data_ptr(), 100, 8, 2.0, ids.data_ptr(), d2.data_ptr(), r2.data_ptr()) == 0
    assert bool((ids == -1).all()) and bool((d2 == 4.0).all()) and bool((r2 == 4.0).all())
    m.close()
