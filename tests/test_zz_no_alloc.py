"""GPU test of vslam_bf_knn2_hamming256_noalloc (include/vslam_b200.h): the allocation-free kNN-2
entry point (static __device__ scratch) must give exactly what vslam_bf_knn2_hamming256 gives --
bit-exact, the bar of every matching test -- on calls in a row that share the scratch.

Collected last on purpose (file name).
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_QUERY = 8192   # VSLAM_NOALLOC_MAX_QUERY


def _lib():
    from orbslam2 import _native
    _native.require_gpu()
    L = _native.lib()
    vp, i32 = C.c_void_p, C.c_int
    L.vslam_bf_knn2_hamming256_noalloc.argtypes = [vp, i32, vp, i32, vp, vp, vp]
    L.vslam_bf_knn2_hamming256_noalloc.restype = i32
    return L


def _both(L, q, t):
    """(idx, dist) of the existing device entry point and of the no-alloc one, on the same device buffers."""
    import torch
    dev = torch.device("cuda:0")
    dq, dt = torch.from_numpy(q).to(dev), torch.from_numpy(t).to(dev)
    nq, nt = len(q), len(t)
    outs = [torch.full((max(nq, 1), 2), -7, dtype=torch.int32, device=dev) for _ in range(4)]
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    p = lambda x: C.c_void_p(x.data_ptr())
    rc = L.vslam_bf_knn2_hamming256(p(dq), nq, nq, p(dt), nt, nt, None, None, 1, p(outs[0]), p(outs[1]), s)
    assert rc == 0, L.vslam_last_error()
    rc = L.vslam_bf_knn2_hamming256_noalloc(p(dq), nq, p(dt), nt, p(outs[2]), p(outs[3]), s)
    assert rc == 0, L.vslam_last_error()
    torch.cuda.synchronize()
    return [o.cpu().numpy()[:nq] for o in outs]


def test_noalloc_knn2_equals_existing_entry_point_on_calls_in_a_row(oracle):
    L = _lib()
    rng = np.random.default_rng(11)
    # sizes in a row on the shared scratch: large -> small (stale partials of the larger call stay behind),
    # split and unsplit train ranges, ragged tails, the maximum, one train row, ties
    cases = [(2000, 2000), (300, 5000), (5, 1), (MAX_QUERY, 4099), (129, 127), (2000, 2000), (40, 70000)]
    for n, (nq, nt) in enumerate(cases):
        q = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
        t = rng.integers(0, 256, (nt, 32), dtype=np.uint8)
        if n == 4:
            t[[3, 9, 100]] = t[3]
            q[0] = t[3]
        idx, dist, idx_na, dist_na = _both(L, q, t)
        assert np.array_equal(idx_na, idx), (nq, nt)
        assert np.array_equal(dist_na, dist), (nq, nt)
        if nq * nt <= 4_000_000:
            ridx, rdist = oracle.bf_knn2(q, t)
            assert np.array_equal(idx_na, ridx) and np.array_equal(dist_na, rdist), (nq, nt)
    # the same inputs twice in a row give the same answer
    q = rng.integers(0, 256, (1500, 32), dtype=np.uint8)
    t = rng.integers(0, 256, (1700, 32), dtype=np.uint8)
    a = _both(L, q, t)
    b = _both(L, q, t)
    assert all(np.array_equal(x, y) for x, y in zip(a, b))
    assert np.array_equal(a[2], a[0]) and np.array_equal(a[3], a[1])


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    import torch
    L = _lib()
    L.vslam_last_error.restype = C.c_char_p
    buf = torch.zeros(64, dtype=torch.uint8, device="cuda:0")
    p = C.c_void_p(buf.data_ptr())
    rc = L.vslam_bf_knn2_hamming256_noalloc(p, MAX_QUERY + 1, p, 1, p, p, None)
    assert rc == -3
    assert b"VSLAM_NOALLOC_MAX_QUERY" in L.vslam_last_error()
    assert L.vslam_bf_knn2_hamming256_noalloc(p, -1, p, 1, p, p, None) == -1
    # empty query set: nothing to do; empty train set: every row (-1, -1)
    assert L.vslam_bf_knn2_hamming256_noalloc(None, 0, None, 0, None, None, None) == 0
    q = torch.zeros((3, 32), dtype=torch.uint8, device="cuda:0")
    oi = torch.zeros((3, 2), dtype=torch.int32, device="cuda:0")
    od = torch.zeros((3, 2), dtype=torch.int32, device="cuda:0")
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = L.vslam_bf_knn2_hamming256_noalloc(C.c_void_p(q.data_ptr()), 3, None, 0, C.c_void_p(oi.data_ptr()),
                                            C.c_void_p(od.data_ptr()), s)
    assert rc == 0
    torch.cuda.synchronize()
    assert (oi.cpu().numpy() == -1).all() and (od.cpu().numpy() == -1).all()
