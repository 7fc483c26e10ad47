"""pcp_knn_bruteforce_noalloc (include/pcp.h): the brute-force kNN entry point that allocates nothing -- its scratch is
a static device array -- against pcp_knn_bruteforce, both through the C ABI.  Collected last on purpose.
Bit-exact: indices, float32 distances and counts, as everywhere in the kNN tests."""
import ctypes
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "point_cloud_process_b200", "libpcp_b200.so")
PCP_HOST = 0
PCP_ERR_UNSUPPORTED = -4
MAX_POINTS = 262144            # PCP_NOALLOC_MAX_POINTS

pytestmark = pytest.mark.gpu


def _lib():
    lib = ctypes.CDLL(LIB)
    lib.pcp_last_error.restype = ctypes.c_char_p
    vp, i64, ci = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int
    lib.pcp_ctx_create.argtypes = [ci, ctypes.POINTER(vp)]
    lib.pcp_ctx_destroy.argtypes = [vp]
    lib.pcp_index_build.argtypes = [vp, vp, i64, i64, ci, ctypes.POINTER(vp)]
    lib.pcp_index_destroy.argtypes = [vp]
    lib.pcp_knn_bruteforce.argtypes = [vp, vp, i64, i64, ci, ci, vp, vp, vp]
    lib.pcp_knn_bruteforce_noalloc.argtypes = [ci, vp, i64, i64, vp, i64, i64, ci, vp, vp, vp, vp]
    return lib


def _existing(lib, pts, qry, k):
    """pcp_knn_bruteforce on host arrays (context + index: the allocating path)."""
    ctx, ix = ctypes.c_void_p(), ctypes.c_void_p()
    assert lib.pcp_ctx_create(0, ctypes.byref(ctx)) == 0, lib.pcp_last_error()
    assert lib.pcp_index_build(ctx, pts.ctypes.data, len(pts), 3, PCP_HOST, ctypes.byref(ix)) == 0, lib.pcp_last_error()
    idx = np.empty((len(qry), k), np.int32)
    dist = np.empty((len(qry), k), np.float32)
    cnt = np.empty(len(qry), np.int32)
    rc = lib.pcp_knn_bruteforce(ix, qry.ctypes.data, len(qry), 3, k, PCP_HOST, idx.ctypes.data, dist.ctypes.data, cnt.ctypes.data)
    assert rc == 0, lib.pcp_last_error()
    lib.pcp_index_destroy(ix)
    lib.pcp_ctx_destroy(ctx)
    return idx, dist, cnt


def _noalloc(lib, torch, d_pts, d_qry, k):
    nq = d_qry.shape[0]
    idx = torch.full((nq, k), -7, dtype=torch.int32, device="cuda:0")
    dist = torch.full((nq, k), -7.0, dtype=torch.float32, device="cuda:0")
    cnt = torch.full((nq,), -7, dtype=torch.int32, device="cuda:0")
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    rc = lib.pcp_knn_bruteforce_noalloc(0, d_pts.data_ptr(), d_pts.shape[0], 3, d_qry.data_ptr(), nq, 3, k,
                                        idx.data_ptr(), dist.data_ptr(), cnt.data_ptr(), None)   # NULL = default stream
    assert rc == 0, lib.pcp_last_error()
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    return idx.cpu().numpy(), dist.cpu().numpy(), cnt.cpu().numpy()


def test_noalloc_matches_bruteforce_two_calls_in_a_row():
    import torch
    lib = _lib()
    rng = np.random.default_rng(20261105)
    # two different clouds back to back: the second call must not see what the first left in the shared scratch
    # (a larger cloud first, then a smaller ragged one whose pads fall where real points were; duplicates for ties)
    cases = []
    a = rng.uniform(-40, 40, (50000, 3)).astype(np.float32)
    cases.append((a, rng.uniform(-45, 45, (3000, 3)).astype(np.float32), 8))
    b = rng.uniform(-5, 5, (1237, 3)).astype(np.float32)
    b[100:140] = b[60]
    cases.append((b, np.concatenate([b[:500], rng.uniform(-6, 6, (777, 3)).astype(np.float32)]), 16))
    cases.append((a[:7], a[:40].copy(), 32))                       # fewer points than k: unfilled slots
    for pts, qry, k in cases:
        want = _existing(lib, pts, qry, k)
        d_pts = torch.from_numpy(pts).to("cuda:0")
        d_qry = torch.from_numpy(qry).to("cuda:0")
        got = _noalloc(lib, torch, d_pts, d_qry, k)
        for g, w, name in zip(got, want, ("idx", "dist", "count")):
            assert np.array_equal(g, w), f"{name} differs (n={len(pts)}, nq={len(qry)}, k={k})"
        again = _noalloc(lib, torch, d_pts, d_qry, k)              # same call again on the warm scratch
        for g, w in zip(again, want):
            assert np.array_equal(g, w)


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    import torch
    lib = _lib()
    one = torch.zeros((1, 3), dtype=torch.float32, device="cuda:0")
    out_i = torch.zeros((1, 8), dtype=torch.int32, device="cuda:0")
    out_d = torch.zeros((1, 8), dtype=torch.float32, device="cuda:0")
    rc = lib.pcp_knn_bruteforce_noalloc(0, one.data_ptr(), MAX_POINTS + 1, 3, one.data_ptr(), 1, 3, 8,
                                        out_i.data_ptr(), out_d.data_ptr(), None, None)
    assert rc == PCP_ERR_UNSUPPORTED
    assert b"PCP_NOALLOC_MAX_POINTS" in lib.pcp_last_error()
    assert lib.pcp_knn_bruteforce_noalloc(0, one.data_ptr(), 1, 3, one.data_ptr(), 1, 3, 33,
                                          out_i.data_ptr(), out_d.data_ptr(), None, None) != 0


def test_noalloc_at_the_maximum_size():
    import torch
    lib = _lib()
    rng = np.random.default_rng(7)
    pts = rng.uniform(0, 100, (MAX_POINTS, 3)).astype(np.float32)
    qry = pts[rng.choice(MAX_POINTS, 2048, replace=False)].copy()
    want = _existing(lib, pts, qry, 8)
    got = _noalloc(lib, torch, torch.from_numpy(pts).to("cuda:0"), torch.from_numpy(qry).to("cuda:0"), 8)
    for g, w in zip(got, want):
        assert np.array_equal(g, w)
