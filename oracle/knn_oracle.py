"""ctypes front end of oracle/knn_oracle.c -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  Mirrors the reference call shapes (kdtree.py:119-208, octree.py:216-328 of
/root/reference/Kdtree_Octree/lesson2) in batched form.
"""
import ctypes
import os

import numpy as np

from . import build_oracle

_lib = None


def lib():
    global _lib
    if _lib is None:
        path = build_oracle.build()
        L = ctypes.CDLL(path)
        vp, i64, i32, f32, f64 = (ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_float,
                                  ctypes.c_double)
        L.oracle_kdtree_construction.restype = vp
        L.oracle_kdtree_construction.argtypes = [vp, i64, i32]
        L.oracle_kdtree_destroy.argtypes = [vp]
        L.oracle_octree_construction.restype = vp
        L.oracle_octree_construction.argtypes = [vp, i64, i32, f64]
        L.oracle_octree_destroy.argtypes = [vp]
        for name in ("oracle_kdtree_knn", "oracle_octree_knn"):
            getattr(L, name).argtypes = [vp, vp, i64, i32, vp, vp, vp, vp, i32]
        for name in ("oracle_kdtree_radius", "oracle_octree_radius"):
            getattr(L, name).argtypes = [vp, vp, i64, f32, vp, vp, vp, i32]
        L.oracle_bruteforce_knn.argtypes = [vp, i64, vp, i64, i32, vp, vp, vp, i32]
        L.oracle_pair_dist.argtypes = [vp, vp, vp, vp, i64, vp]
        _lib = L
    return _lib


def _f32(a):
    a = np.ascontiguousarray(a, dtype=np.float32)
    if a.ndim != 2 or a.shape[1] != 3:
        raise ValueError("expected an (N, 3) array")
    return a


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


class _Tree:
    _ctor = None
    _dtor = None
    _knn = None
    _radius = None

    def __init__(self, db, *args):
        self.db = _f32(db)          # the C side borrows this buffer
        self.h = getattr(lib(), self._ctor)(_p(self.db), self.db.shape[0], *args)

    def __del__(self):
        if getattr(self, "h", None):
            getattr(lib(), self._dtor)(self.h)
            self.h = None

    def knn(self, queries, k, threads=1, with_comparisons=False):
        """-> (dist (Q,k) f32, index (Q,k) i64, count (Q,) i32[, comparisons (Q,) i64]);
        unfilled slots hold (1e10, 0) like KNNResultSet (result_set.py:19-22)."""
        q = _f32(queries)
        nq = q.shape[0]
        dist = np.empty((nq, k), np.float32)
        idx = np.empty((nq, k), np.int64)
        cnt = np.empty(nq, np.int32)
        cmp_ = np.zeros(nq, np.int64)
        getattr(lib(), self._knn)(self.h, _p(q), nq, k, _p(dist), _p(idx), _p(cnt), _p(cmp_), threads)
        return (dist, idx, cnt, cmp_) if with_comparisons else (dist, idx, cnt)

    def radius(self, queries, radius, threads=1):
        """-> CSR (offsets (Q+1,) i64, dist (M,) f32, index (M,) i64), tree-visit order per query."""
        q = _f32(queries)
        nq = q.shape[0]
        offsets = np.zeros(nq + 1, np.int64)
        fn = getattr(lib(), self._radius)
        fn(self.h, _p(q), nq, float(radius), _p(offsets), None, None, threads)
        m = int(offsets[-1])
        dist = np.empty(m, np.float32)
        idx = np.empty(m, np.int64)
        fn(self.h, _p(q), nq, float(radius), _p(offsets), _p(dist), _p(idx), threads)
        return offsets, dist, idx


class KdTree(_Tree):
    """kdtree_construction(db_np, leaf_size) restated (kdtree.py:119-137)."""
    _ctor, _dtor, _knn, _radius = ("oracle_kdtree_construction", "oracle_kdtree_destroy",
                                   "oracle_kdtree_knn", "oracle_kdtree_radius")

    def __init__(self, db, leaf_size=32):
        super().__init__(db, int(leaf_size))


class Octree(_Tree):
    """octree_construction(db_np, leaf_size, min_extent) restated (octree.py:310-328)."""
    _ctor, _dtor, _knn, _radius = ("oracle_octree_construction", "oracle_octree_destroy",
                                   "oracle_octree_knn", "oracle_octree_radius")

    def __init__(self, db, leaf_size=32, min_extent=0.0001):
        super().__init__(db, int(leaf_size), float(min_extent))


def bruteforce_knn(db, queries, k, threads=1):
    """Canonical order: ascending (float32 distance, index)."""
    db = _f32(db)
    q = _f32(queries)
    nq = q.shape[0]
    dist = np.empty((nq, k), np.float32)
    idx = np.empty((nq, k), np.int64)
    cnt = np.empty(nq, np.int32)
    lib().oracle_bruteforce_knn(_p(db), db.shape[0], _p(q), nq, k, _p(dist), _p(idx), _p(cnt), threads)
    return dist, idx, cnt


def pair_dist(db, queries, qidx, pidx):
    db = _f32(db)
    q = _f32(queries)
    qi = np.ascontiguousarray(qidx, np.int64)
    pi = np.ascontiguousarray(pidx, np.int64)
    out = np.empty(qi.shape[0], np.float32)
    lib().oracle_pair_dist(_p(db), _p(q), _p(qi), _p(pi), qi.shape[0], _p(out))
    return out
