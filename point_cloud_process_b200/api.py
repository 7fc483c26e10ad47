"""Batched extension of the reference interface: whole query sets / pair batches per
call.  Thin Python over the pybind11 module, which is thin C++ over the C ABI (include/pcp.h)."""
import threading

import numpy as np

from . import native

_tls = threading.local()
MAX_K = 32                 # PCP_MAX_K: register top-k width of the kNN kernels
EMPTY_DIST = 1e10          # PCP_KNN_EMPTY_DIST (result_set.py:19-22)


def default_context(device=0):
    """One context (GPU + stream + scratch) per host thread and device."""
    ctxs = getattr(_tls, "ctxs", None)
    if ctxs is None:
        ctxs = _tls.ctxs = {}
    if device not in ctxs:
        ctxs[device] = native.Context(device)     # raises RuntimeError when no sm_100 GPU: no CPU fallback
    return ctxs[device]


def _as_cloud(a, name):
    a = np.asarray(a)
    if a.ndim == 1 and a.shape[0] == 3:
        a = a[None, :]
    if a.ndim != 2 or a.shape[1] < 3:
        raise ValueError(f"{name}: expected an (N, 3) row-per-point array, got shape {a.shape}; "
                         "the GPU search structure is 3-D only")
    if a.shape[1] != 3:
        a = a[:, :3]
    # float64 input (kdtree.py:219 demo) is down-cast: bit-parity is guaranteed for float32 only
    return np.ascontiguousarray(a, dtype=np.float32)


class SearchIndex:
    """Replaces the kd-tree / octree root handle (kdtree.py:119-137, octree.py:310-328)."""

    def __init__(self, db_np, device=0, ctx=None):
        self.ctx = ctx or default_context(device)
        self.db = _as_cloud(db_np, "db_np")
        self._ix = native.Index(self.ctx, self.db)

    def __len__(self):
        return self._ix.size()

    def stats(self):
        return self._ix.stats()

    def knn(self, queries, k):
        """-> (idx (Q,k) int32, dist (Q,k) float32 Euclidean, count (Q,) int32); rows ascending by
        (distance, index); slots >= count hold (1e10, 0) like KNNResultSet (result_set.py:19-22)."""
        q, k = _as_cloud(queries, "queries"), int(k)
        if k < 1:
            raise ValueError("k must be >= 1")
        if k <= MAX_K:
            return self._ix.knn(q, k)
        return self._knn_wide(q, k)

    def _knn_wide(self, q, k):
        """k > 32, the register top-k width of the search kernels (PCP_MAX_K): KNNResultSet accepts any capacity
        (result_set.py:16-24), so the same result is assembled from two GPU searches -- the 32-NN distance of a
        query bounds from below the radius that holds k neighbours; a radius search (exact float32 distances,
        rows sorted by (distance, index)) is repeated with a growing radius for the queries still short of
        min(k, n) hits, and the first k entries of each row are the answer.  Host work is bookkeeping only."""
        nq, n = len(q), len(self)
        idx = np.zeros((nq, k), np.int32)
        dist = np.full((nq, k), EMPTY_DIST, np.float32)
        cnt = np.zeros(nq, np.int32)
        kk = min(k, n)
        if nq == 0 or kk == 0:
            return idx, dist, cnt
        _, d32, _ = self._ix.knn(q, MAX_K)
        # k-th distance estimate from the 32nd: volume grows with the cube of the radius
        floor = 1e-4 * float(np.ptp(self.db, axis=0).max()) + 1e-30     # (a query sitting on > 32 duplicates)
        r = np.maximum(d32[:, MAX_K - 1].astype(np.float64), floor) * (kk / MAX_K) ** (1.0 / 3.0) * 1.25
        pending = np.arange(nq)
        while len(pending):
            rad = float(np.float32(r[pending].max()))
            if not np.isfinite(rad):
                rad = float(np.finfo(np.float32).max)
            off, ridx, rdist = self._ix.radius(q[pending], rad, True)
            got = np.diff(off)
            done = got >= kk
            for j in np.nonzero(done)[0]:
                i = pending[j]
                idx[i, :kk] = ridx[off[j]:off[j] + kk]
                dist[i, :kk] = rdist[off[j]:off[j] + kk]
                cnt[i] = kk
            pending = pending[~done]
            r[pending] = np.minimum(2.0 * rad, np.finfo(np.float32).max)
        return idx, dist, cnt

    def knn_bruteforce(self, queries, k):
        """Same contract as knn(), computed by the exhaustive device kernel (the reference's own check,
        np.argsort(np.linalg.norm(db - query, axis=1))[:k], kdtree.py:234-238): every (query, point) pair, no
        spatial pruning.  The on-device exhaustive cross-check of the test-suite; k <= 32."""
        return self._ix.knn_bruteforce(_as_cloud(queries, "queries"), int(k))

    def radius_count_bruteforce(self, queries, radius):
        """Per-query number of points with float32 dist <= radius, by the exhaustive kernel."""
        return self._ix.radius_count_bruteforce(_as_cloud(queries, "queries"), float(np.float32(radius)))

    def knn_self(self, k):
        """kNN of every indexed point against the cloud itself, original order."""
        return self._ix.knn_self(int(k))

    def radius(self, queries, radius, sorted=False):
        """-> CSR (offsets (Q+1,) int64, idx (M,) int32, dist (M,) float32): all points with
        float32 dist <= radius (result_set.py:80)."""
        return self._ix.radius(_as_cloud(queries, "queries"), float(np.float32(radius)), bool(sorted))


def icp_batch(sources, targets, T_init=None, max_iteration=100, R_diff_thres=0.5, t_diff_thres=0.5,
              dist_thres=5.0, first_iter_broadcast=True, device=0, ctx=None, want_log=False):
    """Batched icp_point2point (main.py:97-156) over independent pairs.

    sources / targets: lists of (N_i, 3) arrays.  Returns dict with T_last (P,4,4) -- what the
    reference returns --, T_total (P,4,4) composed pose, iterations (P,), status (P,) and, with
    want_log, diff_log (P, max_iteration, 2)."""
    ctx = ctx or default_context(device)
    src = [_as_cloud(s, "source") for s in sources]
    tgt = [_as_cloud(t, "target") for t in targets]
    if len(src) != len(tgt):
        raise ValueError("sources and targets must have the same length")
    P = len(src)
    so = np.zeros(P + 1, np.int64)
    to = np.zeros(P + 1, np.int64)
    if P:
        so[1:] = np.cumsum([len(s) for s in src])
        to[1:] = np.cumsum([len(t) for t in tgt])
    sp = np.concatenate(src) if P else np.zeros((0, 3), np.float32)
    tp = np.concatenate(tgt) if P else np.zeros((0, 3), np.float32)
    if T_init is None:
        T0 = np.tile(np.eye(4), (P, 1, 1))
    else:
        T0 = np.ascontiguousarray(np.asarray(T_init, np.float64).reshape(P, 4, 4))
    return native.icp_batch(ctx, sp, so, tp, to, T0, int(max_iteration), float(R_diff_thres),
                            float(t_diff_thres), float(dist_thres), bool(first_iter_broadcast), bool(want_log))
