"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the ICP front end used as the checker by
tests/, never imported by the product path.

  voxel_filter   restates Pca_and_Voxel_filter/voxel_filter.py:10-66 (type="centroid")
  pca / normals  restate  Pca_and_Voxel_filter/pca_normal.py:10-36 and the loop at :85-90

Pinned: tests/golden/front_golden.npz holds outputs of the reference's OWN voxel_filter() and PCA() (imported
from /root/reference with stand-ins for the absent open3d / pyntcloud / matplotlib modules, which those two
functions never touch) -- tests/golden/make_golden_front.py; tests/test_oracle_golden.py checks this file
against them.
"""
import numpy as np


def voxel_filter(points, leaf_size):
    """voxel_filter.py:10-66, centroid mode, vectorised: same float64 arithmetic, same stable order, np.mean per
    run (so numpy's own summation order), last run dropped."""
    pc = np.asarray(points, dtype=np.float64)                       # :19 (Open3D points are float64)
    mx = np.max(pc, axis=0)                                         # :20
    mn = np.min(pc, axis=0)                                         # :21
    Dx = (mx[0] - mn[0]) // leaf_size                               # :22
    Dy = (mx[1] - mn[1]) // leaf_size                               # :23
    hx = np.floor((pc[:, 0] - mn[0]) / leaf_size)                   # :29
    hy = np.floor((pc[:, 1] - mn[1]) / leaf_size)                   # :30
    hz = np.floor((pc[:, 2] - mn[2]) / leaf_size)                   # :31
    h = hx + hy * Dx + hz * Dx * Dy                                 # :32
    order = np.argsort(h, kind="stable")                            # :36 (sorted() is stable)
    hs = h[order]
    heads = np.flatnonzero(np.r_[True, hs[1:] != hs[:-1]])
    ends = np.r_[heads[1:], len(hs)]
    out = []
    for s, e in zip(heads[:-1], ends[:-1]):                         # the last run is never appended (:40-50)
        grp = np.ascontiguousarray(pc[order[s:e]].T)                # np.hstack of (3,1) columns -> (3, g)
        out.append(np.mean(grp, axis=1))                            # :48
    return np.array(out, dtype=np.float64)                          # :65


def voxel_runs(points, leaf_size):
    """Sizes of the runs of equal hash, ascending hash, last run included."""
    pc = np.asarray(points, dtype=np.float64)
    mx, mn = np.max(pc, axis=0), np.min(pc, axis=0)
    Dx, Dy = (mx[0] - mn[0]) // leaf_size, (mx[1] - mn[1]) // leaf_size
    h = (np.floor((pc[:, 0] - mn[0]) / leaf_size) + np.floor((pc[:, 1] - mn[1]) / leaf_size) * Dx
         + np.floor((pc[:, 2] - mn[2]) / leaf_size) * Dx * Dy)
    return np.unique(h, return_counts=True)[1]


def pca(data):
    """PCA(data, correlation=False, sort=True) of pca_normal.py:10-36."""
    pc_array = np.transpose(np.asarray(data, dtype=np.float64))     # :21
    pc_mean = np.mean(pc_array, axis=1, keepdims=True)              # :23
    pc_norm = pc_array - pc_mean                                    # :24
    H = np.cov(pc_norm)                                             # :26
    w, v = np.linalg.eigh(H)                                        # :28
    order = w.argsort()[::-1]                                       # :31-33
    return w[order], v[:, order]


def normals(points, nn_idx):
    """Loop of pca_normal.py:85-90 over given neighbour lists (N, k): normal = v[:, 2]; also the eigenvalues."""
    pts = np.asarray(points, dtype=np.float64)
    nrm = np.empty((len(nn_idx), 3))
    ev = np.empty((len(nn_idx), 3))
    for i, idx in enumerate(nn_idx):
        w, v = pca(pts[idx, :])
        nrm[i], ev[i] = v[:, 2], w
    return nrm, ev


def iss_keypoints(points, radius=0.5, lambda21=0.5, lambda32=0.5, non_max_radius=0.5, iss_count=20):
    """The detection loop of Keypoint_detection_ISS/ISS.py:36-73 as a function (the reference keeps it under
    __main__): -> (iss_idx list, cand_idx list, {index: lambda3}).  Pinned by tests/golden/iss_golden.npz, which
    holds the variables of the reference script itself after it was executed on the same clouds
    (tests/golden/make_golden_iss.py)."""
    from scipy import spatial
    pts = np.asarray(points, dtype=np.float64)[:, :3]
    tree = spatial.cKDTree(pts)                                           # :36
    balls = tree.query_ball_point(pts, radius)
    counts = np.array([len(b) for b in balls], dtype=np.float64)
    cand_idx, lambda3 = [], {}
    for i in range(len(pts)):                                             # :41
        nb = np.asarray(balls[i])
        w = 1.0 / counts[nb]                                              # :49
        d = pts[nb] - pts[i]
        nume = np.zeros((3, 3))
        for wj, dj in zip(w, d):                                          # :48-51 (same accumulation order)
            nume += wj * np.outer(dj, dj)
        denom = 0
        for wj in w:
            denom += wj
        with np.errstate(invalid="ignore", divide="ignore"):
            value, _ = np.linalg.eig(nume / denom)                        # :52
            s = np.sort(value)[::-1]                                      # :53
            if s[1] / s[0] < lambda21 and s[2] / s[1] < lambda32:         # :55
                cand_idx.append(i)
                lambda3[i] = s[2]
    all_l3 = dict(lambda3)
    order = dict(sorted(lambda3.items(), key=lambda x: x[1], reverse=True))   # :59
    iss_idx = []
    for idx in order:                                                     # :61-73
        if idx not in lambda3:
            continue
        iss_idx.append(idx)
        for j in tree.query_ball_point(pts[idx], non_max_radius):
            lambda3.pop(j, None)
        if len(iss_idx) > iss_count:
            break
    return iss_idx, cand_idx, all_l3


def dbscan_labels(data, radius=0.5, min_pts=10):
    """DBSCAN.fit + predict of Cluster_dbscan/dbscan.py:10-38 with the same visiting order, using sets where the
    reference scans Python lists (`cur_ind in unvisited`, `unvisited.remove`, :29,:35) -- identical labels, without
    the O(N^2) list work.  Pinned by tests/golden/dbscan_golden.npz (the reference class itself,
    tests/golden/make_golden_dbscan.py)."""
    from scipy.spatial import KDTree
    data = np.asarray(data, dtype=np.float64)
    n = data.shape[0]
    labels = -1 * np.ones(n)                                       # :12
    visited = np.zeros(n, dtype=bool)
    tree = KDTree(data, leafsize=32)                               # :16
    balls = tree.query_ball_point(data, radius)
    label = -1
    for ind in range(n - 1, -1, -1):                               # unvisited.pop() takes the highest index left (:18)
        if visited[ind]:
            continue
        visited[ind] = True
        nb = balls[ind]
        if len(nb) < min_pts:                                      # :20
            labels[ind] = -1
            continue
        label += 1                                                 # :24
        labels[ind] = label
        stack = list(nb)                                           # :26
        while stack:
            cur = stack.pop()                                      # :28
            if not visited[cur]:                                   # :29
                labels[cur] = label
                if len(balls[cur]) > min_pts:                      # :32 (strict)
                    stack.extend(balls[cur])
                visited[cur] = True                                # :35
    return labels.astype(np.int32)
