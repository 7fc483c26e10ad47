"""TEST INFRASTRUCTURE: CPU restatement (numpy / scipy) of the global-initialisation path of Registration/main.py:33-84
and icp_template.py:20-41,56-111 -- hybrid-neighbourhood normals, FPFH, feature matching, 3-point rigid transform and
hypothesis scoring -- as the product implements it on the GPU (pcp_global.cu, pcp_global_ransac.cu).

PARITY UNPINNED: the reference delegates all of this to Open3D (estimate_normals, compute_fpfh_feature,
registration_ransac_based_on_feature_matching), which is absent here, unpinned in the reference, and driven by a global
random generator; no fixture of the reference pins any value on this path.  This file restates the
published algorithm (Rusu, Blodow, Beetz 2009, with Open3D's conventions: Darboux frame anchored at the point whose
normal is closer to the connecting line, bins floor(11 (f1 + pi) / 2 pi), floor(11 (f + 1) / 2), increments
100 / (k - 1), neighbour weights 1 / d^2, each 11-bin block normalised to 100 before SPFH(p) is added).  The tests
hold the GPU path to it and to properties that need no oracle (hand-computed histograms, planted poses).
"""
import numpy as np
from scipy.spatial import cKDTree


def hybrid_neighbors(pts, radius, max_nn):
    """Rows of neighbour indices: the up-to-max_nn nearest points within radius (float32 distances of the product,
    ascending (distance, index), the point itself first)."""
    p32 = np.asarray(pts, np.float32)
    tree = cKDTree(p32.astype(np.float64))
    rows = []
    r32 = np.float32(radius)
    for i, cand in enumerate(tree.query_ball_point(p32.astype(np.float64), float(radius) * 1.0001 + 1e-6)):
        cand = np.asarray(cand, np.int64)
        d = p32[i][None, :] - p32[cand]
        dist = np.sqrt(((d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]).astype(np.float32) + d[:, 2] * d[:, 2]).astype(np.float32))
        keep = ~(dist > r32)
        cand, dist = cand[keep], dist[keep]
        order = np.lexsort((cand, dist))
        rows.append(cand[order][:max_nn])
    return rows


def normals_hybrid(pts, radius, max_nn):
    pts64 = np.asarray(pts, np.float32).astype(np.float64)
    out = np.zeros((len(pts64), 3))
    for i, nb in enumerate(hybrid_neighbors(pts, radius, max_nn)):
        if len(nb) < 3:
            out[i] = (0, 0, 1)
            continue
        q = pts64[nb]
        c = q - q.mean(0)
        w, v = np.linalg.eigh(c.T @ c / len(nb))
        nrm = v[:, 0] / np.linalg.norm(v[:, 0])
        lead = nrm[np.argmax(np.abs(nrm))]
        out[i] = -nrm if lead < 0 else nrm
    return out


def pair_features(p1, n1, p2, n2):
    f = np.zeros(4)
    d = p2 - p1
    ln = np.linalg.norm(d)
    f[3] = ln
    if ln == 0:
        return f
    a1, a2 = n1 @ d / ln, n2 @ d / ln
    if np.arccos(min(1.0, abs(a1))) > np.arccos(min(1.0, abs(a2))) + 1e-9:      # (a clear win only: see pcp_global.cu)
        u, t, d, f[2] = n2, n1, -d, -a2
    else:
        u, t, f[2] = n1, n2, a1
    v = np.cross(d, u)
    vn = np.linalg.norm(v)
    if vn == 0:
        return np.zeros(4)
    v = v / vn
    w = np.cross(u, v)
    f[1] = v @ t
    f[0] = np.arctan2(w @ t, u @ t)
    return f


def _bin(x):
    return int(min(10, max(0, np.floor(x))))


def fpfh(pts, normals, radius, max_nn):
    """-> (n, 33) float64."""
    pts64 = np.asarray(pts, np.float32).astype(np.float64)
    nbrs = hybrid_neighbors(pts, radius, max_nn)
    n = len(pts64)
    spfh = np.zeros((n, 33))
    for i, nb in enumerate(nbrs):
        if len(nb) < 2:
            continue
        incr = 100.0 / (len(nb) - 1)
        for q in nb[1:]:
            f = pair_features(pts64[i], normals[i], pts64[q], normals[q])
            spfh[i, _bin(11 * (f[0] + np.pi) / (2 * np.pi))] += incr
            spfh[i, 11 + _bin(11 * (f[1] + 1) * 0.5)] += incr
            spfh[i, 22 + _bin(11 * (f[2] + 1) * 0.5)] += incr
    out = np.zeros((n, 33))
    for i, nb in enumerate(nbrs):
        acc = np.zeros(33)
        for q in nb[1:]:
            d2 = float(((pts64[q] - pts64[i]) ** 2).sum())
            if d2 == 0:
                continue
            acc += spfh[q] / d2
        for g in range(3):
            s = acc[11 * g:11 * g + 11].sum()
            if s != 0:
                acc[11 * g:11 * g + 11] *= 100.0 / s
        out[i] = acc + spfh[i]
    return out


def feature_match(src, tgt):
    """Nearest row of tgt for every row of src (float64, lowest index wins)."""
    d = ((src[:, None, :] - tgt[None, :, :]) ** 2).sum(-1)
    return d.argmin(1), d.min(1)


def rigid3(s, t):
    """Least-squares rigid transform (proper rotation) mapping the rows of s onto the rows of t: 4x4."""
    sb, tb = s.mean(0), t.mean(0)
    H = (t - tb).T @ (s - sb)
    U, S, Vt = np.linalg.svd(H)
    D = np.diag([1, 1, np.sign(np.linalg.det(U) * np.linalg.det(Vt)) or 1.0])
    R = U @ D @ Vt
    T = np.eye(4)
    T[:3, :3] = R
    T[:3, 3] = tb - R @ sb
    return T


def score(T, src, tgt_tree, max_dist):
    """(fitness, inlier_rmse) of a hypothesis: nearest target point of every transformed source point."""
    q = (np.asarray(src, np.float64) @ T[:3, :3].T + T[:3, 3]).astype(np.float32).astype(np.float64)
    d, _ = tgt_tree.query(q, 1)
    ok = d <= max_dist
    return ok.mean(), (np.sqrt((d[ok] ** 2).mean()) if ok.any() else 0.0)


def synthetic_scene(seed, n=40000, extent=60.0):
    """A cloud with distinctive local geometry: points on random plane patches, spheres and cylinders."""
    rng = np.random.default_rng(seed)
    parts = []
    per = n // 24
    for _ in range(10):                                   # plane patches
        c = rng.uniform(0, extent, 3) * [1, 1, 0.15]
        a, b = np.linalg.qr(rng.normal(size=(3, 3)))[0][:2]
        uv = rng.uniform(-1, 1, (per, 2)) * rng.uniform(4, 10, 2)
        parts.append(c + uv[:, :1] * a + uv[:, 1:] * b)
    for _ in range(8):                                    # spheres
        c = rng.uniform(0, extent, 3) * [1, 1, 0.15]
        v = rng.normal(size=(per, 3))
        parts.append(c + rng.uniform(2, 5) * v / np.linalg.norm(v, axis=1, keepdims=True))
    for _ in range(6):                                    # vertical cylinders
        c = rng.uniform(0, extent, 3) * [1, 1, 0]
        th = rng.uniform(0, 2 * np.pi, per)
        r = rng.uniform(1, 3)
        parts.append(c + np.stack([r * np.cos(th), r * np.sin(th), rng.uniform(0, 8, per)], 1))
    ground = np.stack([rng.uniform(0, extent, n // 4), rng.uniform(0, extent, n // 4), rng.normal(0, 0.02, n // 4)], 1)
    return np.concatenate(parts + [ground]).astype(np.float32)


def planted_pair(seed, n=40000, max_rot_deg=25.0, max_t=6.0, noise=0.02):
    """(target, source, T_gt) with T_gt mapping source onto target; the source is a noisy, re-ordered copy."""
    rng = np.random.default_rng(seed + 1)
    target = synthetic_scene(seed, n)
    ang = np.deg2rad(rng.uniform(-1, 1, 3) * [3.0, 3.0, max_rot_deg])
    cx, cy, cz = np.cos(ang)
    sx, sy, sz = np.sin(ang)
    R = (np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]]) @ np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
         @ np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]]))
    t = rng.uniform(-max_t, max_t, 3) * [1, 1, 0.1]
    c = target.astype(np.float64).mean(0)
    T = np.eye(4)
    T[:3, :3] = R
    T[:3, 3] = c + t - R @ c
    src = (target.astype(np.float64) - T[:3, 3]) @ R + rng.normal(0, noise, target.shape)      # R^T (x - t)
    src = src[rng.permutation(len(src))].astype(np.float32)
    return target, src, T
