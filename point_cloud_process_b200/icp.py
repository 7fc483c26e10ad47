"""Drop-in for the ICP path of Registration/main.py and Registration/icp_template.py.

    icp_point2point(source, target, transformation) -> 4x4 last increment     main.py:97-156
    ICP(src_cloud, tgt_cloud) -> (homo_mat_total 4x4, log)                    icp_template.py:128-200
    find_associations(src_points (3,N), tgt_tree) -> int array (M, 2)         icp_template.py:113-126 / main.py:116-121
    procrustes_transformation(A (3,N), B (3,N)) -> (R, t, cost)               icp_template.py:43-54 / main.py:131-141
    rotmat2quaternion / copysign / homo2tq                                    main.py:158-180
    read_bin_velodyne / read_oxford_bin                                       main.py:10-17 / icp_template.py:11-17
    write_reg_result(rows, path)                                              main.py:220-222 (reg_result.txt format)
    icp_batch(...)                                                            batched extension (api.py)

All arithmetic of the loop runs on the GPU (include/pcp.h: pcp_icp_batch, pcp_procrustes, pcp_knn);
this module only adapts argument types.  `source` / `target` may be Open3D-like objects with a
`.points` attribute (main.py:112-113) or (N, 3) arrays.
"""
import numpy as np

from . import native
from .api import SearchIndex, default_context, icp_batch   # noqa: F401

# constants of main.py:98-103
MAX_ITERATION = 100
R_DIFF_THRES = 0.5
T_DIFF_THRES = 0.5
DIST_THRES = 5.0

FEW_PAIRS = 2      # PCP_ICP_FEW_PAIRS


def _points(cloud):
    pts = np.asarray(cloud.points if hasattr(cloud, "points") else cloud)
    if pts.ndim != 2:
        raise ValueError("expected a point cloud with .points or an (N, 3) array")
    if pts.shape[1] != 3 and pts.shape[0] in (3, 6):       # read_oxford_bin layout: 6 x N / 3 x N
        pts = pts[:3].T
    return np.ascontiguousarray(pts[:, :3], dtype=np.float32)


def icp_point2point(source, target, transformation, max_iteration=MAX_ITERATION, R_diff_thres=R_DIFF_THRES,
                    t_diff_thres=T_DIFF_THRES, dist_thres=DIST_THRES, return_info=False):
    """Same contract as main.py:97-156, including its as-built behaviour: the return value is the LAST
    INCREMENT (main.py:156), iteration 1 uses the broadcast t_diff of main.py:150, associations are kept
    iff squared distance < dist_thres, and an Open3D-like `source` is left transformed by every pose
    that was applied to it (main.py:110 transforms in place)."""
    out = icp_batch([_points(source)], [_points(target)], T_init=np.asarray(transformation, np.float64)[None],
                    max_iteration=max_iteration, R_diff_thres=R_diff_thres, t_diff_thres=t_diff_thres,
                    dist_thres=dist_thres, first_iter_broadcast=True, want_log=return_info)
    T_last, T_total, status = out["T_last"][0], out["T_total"][0], int(out["status"][0])
    if status == FEW_PAIRS:
        print("ICP failed, cannot find enough associations!")     # main.py:126
    applied = T_total if status == FEW_PAIRS else np.linalg.inv(T_last) @ T_total
    if hasattr(source, "transform"):
        source.transform(applied)
    if return_info:
        n = int(out["iterations"][0])
        log = {"R_diff": out["diff_log"][0][:n, 0].tolist(), "t_diff": out["diff_log"][0][:n, 1].tolist()}
        if status == FEW_PAIRS:
            log = {k: v[:n - 1] for k, v in log.items()}
        return T_last, {"total": T_total, "applied": applied, "iterations": n, "status": status, "log": log}
    return T_last


def ICP(src_cloud, tgt_cloud, max_iteration=MAX_ITERATION, R_diff_thres=R_DIFF_THRES, t_diff_thres=T_DIFF_THRES,
        dist_thres=DIST_THRES):
    """icp_template.py:128-200: identity initial pose; returns the ACCUMULATED pose and the
    {"R_diff": [...], "t_diff": [...]} log (plain norms: the template has no (3,1)-(3,) broadcast)."""
    out = icp_batch([_points(src_cloud)], [_points(tgt_cloud)], max_iteration=max_iteration,
                    R_diff_thres=R_diff_thres, t_diff_thres=t_diff_thres, dist_thres=dist_thres,
                    first_iter_broadcast=False, want_log=True)
    if int(out["status"][0]) == FEW_PAIRS:
        print("ICP failed, cannot find enough associations!")     # icp_template.py:170
    n = int(out["iterations"][0]) - (1 if int(out["status"][0]) == FEW_PAIRS else 0)
    log = {"R_diff": out["diff_log"][0][:n, 0].tolist(), "t_diff": out["diff_log"][0][:n, 1].tolist()}
    return out["T_total"][0], log


def find_associations(src_points, tgt_tree=None, dist_thres=DIST_THRES):
    """src_points (3, N); tgt_tree: a SearchIndex / kdtree / octree handle over the target.  Keeps (i, nn(i))
    iff the SQUARED distance is < dist_thres (main.py:119).  float32 search (bit-exact Euclidean distances of
    the lesson2 contract); the ICP loop itself uses the float64-refined search inside pcp_icp_batch."""
    if tgt_tree is None:
        raise ValueError("find_associations needs the target's search handle (SearchIndex(target))")
    index = getattr(tgt_tree, "index", tgt_tree)
    src = np.ascontiguousarray(np.asarray(src_points)[:3].T, dtype=np.float32)
    idx, dist, cnt = index.knn(src, 1)
    d2 = dist[:, 0].astype(np.float64) ** 2
    keep = (cnt > 0) & (d2 < dist_thres)
    return np.stack([np.nonzero(keep)[0], idx[keep, 0]], axis=1).astype(np.int64)


def procrustes_transformation(A, B, ctx=None):
    """A, B (3, N) float64 -> (R (3,3), t (3,1), cost) with R = U V^T of svd((B-mean)(A-mean)^T), no
    determinant fix (main.py:137-141); reduction and 3x3 Jacobi SVD on the GPU (pcp_procrustes)."""
    A = np.ascontiguousarray(A, dtype=np.float64)
    B = np.ascontiguousarray(B, dtype=np.float64)
    return native.procrustes(ctx or default_context(), A, B)


def copysign(v, s):                                           # main.py:176-180
    return -v if v * s < 0 else v


def rotmat2quaternion(m):                                     # main.py:158-168
    H = np.eye(4)
    H[:3, :3] = np.asarray(m, np.float64)[:3, :3]
    tq = native.homo2tq(H)
    return tq[3], tq[4], tq[5], tq[6]


def homo2tq(homo_mat):                                        # main.py:170-174
    return tuple(native.homo2tq(np.ascontiguousarray(homo_mat, np.float64)))


def read_bin_velodyne(path):
    """main.py:10-17: Oxford clouds are 6 float32 per point (x,y,z,nx,ny,nz); returns xyz (N, 3) float32."""
    return np.fromfile(path, dtype=np.float32).reshape(-1, 6)[:, :3].copy()


def read_oxford_bin(bin_path):
    """icp_template.py:11-17: returns [x,y,z,nx,ny,nz] as 6 x N."""
    d = np.fromfile(bin_path, dtype=np.float32)
    return np.transpose(d.reshape(-1, 6))


REG_HEADER = "idx1,idx2,t_x,t_y,t_z,q_w,q_x,q_y,q_z"
REG_FMT = "%i,%i,%f,%f,%f,%f,%f,%f,%f"


def write_reg_result(rows, path):
    """main.py:220-222: rows of [trg, src, tx, ty, tz, qw, qx, qy, qz] -> reg_result.txt."""
    np.savetxt(path, np.asarray(rows, np.float64), delimiter=",", header=REG_HEADER, fmt=REG_FMT)


def read_pair_list(path):
    """Pair list of registration_dataset/reg_result.txt (main.py:186-192): [(trg, src), ...]."""
    pairs = []
    with open(path) as f:
        for line in f.readlines()[1:]:
            parts = line.strip().split(",")
            if len(parts) >= 2:
                pairs.append((int(float(parts[0])), int(float(parts[1]))))
    return pairs


def register_pairs(sources, targets, inits, pair_ids, out_path=None, **kw):
    """Batched form of the main.py:189-218 loop body after the (Open3D, out-of-scope) global
    initialisation: ICP on every pair in one GPU call, rows in the reg_result.txt layout."""
    out = icp_batch(sources, targets, T_init=inits, **kw)
    rows = np.zeros((len(pair_ids), 9))
    for i, (trg, src) in enumerate(pair_ids):
        rows[i] = [trg, src, *homo2tq(out["T_last"][i])]
    if out_path is not None:
        write_reg_result(rows, out_path)
    return rows, out
