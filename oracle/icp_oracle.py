"""oracle/icp_oracle.py -- TEST INFRASTRUCTURE ONLY.

float64 numpy restatement of the reference's point-to-point ICP loop and its pose/IO helpers.
It is the checker for the CUDA ICP kernel and the "port" CPU baseline of bench.py; nothing under
point_cloud_process_b200/ may import it.

Restated (paths relative to /root/reference):
    Registration/main.py:97-156          icp_point2point          -> icp_point2point()
    Registration/main.py:158-180         rotmat2quaternion, homo2tq, copysign
    Registration/main.py:220-222         reg_result.txt writer    -> format_reg_result()
    Registration/icp_template.py:43-54   procrustes_transformation(A, B) -> (R, t, cost)
    Registration/icp_template.py:113-126 find_associations(src_points (3,N), tgt_tree)
    Registration/icp_template.py:128-200 ICP(src, tgt) -> (homo_mat_total, log)

Third-party boundary: the reference calls Open3D (version unpinned in Registration/, absent here)
for KDTreeFlann(target).search_knn_vector_3d(p, 1) (main.py:105,117 -- exact 1-NN, returns the
SQUARED distance) and PointCloud.transform (main.py:110 -- in place).  Exact 1-NN is
implementation independent up to ties, so scipy's cKDTree stands in.  Parity pin:
tests/golden/make_golden_icp.py executes the reference's own icp_point2point source under a
minimal Open3D stand-in and tests/test_oracle_golden.py compares this restatement with it.
"""
import numpy as np
from scipy.spatial import cKDTree


def procrustes_transformation(A, B):
    """main.py:131-141 / icp_template.py:43-54.  A, B: (3, N) float64.  The reference centres with
    the dense N x N matrix L = I - 11^T/N (main.py:134-136); A @ L == A - mean(A), restated so."""
    N = A.shape[1]
    Ap = A - A.mean(axis=1, keepdims=True)
    Bp = B - B.mean(axis=1, keepdims=True)
    mediate = Bp @ Ap.T                                     # main.py:137
    u, s, vt = np.linalg.svd(mediate)                       # main.py:138
    R = u @ vt                                              # main.py:139 (no det(R)=+1 fix)
    t = (1.0 / N) * ((B - R @ A) @ np.ones((N, 1)))         # main.py:140, shape (3,1)
    cost = np.linalg.norm(B - (R @ A + t * np.ones((1, N))))  # main.py:141
    return R, t, cost


def find_associations(src_points, tgt_tree, dist_thres=5.0):
    """icp_template.py:113-126 with main.py:116-121 semantics: src_points (3,N); keep (i, nn(i))
    iff the SQUARED distance is < dist_thres.  Returns int array (M, 2)."""
    d, j = tgt_tree.query(src_points.T, k=1)
    keep = d * d < dist_thres
    return np.stack([np.nonzero(keep)[0], j[keep]], axis=1).astype(np.int64)


def icp_point2point(source, target, transformation, max_iteration=100, R_diff_thres=0.5,
                    t_diff_thres=0.5, dist_thres=5.0, return_info=False):
    """main.py:97-156.  source/target: (N,3) arrays (float32 on disk, float64 inside Open3D).
    Returns the LAST INCREMENT (main.py:156); with return_info also the composed pose
    (the icp_template.py:136,198 contract), iteration count, status and the R_diff/t_diff log."""
    src = np.asarray(source, np.float64).copy()
    tgt = np.asarray(target, np.float64)
    T = np.asarray(transformation, np.float64)
    R_last = T[:3, :3]                                      # main.py:99
    t_last = T[:3, 3]                                       # main.py:100 -- shape (3,)
    tree = cKDTree(tgt)                                     # main.py:105
    total = np.eye(4)
    log = {"R_diff": [], "t_diff": []}
    iters, status = 0, 0
    for _ in range(max_iteration):                          # main.py:107
        src = src @ T[:3, :3].T + T[:3, 3]                  # main.py:110 in place, cumulative
        total = T @ total
        iters += 1
        pairs = find_associations(src.T, tree, dist_thres)  # main.py:116-121
        if pairs.shape[0] < 3:                              # main.py:125-127
            status = 1
            break
        B = tgt[pairs[:, 1]].T                              # main.py:131
        A = src[pairs[:, 0]].T                              # main.py:132
        R, t, _ = procrustes_transformation(A, B)
        T = np.zeros((4, 4))                                # main.py:143-146
        T[:3, :3] = R
        T[:3, 3] = t.squeeze()
        T[3, 3] = 1.0
        R_diff = np.linalg.norm(R - R_last)                 # main.py:149
        t_diff = np.linalg.norm(t - t_last)                 # main.py:150: (3,1)-(3,) broadcasts on iteration 1
        R_last, t_last = R, t                               # main.py:151-152: t_last is (3,1) from now on
        log["R_diff"].append(float(R_diff))
        log["t_diff"].append(float(t_diff))
        if R_diff <= R_diff_thres and t_diff <= t_diff_thres:   # main.py:153-154
            break
    if return_info:
        # pose that maps the ORIGINAL source onto the target, including the last increment
        composed = T @ total if status == 0 else total
        return T, {"total": composed, "applied": total, "iterations": iters, "status": status, "log": log}
    return T


def ICP(src_points, tgt_points, **kw):
    """icp_template.py:128-200 contract: -> (homo_mat_total, log) with identity initial pose."""
    T, info = icp_point2point(src_points, tgt_points, np.eye(4), return_info=True, **kw)
    return info["total"], info["log"]


def copysign(v, s):                                          # main.py:176-180
    if v * s < 0:
        v *= -1
    return v


def rotmat2quaternion(m):                                    # main.py:158-168
    trace = m[0][0] + m[1][1] + m[2][2]
    qw = np.sqrt(max(0, trace + 1)) / 2
    qx = np.sqrt(max(0, 1 + m[0][0] - m[1][1] - m[2][2])) / 2
    qy = np.sqrt(max(0, 1 - m[0][0] + m[1][1] - m[2][2])) / 2
    qz = np.sqrt(max(0, 1 - m[0][0] - m[1][1] + m[2][2])) / 2
    qx = copysign(qx, m[2][1] - m[1][2])
    qy = copysign(qy, m[0][2] - m[2][0])
    qz = copysign(qz, m[1][0] - m[0][1])
    return qw, qx, qy, qz


def homo2tq(homo_mat):                                       # main.py:170-174
    tx, ty, tz = homo_mat[:3, 3]
    qw, qx, qy, qz = rotmat2quaternion(homo_mat[:3, :3])
    return tx, ty, tz, qw, qx, qy, qz


def format_reg_result(rows):
    """main.py:220-222: text np.savetxt(..., fmt='%i,%i,%f,...', header=...) produces."""
    import io
    buf = io.StringIO()
    np.savetxt(buf, np.asarray(rows, np.float64), delimiter=",",
               header="idx1,idx2,t_x,t_y,t_z,q_w,q_x,q_y,q_z", fmt="%i,%i,%f,%f,%f,%f,%f,%f,%f")
    return buf.getvalue()


def make_pair(seed, n=20000, box=(60.0, 60.0, 5.0), max_rot_deg=(2.0, 2.0, 5.0), max_t=0.5, noise=0.01):
    """SURVEY 8(d) C4 recipe: seeded synthetic registration pair (target, source, T_gt) where
    T_gt maps source onto target (x -> R(x-c)+c+t)."""
    rng = np.random.default_rng(seed)
    target = (rng.random((n, 3)) * np.asarray(box)).astype(np.float32)
    c = target.astype(np.float64).mean(0)
    ang = np.deg2rad(rng.uniform(-1, 1, 3) * np.asarray(max_rot_deg))
    cx, cy, cz = np.cos(ang)
    sx, sy, sz = np.sin(ang)
    Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    R = Rz @ Ry @ Rx
    t = rng.uniform(-max_t, max_t, 3)
    source = (target.astype(np.float64) - c - t) @ R + c + rng.normal(0, noise, (n, 3))
    source = source[rng.permutation(n)].astype(np.float32)
    T = np.eye(4)
    T[:3, :3] = R
    T[:3, 3] = c + t - R @ c
    return target, source, T
