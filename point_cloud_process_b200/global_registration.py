"""Global initialisation in front of ICP -- the Open3D calls of Registration/main.py:33-84 and the contract of
icp_template.py:20-41 (find_matchings) and :56-111 (ransac_init), on the GPU behind pcp_normals_hybrid, pcp_fpfh,
pcp_feature_match and pcp_ransac_init (include/pcp.h).  Same function names and argument meaning as main.py; clouds
are (N, 3) arrays (or objects with `.points`).

    preprocess_point_cloud(pcd, voxel_size) -> (pcd_down (M, 3), fpfh (33, M))                  main.py:30-47
    execute_global_registration(source_down, target_down, source_fpfh, target_fpfh, voxel_size) main.py:68-84
        -> RegistrationResult(transformation 4x4, fitness, inlier_rmse, correspondence_set-free)
    find_matchings(src_features (33, Ns), tgt_features (33, Nt)) -> (M, 2) int array            icp_template.py:20-41
    ransac_init(src_cloud, tgt_cloud, voxel_size=2.0) -> (R (3,3), t (3,1))                     icp_template.py:56-111

Open3D is absent here and unpinned in the reference: this restates its published algorithms;
hypotheses come from a counter hash of `seed`, so a run is reproducible -- Open3D's own results are not.
"""
import numpy as np

from . import native
from .api import SearchIndex, default_context
from .voxel_filter import voxel_filter


class RegistrationResult:
    """Stand-in for o3d.pipelines.registration.RegistrationResult: the fields main.py reads (`.transformation`)."""

    def __init__(self, transformation, fitness, inlier_rmse, iterations, evaluated, correspondences):
        self.transformation = transformation
        self.fitness = fitness
        self.inlier_rmse = inlier_rmse
        self.iterations = iterations
        self.hypotheses_evaluated = evaluated
        self.n_correspondences = correspondences

    def __repr__(self):
        return "RegistrationResult with fitness=%e, inlier_rmse=%e" % (self.fitness, self.inlier_rmse)


def _pts(cloud):
    pts = np.asarray(cloud.points if hasattr(cloud, "points") else cloud)
    if pts.ndim != 2 or pts.shape[1] < 3:
        raise ValueError("expected a point cloud with .points or an (N, 3) array")
    return np.ascontiguousarray(pts[:, :3], dtype=np.float32)


def preprocess_point_cloud(pcd, voxel_size, ctx=None):
    """main.py:30-47: voxel down-sample (voxel_size), normals over hybrid neighbourhoods (radius 2 voxels, max_nn 30),
    FPFH (radius 5 voxels, max_nn 100).  Returns the down-sampled points (M, 3) float32 and the features (33, M)."""
    ctx = ctx or default_context()
    down = np.ascontiguousarray(voxel_filter(_pts(pcd), float(voxel_size), "centroid", ctx=ctx), dtype=np.float32)
    ix = SearchIndex(down, ctx=ctx)
    normals = ix._ix.normals_hybrid(float(voxel_size) * 2.0, 30)
    fpfh = ix._ix.fpfh(normals, float(voxel_size) * 5.0, 100)
    return down, np.ascontiguousarray(fpfh.T)


def find_matchings(src_features, tgt_features, src_tree=None, tgt_tree=None, mutual=True, ctx=None):
    """icp_template.py:20-41: features are (length of feature, #points).  Every source feature is paired with its
    nearest target feature; with `mutual` only pairs that are each other's nearest neighbour are kept, unless that
    leaves fewer than 9 of them (Open3D's mutual_filter rule, main.py:75)."""
    ctx = ctx or default_context()
    s = np.ascontiguousarray(np.asarray(src_features, np.float64).T)
    t = np.ascontiguousarray(np.asarray(tgt_features, np.float64).T)
    if len(s) == 0 or len(t) == 0:
        return np.zeros((0, 2), np.int64)
    st, _ = native.feature_match(ctx, s, t)
    pairs = np.stack([np.arange(len(s)), st], axis=1).astype(np.int64)
    if mutual:
        ts, _ = native.feature_match(ctx, t, s)
        keep = ts[st] == np.arange(len(s))
        if int(keep.sum()) >= 9:
            pairs = pairs[keep]
    return pairs


def execute_global_registration(source_down, target_down, source_fpfh, target_fpfh, voxel_size, max_iteration=100000,
                                confidence=0.999, seed=0, ctx=None):
    """main.py:68-84: feature-matching RANSAC, ransac_n = 3, edge-length checker 0.9, distance checker and
    correspondence distance 1.5 voxels, RANSACConvergenceCriteria(100000, 0.999)."""
    ctx = ctx or default_context()
    src, tgt = _pts(source_down), _pts(target_down)
    corr = find_matchings(source_fpfh, target_fpfh, mutual=True, ctx=ctx)
    tix = SearchIndex(tgt, ctx=ctx)
    out = native.ransac_init(ctx, src, tix._ix, np.ascontiguousarray(corr, np.int32), float(voxel_size) * 1.5, 0.9,
                             int(max_iteration), float(confidence), int(seed))
    return RegistrationResult(np.asarray(out["transformation"]), float(out["fitness"]), float(out["inlier_rmse"]),
                              int(out["iterations"]), int(out["evaluated"]), len(corr))


def ransac_init(src_cloud, tgt_cloud, voxel_size=2.0, seed=0, ctx=None):
    """icp_template.py:56-111: feature detection / description / correspondence / RANSAC -> best (R, t)."""
    sd, sf = preprocess_point_cloud(src_cloud, voxel_size, ctx=ctx)
    td, tf = preprocess_point_cloud(tgt_cloud, voxel_size, ctx=ctx)
    res = execute_global_registration(sd, td, sf, tf, voxel_size, seed=seed, ctx=ctx)
    T = res.transformation
    return T[:3, :3].copy(), T[:3, 3:4].copy()
