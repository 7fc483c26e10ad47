"""Host-side mirror of the two SEARCH ops of the reference's PointNet++ extension
(Classification_ModelNet40/pointnet2_ops_lib/pointnet2_ops/pointnet2_utils.py), same names and argument order,
numpy arrays in place of torch tensors:

    ball_query(radius, nsample, xyz (B, N, 3), new_xyz (B, npoint, 3)) -> idx (B, npoint, nsample) int32   :243-276
    three_nn(unknown (B, n, 3), known (B, m, 3)) -> (dist (B, n, 3) = sqrt of the squared distances, idx)   :104-136

Both run on the GPU behind pcp_ball_query / pcp_three_nn (include/pcp.h) with the selection rules of the
reference's kernels (first nsample hits in index order; strict '<', lowest index wins).  The rest of that
extension (grouping, interpolation, sampling, the networks) is deep-learning code outside this path.
"""
import numpy as np

from . import native
from .api import default_context


def _f32(a, name):
    a = np.ascontiguousarray(a, dtype=np.float32)
    if a.ndim != 3 or a.shape[2] != 3:
        raise ValueError(f"{name}: expected a (B, N, 3) array, got shape {a.shape}")
    return a


def ball_query(radius, nsample, xyz, new_xyz, ctx=None):
    return native.ball_query(ctx or default_context(), float(radius), int(nsample), _f32(xyz, "xyz"), _f32(new_xyz, "new_xyz"))


def three_nn(unknown, known, ctx=None):
    dist2, idx = native.three_nn(ctx or default_context(), _f32(unknown, "unknown"), _f32(known, "known"))
    return np.sqrt(dist2), idx
