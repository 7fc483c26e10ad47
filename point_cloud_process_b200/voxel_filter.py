"""Host-side mirror of Pca_and_Voxel_filter/voxel_filter.py: same name, arguments and result.

    voxel_filter(point_cloud, leaf_size, type) -> (M, 3) float64        voxel_filter.py:10-66

The whole computation (bounding box, voxel hash, stable radix sort, per-voxel centroid in numpy's summation
order) runs on the GPU behind pcp_voxel_filter (include/pcp.h); there is no CPU path.  Reference behaviour that
is kept on purpose: D = (max - min) // leaf_size without +1 (colliding cells merge, :22-24,33), output in
ascending hash order, the last voxel is not emitted (:40-50).  The points are read as float32 rows and promoted
to float64 (what Open3D does with a float32 file); a float64 cloud is rounded to float32 first.
"type='random'" picks one point per voxel from a counter hash of `seed` (the reference draws from Python's
global `random`, :60, which nothing can reproduce); only the voxel structure is comparable.
"""
import numpy as np

from . import native
from .api import default_context


def voxel_filter(point_cloud, leaf_size, type="centroid", seed=0, device=0, ctx=None):
    pts = np.asarray(getattr(point_cloud, "points", point_cloud))
    if pts.ndim != 2 or pts.shape[1] < 3:
        raise ValueError(f"point_cloud: expected (N, 3) rows, got shape {pts.shape}")
    if pts.shape[0] == 0:
        raise ValueError("zero-size array to reduction operation maximum which has no identity")   # np.max, :20
    pts = np.ascontiguousarray(pts[:, :3], dtype=np.float32)
    return native.voxel_filter(ctx or default_context(device), pts, float(leaf_size), str(type), int(seed))
