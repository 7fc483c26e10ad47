"""Host-side mirror of the normal estimation of Pca_and_Voxel_filter/pca_normal.py:85-90:

    for every point: search_knn_vector_3d(p, k) -> PCA(pc_array[idx, :]) -> normal = v[:, 2]

as one batched call: exact kNN of every point (the library's hot path) followed by a per-point kernel that
restates PCA() (pca_normal.py:10-36: mean, np.cov with 1/(k-1), symmetric eigen-decomposition, eigenvalues
descending) in float64.  No CPU path.  Signs: numpy leaves the eigenvector sign to LAPACK; here the component
of largest magnitude is positive.
"""
import numpy as np

from . import native
from .api import SearchIndex, default_context


def PCA(data, correlation=False, sort=True, device=0, ctx=None):
    """PCA(data, correlation=False, sort=True) of pca_normal.py:10-36 on the GPU (pcp_pca): data (N, 3) ->
    (eigenvalues (3,), eigenvectors (3, 3) as columns).  `correlation` is accepted and unused, as in the
    reference (:10,26); sort=False gives np.linalg.eigh's ascending order (:28)."""
    pts = np.asarray(getattr(data, "points", data))
    if pts.ndim != 2 or pts.shape[1] < 3:
        raise ValueError(f"data: expected (N, 3) rows, got shape {pts.shape}")
    w, v = native.pca(ctx or default_context(device), np.ascontiguousarray(pts[:, :3], dtype=np.float32))
    return (w, v) if sort else (w[::-1].copy(), v[:, ::-1].copy())


def estimate_normals(points, k=5, device=0, ctx=None, index=None, return_eigenvalues=False):
    """-> normals (N, 3) float64 [, eigenvalues (N, 3) float64 descending]."""
    ix = index if index is not None else SearchIndex(getattr(points, "points", points), device=device, ctx=ctx)
    nrm, ev = ix._ix.normals(int(k))
    return (nrm, ev) if return_eigenvalues else nrm
