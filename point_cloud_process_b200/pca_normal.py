"""Host-side mirror of the normal estimation of Pca_and_Voxel_filter/pca_normal.py:85-90:

    for every point: search_knn_vector_3d(p, k) -> PCA(pc_array[idx, :]) -> normal = v[:, 2]

as one batched call: exact kNN of every point (the library's hot path) followed by a per-point kernel that
restates PCA() (pca_normal.py:10-36: mean, np.cov with 1/(k-1), symmetric eigen-decomposition, eigenvalues
descending) in float64.  No CPU path.  Signs: numpy leaves the eigenvector sign to LAPACK; here the component
of largest magnitude is positive.
"""
import numpy as np

from .api import SearchIndex


def estimate_normals(points, k=5, device=0, ctx=None, index=None, return_eigenvalues=False):
    """-> normals (N, 3) float64 [, eigenvalues (N, 3) float64 descending]."""
    ix = index if index is not None else SearchIndex(getattr(points, "points", points), device=device, ctx=ctx)
    nrm, ev = ix._ix.normals(int(k))
    return (nrm, ev) if return_eigenvalues else nrm
