"""Host-side mirror of Keypoint_detection_ISS/ISS.py: the script's detection loop (ISS.py:36-73, which the
reference keeps under `if __name__ == "__main__"`) as one function with the script's hyper-parameter names and
defaults (ISS.py:19-27).  Everything -- neighbour counts, weighted scatter matrices, eigenvalues, the ratio tests
and the non-maximum suppression -- runs on the GPU behind pcp_iss_keypoints (include/pcp.h); no CPU path.
"""
import numpy as np

from .api import SearchIndex


def readModelNet40(path):
    """ISS.py:7-13: comma-separated rows -> (N, C) float64."""
    return np.loadtxt(path, delimiter=",", dtype=np.float64, ndmin=2)


def iss_keypoints(points, radius=0.5, lambda21=0.5, lambda32=0.5, non_max_radius=0.5, iss_count=20, device=0, ctx=None,
                  index=None, return_lambda3=False):
    """-> iss_idx (list of point indices in selection order; at most iss_count + 1 of them, ISS.py:72)
    [, lambda3 (N,) float64: the smallest eigenvalue of every candidate, NaN elsewhere]."""
    ix = index
    if ix is None:
        pts = np.asarray(points)[:, :3]
        ix = SearchIndex(pts, device=device, ctx=ctx)
        if pts.dtype == np.float64:
            # the reference searches the float64 rows with scipy's cKDTree (ISS.py:38-40): borderline neighbour pairs
            # of the float32 search are decided on these originals (pcp_index_attach_f64)
            ix._ix.attach_f64(np.ascontiguousarray(pts))
    kp, l3 = ix._ix.iss_keypoints(float(radius), float(lambda21), float(lambda32), float(non_max_radius), int(iss_count))
    kp = [int(i) for i in kp]
    return (kp, l3) if return_lambda3 else kp
