"""Host-side mirror of Cluster_dbscan/dbscan.py: the same class, constructor arguments and fit / predict methods.

    DBSCAN(radius=0.5, Min_Pts=10).fit(data); labels = .predict()          dbscan.py:4-41

fit() runs entirely on the GPU behind pcp_dbscan (include/pcp.h): radius search of the cloud against itself
(the hot path's CSR), connected components of the expanding points and a rank-ordered fixed point that reproduces
the labels of the reference's SEQUENTIAL loop (cluster numbering, permanent noise, strict `> Min_Pts` expansion --
DESIGN.md section 8).  No CPU path.  data: (N, 3) rows (or (N, 2): a zero z column is appended).
"""
import numpy as np

from .api import SearchIndex


class DBSCAN(object):
    def __init__(self, radius=0.5, Min_Pts=10, device=0, ctx=None):
        self.radius = radius
        self.Min_Pts = Min_Pts
        self.labels_ = None
        self.n_clusters_ = None
        self._device, self._ctx = device, ctx

    def fit(self, data):
        data = np.asarray(data)
        if data.ndim != 2 or data.shape[1] not in (2, 3):
            raise ValueError(f"data: expected (N, 2) or (N, 3) rows, got shape {data.shape}; the GPU search structure is 3-D")
        if data.shape[1] == 2:
            data = np.concatenate([data, np.zeros((len(data), 1), data.dtype)], axis=1)
        ix = SearchIndex(data, device=self._device, ctx=self._ctx)
        if data.dtype == np.float64:
            # the reference searches the float64 rows with scipy's KDTree (dbscan.py:13): the float32 search structure
            # finds the candidates, borderline pairs are decided on these originals (pcp_index_attach_f64)
            ix._ix.attach_f64(np.ascontiguousarray(data))
        labels, self.n_clusters_ = ix._ix.dbscan(float(self.radius), int(self.Min_Pts))
        self.labels_ = np.asarray(labels, dtype=np.int32)

    def predict(self):
        return self.labels_
