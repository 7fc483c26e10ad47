"""Drop-in for Kdtree_Octree/lesson2/kdtree.py: same function names and argument meaning, the
recursive median-split tree replaced by the GPU bucket hierarchy (include/pcp.h).

    kdtree_construction(db_np, leaf_size) -> root handle          kdtree.py:119-137
    kdtree_knn_search(root, db, result_set, query) -> False       kdtree.py:141-172
    kdtree_radius_search(root, db, result_set, query) -> False    kdtree.py:176-208

`leaf_size` is accepted and ignored (the GPU structure always uses 32-point buckets).  db_np must be
(N, 3): the GPU structure is 3-D only, and a (3, N) array -- what the shipped benchmark.py:27 passes
because of its `.T` -- is rejected with a ValueError instead of being silently mis-read as three
points of dimension N.
Batched extension (one call for many queries): root.knn(queries, k), root.radius(queries, r).
"""
import numpy as np

from .api import SearchIndex


class Node:
    """Handle returned by kdtree_construction (stands in for the reference's root Node, kdtree.py:10-31).
    The tree's nodes live on the GPU; the attributes below describe the whole cloud."""

    def __init__(self, index, leaf_size):
        self.index = index
        self.leaf_size = leaf_size
        self.axis = 0
        self.value = None
        self.left = None
        self.right = None
        self.point_indices = np.arange(len(index))

    def is_leaf(self):
        return self.value is None

    # batched extension
    def knn(self, queries, k):
        return self.index.knn(queries, k)

    def radius(self, queries, radius, sorted=False):
        return self.index.radius(queries, radius, sorted=sorted)

    def __str__(self):
        nb, nl = self.index.stats()
        return "GPU bucket hierarchy: %d points, %d buckets, %d levels" % (len(self.index), nb, nl)


def _check_db(db_np):
    db_np = np.asarray(db_np)
    if db_np.ndim != 2 or db_np.shape[1] != 3:
        raise ValueError(
            "expected db_np with shape (N, 3), got %r. (benchmark.py:27 returns the scan transposed to (3, N); "
            "pass read_velodyne_bin(path).T or load the file as N x 3.)" % (db_np.shape,))
    return db_np


def kdtree_construction(db_np, leaf_size):
    """Build the search structure over db_np (N, 3) on the GPU and return its handle."""
    return Node(SearchIndex(_check_db(db_np)), leaf_size)


def traverse_kdtree(root, depth, max_depth):
    """kdtree.py:100-116 prints the tree; here: the structure's summary.  depth/max_depth are lists."""
    nb, nl = root.index.stats()
    depth[0] = max_depth[0] = max(max_depth[0], nl)
    print(root)


def _query_row(query):
    q = np.asarray(query, dtype=np.float32).reshape(-1)
    if q.shape[0] != 3:
        raise ValueError("query must have 3 components, got shape %r" % (np.asarray(query).shape,))
    return q[None, :]


def knn_into(root, result_set, query):
    idx, dist, cnt = root.index.knn(_query_row(query), result_set.capacity)
    result_set._load(dist[0], idx[0], cnt[0])


def radius_into(root, result_set, query):
    off, idx, dist = root.index.radius(_query_row(query), result_set.radius, sorted=True)
    result_set._load(dist, idx)


def kdtree_knn_search(root, db, result_set, query):
    """Fill result_set (KNNResultSet) with the capacity nearest neighbours of query; returns False
    like the reference (kdtree.py:172).  `db` is unused: the handle owns a copy of the cloud."""
    if root is None:
        return False
    knn_into(root, result_set, query)
    return False


def kdtree_radius_search(root, db, result_set, query):
    """Fill result_set (RadiusNNResultSet) with every point at float32 distance <= radius; returns False."""
    if root is None:
        return False
    radius_into(root, result_set, query)
    return False
