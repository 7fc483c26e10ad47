"""Drop-in for Kdtree_Octree/lesson2/octree.py: same function names and argument meaning, served by
the same GPU bucket hierarchy as kdtree.py (one structure replaces both trees).

    octree_construction(db_np, leaf_size, min_extent) -> root handle        octree.py:310-328
    octree_knn_search(root, db, result_set, query) -> bool                  octree.py:262-306
    octree_radius_search(root, db, result_set, query) -> bool               octree.py:216-259
    octree_radius_search_fast(root, db, result_set, query) -> bool          octree.py:166-212

The reference's searches return "the query ball lies inside this octant, the caller may stop"; at the
root that is inside(query, worstDist, root) (octree.py:106-117), which the handle reproduces from the
root cube (centre = mean of the points, extent = half the largest bounding-box side, octree.py:318-322).
"""
import numpy as np

from .api import SearchIndex
from .kdtree import _check_db, knn_into, radius_into


class Octant:
    """Handle returned by octree_construction (stands in for the root Octant, octree.py:11-32)."""

    def __init__(self, index, db_np, leaf_size, min_extent):
        self.index = index
        self.leaf_size, self.min_extent = leaf_size, min_extent
        self.children = [None] * 8
        self.is_leaf = False
        self.point_indices = np.arange(len(index))
        if len(db_np):
            self.center = np.mean(db_np, axis=0)
            self.extent = np.max(np.amax(db_np, axis=0) - np.amin(db_np, axis=0)) * 0.5
        else:
            self.center, self.extent = np.zeros(3, np.float32), 0.0

    def knn(self, queries, k):
        return self.index.knn(queries, k)

    def radius(self, queries, radius, sorted=False):
        return self.index.radius(queries, radius, sorted=sorted)

    def __str__(self):
        nb, nl = self.index.stats()
        return "GPU bucket hierarchy: center %s, extent %.2f, %d points, %d buckets, %d levels" % (
            np.array2string(np.asarray(self.center), precision=2), self.extent, len(self.index), nb, nl)


def inside(query, radius, octant):
    """Ball (query, radius) strictly inside the octant's cube (octree.py:106-117)."""
    offset = np.fabs(np.asarray(query) - octant.center)
    return bool(np.all(offset + radius < octant.extent))


def octree_construction(db_np, leaf_size, min_extent):
    db_np = _check_db(db_np)
    return Octant(SearchIndex(db_np), db_np, leaf_size, min_extent)


def octree_knn_search(root, db, result_set, query):
    if root is None:
        return False
    knn_into(root, result_set, query)
    return inside(query, result_set.worstDist(), root)


def octree_radius_search(root, db, result_set, query):
    if root is None:
        return False
    radius_into(root, result_set, query)
    return inside(query, result_set.worstDist(), root)


def octree_radius_search_fast(root, db, result_set, query):
    """Same result set as octree_radius_search (the reference's variant only adds a contains()
    shortcut, octree.py:150-164,186-190)."""
    return octree_radius_search(root, db, result_set, query)
