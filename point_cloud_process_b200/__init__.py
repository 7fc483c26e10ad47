"""point_cloud_process_b200 -- B200-native exact kNN / radius search and point-to-point ICP.

Drop-in for the one hot path of ruanyyyyyyy/point-cloud-process (Kdtree_Octree/lesson2 search +
Registration ICP loop).  All compute runs in hand-written CUDA (sm_100a) behind the C ABI of
include/pcp.h; this package is the host-side mirror of the reference's Python interface:

    point_cloud_process_b200.result_set   KNNResultSet / RadiusNNResultSet / DistIndex
    point_cloud_process_b200.kdtree       kdtree_construction / kdtree_knn_search / kdtree_radius_search
    point_cloud_process_b200.octree       octree_construction / octree_knn_search / octree_radius_search(_fast)
    point_cloud_process_b200.icp          icp_point2point / ICP / homo2tq / write_reg_result ...
    point_cloud_process_b200.api          batched extension: SearchIndex.knn / .radius, icp_batch

There is no CPU fallback: importing the native module fails loudly when it has not been built,
and creating a context fails loudly when no sm_100 GPU is present.
"""
import os as _os

_HERE = _os.path.dirname(_os.path.abspath(__file__))


def _load_native():
    try:
        from . import _pcp   # noqa: F401
    except ImportError as e:   # pragma: no cover
        raise ImportError(
            "point_cloud_process_b200: the native CUDA module _pcp is not built "
            "(run `python -m point_cloud_process_b200.build`); there is no CPU fallback. "
            f"Original error: {e}") from e
    return _pcp


native = _load_native()
__version__ = native.__version__

from .api import SearchIndex, default_context, icp_batch   # noqa: E402,F401
from . import result_set, kdtree, octree, icp   # noqa: E402,F401
