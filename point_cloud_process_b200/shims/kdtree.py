"""Top-level module named like Kdtree_Octree/lesson2/kdtree.py so that the reference's unchanged drivers
import the GPU implementation:  cd Kdtree_Octree/lesson2 && PYTHONPATH=<repo>/point_cloud_process_b200/shims:<repo> python -P benchmark.py"""
from point_cloud_process_b200.kdtree import *          # noqa: F401,F403
from point_cloud_process_b200.kdtree import __doc__ as _doc   # noqa: F401
