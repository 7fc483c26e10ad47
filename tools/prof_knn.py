"""Profiling target: a few launches of the kNN kernel on the KITTI scan (or the uniform 1M cloud)."""
import os, sys
import numpy as np
sys.path.insert(0, os.environ.get("PCP_ROOT") or os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import point_cloud_process_b200 as pcp
which = sys.argv[1] if len(sys.argv) > 1 else "kitti"
k = int(sys.argv[2]) if len(sys.argv) > 2 else 8
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
if which == "kitti":
    pts = np.fromfile(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "kitti_000000_xyz.f32"), np.float32).reshape(-1, 3)
else:
    pts = (np.random.default_rng(1234).random((1_000_000, 3)) * 25.4).astype(np.float32)
ix = pcp.SearchIndex(pts)
for _ in range(reps):
    idx, dist, cnt = ix.knn_self(k)
print("ok", idx[0])
