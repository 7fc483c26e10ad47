"""Profiling target: build + general-query kNN (the bench step) on the KITTI scan or the uniform 1M cloud."""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import point_cloud_process_b200 as pcp
native = pcp.native
which = sys.argv[1] if len(sys.argv) > 1 else "kitti"
k = int(sys.argv[2]) if len(sys.argv) > 2 else 8
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
if which == "kitti":
    pts = np.fromfile(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "kitti_000000_xyz.f32"), np.float32).reshape(-1, 3)
else:
    pts = (np.random.default_rng(1234).random((1_000_000, 3)) * 25.4).astype(np.float32)
ctx = native.Context(0)
d = torch.from_numpy(pts).cuda()
n = len(pts)
ix = native.Index(ctx, d.data_ptr(), n, 3)
idx = torch.empty((n, k), dtype=torch.int32, device="cuda")
dist = torch.empty((n, k), dtype=torch.float32, device="cuda")
cnt = torch.empty(n, dtype=torch.int32, device="cuda")
for _ in range(reps):
    ix.rebuild_device(d.data_ptr(), n, 3)
    ix.knn_device(d.data_ptr(), n, 3, k, idx.data_ptr(), dist.data_ptr(), cnt.data_ptr())
ctx.synchronize()
print("ok", idx[0].tolist())
