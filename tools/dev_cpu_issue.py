"""Developer probe: host time to ISSUE one build + kNN step (no synchronisation inside the loop)."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import point_cloud_process_b200 as pcp
from tools.dev_timing import kitti
native = pcp.native
ctx = native.Context(0)
n, k = len(kitti), 8
d = torch.from_numpy(kitti).cuda()
ix = native.Index(ctx, d.data_ptr(), n, 3)
idx = torch.empty((n, k), dtype=torch.int32, device="cuda"); dist = torch.empty((n, k), dtype=torch.float32, device="cuda")
cnt = torch.empty(n, dtype=torch.int32, device="cuda")
for _ in range(10):
    ix.rebuild_device(d.data_ptr(), n, 3); ix.knn_device(d.data_ptr(), n, 3, k, idx.data_ptr(), dist.data_ptr(), cnt.data_ptr())
ctx.synchronize()
for name, fn in (("rebuild_device", lambda: ix.rebuild_device(d.data_ptr(), n, 3)),
                 ("knn_device", lambda: ix.knn_device(d.data_ptr(), n, 3, k, idx.data_ptr(), dist.data_ptr(), cnt.data_ptr())),
                 ("knn_self_device", lambda: ix.knn_self_device(k, idx.data_ptr(), dist.data_ptr(), cnt.data_ptr()))):
    ts = []
    for _ in range(30):
        ctx.synchronize()
        t0 = time.perf_counter(); fn(); ts.append(time.perf_counter() - t0)
    ctx.synchronize()
    l0 = ctx.launch_count(); fn(); l1 = ctx.launch_count()
    print(f"{name:16s} host issue time median {np.median(ts)*1e6:7.1f} us  min {min(ts)*1e6:7.1f} us  launches {l1-l0}")
