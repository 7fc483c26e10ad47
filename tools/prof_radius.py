"""Profiling target: C3 radius count + fill (r = 0.5) on the uniform 1M cloud, device-resident."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import point_cloud_process_b200 as pcp
native = pcp.native
n = 1_000_000
pts = (np.random.default_rng(1234).random((n, 3)) * 25.4).astype(np.float32)
ctx = native.Context(0)
d = torch.from_numpy(pts).cuda()
ix = native.Index(ctx, d.data_ptr(), n, 3)
off = torch.empty(n + 1, dtype=torch.int64, device="cuda")
for _ in range(2):
    ix.radius_count_device(d.data_ptr(), n, 3, 0.5, off.data_ptr())
ctx.synchronize()
m = int(off[-1].item())
ridx = torch.empty(m, dtype=torch.int32, device="cuda"); rdist = torch.empty(m, dtype=torch.float32, device="cuda")
ix.radius_fill_device(d.data_ptr(), n, 3, 0.5, off.data_ptr(), ridx.data_ptr(), rdist.data_ptr(), False)
ctx.synchronize()
print("pairs", m)
