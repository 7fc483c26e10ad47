"""Workload for the ncu launch list / captures of the ICP front end: voxel filter on 16 M and 1 M points, normals k=5 on 1 M.
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/front_launches.csv python tools/prof_front.py"""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import point_cloud_process_b200 as pcp
native = pcp.native
ctx = native.Context(0)
n_big = int(os.environ.get("PCP_PROF_N", 16_000_000))
big = torch.from_numpy((np.random.default_rng(99).random((n_big, 3)) * 64.0).astype(np.float32)).cuda()
uni_h = (np.random.default_rng(1234).random((1_000_000, 3)) * 25.4).astype(np.float32)
uni = torch.from_numpy(uni_h).cuda()
for _ in range(2):
    native.voxel_filter_device(ctx, big.data_ptr(), n_big, 3, 0.4, 0, 0)
    native.voxel_filter_device(ctx, uni.data_ptr(), 1_000_000, 3, 0.4, 0, 0)
ix = native.Index(ctx, uni.data_ptr(), 1_000_000, 3)
nrm = torch.empty((1_000_000, 3), dtype=torch.float64, device="cuda"); ev = torch.empty_like(nrm)
for _ in range(2):
    ix.normals_device(5, nrm.data_ptr(), ev.data_ptr())
ctx.synchronize()
print("done")
