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
if os.environ.get("PCP_PROF_CONSUMERS", "1") == "1":      # ISS keypoints + DBSCAN on the KITTI scan (r = 1.0: 80 M pairs)
    import bench
    kitti = torch.from_numpy(np.fromfile(bench.KITTI, np.float32).reshape(-1, 3)).cuda()
    kx = native.Index(ctx, kitti.data_ptr(), kitti.shape[0], 3)
    lab = torch.empty(kitti.shape[0], dtype=torch.int32, device="cuda")
    for _ in range(2):
        kx.iss_keypoints(1.0, 0.5, 0.5, 1.0, 20)
        kx.dbscan_device(1.0, 10, lab.data_ptr())
ctx.synchronize()
print("done")
