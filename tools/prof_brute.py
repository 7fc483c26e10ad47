"""Profiling target: the exhaustive kNN kernel (pcp_brute.cu) over the whole KITTI scan, device-resident.
Prints the time per launch and the fraction of the FP32 peak (8 unfused ops per evaluation, every pair evaluated)."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.environ.get("PCP_ROOT") or os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import point_cloud_process_b200 as pcp
native = pcp.native
pts = np.fromfile(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden", "kitti_000000_xyz.f32"), np.float32).reshape(-1, 3)
n = len(pts)
ctx = native.Context(0); stream = torch.cuda.ExternalStream(ctx.stream())
d = torch.from_numpy(pts).cuda()
ix = native.Index(ctx, d.data_ptr(), n, 3)
for k in (8, 16, 32):
    idx = torch.empty((n, k), dtype=torch.int32, device="cuda"); dist = torch.empty((n, k), dtype=torch.float32, device="cuda")
    cnt = torch.empty(n, dtype=torch.int32, device="cuda")
    ms = []
    for i in range(4):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        ix.knn_bruteforce_device(d.data_ptr(), n, 3, k, idx.data_ptr(), dist.data_ptr(), cnt.data_ptr())
        e1.record(stream); ctx.synchronize()
        ms.append(e0.elapsed_time(e1))
    t = min(ms[1:]) * 1e-3
    slots = (n + 31) // 32 * 32
    evals = n * slots
    peak = 148 * 128 * 1.965e9
    print({"k": k, "ms": round(t * 1e3, 3), "evaluations": evals, "T_op_per_s": round(8 * evals / t / 1e12, 2),
           "fp32_frac": round(8 * evals / t / peak, 4), "queries_per_s": round(n / t)})
