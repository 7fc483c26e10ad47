"""Profiling target: one batched ICP call (mode A) on P synthetic pairs, device-resident."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.environ.get("PCP_ROOT") or os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import point_cloud_process_b200 as pcp
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
native = pcp.native
P = int(sys.argv[1]) if len(sys.argv) > 1 else 512
mode = sys.argv[2] if len(sys.argv) > 2 else "A"
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
h_src, h_tgt = bench.make_pairs(0, P, 20000, os.cpu_count() or 1)
ctx = native.Context(0); stream = torch.cuda.ExternalStream(ctx.stream())
off = np.arange(P + 1, dtype=np.int64) * 20000
d_src, d_tgt = torch.from_numpy(h_src).cuda(), torch.from_numpy(h_tgt).cuda()
prm = bench.ICP_MODES[mode]
a = (prm["max_iteration"], prm["R_diff_thres"], prm["t_diff_thres"], 5.0, True)
torch.cuda.synchronize()
ms = []
for i in range(reps + 1):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    out = native.icp_batch_device(ctx, d_src.data_ptr(), off, d_tgt.data_ptr(), off, 3, *a)
    e1.record(stream); ctx.synchronize()
    ms.append(e0.elapsed_time(e1))
print({"pairs": P, "ms": ms, "pairs_per_s": P / (min(ms[1:]) * 1e-3), "iterations_hist": np.bincount(out["iterations"]).tolist()})
its = np.asarray(out["iterations"]); print({"pair_iterations": int(its.sum()), "pair_iterations_per_s": float(its.sum() / (min(ms[1:]) * 1e-3)), "mean_iterations": float(its.mean())})
