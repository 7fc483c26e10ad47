"""Developer probe: KITTI step (rebuild + kNN) as stream launches vs one CUDA-graph launch: device time and host issue time."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.environ.get("PCP_ROOT") or os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import point_cloud_process_b200 as pcp
native = pcp.native
kitti = np.fromfile(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "kitti_000000_xyz.f32"), np.float32).reshape(-1, 3)
ctx = native.Context(0); stream = torch.cuda.ExternalStream(ctx.stream())
n, k = len(kitti), 8
d = torch.from_numpy(kitti).cuda()
idx = torch.empty((n, k), dtype=torch.int32, device="cuda"); dst = torch.empty((n, k), dtype=torch.float32, device="cuda"); cnt = torch.empty(n, dtype=torch.int32, device="cuda")
torch.cuda.synchronize()
ix = native.Index(ctx, d.data_ptr(), n, 3)
def step():
    ix.rebuild_device(d.data_ptr(), n, 3)
    ix.knn_device(d.data_ptr(), n, 3, k, idx.data_ptr(), dst.data_ptr(), cnt.data_ptr())
for _ in range(5): step()
ctx.synchronize()
ctx.graph_begin(); step(); g = ctx.graph_end()
def bench(fn, reps=50):
    for _ in range(5): fn()
    ctx.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record(stream)
    for _ in range(reps): fn()
    t_issue = time.perf_counter() - t0
    e1.record(stream); ctx.synchronize()
    shots = []
    for _ in range(20):
        ctx.synchronize(); t0 = time.perf_counter(); fn(); ctx.synchronize(); shots.append(time.perf_counter() - t0)
    return e0.elapsed_time(e1) / reps * 1e3, t_issue / reps * 1e6, sorted(shots)[10] * 1e6
print("stream launches: device %.1f us/step, host issue %.1f us/step, single shot %.1f us" % bench(step))
print("graph (%d nodes): device %.1f us/step, host issue %.1f us/step, single shot %.1f us" % ((g.nodes(),) + bench(g.launch)))
