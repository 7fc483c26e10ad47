"""Developer probe: device-resident kNN timing for the current env settings (one line per workload)."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.environ.get("PCP_ROOT") or os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import point_cloud_process_b200 as pcp
native = pcp.native
which = sys.argv[1:] or ["kitti"]
kitti = np.fromfile(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "kitti_000000_xyz.f32"), np.float32).reshape(-1, 3)
def timeit(stream, fn, reps=20, warm=3):
    for _ in range(warm): fn()
    stream.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for a, b in evs:
        a.record(stream); fn(); b.record(stream)
    stream.synchronize()
    return sorted(a.elapsed_time(b) for a, b in evs)[reps // 2]
for w in which:
    pts, k = (kitti, 8) if w == "kitti" else ((np.random.default_rng(1234).random((1_000_000, 3)) * 25.4).astype(np.float32), 16)
    ctx = native.Context(0); stream = torch.cuda.ExternalStream(ctx.stream())
    d = torch.from_numpy(pts).cuda(); n = len(pts)
    ix = native.Index(ctx, d.data_ptr(), n, 3)
    idx = torch.empty((n, k), dtype=torch.int32, device="cuda"); dist = torch.empty((n, k), dtype=torch.float32, device="cuda"); cnt = torch.empty(n, dtype=torch.int32, device="cuda")
    ts = timeit(stream, lambda: ix.knn_self_device(k, idx.data_ptr(), dist.data_ptr(), cnt.data_ptr()))
    tg = timeit(stream, lambda: ix.knn_device(d.data_ptr(), n, 3, k, idx.data_ptr(), dist.data_ptr(), cnt.data_ptr()))
    ctx.enable_counters(True); ix.knn_self_device(k, idx.data_ptr(), dist.data_ptr(), cnt.data_ptr()); c = ctx.read_counters(); ctx.enable_counters(False)
    print(f"{w} k={k} env scans={os.environ.get('PCP_TILE_MAX_SCANS')} tests={os.environ.get('PCP_TILE_MAX_TESTS')} kern={os.environ.get('PCP_KNN_KERNEL')}: self {ts*1e3:.0f} us  general {tg*1e3:.0f} us  evals/q {c[0]/n:.0f} overflow_groups {c[6]}")
