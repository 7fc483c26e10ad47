"""Developer probe: distribution of per-group work in the tiled kNN kernel."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["PCP_KNN_KERNEL"] = "tile"
os.environ["PCP_DUMP_GROUPS"] = "/tmp/groups.bin"
import point_cloud_process_b200 as pcp
kitti = np.fromfile("tests/golden/kitti_000000_xyz.f32", np.float32).reshape(-1, 3)
for nm, pts, kk in (("kitti", kitti, 8), ("uni1M", (np.random.default_rng(1234).random((1000000, 3)) * 25.4).astype(np.float32), 16)):
    ctx = pcp.native.Context(0)
    ctx.enable_counters(True)
    ix = pcp.native.Index(ctx, pts)
    ix.knn_self(kk)
    ctx.read_counters()
    g = np.fromfile("/tmp/groups.bin", np.uint32).reshape(-1, 4)
    for j, name in enumerate(("scans", "tests", "rounds", "opens")):
        c = g[:, j]
        print(nm, name, "mean %.1f" % c.mean(), "p50 %d p90 %d p99 %d p99.9 %d max %d" % tuple(np.percentile(c, [50, 90, 99, 99.9, 100])), "sum", c.sum())
    cost = g[:, 0] * 352.0 + g[:, 1] * 30 + g[:, 2] * (60 if kk == 8 else 110) + g[:, 3] * 80
    print(nm, "est instr/group mean %.0f p99 %.0f max %.0f; share of groups>4x mean: %.4f, their cost share %.3f" % (
        cost.mean(), np.percentile(cost, 99), cost.max(), (cost > 4 * cost.mean()).mean(), cost[cost > 4 * cost.mean()].sum() / cost.sum()))
