"""Developer probe: distribution of per-group work in the tiled kNN kernel."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["PCP_KNN_KERNEL"] = "tile"
os.environ["PCP_DUMP_GROUPS"] = "/tmp/groups.bin"
import point_cloud_process_b200 as pcp
kitti = np.fromfile("tests/golden/kitti_000000_xyz.f32", np.float32).reshape(-1, 3)
uni = (np.random.default_rng(1234).random((1000000, 3)) * 25.4).astype(np.float32)
for nm, pts, kk in (("kitti", kitti, 8), ("uni1M_k8", uni, 8), ("uni1M_k16", uni, 16)):
    ctx = pcp.native.Context(0)
    ctx.enable_counters(True)
    ix = pcp.native.Index(ctx, pts)
    ix.knn_self(kk)
    ctx.read_counters()
    g = np.fromfile("/tmp/groups.bin", np.uint32).reshape(-1, 8)
    for j, name in enumerate(("scans", "tests", "rounds", "opens")):
        c = g[:, j]
        print(nm, name, "mean %.1f" % c.mean(), "p50 %d p90 %d p99 %d p99.9 %d max %d" % tuple(np.percentile(c, [50, 90, 99, 99.9, 100])), "sum", c.sum())
    cost = g[:, 0] * 352.0 + g[:, 1] * 30 + g[:, 2] * (60 if kk == 8 else 110) + g[:, 3] * 80
    print(nm, "est instr/group mean %.0f p99 %.0f max %.0f; share of groups>4x mean: %.4f, their cost share %.3f" % (
        cost.mean(), np.percentile(cost, 99), cost.max(), (cost > 4 * cost.mean()).mean(), cost[cost > 4 * cost.mean()].sum() / cost.sum()))
    # timeline: 32 ns units
    t0 = g[:, 4].min()
    beg = (g[:, 4] - t0).astype(np.int64) * 0.032
    end = (g[:, 5] - t0).astype(np.int64) * 0.032
    dur = end - beg
    dry = beg.max()
    print(nm, "kernel span %.0f us; last group handed out at %.0f us; group duration mean %.0f p50 %.0f p99 %.0f max %.0f us" % (
        end.max(), dry, dur.mean(), np.percentile(dur, 50), np.percentile(dur, 99), dur.max()))
    late = end > dry
    print(nm, "groups running when the queue ran dry: %d; of them ending >50/100/200/400 us later: %s" % (
        late.sum(), [(end > dry + x).sum() for x in (50, 100, 200, 400)]))
    sm_end = np.zeros(g[:, 6].max() + 1)
    np.maximum.at(sm_end, g[:, 6], end)
    print(nm, "per-SM last end: min %.0f p50 %.0f p90 %.0f max %.0f us; SMs used %d" % (
        sm_end.min(), np.percentile(sm_end, 50), np.percentile(sm_end, 90), sm_end.max(), (sm_end > 0).sum()))
    worst = np.argsort(-end)[:8]
    for w in worst:
        print("   group %d sm %d beg %.0f end %.0f dur %.0f scans %d tests %d rounds %d opens %d aborted %d" % (
            w, g[w, 6], beg[w], end[w], dur[w], g[w, 0], g[w, 1], g[w, 2], g[w, 3], g[w, 7]))
    # concurrency over time
    edges = np.linspace(0, end.max(), 21)
    conc = [((beg <= t) & (end > t)).sum() for t in edges[:-1]]
    print(nm, "groups in flight at 5% steps:", conc)
    np.save("/root/repo/gpurun_out/groups_%s.npy" % nm, g)
