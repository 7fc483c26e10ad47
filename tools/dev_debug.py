import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import point_cloud_process_b200 as pcp
from oracle import knn_oracle
g = np.load("tests/golden/knn_golden.npz")
name = "clustered"
db, q, k = g[name + "/db"], g[name + "/queries"], int(g[name + "/k"])
idx, dist, cnt = pcp.SearchIndex(db).knn(q, k)
gd, gi = g[name + "/oc_knn_dist"], g[name + "/oc_knn_idx"]
bad = np.nonzero((dist != gd).any(1))[0]
print("bad rows", bad)
for r in bad[:5]:
    print(r, q[r]); print(" got ", idx[r], dist[r]); print(" want", gi[r], gd[r])
    print(" true dist of got", knn_oracle.pair_dist(db, q, np.full(k, r), idx[r]))
kitti = np.fromfile("tests/golden/kitti_000000_xyz.f32", np.float32).reshape(-1, 3)
for nm, pts, kk, r in (("kitti", kitti, 8, 1.0), ("uni1M", (np.random.default_rng(1234).random((1000000, 3)) * 25.4).astype(np.float32), 16, 0.5)):
    ctx = pcp.native.Context(0)
    ctx.enable_counters(True)
    ix = pcp.native.Index(ctx, pts)
    ix.knn_self(kk)
    c = ctx.read_counters(); ng = c[4]
    print(nm, "knn_self: evals/query", c[0] / len(pts), "tests/group", c[1] / ng, "scans/group", c[5] / ng, "opens/group", c[2] / ng, "drains/group", c[3] / ng)
    ix.knn(pts, kk)
    c = ctx.read_counters(); ng = c[4]
    print(nm, "knn     : evals/query", c[0] / len(pts), "tests/group", c[1] / ng, "scans/group", c[5] / ng, "opens/group", c[2] / ng, "drains/group", c[3] / ng)
    ix.radius(pts[::16], r)
    c = ctx.read_counters(); nq = c[4]
    print(nm, "radius  : evals/query", c[0] / nq, "boxtests/query", c[1] / nq, "opens/query", c[2] / nq, "(two passes)")
