"""Generate golden kNN / radius vectors with the REFERENCE's own Python code.

Run in the build container only (needs /root/reference):

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden_knn.py

Imports Kdtree_Octree/lesson2/{kdtree,octree,result_set}.py unmodified from /root/reference and
records, for seeded inputs, exactly what kdtree_knn_search / kdtree_radius_search /
octree_knn_search / octree_radius_search(_fast) leave in the result sets
(dist_index_list order, count, worst_dist, comparison_counter).  Outputs:

    tests/golden/kitti_000000_xyz.f32   the bundled KITTI scan as (124668, 3) float32 (x, y, z)
    tests/golden/knn_golden.npz         all cases

Nothing under /root/reference is copied: the .f32 file is the xyz columns of the public KITTI
scan the reference ships as data, the .npz holds computed outputs.
"""
import os
import sys

import numpy as np

REF = "/root/reference/Kdtree_Octree/lesson2"
sys.dont_write_bytecode = True
sys.path.insert(0, REF)
import kdtree as ref_kdtree            # noqa: E402
import octree as ref_octree            # noqa: E402
from result_set import KNNResultSet, RadiusNNResultSet   # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def run_knn(search, root, db, queries, k):
    nq = len(queries)
    dist = np.empty((nq, k), np.float32)
    idx = np.empty((nq, k), np.int64)
    cnt = np.empty(nq, np.int32)
    cmpc = np.empty(nq, np.int64)
    for i, q in enumerate(queries):
        rs = KNNResultSet(capacity=k)
        search(root, db, rs, q)
        dist[i] = [d.distance for d in rs.dist_index_list]
        idx[i] = [d.index for d in rs.dist_index_list]
        cnt[i] = rs.count
        cmpc[i] = rs.comparison_counter
    return dist, idx, cnt, cmpc


def run_radius(search, root, db, queries, r):
    offsets = [0]
    dist, idx, cmpc = [], [], []
    for q in queries:
        rs = RadiusNNResultSet(radius=r)
        search(root, db, rs, q)
        dist += [d.distance for d in rs.dist_index_list]      # visit order
        idx += [d.index for d in rs.dist_index_list]
        offsets.append(offsets[-1] + rs.count)
        cmpc.append(rs.comparison_counter)
    return (np.asarray(offsets, np.int64), np.asarray(dist, np.float32), np.asarray(idx, np.int64),
            np.asarray(cmpc, np.int64))


def case(out, name, db, queries, k, r, leaf_size=32, min_extent=0.0001):
    db = np.ascontiguousarray(db, np.float32)
    queries = np.ascontiguousarray(queries, np.float32)
    kd = ref_kdtree.kdtree_construction(db, leaf_size)
    oc = ref_octree.octree_construction(db, leaf_size, min_extent)
    out[name + "/db"] = db
    out[name + "/queries"] = queries
    out[name + "/k"] = np.int64(k)
    out[name + "/r"] = np.float64(r)
    out[name + "/leaf_size"] = np.int64(leaf_size)
    for tag, fn, root in (("kd", ref_kdtree.kdtree_knn_search, kd), ("oc", ref_octree.octree_knn_search, oc)):
        d, i, c, m = run_knn(fn, root, db, queries, k)
        out[f"{name}/{tag}_knn_dist"] = d
        out[f"{name}/{tag}_knn_idx"] = i
        out[f"{name}/{tag}_knn_cnt"] = c
        out[f"{name}/{tag}_knn_cmp"] = m
    for tag, fn, root in (("kd", ref_kdtree.kdtree_radius_search, kd),
                          ("oc", ref_octree.octree_radius_search, oc),
                          ("ocf", ref_octree.octree_radius_search_fast, oc)):
        o, d, i, m = run_radius(fn, root, db, queries, r)
        out[f"{name}/{tag}_rad_off"] = o
        out[f"{name}/{tag}_rad_dist"] = d
        out[f"{name}/{tag}_rad_idx"] = i
        out[f"{name}/{tag}_rad_cmp"] = m
    print(name, "db", db.shape, "queries", queries.shape, "k", k, "r", r)


def main():
    out = {}
    kitti = np.fromfile("/root/reference/Kdtree_Octree/000000.bin", np.float32).reshape(-1, 4)[:, :3].copy()
    kitti.tofile(os.path.join(HERE, "kitti_000000_xyz.f32"))
    rng = np.random.default_rng(20240001)

    # C1: bundled KITTI scan, benchmark.py:31-34 constants, seeded in-cloud queries (+ query 0,
    # the one benchmark.py:122 issues).  The db itself is NOT stored in the npz (it is the .f32).
    qi = np.concatenate([[0], rng.choice(len(kitti), 255, replace=False)])
    tmp = {}
    case(tmp, "kitti", kitti, kitti[qi], 8, 1.0)
    del tmp["kitti/db"]
    tmp["kitti/query_index"] = qi.astype(np.int64)
    out.update(tmp)
    tmp = {}
    case(tmp, "kitti_k16", kitti, kitti[qi[:64]], 16, 0.5)
    del tmp["kitti_k16/db"]
    tmp["kitti_k16/query_index"] = qi[:64].astype(np.int64)
    out.update(tmp)

    # uniform cloud, C3 density (61 pts/m^3), queries in and around the cloud
    n = 4000
    side = (n / 61.0) ** (1 / 3)
    uni = rng.random((n, 3), np.float32) * np.float32(side)
    q = np.concatenate([uni[:100], rng.random((60, 3), np.float32) * np.float32(side * 1.5) - np.float32(0.25 * side)])
    case(out, "uniform", uni, q, 16, 0.5)

    # clustered cloud (very uneven density)
    centres = rng.normal(0, 20, (12, 3))
    clu = np.concatenate([rng.normal(c, s, (300, 3)) for c, s in zip(centres, rng.uniform(0.05, 3, 12))]).astype(np.float32)
    case(out, "clustered", clu, np.concatenate([clu[::37], (rng.normal(0, 25, (40, 3))).astype(np.float32)]), 8, 1.0, leaf_size=8)

    # duplicates and exact distance ties: integer lattice with every point repeated twice
    g = np.stack(np.meshgrid(np.arange(6), np.arange(6), np.arange(4), indexing="ij"), -1).reshape(-1, 3)
    lat = np.concatenate([g, g]).astype(np.float32)
    lat = lat[rng.permutation(len(lat))]
    case(out, "lattice_ties", lat, np.concatenate([lat[:20], np.float32([[2.5, 2.5, 1.5], [0.5, 0.0, 0.0], [-3, -3, -3]])]), 8, 1.0, leaf_size=4)

    # fewer points than k; single point; far queries; r = 0 (self only); r on an exact boundary
    few = rng.random((5, 3), np.float32)
    case(out, "fewer_than_k", few, np.concatenate([few, np.float32([[10, 10, 10]])]), 8, 0.5, leaf_size=2)
    one = np.float32([[1.0, 2.0, 3.0]])
    case(out, "single_point", one, np.float32([[1, 2, 3], [0, 0, 0]]), 3, 0.0, leaf_size=32)
    line = np.zeros((9, 3), np.float32)
    line[:, 0] = np.arange(9) * 0.25          # distances 0.25*i are exact in float32
    case(out, "radius_boundary", line, np.float32([[0, 0, 0], [1.0, 0, 0]]), 4, 0.5, leaf_size=2)

    np.savez_compressed(os.path.join(HERE, "knn_golden.npz"), **out)
    print("wrote", os.path.join(HERE, "knn_golden.npz"))


if __name__ == "__main__":
    main()
