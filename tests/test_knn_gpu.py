"""GPU parity: CUDA kNN / radius (through pybind11 -> C ABI) vs the CPU oracle and the golden vectors."""
import numpy as np
import pytest

from conftest import csr_sets

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pcp():
    import point_cloud_process_b200 as m
    return m


@pytest.fixture(scope="module")
def kitti_index(pcp, kitti):
    return pcp.SearchIndex(kitti)


def _check_knn_against_oracle(idx, dist, cnt, o_dist, o_idx, o_cnt, db, queries, k):
    """SURVEY 8(c) criteria: bit-equal distances everywhere; identical index sequence wherever the
    oracle's distances around the row are free of ties; otherwise every returned index must carry
    exactly the distance reported for it."""
    from oracle import knn_oracle
    assert np.array_equal(cnt, o_cnt)
    assert np.array_equal(dist, o_dist), "float32 distances must be bit-equal"
    same = (idx == o_idx).all(axis=1)
    if not same.all():
        bad = np.nonzero(~same)[0]
        for r in bad:
            c = int(cnt[r])
            d = o_dist[r, :c]
            assert (np.diff(d) == 0).any() or c == k, f"row {r} differs without a tie"
            got = knn_oracle.pair_dist(db, queries, np.full(c, r), idx[r, :c])
            assert np.array_equal(got, dist[r, :c])
            assert len(set(idx[r, :c].tolist())) == c
    return int((~same).sum())


@pytest.mark.parametrize("name", ["kitti", "kitti_k16", "uniform", "clustered", "lattice_ties", "fewer_than_k",
                                  "single_point", "radius_boundary"])
def test_knn_golden(pcp, knn_golden, kitti, name):  # kd-tree golden: the reference octree kNN is inexact on "clustered"
    """Against vectors produced by the reference's own kdtree.py / octree.py."""
    g = knn_golden
    db = kitti if name.startswith("kitti") else g[name + "/db"]
    q, k = g[name + "/queries"], int(g[name + "/k"])
    idx, dist, cnt = pcp.SearchIndex(db).knn(q, k)
    assert np.array_equal(cnt, g[name + "/kd_knn_cnt"])
    assert np.array_equal(dist, g[name + "/kd_knn_dist"])          # bit-exact float32 Euclidean distances
    gi = g[name + "/kd_knn_idx"]
    d = g[name + "/kd_knn_dist"]
    for r in range(len(q)):
        c = int(cnt[r])
        if c == 0:
            continue
        # positions not involved in an exact distance tie must agree index for index
        tie = np.zeros(c, bool)
        eq = d[r, 1:c] == d[r, :c - 1]
        tie[1:] |= eq
        tie[:-1] |= eq
        if c == k:
            tie[-1] = True     # the k-th slot may tie with the unseen (k+1)-th
        assert np.array_equal(idx[r, :c][~tie], gi[r, :c][~tie])
        assert (idx[r, c:] == 0).all() and (dist[r, c:] == np.float32(1e10)).all()


@pytest.mark.parametrize("name", ["kitti", "kitti_k16", "uniform", "clustered", "lattice_ties", "fewer_than_k",
                                  "single_point", "radius_boundary"])
def test_radius_golden(pcp, knn_golden, kitti, name):
    """Both reference trees must agree with the product wherever they are exact.  They are NOT exact
    everywhere: the octree's mean-centred root cube (octree.py:322) does not contain every point of
    "clustered" and prunes real neighbours, and both trees prune with strict '<'
    (kdtree.py:201,205; octree.py:147) and so drop dist == radius points that sit exactly on a
    split plane / octant face ("lattice_ties", "radius_boundary").  The product implements the
    result-set rule itself -- keep iff NOT dist > radius (result_set.py:80) -- so: golden results
    are always a subset, equal on the tie-free cases, and the definition is checked directly."""
    g = knn_golden
    db = kitti if name.startswith("kitti") else g[name + "/db"]
    q, r = g[name + "/queries"], float(g[name + "/r"])
    off, idx, dist = pcp.SearchIndex(db).radius(q, r)
    mine = csr_sets(off, idx)
    exact_tags = {"clustered": ["kd"], "lattice_ties": [], "radius_boundary": []}.get(name, ["kd", "oc", "ocf"])
    for tag in ("kd", "oc", "ocf"):
        go, gi, gd = g[f"{name}/{tag}_rad_off"], g[f"{name}/{tag}_rad_idx"], g[f"{name}/{tag}_rad_dist"]
        ref = csr_sets(go, gi)
        assert all(set(a) <= set(b) for a, b in zip(ref, mine))
        if tag in exact_tags:
            assert np.array_equal(off, go) and ref == mine
        for i in range(len(q)):      # distances bit-equal for every pair the reference reports
            a = dict(zip(idx[off[i]:off[i + 1]].tolist(), dist[off[i]:off[i + 1]].tolist()))
            b = dict(zip(gi[go[i]:go[i + 1]].tolist(), gd[go[i]:go[i + 1]].tolist()))
            assert all(a[j] == b[j] for j in b)
    if len(db) < 5000:   # the definition itself (result_set.py:80 on kdtree.py:158 distances)
        d = np.linalg.norm(q[:, None, :] - db[None, :, :], axis=2)
        truth = [tuple(np.nonzero(~(d[i] > np.float32(r)))[0].tolist()) for i in range(len(q))]
        assert mine == truth
    # sorted variant: ascending (distance, index)
    off2, idx2, dist2 = pcp.SearchIndex(db).radius(q, r, sorted=True)
    assert np.array_equal(off2, off)
    for i in range(len(q)):
        seg = list(zip(dist2[off[i]:off[i + 1]].tolist(), idx2[off[i]:off[i + 1]].tolist()))
        assert seg == sorted(seg)
        assert sorted(idx2[off[i]:off[i + 1]].tolist()) == sorted(idx[off[i]:off[i + 1]].tolist())


@pytest.mark.parametrize("k", [1, 8, 16])
def test_knn_full_kitti_vs_oracle(pcp, kitti, kitti_index, k):
    """C1 at full size: all 124 668 points as queries (self included), canonical brute-force-equal
    oracle = reference kd-tree restatement; ties checked by distance."""
    from oracle import knn_oracle
    idx, dist, cnt = kitti_index.knn(kitti, k)
    o_dist, o_idx, o_cnt = knn_oracle.KdTree(kitti, 32).knn(kitti, k, threads=8)
    ndiff = _check_knn_against_oracle(idx, dist, cnt, o_dist, o_idx.astype(np.int32), o_cnt, kitti, kitti, k)
    assert ndiff < 100          # ~22 tie rows at k=8 on this scan
    # the self-query entry point must give the same answer
    idx2, dist2, cnt2 = kitti_index.knn_self(k)
    assert np.array_equal(idx2, idx) and np.array_equal(dist2, dist) and np.array_equal(cnt2, cnt)


def test_knn_canonical_order_vs_bruteforce(pcp, kitti, kitti_index):
    """Index-exact including ties: the product's tie rule is ascending (distance, index)."""
    from oracle import knn_oracle
    rng = np.random.default_rng(7)
    qi = rng.choice(len(kitti), 4096, replace=False)
    idx, dist, cnt = kitti_index.knn(kitti[qi], 8)
    b_dist, b_idx, b_cnt = knn_oracle.bruteforce_knn(kitti, kitti[qi], 8, threads=8)
    assert np.array_equal(dist, b_dist) and np.array_equal(idx, b_idx.astype(np.int32)) and np.array_equal(cnt, b_cnt)


def test_radius_full_kitti_vs_oracle(pcp, kitti, kitti_index):
    """C1 radius r=1.0 on every 8th point (15 584 queries, ~10 M pairs) vs the kd-tree restatement."""
    from oracle import knn_oracle
    q = kitti[::8]
    off, idx, dist = kitti_index.radius(q, 1.0, sorted=True)
    o_off, o_dist, o_idx = knn_oracle.KdTree(kitti, 32).radius(q, 1.0, threads=8)
    assert np.array_equal(off, o_off)
    # sort the oracle's visit-order segments the same way and compare everything bit for bit
    seg = np.repeat(np.arange(len(q)), np.diff(o_off))
    order = np.lexsort((o_idx, o_dist, seg))
    assert np.array_equal(idx, o_idx[order].astype(np.int32))
    assert np.array_equal(dist, o_dist[order])


def test_queries_outside_and_degenerate(pcp):
    from oracle import knn_oracle
    rng = np.random.default_rng(3)
    db = rng.normal(size=(1000, 3)).astype(np.float32)
    q = np.concatenate([rng.normal(size=(50, 3)) * 30, [[1e6, -1e6, 0]], db[:5]]).astype(np.float32)
    idx, dist, cnt = pcp.SearchIndex(db).knn(q, 5)
    b_dist, b_idx, b_cnt = knn_oracle.bruteforce_knn(db, q, 5)
    assert np.array_equal(idx, b_idx.astype(np.int32)) and np.array_equal(dist, b_dist)
    # all points identical (zero-extent bounding box)
    same = np.ones((100, 3), np.float32)
    idx, dist, cnt = pcp.SearchIndex(same).knn(same[:3], 4)
    assert (dist == 0).all() and idx.tolist() == [[0, 1, 2, 3]] * 3
    # empty cloud / empty query set
    idx, dist, cnt = pcp.SearchIndex(np.zeros((0, 3), np.float32)).knn(q[:4], 3)
    assert (cnt == 0).all() and (dist == np.float32(1e10)).all() and (idx == 0).all()
    off, i2, d2 = pcp.SearchIndex(np.zeros((0, 3), np.float32)).radius(q[:4], 1.0)
    assert off.tolist() == [0] * 5 and len(i2) == 0
    idx, dist, cnt = pcp.SearchIndex(db).knn(np.zeros((0, 3), np.float32), 3)
    assert idx.shape == (0, 3)


def test_shape_errors(pcp, kitti):
    with pytest.raises(ValueError):
        pcp.SearchIndex(kitti[:100].T.copy()[:, :2])       # dim != 3
    with pytest.raises(ValueError):
        pcp.SearchIndex(kitti[:100]).knn(kitti[:4], 0)
    with pytest.raises(RuntimeError):
        pcp.native.Index(pcp.default_context(), kitti[:100]).knn(kitti[:4], 64)   # C ABI: k > PCP_MAX_K is unsupported, loudly
    idx, dist, cnt = pcp.SearchIndex(kitti[:100]).knn(kitti[:4], 64)   # the shim assembles wide rows (tests/test_shims_gpu.py)
    assert idx.shape == (4, 64) and (cnt == 64).all()


def test_knn_uniform_1m_properties(pcp):
    """C3 size: properties that need no oracle -- sorted rows, self at distance 0, symmetric
    sanity vs an independent radius search."""
    rng = np.random.default_rng(1234)
    db = (rng.random((1_000_000, 3)) * 25.4).astype(np.float32)
    ix = pcp.SearchIndex(db)
    idx, dist, cnt = ix.knn_self(16)
    assert (cnt == 16).all()
    assert (np.diff(dist, axis=1) >= 0).all()
    assert (dist[:, 0] == 0).all()
    # k-th distance consistency with radius search on a sample: count(dist <= d_k) >= k and count(dist < d_k) < k
    qi = rng.choice(len(db), 2000, replace=False)
    off, ridx, rdist = ix.radius(db[qi], 0.5)
    for j, i in enumerate(qi[:200]):
        seg = rdist[off[j]:off[j + 1]]
        dk = dist[i, 15]
        if dk <= 0.5:
            assert (seg <= dk).sum() >= 16 and (seg < dk).sum() < 16
    from oracle import knn_oracle
    b_dist, b_idx, _ = knn_oracle.bruteforce_knn(db, db[qi[:256]], 16, threads=8)
    assert np.array_equal(dist[qi[:256]], b_dist) and np.array_equal(idx[qi[:256]], b_idx.astype(np.int32))


def test_knn_mass_ties_tile_kernel(pcp):
    """Tiled kernel with masses of exactly tied distances (every point repeated 40 times): the candidate
    columns stay full after a compaction and the group must take the overflow path; canonical
    (distance, index) order vs brute force."""
    from oracle import knn_oracle
    rng = np.random.default_rng(11)
    base = (rng.random((400, 3)) * 10).astype(np.float32)
    owner = np.repeat(np.arange(400), 40)
    order = rng.permutation(len(owner))
    owner = owner[order]
    db = base[owner]
    assert len(db) >= 8192                                   # large enough for the tiled path
    idx, dist, cnt = pcp.SearchIndex(db).knn(db, 8)
    sel = rng.choice(len(db), 1024, replace=False)
    b_dist, b_idx, b_cnt = knn_oracle.bruteforce_knn(db, db[sel], 8, threads=8)
    assert np.array_equal(dist[sel], b_dist) and np.array_equal(idx[sel], b_idx.astype(np.int32))
    assert (cnt == 8).all() and (dist == 0).all()
    # every query: the 8 lowest row numbers among its 40 copies (a group that gives up must not disturb the
    # candidate columns of the groups next to it)
    rows = np.argsort(owner, kind="stable").reshape(400, 40)[:, :8]      # ascending rows per base point
    assert np.array_equal(idx, rows[owner].astype(np.int32))


def test_result_arrays_outlive_index_and_pool_reuse(pcp, kitti):
    """Results come back in page-locked pool buffers owned by the numpy arrays: they stay valid after the
    index is gone, and a collected result hands its buffer back (a second call must not corrupt a live one)."""
    ix = pcp.SearchIndex(kitti)
    idx1, dist1, cnt1 = ix.knn(kitti, 8)
    keep = idx1.copy(), dist1.copy()
    idx2, dist2, _ = ix.knn(kitti[::-1].copy(), 8)           # new buffers while the first results are alive
    del ix
    assert np.array_equal(idx1, keep[0]) and np.array_equal(dist1, keep[1])
    assert np.array_equal(idx2[::-1], idx1)
    del idx2, dist2
    idx3, dist3, _ = pcp.SearchIndex(kitti).knn(kitti, 8)    # reuses the freed pool buffers
    assert np.array_equal(idx3, keep[0]) and np.array_equal(idx1, keep[0])


def test_kernel_timing_api(pcp, kitti):
    ctx = pcp.native.Context(0)
    ix = pcp.native.Index(ctx, kitti)
    ctx.enable_kernel_timing(True)
    ix.knn(kitti, 8)
    ix.knn(kitti, 8)
    ms, n = ctx.read_kernel_timing()
    assert n == 2 and 0.0 < ms < 50.0
    ms, n = ctx.read_kernel_timing()
    assert n == 0 and ms == 0.0
    ctx.enable_kernel_timing(False)


def test_radius_tiled_fill_matches(kitti):
    """PCP_RADIUS_FILL=tile (the tiled FILL pass, off by default) gives the same CSR as the default path."""
    import os, subprocess, sys, tempfile
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = ("import sys, numpy as np; sys.path.insert(0, %r); import point_cloud_process_b200 as pcp;"
            "k = np.fromfile(%r, np.float32).reshape(-1, 3); q = k[::8];"
            "off, idx, dist = pcp.SearchIndex(k).radius(q, 1.0, sorted=True);"
            "np.savez(sys.argv[1], off=off, idx=idx, dist=dist)") % (root, os.path.join(root, "tests", "golden", "kitti_000000_xyz.f32"))
    outs = []
    with tempfile.TemporaryDirectory() as tmp:
        for mode in ("warp", "tile"):
            path = os.path.join(tmp, mode + ".npz")
            env = dict(os.environ, PCP_RADIUS_FILL=mode, PCP_RADIUS_KERNEL="tile")
            r = subprocess.run([sys.executable, "-c", code, path], env=env, capture_output=True, text=True, timeout=300)
            assert r.returncode == 0, r.stderr[-2000:]
            outs.append(dict(np.load(path)))
    for key in ("off", "idx", "dist"):
        assert np.array_equal(outs[0][key], outs[1][key]), key


@pytest.mark.timeout(120)
def test_work_counters_all_search_kernels(pcp, kitti):
    """pcp_ctx_enable_counters on every search kernel (tiled and warp-per-query, kNN and radius): the counted
    runs must finish (a divergent full-mask vote in a counter branch once hung the tiled radius kernel) and
    give the same results as the uncounted ones."""
    ctx = pcp.native.Context(0)
    ix = pcp.native.Index(ctx, kitti)
    q = kitti[::4].copy()                                   # 31 167 queries: tiled kernels
    ref_knn = ix.knn(q, 8)
    ref_rad = ix.radius(q, 0.5, True)
    small = kitti[:100].copy()                              # warp-per-query kernels
    ref_small = ix.knn(small, 8), ix.radius(small, 0.5, True)
    ctx.enable_counters(True)
    got_knn, got_rad = ix.knn(q, 8), ix.radius(q, 0.5, True)
    c_big = ctx.read_counters()
    got_small = ix.knn(small, 8), ix.radius(small, 0.5, True)
    c_small = ctx.read_counters()
    ix.knn_self(8)
    c_self = ctx.read_counters()
    ctx.enable_counters(False)
    for a, b in zip(ref_knn + ref_rad, got_knn + got_rad):
        assert np.array_equal(a, b)
    for a, b in zip(ref_small[0] + ref_small[1], got_small[0] + got_small[1]):
        assert np.array_equal(a, b)
    assert c_big[0] > 0 and c_small[0] > 0 and c_self[0] > 0


def _run_variants(code, variants, timeout=600):
    """Run `code` (writes an .npz to sys.argv[1]) once per environment variant; returns the loaded arrays."""
    import os, subprocess, sys, tempfile
    outs = []
    with tempfile.TemporaryDirectory() as tmp:
        for i, extra in enumerate(variants):
            path = os.path.join(tmp, "v%d.npz" % i)
            env = dict(os.environ, **extra)
            r = subprocess.run([sys.executable, "-c", code, path], env=env, capture_output=True, text=True, timeout=timeout)
            assert r.returncode == 0, r.stderr[-2000:]
            outs.append(dict(np.load(path)))
    return outs


ROOT = __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__)))


@pytest.mark.timeout(900)
def test_knn_tail_handover_same_results():
    """400 000 uniform points (12 500 groups: the unbudgeted large-grid path): the tiled kernel with the tail
    hand-over (default), without it (PCP_TILE_TAIL=0) and the warp-per-query kernel give identical arrays for
    EVERY query -- the stragglers that move to the overflow kernel, seeded with their gates, included."""
    code = ("import sys, numpy as np; sys.path.insert(0, %r); import point_cloud_process_b200 as pcp;"
            "db = (np.random.default_rng(77).random((400000, 3)) * 20).astype(np.float32);"
            "ix = pcp.SearchIndex(db); i1, d1, c1 = ix.knn_self(8); i2, d2, c2 = ix.knn(db[::3].copy(), 16);"
            "np.savez(sys.argv[1], i1=i1, d1=d1, c1=c1, i2=i2, d2=d2, c2=c2)") % ROOT
    outs = _run_variants(code, [{}, {"PCP_TILE_TAIL": "0"}, {"PCP_TILE_TAIL": "0.5"}, {"PCP_KNN_KERNEL": "warp"}])
    for o in outs[1:]:
        for key in outs[0]:
            assert np.array_equal(outs[0][key], o[key]), key


@pytest.mark.timeout(900)
def test_radius_large_query_set_paths_agree():
    """300 000 queries (>= 262 144: tiled COUNT with hand-over + tiled staged FILL by default) against the
    warp-per-query FILL, the tiled kernels without hand-over, and the all-warp path: same sorted CSR."""
    code = ("import sys, numpy as np; sys.path.insert(0, %r); import point_cloud_process_b200 as pcp;"
            "rng = np.random.default_rng(78); db = (rng.random((300000, 3)) * 15).astype(np.float32);"
            "db[:40000] = db[:40000] * 0.1 + 3;"                      # a dense clump: long rows, many flushes
            "off, idx, dist = pcp.SearchIndex(db).radius(db, 0.4, sorted=True);"
            "np.savez(sys.argv[1], off=off, idx=idx, dist=dist)") % ROOT
    outs = _run_variants(code, [{}, {"PCP_RADIUS_FILL": "warp"}, {"PCP_TILE_TAIL": "0"}, {"PCP_RADIUS_KERNEL": "warp"}])
    assert outs[0]["off"][-1] > 300000
    for o in outs[1:]:
        for key in outs[0]:
            assert np.array_equal(outs[0][key], o[key]), key


def test_bruteforce_kernel_vs_cpu_oracle(pcp):
    """The exhaustive device kernel (second oracle, kdtree.py:234-238) against the CPU brute force: every list
    width, k below the width, fewer points than k, duplicates, queries outside the cloud, tile boundaries."""
    from oracle import knn_oracle
    rng = np.random.default_rng(31)
    for n in (1, 31, 33, 1024, 1025, 5000):
        db = (rng.random((n, 3)) * np.float32([30, 30, 4])).astype(np.float32)
        if n > 100:
            db[7:47] = db[7]                                    # 40 exact duplicates
        q = np.concatenate([db[rng.choice(n, min(n, 150))], (rng.random((50, 3)) * 60 - 15).astype(np.float32)])
        ix = pcp.SearchIndex(db)
        for k in (1, 3, 8, 11, 16, 32):
            idx, dist, cnt = ix.knn_bruteforce(q, k)
            b_dist, b_idx, b_cnt = knn_oracle.bruteforce_knn(db, q, k, threads=8)
            assert np.array_equal(dist, b_dist) and np.array_equal(idx, b_idx.astype(np.int32)) and np.array_equal(cnt, b_cnt), (n, k)
        for r in (0.0, 0.7, 3.0):
            c = ix.radius_count_bruteforce(q, r)
            off, _, _ = ix.radius(q, r)
            assert np.array_equal(c, np.diff(off)), (n, r)


@pytest.mark.parametrize("name", ["kitti", "uniform1m"])
def test_full_set_parity_vs_bruteforce_kernel(pcp, kitti, name):
    """FULL-set parity at config size (C1: every KITTI point, k = 8 and r = 1.0; C3: 1 M uniform points, k = 16 and
    r = 0.5): the tiled kernels against the exhaustive device kernel, every query, bit-equal indices and float32
    distances -- where round 1 could only sample 256 queries against the CPU."""
    if name == "kitti":
        db, k, r = kitti, 8, 1.0
    else:
        db, k, r = (np.random.default_rng(1234).random((1_000_000, 3)) * 25.4).astype(np.float32), 16, 0.5
    ix = pcp.SearchIndex(db)
    idx, dist, cnt = ix.knn(db, k)
    b_idx, b_dist, b_cnt = ix.knn_bruteforce(db, k)
    assert np.array_equal(dist, b_dist) and np.array_equal(cnt, b_cnt)
    assert np.array_equal(idx, b_idx)
    s_idx, s_dist, _ = ix.knn_self(k)
    assert np.array_equal(s_idx, b_idx) and np.array_equal(s_dist, b_dist)
    off, _, _ = ix.radius(db, r)
    assert np.array_equal(np.diff(off), ix.radius_count_bruteforce(db, r))


def test_multi_single_rank_and_two_rank_c5(pcp):
    """pcp_multi_* (C ABI): world 1 needs no NCCL and must equal the plain index; with >= 2 GPUs a scaled config 5
    runs under torchrun (ncclAllGather of the reference shards inside the library, queries sharded) and every
    sampled query equals the exhaustive device kernel."""
    import json
    import os
    import subprocess
    import sys
    from point_cloud_process_b200 import multi
    rng = np.random.default_rng(17)
    ref = (rng.random((50_000, 3)) * 20).astype(np.float32)
    q = (rng.random((9000, 3)) * 20).astype(np.float32)
    m = multi.create(rank=0, world=1, device=0)
    assert m.set_reference(ref) == len(ref) and m.info()["comm_nranks"] == 1
    idx, dist, cnt = m.knn(q, 16)
    i2, d2, c2 = pcp.SearchIndex(ref).knn(q, 16)
    assert np.array_equal(idx, i2) and np.array_equal(dist, d2) and np.array_equal(cnt, c2)
    lo, hi, out = multi.icp_batch_sharded([ref[:3000], ref[3000:7000]], [ref[:3000] + 0.01, ref[3000:7000] + 0.01], multi=m)
    assert (lo, hi) == (0, 2) and out["T_last"].shape == (2, 4, 4)
    if pcp.native.device_count() < 2:
        return
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                        "127.0.0.1", "--master-port", "29641", os.path.join(root, "tools", "run_c5.py"), "--ref", "2000001",
                        "--queries", "4000000", "--check", "2048"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    rep = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])
    assert rep["comm_nranks"] == 2 and rep["reference_rows_gathered"] == 2000001 and all(rep["parity_sampled_vs_device_bruteforce"])


def test_cuda_graph_of_build_and_knn(pcp, kitti):
    """pcp_graph_*: index rebuild + kNN of one frame captured once and replayed; the data behind the captured
    device pointers changes between launches and every replay equals the plain calls.  A capture that needs a
    buffer to grow is refused."""
    import torch
    native = pcp.native
    ctx = native.Context(0)
    n, k = len(kitti), 8
    d = torch.from_numpy(kitti).cuda()
    idx = torch.empty((n, k), dtype=torch.int32, device="cuda")
    dst = torch.empty((n, k), dtype=torch.float32, device="cuda")
    cnt = torch.empty(n, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    ix = native.Index(ctx, d.data_ptr(), n, 3)

    def step():
        ix.rebuild_device(d.data_ptr(), n, 3)
        ix.knn_device(d.data_ptr(), n, 3, k, idx.data_ptr(), dst.data_ptr(), cnt.data_ptr())
    step()
    ctx.synchronize()
    want = [(idx.clone(), dst.clone())]
    ctx.graph_begin()
    step()
    g = ctx.graph_end()
    assert g.nodes() >= 5
    frames = [kitti, (kitti[::-1] * np.float32(1.01)).copy(), kitti + np.float32(0.25)]
    for f in frames[1:]:
        d.copy_(torch.from_numpy(f).cuda())
        torch.cuda.synchronize()
        step()
        ctx.synchronize()
        want.append((idx.clone(), dst.clone()))
    for f, (wi, wd) in zip(frames, want):
        d.copy_(torch.from_numpy(f).cuda())
        idx.zero_(); dst.zero_()
        torch.cuda.synchronize()
        g.launch()
        ctx.synchronize()
        assert torch.equal(idx, wi) and torch.equal(dst, wd)
    # a capture whose calls were never run at that size must be refused, and the context stays usable
    big = torch.from_numpy(np.concatenate([kitti, kitti + np.float32(200.0)])).cuda()
    torch.cuda.synchronize()
    ctx.graph_begin()
    ix.rebuild_device(big.data_ptr(), len(big), 3)
    with pytest.raises(ValueError):
        ctx.graph_end()
    step()
    ctx.synchronize()
    assert torch.equal(idx, want[-1][0])


def test_two_devices_in_one_process(kitti):
    """Contexts on two GPUs of one process: kernels that opt into > 48 KB of dynamic shared memory (tiled kNN,
    brute force, per-cloud sort of the ICP builds) must set their attributes on EVERY device, not once per process."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 visible GPUs")
    import point_cloud_process_b200 as pcp
    native = pcp.native
    from oracle import icp_oracle
    pts = np.ascontiguousarray(kitti[:40000])
    tgt, src, _ = icp_oracle.make_pair(77, n=3000, box=(30.0, 30.0, 5.0))
    so = np.array([0, len(src), 2 * len(src)], np.int64)
    out = []
    for dev in (0, 1):
        ctx = native.Context(dev)
        ix = native.Index(ctx, pts)
        a = ix.knn(pts, 8)
        b = ix.knn_bruteforce(pts[:4096], 8)
        icp = native.icp_batch(ctx, np.concatenate([src, src]), so, np.concatenate([tgt, tgt]), so,
                               np.tile(np.eye(4), (2, 1, 1)), 30, 1e-5, 1e-5, 5.0, True, False)
        assert np.array_equal(a[0][:4096], b[0]) and np.array_equal(a[1][:4096], b[1])
        out.append((a, icp))
    assert np.array_equal(out[0][0][0], out[1][0][0]) and np.array_equal(out[0][0][1], out[1][0][1])
    assert np.array_equal(out[0][1]["iterations"], out[1][1]["iterations"])
    assert np.allclose(out[0][1]["T_total"], out[1][1]["T_total"], rtol=0, atol=1e-12)


@pytest.mark.parametrize("n", [1, 2, 31, 32, 33, 1000, 1024, 1025, 4096, 20000, 37888, 37889, 131071, 262144, 262145, 1100000])
def test_build_paths_by_cloud_size(n):
    """Every build path by size -- the one-kernel sort (<= 2^18 points; grids of 1 .. 148 CTAs, 1 .. 7 keys per thread),
    cub's device sort beyond it, the upper hierarchy levels fused into the refinement (<= 1024 level-1 nodes) or
    launched separately -- gives an index whose searches equal the exhaustive kernel's, for query sets on both the
    warp-per-query and the tiled path (whose regrouping uses the same sort)."""
    import point_cloud_process_b200 as pcp
    rng = np.random.default_rng(n)
    pts = (rng.random((n, 3)) * np.array([40.0, 30.0, 6.0])).astype(np.float32)
    if n > 64:
        pts[n // 3:n // 3 + 40] = pts[7]                          # a mass of identical points
    ctx = pcp.native.Context(0)
    ix = pcp.native.Index(ctx, pts)
    k = 4
    for nq in (min(n, 500), 9000):
        q = np.concatenate([pts[rng.integers(0, n, nq // 2 + 1)],
                            (rng.random((nq // 2 + 1, 3)) * np.array([44.0, 33.0, 7.0]) - 2.0).astype(np.float32)])
        got = ix.knn(q, k)
        want = ix.knn_bruteforce(q, k)
        for g, w, name in zip(got, want, ("idx", "dist", "count")):
            assert np.array_equal(g, w), f"n={n} nq={len(q)}: {name} differs"
    if n >= 1000:
        sel = rng.integers(0, n, 3000)
        got = ix.knn_self(k)
        want = ix.knn_bruteforce(pts[sel], k)
        assert np.array_equal(got[0][sel], want[0]) and np.array_equal(got[1][sel], want[1])


@pytest.mark.parametrize("n", [3000, 60000])
def test_rows_with_extra_columns(n):
    """(N, 4) KITTI rows (x, y, z, intensity) and (Q, 6) query rows: only x, y, z are read (benchmark.py:16-27 keeps
    the first three columns; the C ABI takes a row stride): same results as the packed (N, 3) arrays, on the
    warp-per-query and the tiled path, for kNN, self-kNN and radius search."""
    import point_cloud_process_b200 as pcp
    rng = np.random.default_rng(n)
    p3 = (rng.random((n, 3)) * 30).astype(np.float32)
    q3 = (rng.random((n // 2, 3)) * 30).astype(np.float32)
    p4 = np.concatenate([p3, rng.random((n, 1)).astype(np.float32) * 1e6], 1)
    q6 = np.concatenate([q3, rng.normal(size=(len(q3), 3)).astype(np.float32) * 1e3], 1)
    ctx = pcp.native.Context(0)                  # (the native layer passes the row stride on; api.py would slice)
    a, b = pcp.native.Index(ctx, p3), pcp.native.Index(ctx, p4)
    for x, y in zip(a.knn(q3, 8), b.knn(q6, 8)):
        assert np.array_equal(x, y)
    for x, y in zip(a.knn_self(5), b.knn_self(5)):
        assert np.array_equal(x, y)
    ra, rb = a.radius(q3[:3000], 1.5, True), b.radius(q6[:3000], 1.5, True)
    for x, y in zip(ra, rb):
        assert np.array_equal(x, y)
