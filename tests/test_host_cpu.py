"""CPU-only checks of the host side: the C-ABI library loads and exports every symbol include/pcp.h
declares, argument validation that needs no GPU, the result-set containers' incremental semantics
against the reference-pinned oracle, and loud failure (no CPU fallback) without a device."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "point_cloud_process_b200", "libpcp_b200.so")


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(LIB):
        import __graft_entry__
        __graft_entry__.build()
    return ctypes.CDLL(LIB)


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "pcp.h")).read()
    return sorted(set(re.findall(r"PCP_API\s+[\w\s\*]+?\b(pcp_\w+)\s*\(", text)))


def test_header_declares_the_documented_surface():
    syms = declared_symbols()
    for s in ("pcp_ctx_create", "pcp_index_build", "pcp_knn", "pcp_knn_self", "pcp_radius_count", "pcp_radius_fill",
              "pcp_icp_batch", "pcp_procrustes", "pcp_homo2tq", "pcp_last_error", "pcp_voxel_filter",
              "pcp_voxel_filter_read", "pcp_normals", "pcp_pca", "pcp_iss_keypoints", "pcp_dbscan"):
        assert s in syms


def test_library_exports_every_declared_symbol(lib):
    missing = [s for s in declared_symbols() if not hasattr(lib, s)]
    assert not missing, missing


def test_version_and_argument_errors_without_gpu(lib):
    assert lib.pcp_version() == 100
    lib.pcp_last_error.restype = ctypes.c_char_p
    assert lib.pcp_ctx_create(0, None) == -1                      # PCP_ERR_INVALID
    assert b"NULL" in lib.pcp_last_error()
    assert lib.pcp_ctx_synchronize(None) == -1
    assert lib.pcp_knn(None, None, 0, 3, 8, 0, None, None, None) == -1
    tq = (ctypes.c_double * 7)()
    eye = (ctypes.c_double * 16)(*np.eye(4).ravel())
    assert lib.pcp_homo2tq(eye, tq) == 0 and list(tq) == [0, 0, 0, 1, 0, 0, 0]
    # the widened entry points (voxel filter, normals, PCA, ISS, DBSCAN) reject NULL handles the same way
    lib.pcp_voxel_filter.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int64, ctypes.c_double,
                                     ctypes.c_int, ctypes.c_uint64, ctypes.c_int, ctypes.c_void_p]
    assert lib.pcp_voxel_filter(None, None, 0, 3, 0.5, 0, 0, 0, None) == -1
    assert lib.pcp_voxel_filter_read(None, 0, None) == -1
    assert lib.pcp_normals(None, 5, 0, None, None) == -1
    lib.pcp_pca.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int64, ctypes.c_int,
                            ctypes.c_void_p, ctypes.c_void_p]
    assert lib.pcp_pca(None, None, 0, 3, 0, None, None) == -1
    lib.pcp_iss_keypoints.argtypes = [ctypes.c_void_p, ctypes.c_float, ctypes.c_double, ctypes.c_double, ctypes.c_float,
                                      ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    assert lib.pcp_iss_keypoints(None, 0.5, 0.5, 0.5, 0.5, 20, None, None, None) == -1
    lib.pcp_dbscan.argtypes = [ctypes.c_void_p, ctypes.c_float, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
    assert lib.pcp_dbscan(None, 0.5, 10, 0, None, None) == -1
    assert b"pcp_dbscan" in lib.pcp_last_error()
    prm = (ctypes.c_byte * 64)()
    assert lib.pcp_icp_default_params(prm) == 0
    assert np.frombuffer(prm, np.int32, 1)[0] == 100              # main.py:98


def test_no_cpu_fallback_without_device(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    ctx = ctypes.c_void_p()
    rc = lib.pcp_ctx_create(0, ctypes.byref(ctx))
    assert rc == -2 and not ctx.value                             # PCP_ERR_CUDA: fails loudly
    import point_cloud_process_b200 as pcp
    with pytest.raises(RuntimeError):
        pcp.SearchIndex(np.zeros((10, 3), np.float32))
    with pytest.raises(RuntimeError):
        pcp.icp_batch([np.zeros((10, 3), np.float32)], [np.zeros((10, 3), np.float32)])
    from point_cloud_process_b200 import dbscan, iss, pca_normal, voxel_filter
    pts = np.random.default_rng(0).random((50, 3)).astype(np.float32)
    for call in (lambda: voxel_filter.voxel_filter(pts, 0.1, "centroid"), lambda: pca_normal.PCA(pts),
                 lambda: pca_normal.estimate_normals(pts), lambda: iss.iss_keypoints(pts),
                 lambda: dbscan.DBSCAN(0.2, 3).fit(pts)):
        with pytest.raises(RuntimeError):
            call()


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "point_cloud_process_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.replace("# oracle", ""), f"{f} mentions the oracle"


def test_knn_result_set_incremental_semantics():
    """SURVEY Appendix A.3 sequence + random streams against the reference-pinned C restatement
    (oracle_knn_add via a 1-leaf kd-tree visiting points in index order)."""
    from point_cloud_process_b200.result_set import KNNResultSet, RadiusNNResultSet
    rs = KNNResultSet(3)
    seq = [(1.0, 10), (2.0, 11), (2.0, 12), (2.0, 13), (0.5, 14), (2.0, 15)]
    states = []
    for d, i in seq:
        rs.add_point(d, i)
        states.append([(x.distance, x.index) for x in rs.dist_index_list[:rs.count]])
    assert states[2] == [(1.0, 10), (2.0, 11), (2.0, 12)]
    assert states[3] == [(1.0, 10), (2.0, 11), (2.0, 13)]
    assert states[4] == [(0.5, 14), (1.0, 10), (2.0, 11)]
    assert states[5] == [(0.5, 14), (1.0, 10), (2.0, 15)]
    assert rs.comparison_counter == 6 and rs.full() and rs.worstDist() == 2.0
    # fresh set: worst 1e10, slots (1e10, 0)
    e = KNNResultSet(4)
    assert e.worst_dist == 1e10 and [(x.distance, x.index) for x in e.dist_index_list] == [(1e10, 0)] * 4
    assert str(e).endswith("In total 0 comparison operations.")
    # random streams with many ties vs a straightforward restatement of result_set.py:35-53
    rng = np.random.default_rng(0)
    for _ in range(50):
        k = int(rng.integers(1, 9))
        a = KNNResultSet(k)
        dist, index, count, worst = [1e10] * k, [0] * k, 0, 1e10
        for j in range(60):
            d = float(rng.integers(0, 12)) / 4
            a.add_point(d, j)
            if not d > worst:
                if count < k:
                    count += 1
                i = count - 1
                while i > 0 and dist[i - 1] > d:
                    dist[i], index[i] = dist[i - 1], index[i - 1]
                    i -= 1
                dist[i], index[i] = d, j
                worst = dist[k - 1]
            assert [(x.distance, x.index) for x in a.dist_index_list] == list(zip(dist, index))
            assert a.count == count and a.worst_dist == worst
    r = RadiusNNResultSet(1.0)
    for d, i in [(0.5, 1), (1.0, 2), (1.0000001, 3), (0.0, 4)]:
        r.add_point(d, i)
    assert r.size() == 3 and r.worstDist() == 1.0
    assert str(r).startswith("4 - 0.00\n1 - 0.50\n2 - 1.00\nIn total 3 neighbors within 1.000000.")


def test_shim_modules_mirror_reference_names():
    import point_cloud_process_b200 as pcp
    for name in ("kdtree_construction", "kdtree_knn_search", "kdtree_radius_search", "traverse_kdtree", "Node"):
        assert hasattr(pcp.kdtree, name)
    for name in ("octree_construction", "octree_knn_search", "octree_radius_search", "octree_radius_search_fast",
                 "inside", "Octant"):
        assert hasattr(pcp.octree, name)
    for name in ("icp_point2point", "ICP", "find_associations", "procrustes_transformation", "homo2tq",
                 "rotmat2quaternion", "copysign", "read_bin_velodyne", "read_oxford_bin", "write_reg_result"):
        assert hasattr(pcp.icp, name)
    with pytest.raises(ValueError):
        pcp.kdtree.kdtree_construction(np.zeros((3, 1000), np.float32), 32)   # the shipped benchmark.py:27 layout


def test_reg_result_writer_format(tmp_path):
    """reg_result.txt layout of main.py:220-222 (header line prefixed '# ' by np.savetxt, %f columns)."""
    from oracle import icp_oracle
    from point_cloud_process_b200 import icp
    rows = [[643, 456, 1.5, -2.25, 0.125, 0.5437, 0.1, -0.2, 0.8], [0, 456, 0, 0, 0, 1, 0, 0, 0]]
    p = tmp_path / "reg_result.txt"
    icp.write_reg_result(rows, p)
    text = p.read_text()
    assert text == icp_oracle.format_reg_result(rows)
    assert text.splitlines()[0] == "# idx1,idx2,t_x,t_y,t_z,q_w,q_x,q_y,q_z"
    assert text.splitlines()[1] == "643,456,1.500000,-2.250000,0.125000,0.543700,0.100000,-0.200000,0.800000"
    assert icp.read_pair_list(p) == [(643, 456), (0, 456)]


def test_evaluate_rt_matches_reference(tmp_path):
    """the evaluator mirror against outputs of the reference's own evaluate_rt.py on the two
    reg_result.txt files it ships (tests/golden/eval_golden.npz <- tests/golden/make_golden_eval.py); the files
    are re-written through this package's writer (main.py:220-222 format) first."""
    from point_cloud_process_b200 import evaluate_rt as ev
    from point_cloud_process_b200.icp import write_reg_result
    g = np.load(os.path.join(ROOT, "tests", "golden", "eval_golden.npz"))
    a, b = str(tmp_path / "a.txt"), str(tmp_path / "b.txt")
    write_reg_result(g["rows_a"], a)
    write_reg_result(g["rows_b"], b)
    assert open(a).readline().rstrip("\n") == str(g["header_a"])          # numpy's "# " prefix, as shipped
    for (x, y), key in (((a, b), "eval_ab"), ((b, a), "eval_ba"), ((a, a), "eval_aa")):
        got = np.array(ev.evaluate_rt(x, y))
        assert np.allclose(got, g[key], rtol=1e-12, atol=1e-15), (key, got, g[key])
    ra, rb = ev.read_reg_results(a), ev.read_reg_results(b)
    assert len(ra) == 343 and ra[1][0] == "643"
    for i in range(1, 40):
        _, _, ta, qa = ev.reg_result_row_to_array(ra[i])
        _, _, tb, qb = ev.reg_result_row_to_array(rb[i])
        d = ev.get_P_diff(ev.get_P_from_Rt(qa.as_matrix(), ta), ev.get_P_from_Rt(qb.as_matrix(), tb))
        assert np.allclose(d, g["p_diff_first39"][i - 1], rtol=1e-12, atol=1e-15)
    with pytest.raises(AssertionError):
        open(b, "a").write("1,2,0,0,0,1,0,0,0\n")
        ev.evaluate_rt(a, b)


def test_front_end_modules_mirror_reference_names():
    from point_cloud_process_b200 import dbscan, evaluate_rt, iss, pca_normal, voxel_filter
    assert callable(voxel_filter.voxel_filter) and callable(pca_normal.estimate_normals) and callable(pca_normal.PCA)
    assert callable(iss.iss_keypoints) and callable(iss.readModelNet40)
    d = dbscan.DBSCAN(radius=0.3, Min_Pts=7)                    # dbscan.py:5-8
    assert (d.radius, d.Min_Pts, d.labels_) == (0.3, 7, None) and d.predict() is None
    for name in ("get_P_from_Rt", "is_registration_successful", "get_P_diff", "read_oxford_bin", "read_reg_results",
                 "reg_result_row_to_array", "evaluate_rt"):
        assert hasattr(evaluate_rt, name)


def test_knn_result_set_load_merges_like_add_point():
    """KNNResultSet._load on a set that already holds candidates merges (result_set.py:35-53 semantics: a later
    arrival goes behind equal distances, the list stays bounded)."""
    from point_cloud_process_b200.result_set import KNNResultSet
    rng = np.random.default_rng(3)
    for cap in (1, 4, 8):
        rows = [np.sort(rng.integers(0, 12, cap).astype(np.float32)) for _ in range(3)]
        ids = [rng.integers(0, 1000, cap) for _ in range(3)]
        a, b = KNNResultSet(cap), KNNResultSet(cap)
        for d, i in zip(rows, ids):
            a._load(d, i, cap)
            for x, y in zip(d, i):
                b.add_point(x, int(y))
        assert [(x.distance, x.index) for x in a.dist_index_list] == [(x.distance, x.index) for x in b.dist_index_list]
        assert a.count == b.count and a.worst_dist == b.worst_dist
    part = KNNResultSet(4)
    part._load(np.float32([1, 2, 1e10, 1e10]), [7, 8, 0, 0], 2)
    assert part.count == 2 and part.worst_dist == 1e10 and part.dist_index_list[2].index == 0
