"""GPU parity: batched CUDA ICP (pybind11 -> C ABI -> pcp_icp.cu) vs the reference-generated golden
vectors (Registration/main.py:icp_point2point run under an Open3D stand-in) and the float64 oracle.

Tolerances (BASELINE.json north_star): rotation within 1e-4 rad, translation within 1e-4 m, identical
iteration counts / status, for BOTH the last increment (what main.py:156 returns) and the composed pose
(icp_template.py:198)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROT_TOL = 1e-4      # rad
TRANS_TOL = 1e-4    # m


@pytest.fixture(scope="module")
def pcp():
    import point_cloud_process_b200 as m
    return m


def pose_err(Ta, Tb):
    dR = Ta[:3, :3] @ Tb[:3, :3].T
    ang = float(np.arccos(np.clip((np.trace(dR) - 1.0) / 2.0, -1.0, 1.0)))
    return ang, float(np.linalg.norm(Ta[:3, 3] - Tb[:3, 3]))


def assert_pose(Ta, Tb, what=""):
    ang, dt = pose_err(np.asarray(Ta), np.asarray(Tb))
    assert ang <= ROT_TOL and dt <= TRANS_TOL, f"{what}: rotation {ang:.3e} rad, translation {dt:.3e} m"


MODES = {"A": {}, "B": dict(max_iteration=50, R_diff_thres=1e-6, t_diff_thres=1e-6)}


@pytest.mark.parametrize("name", ["a0", "a1", "a2", "a3", "a4", "a5"])
@pytest.mark.parametrize("mode", ["A", "B"])
def test_icp_golden(pcp, icp_golden, name, mode):
    """Against the reference's own icp_point2point: last increment + final in-place source points."""
    g = icp_golden
    src, tgt, T0 = g[name + "/source"], g[name + "/target"], g[name + "/T_init"]
    out = pcp.icp_batch([src], [tgt], T_init=T0[None], **MODES[mode])
    assert_pose(out["T_last"][0], g[f"{name}/{mode}_T_last"], "last increment")
    # the reference leaves `source` transformed by every pose it applied (main.py:110): T_total minus the last increment
    applied = np.linalg.inv(out["T_last"][0]) @ out["T_total"][0] if out["status"][0] != 2 else out["T_total"][0]
    moved = src.astype(np.float64) @ applied[:3, :3].T + applied[:3, 3]
    assert np.abs(moved - g[f"{name}/{mode}_source_final"]).max() <= 2e-4


def test_icp_failure_case(pcp, icp_golden):
    g = icp_golden
    out = pcp.icp_batch([g["fail/source"]], [g["fail/target"]], T_init=g["fail/T_init"][None])
    assert out["status"][0] == 2 and out["iterations"][0] == 1      # PCP_ICP_FEW_PAIRS after one pass (main.py:125-127)
    assert np.array_equal(out["T_last"][0], g["fail/A_T_last"])
    assert np.array_equal(out["T_total"][0], g["fail/T_init"])


@pytest.mark.parametrize("mode", ["A", "B"])
def test_icp_batch_vs_oracle(pcp, mode):
    """A ragged batch (different sizes, one perturbed init, one hopeless pair) in ONE call vs the float64
    oracle pair by pair: iterations, status, both poses and the R_diff / t_diff log."""
    from oracle import icp_oracle
    specs = [(1000, 4000), (1001, 2500), (1002, 6000), (1003, 777), (1004, 3000), (1005, 5000)]
    srcs, tgts, inits = [], [], []
    for i, (seed, n) in enumerate(specs):
        tgt, src, T_gt = icp_oracle.make_pair(seed, n=n, box=(30.0, 30.0, 5.0))
        T0 = np.eye(4)
        if i == 4:
            T0 = T_gt.copy()
            T0[:3, 3] += 0.05
        srcs.append(src), tgts.append(tgt), inits.append(T0)
    tgt, src, _ = icp_oracle.make_pair(1100, n=300, box=(5.0, 5.0, 2.0))
    srcs.append(src + np.float32(100.0)), tgts.append(tgt), inits.append(np.eye(4))
    kw = MODES[mode]
    out = pcp.icp_batch(srcs, tgts, T_init=np.asarray(inits), want_log=True, **kw)
    for p in range(len(srcs)):
        T, info = icp_oracle.icp_point2point(srcs[p], tgts[p], inits[p], return_info=True, **kw)
        assert out["iterations"][p] == info["iterations"], f"pair {p}"
        want_status = 2 if info["status"] == 1 else (1 if (info["iterations"] == kw.get("max_iteration", 100) and not (
            info["log"]["R_diff"][-1] <= kw.get("R_diff_thres", 0.5) and info["log"]["t_diff"][-1] <= kw.get("t_diff_thres", 0.5))) else 0)
        assert out["status"][p] == want_status, f"pair {p}"
        assert_pose(out["T_last"][p], T, f"pair {p} last increment")
        assert_pose(out["T_total"][p], info["total"], f"pair {p} composed pose")
        nlog = len(info["log"]["R_diff"])
        log = out["diff_log"][p]
        assert np.allclose(log[:nlog, 0], info["log"]["R_diff"], rtol=0, atol=2e-4)
        assert np.allclose(log[:nlog, 1], info["log"]["t_diff"], rtol=0, atol=2e-4)
        assert np.isnan(log[nlog:]).all()


def test_icp_first_iteration_broadcast_switch(pcp):
    """main.py:150: on iteration 1 t (3,1) - t_last (3,) broadcasts to 3x3; the flag reproduces / removes it."""
    from oracle import icp_oracle
    tgt, src, _ = icp_oracle.make_pair(1010, n=3000, box=(25.0, 25.0, 5.0))
    a = pcp.icp_batch([src], [tgt], want_log=True)
    b = pcp.icp_batch([src], [tgt], want_log=True, first_iter_broadcast=False)
    T, info = icp_oracle.icp_point2point(src, tgt, np.eye(4), return_info=True)
    assert abs(a["diff_log"][0][0, 1] - info["log"]["t_diff"][0]) < 1e-6
    t = a["T_last"][0][:3, 3] if a["iterations"][0] == 1 else None
    # without the quirk the first t_diff is the plain norm, which is sqrt(3) times smaller for t_last = 0
    assert abs(b["diff_log"][0][0, 1] * np.sqrt(3.0) - a["diff_log"][0][0, 1]) < 1e-9
    assert t is None or abs(np.linalg.norm(t) - b["diff_log"][0][0, 1]) < 1e-9


def test_icp_pose_accuracy_mode_b(pcp):
    """SURVEY 8(d) C4 Mode B at the full 20 000-point size: converges to the ground truth."""
    from oracle import icp_oracle
    tgt, src, T_gt = icp_oracle.make_pair(1000, n=20000)
    out = pcp.icp_batch([src], [tgt], max_iteration=50, R_diff_thres=1e-6, t_diff_thres=1e-6)
    ang, dt = pose_err(out["T_total"][0], T_gt)
    assert ang < 2e-4 and dt < 2e-3, (ang, dt)
    T, info = icp_oracle.icp_point2point(src, tgt, np.eye(4), max_iteration=50, R_diff_thres=1e-6, t_diff_thres=1e-6,
                                         return_info=True)
    assert out["iterations"][0] == info["iterations"]
    assert_pose(out["T_total"][0], info["total"], "composed pose")


def test_icp_edge_cases(pcp):
    e = np.zeros((0, 3), np.float32)
    out = pcp.icp_batch([], [])
    assert out["T_last"].shape == (0, 4, 4)
    rng = np.random.default_rng(0)
    pts = rng.random((100, 3)).astype(np.float32)
    out = pcp.icp_batch([e, pts], [pts, e])          # empty source / empty target -> too few associations
    assert out["status"].tolist() == [2, 2] and out["iterations"].tolist() == [1, 1]
    out = pcp.icp_batch([pts], [pts])                # identical clouds -> identity in one pass
    assert out["status"][0] == 0 and np.allclose(out["T_last"][0], np.eye(4), atol=1e-12)
    with pytest.raises(ValueError):
        pcp.icp_batch([pts], [pts, pts])


def test_homo2tq_golden(pcp, icp_golden):
    g = icp_golden
    got = np.asarray([pcp.native.homo2tq(M) for M in g["quat/homo"]])
    assert np.array_equal(got, g["quat/tq"])


@pytest.mark.timeout(600)
def test_icp_batch_key_width_same_results():
    """The batched build packs (segment id, significant Morton bits) into 32-bit sort keys when they fit;
    PCP_BUILD_KEYS=64 forces the 64-bit layout.  Same order, so bit-identical poses and iteration counts."""
    import os, subprocess, sys, tempfile
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = ("import sys, numpy as np; sys.path.insert(0, %r); import point_cloud_process_b200 as pcp;"
            "from oracle import icp_oracle;"
            "prs = [icp_oracle.make_pair(2000 + i, n=1500 + 331 * i, box=(30.0, 30.0, 5.0)) for i in range(9)];"
            "out = pcp.icp_batch([p[1] for p in prs], [p[0] for p in prs], max_iteration=30, R_diff_thres=1e-6, t_diff_thres=1e-6);"
            "np.savez(sys.argv[1], T=out['T_total'], L=out['T_last'], it=out['iterations'], st=out['status'])") % root
    outs = []
    with tempfile.TemporaryDirectory() as tmp:
        for i, extra in enumerate(({}, {"PCP_BUILD_KEYS": "64"})):
            path = os.path.join(tmp, "v%d.npz" % i)
            r = subprocess.run([sys.executable, "-c", code, path], env=dict(os.environ, **extra), capture_output=True,
                               text=True, timeout=300, cwd=root)
            assert r.returncode == 0, r.stderr[-2000:]
            outs.append(dict(np.load(path)))
    for key in outs[0]:
        assert np.array_equal(outs[0][key], outs[1][key]), key


def _oracle_pair_job(args):
    seed, n, kw = args
    from oracle import icp_oracle
    tgt, src, _ = icp_oracle.make_pair(seed, n=n)
    T, info = icp_oracle.icp_point2point(src, tgt, np.eye(4), return_info=True, **kw)
    return T, info["total"], info["iterations"], info.get("status", 0)


def _oracle_pairs(seeds, n, kw):
    import multiprocessing
    with multiprocessing.get_context("spawn").Pool(min(8, len(seeds))) as pool:
        return pool.map(_oracle_pair_job, [(s, n, kw) for s in seeds], chunksize=4)


def _pose_err(T, T_ref):
    dR = T[:3, :3] @ T_ref[:3, :3].T
    return float(np.arccos(np.clip((np.trace(dR) - 1) / 2, -1, 1))), float(np.linalg.norm(T[:3, 3] - T_ref[:3, 3]))


def test_c2_full_pair_list_to_reg_result(pcp, tmp_path):
    """BASELINE config 2 at its stated size: all 342 rows of the reference's pair list
    (registration_dataset/reg_result.txt:2-343), ~20 000-point clouds in the Oxford 6 x float32 layout written
    to disk and read back through read_bin_velodyne (main.py:10-17), one batched ICP call, reg_result.txt written
    in the reference's format -- every row equal to the oracle's after %f formatting, and the file round-trips
    through the evaluator (evaluate_rt.py)."""
    import os
    from oracle import icp_oracle
    from point_cloud_process_b200 import evaluate_rt, icp, registration
    here = os.path.dirname(os.path.abspath(__file__))
    pair_list = os.path.join(here, "golden", "reg_pairs.txt")
    pair_ids = icp.read_pair_list(pair_list)
    assert len(pair_ids) == 342 and pair_ids[:3] == [(643, 456), (0, 456), (645, 189)]
    cloud_dir = tmp_path / "point_clouds"
    registration.write_synthetic_dataset(pair_ids, str(cloud_dir), lambda seed, n: icp_oracle.make_pair(seed, n=n), 20000)
    out_path = tmp_path / "reg_result.txt"
    rows, out = registration.run(pair_list, str(cloud_dir), str(out_path))
    want = _oracle_pairs([7000 + j for j in range(342)], 20000, {})
    rows_o = [[t, s, *icp_oracle.homo2tq(w[0])] for (t, s), w in zip(pair_ids, want)]
    assert np.array_equal(out["iterations"], [w[2] for w in want])
    got_txt, want_txt = out_path.read_text().splitlines(), icp_oracle.format_reg_result(rows_o).splitlines()
    assert len(got_txt) == 343 and got_txt[0] == want_txt[0] == "# idx1,idx2,t_x,t_y,t_z,q_w,q_x,q_y,q_z"
    for a, b in zip(got_txt[1:], want_txt[1:]):
        fa, fb = np.float64(a.split(",")), np.float64(b.split(","))
        assert np.abs(fa - fb).max() <= 1.01e-6            # equal after %f formatting up to one rounding step
    # round trip: the evaluator reads the file back; scored against the oracle's own rows every pair succeeds
    gt_path = tmp_path / "groundtruths.txt"
    gt_path.write_text(icp_oracle.format_reg_result(rows_o))
    rate, rte, rre = evaluate_rt.evaluate_rt(str(gt_path), str(out_path))
    assert rate == 342 / 343 and rte < 1e-5 and rre < 1e-3      # (the reference divides by the line count incl. the header)


@pytest.mark.parametrize("mode", ["A", "B"])
def test_c4_64_pairs_of_20000_points_vs_oracle(pcp, mode):
    """Config 4 at its per-pair size: 64 pairs x 20 000 points in ONE batch, reference constants (mode A) and the
    converging variant (mode B): iteration counts and status equal to the oracle's, last increment and composed
    pose within 1e-4 rad / 1e-4 m."""
    from oracle import icp_oracle
    kw = {} if mode == "A" else dict(max_iteration=50, R_diff_thres=1e-6, t_diff_thres=1e-6)
    seeds = [1000 + i for i in range(64)]
    pairs = [icp_oracle.make_pair(s, n=20000) for s in seeds]
    out = pcp.icp_batch([p[1] for p in pairs], [p[0] for p in pairs], **kw)
    want = _oracle_pairs(seeds, 20000, kw)
    assert np.array_equal(out["iterations"], [w[2] for w in want])
    assert ((out["status"] == 2) == np.asarray([w[3] == 1 for w in want])).all()      # "fewer than 3 associations" agrees (none here)
    limit = 100 if mode == "A" else 50
    assert ((out["status"] == 1) == (np.asarray([w[2] for w in want]) == limit)).all() or (out["status"] <= 1).all()
    for i, w in enumerate(want):
        a1, t1 = _pose_err(out["T_last"][i], w[0])
        a2, t2 = _pose_err(out["T_total"][i], w[1])
        assert max(a1, a2) <= 1e-4 and max(t1, t2) <= 1e-4, (i, a1, t1, a2, t2)
    if mode == "A":
        assert set(np.unique(out["iterations"]).tolist()) <= {1, 2}      # 1 iteration on ~80 %, 2 on the rest


def test_icp_batch_pinned_host_pipeline_same_results(pcp, monkeypatch):
    """pcp_icp_batch(PCP_HOST) with page-locked clouds copies and registers the batch in runs of pairs (copy of run
    c + 1 beside the registration of run c): same iterations, poses and log (1e-9) as the single-shot pageable path.
    Ragged batch, large enough for several runs (kIcpChunkPoints = 12 Mi points)."""
    import torch
    from oracle import icp_oracle
    native = pcp.native
    rng = np.random.default_rng(5)
    sizes = [int(s) for s in rng.integers(150_000, 420_000, 48)] + [0, 17]
    srcs, tgts = [], []
    for i, n in enumerate(sizes):
        if n == 0:
            srcs.append(np.zeros((0, 3), np.float32)), tgts.append(np.zeros((0, 3), np.float32))
            continue
        tgt, src, _ = icp_oracle.make_pair(4000 + i, n=n, box=(60.0, 60.0, 8.0))
        srcs.append(src), tgts.append(tgt)
    so = np.concatenate([[0], np.cumsum([len(s) for s in srcs])]).astype(np.int64)
    to = np.concatenate([[0], np.cumsum([len(t) for t in tgts])]).astype(np.int64)
    assert so[-1] + to[-1] > 2 * (12 << 20)                       # at least three runs
    sp, tp = np.concatenate(srcs), np.concatenate(tgts)
    T0 = np.tile(np.eye(4), (len(srcs), 1, 1))
    ctx = native.Context(0)
    args = (T0, 30, 1e-5, 1e-5, 5.0, True, True)
    ref = native.icp_batch(ctx, sp, so, tp, to, *args)            # pageable: one shot
    pin_s, pin_t = torch.from_numpy(sp).pin_memory().numpy(), torch.from_numpy(tp).pin_memory().numpy()
    got = native.icp_batch(ctx, pin_s, so, pin_t, to, *args)      # page-locked: pipelined
    for key in ("iterations", "status"):
        assert np.array_equal(ref[key], got[key]), key
    # (the Morton key width of a build follows its largest cloud, so a run may bucket a cloud differently from the
    # whole batch: float64 sums in another order)
    for key in ("T_last", "T_total", "diff_log"):
        assert np.allclose(np.asarray(ref[key]), np.asarray(got[key]), rtol=0, atol=1e-9, equal_nan=True), key


def test_icp_rows_with_extra_columns(pcp):
    """Oxford-layout clouds (x, y, z, nx, ny, nz per row, icp_template.py:11-17): only the first three columns are
    read; same result as the packed clouds."""
    from oracle import icp_oracle
    rng = np.random.default_rng(3)
    srcs, tgts, srcs6, tgts6 = [], [], [], []
    for i, n in enumerate((5000, 20000, 333)):
        tgt, src, _ = icp_oracle.make_pair(900 + i, n=n, box=(40.0, 40.0, 6.0))
        srcs.append(src), tgts.append(tgt)
        srcs6.append(np.concatenate([src, rng.normal(size=(n, 3)).astype(np.float32)], 1))
        tgts6.append(np.concatenate([tgt, rng.normal(size=(n, 3)).astype(np.float32)], 1))
    off = np.concatenate([[0], np.cumsum([len(s) for s in srcs])]).astype(np.int64)
    T0 = np.tile(np.eye(4), (len(srcs), 1, 1))
    ctx = pcp.native.Context(0)                  # (the native layer passes the row stride on; api.py would slice)
    args = (T0, 30, 1e-6, 1e-6, 5.0, True, False)
    a = pcp.native.icp_batch(ctx, np.concatenate(srcs), off, np.concatenate(tgts), off, *args)
    b = pcp.native.icp_batch(ctx, np.concatenate(srcs6), off, np.concatenate(tgts6), off, *args)
    assert np.array_equal(a["iterations"], b["iterations"]) and np.array_equal(a["status"], b["status"])
    assert np.array_equal(np.asarray(a["T_total"]), np.asarray(b["T_total"]))
