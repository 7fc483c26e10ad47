#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200-native nearest-neighbour + ICP hot path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

BASELINE.json metric: "kNN queries/s (k=8, 120k-pt KITTI scan); ICP pairs/s at 1/2/4/8 B200".

N = 1   headline = kNN queries/s, k = 8, on the bundled KITTI scan 000000.bin (config 1: 124 668 points, every
        point a query).  One STEP = one pass of the hot path over one scan: index build (bounding box -> Morton
        keys -> radix sort -> kd-refined buckets -> AABB hierarchy) + batched exact kNN, replayed as ONE CUDA graph
        (pcp_graph_*).  `value` = queries / device time of the K steps (CUDA events on the library's stream),
        inputs resident in HBM, L2 flushed between steps.  `e2e` = the same step through the public API
        (pybind11 -> C ABI, PCP_HOST) from pinned HOST buffers, H2D and D2H inside the timed region.
        `also.icp`    = config 4: 4096 synthetic pairs x 20 000 points, reference constants (main.py:98-103).
        `also.radius` = config 1, second half: radius search r = 1.0 (count + fill of 80.2 M result pairs).
N > 1   headline = ICP pairs/s on config 4 AS WRITTEN: the same 4096 pairs split over the N ranks by points
        (pcp_multi_pair_range), no collective -> "strong" scaling.  `also.knn` = one independent KITTI scan per
        rank (replicas: the single scan does not shard), `also.c5` = config 5 scaled to N/8 of its size (query-
        sharded kNN k = 16 over a reference cloud replicated by ONE ncclAllGather inside the C ABI).
        The two series across N: kNN = `value` at N = 1 and `also.knn.value` at N > 1; ICP = `also.icp.value` at
        N = 1 and `value` at N > 1.

--impl reference: the CPU arm on the same metric -- oracle/ (C restatement of the reference kd-tree for kNN; float64
numpy + cKDTree restatement of icp_point2point for ICP; the reference itself is pure Python and Open3D-bound).
torch is used here for device buffers, CUDA events and the process group (barrier, max over ranks) only.
"""
import argparse
import json
import multiprocessing
import os
import statistics
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
KITTI = os.path.join(ROOT, "tests", "golden", "kitti_000000_xyz.f32")
FP32_LANES_PER_SM = 128
ICP_MODES = {"A": dict(max_iteration=100, R_diff_thres=0.5, t_diff_thres=0.5),          # main.py:98-102 as shipped
             "B": dict(max_iteration=50, R_diff_thres=1e-6, t_diff_thres=1e-6)}        # converging variant
C4_PAIRS, C4_POINTS = 4096, 20000


def knn_config(workload, desc):
    """`config` of the kNN line -- the same dict on the GPU arm and the CPU (reference) arm."""
    return {"workload": workload, "desc": desc, "step": "index (GPU: bucket hierarchy; CPU arm: kd-tree, leaf 32) build + kNN of every point",
            "l2": "GPU arm: flushed between timed steps (256 MB write)", "multi_gpu": "one independent scan per rank, no collective (replicas)"}


def icp_config(pairs, points, mode):
    """`config` of the ICP line -- the same dict on both arms."""
    return {"workload": "icp_c4", "desc": f"{pairs} synthetic pairs x {points} pts (SURVEY C4 recipe), mode {mode} ({ICP_MODES[mode]})",
            "step": "GPU arm: the whole batch, pairs split over the ranks by points, no collective; CPU arm: a bounded sample of the batch",
            "l2": "inputs (GBs of clouds) exceed the 126 MB L2"}


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f), "measured"
    return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0}, "fallback"


def scan_for_rank(rank):
    """Rank 0: the bundled scan.  Rank r > 0: an independent frame derived from it (seeded 5 mm
    jitter + a yaw of 7 degrees * r), so that no two ranks do identical work."""
    pts = np.fromfile(KITTI, np.float32).reshape(-1, 3)
    if rank == 0:
        return pts
    rng = np.random.default_rng(9000 + rank)
    a = np.deg2rad(7.0 * rank)
    R = np.array([[np.cos(a), -np.sin(a), 0], [np.sin(a), np.cos(a), 0], [0, 0, 1]])
    return (pts.astype(np.float64) @ R.T + rng.normal(0, 0.005, pts.shape)).astype(np.float32)


def make_workload(name, rank):
    if name == "kitti_knn8":
        return dict(pts=scan_for_rank(rank), k=8, desc="KITTI 000000.bin 124668 pts, k=8, queries = all points")
    if name == "uniform1m_knn16":
        rng = np.random.default_rng(1234 + rank)
        return dict(pts=(rng.random((1_000_000, 3)) * 25.4).astype(np.float32), k=16,
                    desc="uniform 1M pts in [0,25.4)^3, k=16, queries = all points")
    raise SystemExit(f"unknown workload {name}")


def make_pair(seed, n=C4_POINTS, box=(60.0, 60.0, 5.0), max_rot_deg=(2.0, 2.0, 5.0), max_t=0.5, noise=0.01):
    """SURVEY 8(d) C4 recipe (same stream as the test-suite generator): seeded synthetic registration pair."""
    rng = np.random.default_rng(seed)
    target = (rng.random((n, 3)) * np.asarray(box)).astype(np.float32)
    c = target.astype(np.float64).mean(0)
    ang = np.deg2rad(rng.uniform(-1, 1, 3) * np.asarray(max_rot_deg))
    cx, cy, cz = np.cos(ang)
    sx, sy, sz = np.sin(ang)
    Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    R = Rz @ Ry @ Rx
    t = rng.uniform(-max_t, max_t, 3)
    source = (target.astype(np.float64) - c - t) @ R + c + rng.normal(0, noise, (n, 3))
    source = source[rng.permutation(n)].astype(np.float32)
    return target, source


def _pair_job(args):
    seed, n = args
    return make_pair(seed, n)


def make_pairs(first, count, n_pts, procs):
    """Pairs first .. first + count - 1 of the C4 batch (seed 1000 + i), generated on `procs` host processes."""
    jobs = [(1000 + first + i, n_pts) for i in range(count)]
    if count == 0:
        return np.zeros((0, 3), np.float32), np.zeros((0, 3), np.float32)
    if procs > 1 and count >= 8:
        with multiprocessing.get_context("fork").Pool(procs) as pool:
            out = pool.map(_pair_job, jobs, chunksize=max(1, count // (procs * 4)))
    else:
        out = [_pair_job(j) for j in jobs]
    return np.concatenate([o[1] for o in out]), np.concatenate([o[0] for o in out])      # sources, targets


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index, period_s=0.002):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag = index, [], set(), False
        self.max_mhz = None
        self.period_s = period_s
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:   # noqa: BLE001
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
            getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80): "hw_power_brake",
        }
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:   # noqa: BLE001
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, nm in names.items():
                    if r & bit:
                        self.reasons.add(nm)
            except Exception:   # noqa: BLE001
                pass
            time.sleep(self.period_s)

    def result(self):
        self.stop_flag = True
        self.join(timeout=1.0)
        return {"sm_mhz": (statistics.median(self.samples) if self.samples else None),
                "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ---------------------------------------------------------------------------------------------- CPU arm (oracle/)

def cpu_knn_port(pts, k, threads, min_seconds, max_reps=50):
    """oracle port: kd-tree build + kNN over ALL queries of the scan, repeated for >= min_seconds; plus one
    single-thread search over a sample of the queries."""
    from oracle import knn_oracle
    reps, t_build, t_search = 0, 0.0, 0.0
    t_start = time.perf_counter()
    while True:
        t0 = time.perf_counter()
        kd = knn_oracle.KdTree(pts, 32)
        t1 = time.perf_counter()
        kd.knn(pts, k, threads=threads)
        t2 = time.perf_counter()
        t_build += t1 - t0
        t_search += t2 - t1
        reps += 1
        if time.perf_counter() - t_start >= min_seconds or reps >= max_reps:
            break
    sample = pts[:: max(1, len(pts) // 20000)]
    t0 = time.perf_counter()
    kd.knn(sample, k, threads=1)
    t_single = (time.perf_counter() - t0) / len(sample)
    return reps, t_build / reps, t_search / reps, 1.0 / t_single


def _icp_job(args):
    seed, n_pts, mode = args
    from oracle import icp_oracle
    t, s = make_pair(seed, n_pts)
    t0 = time.perf_counter()
    T, info = icp_oracle.icp_point2point(s, t, np.eye(4), return_info=True, **ICP_MODES[mode])
    return time.perf_counter() - t0, info["iterations"]


def cpu_icp_port(n_pairs, n_pts, mode, procs, first_pair=0):
    """oracle port of icp_point2point (float64 numpy + cKDTree restatement of main.py:97-156): `n_pairs` pairs of
    the C4 batch on a pool of `procs` processes (one pair per process at a time).  Returns pairs/s (wall clock of
    the pool), the per-pair single-core seconds and the iteration counts."""
    jobs = [(1000 + first_pair + i, n_pts, mode) for i in range(n_pairs)]
    procs = max(1, min(procs, n_pairs))
    t0 = time.perf_counter()
    if procs > 1:
        with multiprocessing.get_context("fork").Pool(procs) as pool:
            pool.map(_icp_job, jobs[:procs], chunksize=1)                  # warm the workers (imports, first touch)
            t0 = time.perf_counter()
            out = pool.map(_icp_job, jobs, chunksize=1)
    else:
        out = [_icp_job(j) for j in jobs]
    wall = time.perf_counter() - t0
    return n_pairs / wall, [o[0] for o in out], [o[1] for o in out]


def pose_error(T, T_ref):
    """(rotation angle in rad, translation distance) between two 4x4 poses."""
    dR = T[:3, :3] @ T_ref[:3, :3].T
    ang = float(np.arccos(np.clip((np.trace(dR) - 1.0) / 2.0, -1.0, 1.0)))
    return ang, float(np.linalg.norm(T[:3, 3] - T_ref[:3, 3]))


# ---------------------------------------------------------------------------------------------- kNN (config 1)

def measure_knn(torch, pcp, ctx, stream, w, args, dist, rank, world, local_rank, peaks, peaks_kind):
    """KITTI step (index build + kNN of every point) on this rank's scan: CUDA-graph replay for the throughput
    figure, plain stream launches with per-launch events for the dominant kernel's own time, work counters,
    single-shot latency, end to end from pinned host buffers."""
    native = pcp.native
    pts, k = w["pts"], w["k"]
    n = len(pts)
    d_pts = torch.from_numpy(pts).cuda()
    idx = torch.empty((n, k), dtype=torch.int32, device="cuda")
    dst = torch.empty((n, k), dtype=torch.float32, device="cuda")
    cnt = torch.empty(n, dtype=torch.int32, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")   # 2x the 126 MB L2
    torch.cuda.synchronize()
    ix = native.Index(ctx, d_pts.data_ptr(), n, 3)

    def step():
        ix.rebuild_device(d_pts.data_ptr(), n, 3)
        ix.knn_device(d_pts.data_ptr(), n, 3, k, idx.data_ptr(), dst.data_ptr(), cnt.data_ptr())

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    ctx.synchronize()
    l0 = ctx.launch_count()
    step()
    launches_per_step = ctx.launch_count() - l0
    ctx.synchronize()
    ctx.graph_begin()
    step()
    graph = ctx.graph_end()
    for _ in range(max(args.warmup, 3)):
        graph.launch()
    ctx.synchronize()

    # ---- timed region: K graph replays, L2 flushed (untimed) before each
    K = args.steps
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(2)] for _ in range(K)]
    sampler = ClockSampler(local_rank, 0.002 if rank == 0 else 0.05)   # NVML polls serialise with launches in the driver
    barrier()
    sampler.start()
    t_wall0 = time.perf_counter()
    for i in range(K):
        with torch.cuda.stream(stream):
            flush.zero_()                       # untimed: evicts the previous step's lines from L2
        evs[i][0].record(stream)
        graph.launch()
        evs[i][1].record(stream)
    ctx.synchronize()
    barrier()
    t_wall = time.perf_counter() - t_wall0
    clocks = sampler.result()
    step_ms = [e[0].elapsed_time(e[1]) for e in evs]
    total_ms_local = sum(step_ms)
    total_ms = total_ms_local
    per_rank = None
    if dist is not None:
        t = torch.tensor([total_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t.item())
        seen = [None] * world
        dist.all_gather_object(seen, {"reasons": clocks["reasons"], "ms_per_step": total_ms_local / K,
                                      "median_step_ms": statistics.median(step_ms), "max_step_ms": max(step_ms),
                                      "sm_mhz": clocks.get("sm_mhz"), "clock_samples": clocks.get("samples"),
                                      "wall_ms_per_step": 1e3 * t_wall / K})
        clocks["reasons"] = sorted(set(r for s in seen for r in s["reasons"]))
        clocks["note"] = "sm_mhz: rank 0 (500 Hz); reasons: union over ranks (other ranks polled at 20 Hz)"
        per_rank = [{k2: v for k2, v in s.items() if k2 != "reasons"} for s in seen]
    value = world * n * K / (total_ms * 1e-3)

    # ---- the dominant kernel's own duration: stream launches, an event pair around every knn_tile_kernel launch
    Kk = max(5, min(K, 20))
    ev3 = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(Kk)]
    ctx.enable_kernel_timing(True)
    for i in range(Kk):
        with torch.cuda.stream(stream):
            flush.zero_()
        ev3[i][0].record(stream)
        ix.rebuild_device(d_pts.data_ptr(), n, 3)
        ev3[i][1].record(stream)
        ix.knn_device(d_pts.data_ptr(), n, 3, k, idx.data_ptr(), dst.data_ptr(), cnt.data_ptr())
        ev3[i][2].record(stream)
    ctx.synchronize()
    tile_ms_total, tile_launches = ctx.read_kernel_timing()
    ctx.enable_kernel_timing(False)
    build_ms = statistics.mean(e[0].elapsed_time(e[1]) for e in ev3)
    knn_ms = statistics.mean(e[1].elapsed_time(e[2]) for e in ev3)
    tile_s = tile_ms_total / max(tile_launches, 1) * 1e-3

    # ---- work counters (untimed pass): distance evaluations C of the kNN kernel
    ctx.enable_counters(True)
    ix.knn_device(d_pts.data_ptr(), n, 3, k, idx.data_ptr(), dst.data_ptr(), cnt.data_ptr())
    ctrs = ctx.read_counters()
    ctx.enable_counters(False)
    evals_all = int(ctrs[0])
    evals = evals_all - int(ctrs[7])            # evaluated by the tiled kernel itself (the rest: overflow groups)
    sm_count = torch.cuda.get_device_properties(local_rank).multi_processor_count
    fp32_peak = sm_count * FP32_LANES_PER_SM * float(peaks.get("sm_max_mhz", 1965.0)) * 1e6 / 1e12   # T unfused FP32 op/s
    achieved = 8.0 * evals / tile_s / 1e12
    hbm_bytes = 16 * n + 12 * n + 8 * k * n + 4 * n
    roofline = {
        "bound": "fp32", "kernel": "knn_tile_kernel", "achieved": achieved, "peak": fp32_peak,
        "unit": "TFLOP/s (unfused FP32 ops; 8 per distance evaluation: 3 sub + 3 mul + 2 add, issued as 3 FADD2 + 3 FMUL2 + 4 FADD per pair of evaluations)",
        "frac": achieved / fp32_peak, "frac_fma2_convention": achieved / fp32_peak / 2.0,
        "traffic": ncu_traffic_bytes(args.workload),
        "traffic_note": "bytes per launch, dram__bytes_read.sum + dram__bytes_write.sum of the committed ncu --set full capture of this kernel on this workload (profiles/r2_ncu_knn_tile_*.txt); not re-measured inside this run (needs ncu); null when there is no capture for the workload",
        "peak_source": f"{sm_count} SMs x 128 FP32 lanes x sm_max_mhz ({peaks_kind} MEASURED_PEAKS.json clock); no FP32 peak is measured by the driver",
        "distance_evaluations_per_launch": evals, "evals_per_query": evals / n,
        "kernel_ms": tile_s * 1e3, "kernel_launches_timed": int(tile_launches),
        "kernel_share_of_step": tile_s * 1e3 / (build_ms + knn_ms),
        "timer": "CUDA events recorded by the library on its own stream around every knn_tile_kernel launch of a second timed loop (stream launches; event records cannot sit inside the replayed graph)",
        "whole_search": {"ms": knn_ms, "evals": evals_all, "fp32_frac": 8.0 * evals_all / (knn_ms * 1e-3) / 1e12 / fp32_peak,
                         "note": "query Morton sort + kd regrouping + tiled kernel + overflow kernel"},
        "hbm": {"algorithmic_bytes": hbm_bytes, "achieved_gbs": hbm_bytes / tile_s / 1e9, "peak_gbs": peaks.get("hbm_gbs"),
                "frac": hbm_bytes / tile_s / 1e9 / peaks.get("hbm_gbs", 6533.8)},
    }
    if rank == 0:
        try:     # SURVEY 8(d) companion: work efficiency eta = C_min / C on a query sample
            from scipy.spatial import cKDTree
            ctx.synchronize()
            d_k = dst[:, k - 1].cpu().numpy().astype(np.float64)          # true k-th distances (exact search)
            sel = np.random.default_rng(5).choice(n, min(n, 4000), replace=False)
            tree = cKDTree(pts.astype(np.float64))
            c_min = sum(len(tree.query_ball_point(pts[i].astype(np.float64), d_k[i], p=np.inf)) for i in sel)
            roofline["work_efficiency"] = {"eta": (c_min / len(sel)) / (evals_all / n), "c_min_per_query": c_min / len(sel),
                                           "c_per_query": evals_all / n, "sample_queries": int(len(sel)),
                                           "note": "C_min = points with |p - q|_inf <= true k-th distance (host, scipy)"}
        except Exception as e:                                             # a reporting extra, never fatal
            roofline["work_efficiency"] = {"error": repr(e)[:200]}

    # ---- single-shot latency: one step issued on an idle stream and waited for (host clock)
    shots_g, shots_s = [], []
    for _ in range(20):
        ctx.synchronize()
        t0 = time.perf_counter()
        graph.launch()
        ctx.synchronize()
        shots_g.append(time.perf_counter() - t0)
        t0 = time.perf_counter()
        step()
        ctx.synchronize()
        shots_s.append(time.perf_counter() - t0)
    latency = {"single_shot_ms": 1e3 * statistics.median(shots_g), "min_ms": 1e3 * min(shots_g),
               "stream_launches_single_shot_ms": 1e3 * statistics.median(shots_s),
               "note": "one build + kNN step issued on an idle stream and waited for (host clock, launch included): one "
                       f"CUDA-graph launch of {graph.nodes()} nodes vs {launches_per_step} stream launches"}

    # ---- end to end through the public API with pinned host buffers (H2D + D2H inside)
    h_np = torch.from_numpy(pts).pin_memory().numpy()
    six = pcp.SearchIndex.__new__(pcp.SearchIndex)
    six.ctx, six.db = ctx, h_np
    six._ix = native.Index(ctx, h_np)
    for _ in range(4):                             # results are held like in the timed loop, so that the pinned
        six._ix.rebuild(h_np)                      # result-buffer pool reaches its steady state (two sets) here
        r_idx, r_dist, r_cnt = six.knn(h_np, k)
    barrier()
    Ke = max(5, min(K, 30))
    t0 = time.perf_counter()
    for _ in range(Ke):
        six._ix.rebuild(h_np)                      # H2D of the scan + build
        r_idx, r_dist, r_cnt = six.knn(h_np, k)    # H2D of the queries, search, D2H of the results
    t_e2e = time.perf_counter() - t0
    if dist is not None:
        t = torch.tensor([t_e2e], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        t_e2e = float(t.item())
    e2e = {"value": world * n * Ke / t_e2e, "unit": "queries/s", "h2d_bytes_per_step": 2 * n * 12,
           "d2h_bytes_per_step": n * k * 8 + n * 4, "steps": Ke, "ms_per_step": 1e3 * t_e2e / Ke,
           "timer": "host perf_counter around synchronous C-ABI calls (PCP_HOST mode)"}
    return {"value": value, "ms_per_step": total_ms / K, "clocks": clocks, "per_rank": per_rank, "roofline": roofline,
            "latency": latency, "e2e": e2e, "gpu_launches": int(K * graph.nodes()), "graph_nodes": int(graph.nodes()),
            "stream_launches_per_step": int(launches_per_step),
            "breakdown_ms": {"build": build_ms, "knn": knn_ms, "graph_step_median": statistics.median(step_ms),
                             "stream_step": build_ms + knn_ms, "wall_region_ms_per_step": 1e3 * t_wall / K,
                             "note": "build / knn: the second loop (stream launches); graph_step: the replayed graph"},
            "n": n, "k": k}


# ---------------------------------------------------------------------------------------------- ICP (config 4)

def measure_icp(torch, pcp, ctx, stream, h_src, h_tgt, first_pair, n_pairs, n_pts, mode, steps, warmup, peaks,
                sm_count, parity_pairs=16):
    """Batched ICP over this rank's `n_pairs` pairs (global ids first_pair ...): device-resident pairs/s, end to
    end from pinned host buffers, iteration histogram, launches / host synchronisations per batch, work counters
    -> HBM and FP32 figures, and a sampled parity check against the oracle restatement of icp_point2point."""
    native = pcp.native
    off = np.arange(n_pairs + 1, dtype=np.int64) * n_pts
    prm = ICP_MODES[mode]
    a = (prm["max_iteration"], prm["R_diff_thres"], prm["t_diff_thres"], 5.0, True)
    res = {"pairs": int(n_pairs), "first_pair": int(first_pair), "points_per_cloud": int(n_pts), "mode": mode}
    if n_pairs == 0:
        res.update(device_total_ms=0.0, e2e_total_ms=0.0, steps=steps)
        return res
    d_src, d_tgt = torch.from_numpy(h_src).cuda(), torch.from_numpy(h_tgt).cuda()
    torch.cuda.synchronize()

    def run():
        return native.icp_batch_device(ctx, d_src.data_ptr(), off, d_tgt.data_ptr(), off, 3, *a)
    for _ in range(max(1, warmup)):
        out = run()
    ctx.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for e0, e1 in evs:
        e0.record(stream)
        out = run()
        e1.record(stream)
    ctx.synchronize()
    ms = [e0.elapsed_time(e1) for e0, e1 in evs]
    launches, syncs, enq = ctx.icp_last_stats()
    its = np.asarray(out["iterations"])
    res.update(device_ms_per_batch=float(np.mean(ms)), device_total_ms=float(np.sum(ms)), steps=steps,
               iterations_hist=np.bincount(its).tolist(), status_hist=np.bincount(out["status"], minlength=3).tolist(),
               pair_iterations=int(its.sum()), iteration_loop_launches_per_batch=int(launches), host_syncs_per_batch=int(syncs),
               iterations_enqueued=int(enq),
               host_sync_note="the one synchronisation is the copy of the results: no host read between iterations")
    # work counters (untimed): distance evaluations of the association kernel
    ctx.enable_counters(True)
    run()
    c = ctx.read_counters()
    ctx.enable_counters(False)
    evals = int(c[0])
    t_s = res["device_ms_per_batch"] * 1e-3
    build_bytes = 92 * 2 * n_pts * n_pairs                       # two segmented bucket builds (sources, targets)
    iter_bytes = 28 * n_pts * res["pair_iterations"]             # 12 B source row + 16 B matched target row
    fp32_peak = sm_count * FP32_LANES_PER_SM * float(peaks.get("sm_max_mhz", 1965.0)) * 1e6 / 1e12
    res["roofline"] = {
        "bound": "hbm", "kernel": "pcp_icp_batch (2 segmented builds + association / solve per iteration)",
        "achieved": (build_bytes + iter_bytes) / t_s / 1e9, "peak": peaks.get("hbm_gbs"), "unit": "GB/s",
        "frac": (build_bytes + iter_bytes) / t_s / 1e9 / peaks.get("hbm_gbs", 6533.8), "traffic": None,
        "algorithmic_bytes": {"builds_92B_per_point": build_bytes, "iterations_28B_per_source_point": iter_bytes},
        "fp32": {"distance_evaluations_per_batch": evals, "evals_per_source_point_iteration": evals / max(1, n_pts * res["pair_iterations"]),
                 "achieved_tflops_over_whole_batch": 8.0 * evals / t_s / 1e12, "frac_over_whole_batch": 8.0 * evals / t_s / 1e12 / fp32_peak,
                 "note": "8 unfused FP32 ops per float32 candidate evaluation of icp_assoc_tile_kernel, divided by the time of the WHOLE batch (builds included)"},
        "timer": "CUDA events on the library's stream around each pcp_icp_batch call (device-resident clouds)"}
    # sampled parity against the oracle (untimed)
    from oracle import icp_oracle
    sel = np.linspace(0, n_pairs - 1, min(parity_pairs, n_pairs)).astype(int)
    worst = [0.0, 0.0]
    ok = True
    for p in sel:
        s, t = h_src[p * n_pts:(p + 1) * n_pts], h_tgt[p * n_pts:(p + 1) * n_pts]
        T, info = icp_oracle.icp_point2point(s, t, np.eye(4), return_info=True, **prm)
        e1 = pose_error(np.asarray(out["T_last"][p]), T)
        e2 = pose_error(np.asarray(out["T_total"][p]), info["total"])
        worst = [max(worst[0], e1[0], e2[0]), max(worst[1], e1[1], e2[1])]
        ok = ok and int(its[p]) == info["iterations"] and max(e1[0], e2[0]) <= 1e-4 and max(e1[1], e2[1]) <= 1e-4
    res["parity"] = {"pairs_checked": int(len(sel)), "ok": bool(ok), "max_rot_err_rad": worst[0], "max_trans_err_m": worst[1],
                     "criterion": "iterations equal; last increment and composed pose within 1e-4 rad / 1e-4 m of the oracle"}
    # end to end: pinned host clouds in, poses out
    T0 = np.tile(np.eye(4), (n_pairs, 1, 1))
    p_src, p_tgt = torch.from_numpy(h_src).pin_memory().numpy(), torch.from_numpy(h_tgt).pin_memory().numpy()
    native.icp_batch(ctx, p_src, off, p_tgt, off, T0, *a, False)
    reps = max(1, min(steps, 3))
    t0 = time.perf_counter()
    for _ in range(reps):
        native.icp_batch(ctx, p_src, off, p_tgt, off, T0, *a, False)
    res["e2e_total_ms"] = 1e3 * (time.perf_counter() - t0)
    res["e2e_steps"] = reps
    res["e2e_h2d_bytes_per_batch"] = int(h_src.nbytes + h_tgt.nbytes)
    res["e2e_d2h_bytes_per_batch"] = int(n_pairs * (2 * 128 + 8))
    del d_src, d_tgt
    return res


# ---------------------------------------------------------------------------------------------- exhaustive kernel (K5)

def measure_bruteforce(torch, pcp, ctx, stream, pts, k, reps, sm_count, sm_max_mhz):
    """The exhaustive kNN kernel (pcp_brute.cu) over the whole scan, device resident: every (query, point) pair is
    evaluated, so the evaluation count is exact (no counters) and the FP32 fraction is the evaluation loop's own."""
    native = pcp.native
    n = len(pts)
    d = torch.from_numpy(pts).cuda()
    idx = torch.empty((n, k), dtype=torch.int32, device="cuda")
    dist = torch.empty((n, k), dtype=torch.float32, device="cuda")
    cnt = torch.empty(n, dtype=torch.int32, device="cuda")
    ix = native.Index(ctx, d.data_ptr(), n, 3)
    ms = []
    for _ in range(reps + 1):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        ix.knn_bruteforce_device(d.data_ptr(), n, 3, k, idx.data_ptr(), dist.data_ptr(), cnt.data_ptr())
        e1.record(stream)
        ctx.synchronize()
        ms.append(e0.elapsed_time(e1))
    t = statistics.median(ms[1:]) * 1e-3
    evals = n * ((n + 31) // 32 * 32)
    peak = sm_count * 128 * sm_max_mhz * 1e6
    return {"workload": f"kitti_bruteforce k={k}", "queries": n, "ms": t * 1e3, "queries_per_s": n / t,
            "roofline": {"bound": "fp32", "kernel": "brute_kernel", "achieved": 8 * evals / t / 1e12, "peak": peak / 1e12,
                         "unit": "TFLOP/s (unfused FP32 ops; 8 per distance evaluation)", "frac": 8 * evals / t / peak,
                         "distance_evaluations_per_launch": evals, "traffic": None,
                         "note": "every pair of the 124 668 x 124 672 slots is evaluated (no pruning): the exhaustive "
                                 "cross-check kernel of SURVEY K5, the packed-FP32 evaluation loop on its own"}}


# ---------------------------------------------------------------------------------------------- radius (config 1)

def measure_radius(torch, pcp, ctx, stream, pts, radius, reps, peaks):
    """Radius search r on the KITTI scan, every point a query: count pass (offsets) + fill pass (indices and
    distances of all result pairs), device resident.  HBM roofline on the fill."""
    native = pcp.native
    n = len(pts)
    d = torch.from_numpy(pts).cuda()
    off = torch.empty(n + 1, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    ix = native.Index(ctx, d.data_ptr(), n, 3)
    ix.radius_count_device(d.data_ptr(), n, 3, radius, off.data_ptr())
    ctx.synchronize()
    m = int(off[-1].item())
    ridx = torch.empty(m, dtype=torch.int32, device="cuda")
    rdst = torch.empty(m, dtype=torch.float32, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    tc, tf = [], []
    for i in range(reps + 2):
        with torch.cuda.stream(stream):
            flush.zero_()
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        e[0].record(stream)
        ix.radius_count_device(d.data_ptr(), n, 3, radius, off.data_ptr())
        e[1].record(stream)
        ix.radius_fill_device(d.data_ptr(), n, 3, radius, off.data_ptr(), ridx.data_ptr(), rdst.data_ptr(), False)
        e[2].record(stream)
        ctx.synchronize()
        if i >= 2:
            tc.append(e[0].elapsed_time(e[1]))
            tf.append(e[1].elapsed_time(e[2]))
    ctx.enable_counters(True)                   # ONE counted pass, counters off while timing
    ix.radius_count_device(d.data_ptr(), n, 3, radius, off.data_ptr())
    c = ctx.read_counters()
    ctx.enable_counters(False)
    fill_bytes = 16 * n + 12 * n + 8 * (n + 1) + 8 * m
    fill_s = statistics.median(tf) * 1e-3
    return {"workload": f"kitti_radius r={radius}", "queries": n, "result_pairs": m, "count_ms": statistics.median(tc),
            "fill_ms": statistics.median(tf), "queries_per_s": n / ((statistics.median(tc) + statistics.median(tf)) * 1e-3),
            "evals_per_query_count_pass": int(c[0]) / n,
            "roofline": {"bound": "hbm", "kernel": "radius fill pass", "achieved": fill_bytes / fill_s / 1e9, "peak": peaks.get("hbm_gbs"),
                         "unit": "GB/s", "frac": fill_bytes / fill_s / 1e9 / peaks.get("hbm_gbs", 6533.8), "traffic": None,
                         "algorithmic_bytes": fill_bytes, "note": "16 N + 12 Q + 8 (Q + 1) + 8 M bytes (SURVEY 8d); L2 flushed before every repetition"}}


# ---------------------------------------------------------------------------------------------- reference arm

def run_reference(args):
    """CPU arm on the metric the GPU arm reports at this N (module docstring): N = 1 the kd-tree kNN (oracle/
    knn_oracle.c, all host threads, build + search of the whole scan per step); N > 1 the ICP batch (oracle/
    icp_oracle.py on a pool over all cores, a bounded sample of config 4 per step).  Rank 0 alone runs it."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    if args.gpus > 1:
        n_pairs = max(8, 2 * threads)
        vals, its_all = [], []
        for i in range(args.warmup + args.steps):
            v, _, its = cpu_icp_port(n_pairs, args.icp_points, args.icp_mode, threads)
            if i >= args.warmup:
                vals.append(v)
                its_all = its
        value = statistics.mean(vals)
        line = {
            "impl": "reference", "metric": "ICP pairs/s", "value": value, "unit": "pairs/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * n_pairs / value, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": icp_config(args.icp_pairs, args.icp_points, args.icp_mode),
            "cpu_baseline": {"value": value, "unit": "pairs/s", "cores": threads, "kind": "port",
                             "sample": f"{n_pairs} pairs per step on a pool of {threads} processes; float64 numpy + cKDTree restatement of "
                                       f"main.py:97-156 (the reference needs Open3D: absent); iterations {sorted(set(its_all))}"},
            "e2e": {"value": value, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
        emit_line(line)
        return
    w = make_workload(args.workload, 0)
    pts, k = w["pts"], w["k"]
    from oracle import knn_oracle
    times = []
    for i in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        kd = knn_oracle.KdTree(pts, 32)
        kd.knn(pts, k, threads=threads)
        dt = time.perf_counter() - t0
        if i >= args.warmup:
            times.append(dt)
    total = sum(times)
    value = len(pts) * len(times) / total
    line = {
        "impl": "reference", "metric": "kNN queries/s", "value": value, "unit": "queries/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "KITTI scan bundled with the reference (tests/golden copy)",
        "config": knn_config(args.workload, w["desc"]),
        "cpu_baseline": {"value": value, "unit": "queries/s", "cores": threads, "kind": "port",
                         "sample": f"whole scan ({len(pts)} queries) per step, {len(times)} steps; C restatement of "
                                   "kdtree.py/result_set.py (the reference itself is pure Python: ~830 queries/s/core)"},
        "e2e": {"value": value, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit_line(line)


# ---------------------------------------------------------------------------------------------- main

def ncu_traffic_bytes(workload):
    """DRAM bytes per launch of the tiled kNN kernel from the committed ncu capture of that workload (MB in the file)."""
    name = {"kitti_knn8": "r2_ncu_knn_tile_kitti_k8.txt"}.get(workload)
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", name) if name else None
    if not path or not os.path.exists(path):
        return None
    total = 0.0
    for line in open(path):
        if line.startswith("dram__bytes_read.sum") or line.startswith("dram__bytes_write.sum"):
            total += float(line.split("=")[1])
    return int(total * 1e6)


_STDOUT_FD = None


def quiet_stdout():
    """Libraries below us (NCCL prints its version banner on stdout) must not add lines to the one-JSON-line contract:
    everything written to fd 1 from here on goes to stderr, the JSON line goes to the saved descriptor."""
    global _STDOUT_FD
    if _STDOUT_FD is None:
        sys.stdout.flush()
        _STDOUT_FD = os.dup(1)
        os.dup2(2, 1)


def emit_line(line):
    data = (json.dumps(line) + "\n").encode()
    if _STDOUT_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_STDOUT_FD, data)


def main():
    quiet_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="kitti_knn8")
    ap.add_argument("--icp-pairs", type=int, default=C4_PAIRS, help="pairs of the config-4 batch, split over the ranks (0 = skip)")
    ap.add_argument("--icp-points", type=int, default=C4_POINTS)
    ap.add_argument("--icp-mode", default="A", choices=["A", "B"])
    ap.add_argument("--icp-steps", type=int, default=3)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-radius", action="store_true")
    ap.add_argument("--no-c5", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=10.0)
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
        return

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    threads = os.cpu_count() or 1
    # ---- host data first (worker processes are forked before CUDA exists in this process)
    P = args.icp_pairs
    lo, hi = (rank * P) // world, ((rank + 1) * P) // world      # equal clouds: pcp_multi_pair_range gives these bounds
    h_src = h_tgt = None
    if P > 0:
        h_src, h_tgt = make_pairs(lo, hi - lo, args.icp_points, max(1, threads // world))

    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        # NCCL_DEBUG=VERSION (the box default) prints "NCCL version ..." on STDOUT, in front of the one JSON line
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    import point_cloud_process_b200 as pcp
    native = pcp.native
    peaks, peaks_kind = load_peaks()
    ctx = native.Context(local_rank)
    stream = torch.cuda.ExternalStream(ctx.stream())
    sm_count = torch.cuda.get_device_properties(local_rank).multi_processor_count

    def max_over_ranks(*vals):
        if dist is None:
            return [float(v) for v in vals]
        t = torch.tensor(list(vals), dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return [float(x) for x in t.tolist()]

    # ---- kNN block (config 1)
    w = make_workload(args.workload, rank)
    knn = measure_knn(torch, pcp, ctx, stream, w, args, dist, rank, world, local_rank, peaks, peaks_kind)
    knn_cfg = knn_config(args.workload, w["desc"])

    # ---- ICP block (config 4, pairs split over the ranks)
    icp = None
    if P > 0:
        if world > 1:
            want = native.multi_pair_range(np.arange(P + 1, dtype=np.int64) * args.icp_points,
                                           np.arange(P + 1, dtype=np.int64) * args.icp_points, rank, world)
            assert tuple(want) == (lo, hi), (want, lo, hi)
        r = measure_icp(torch, pcp, ctx, stream, h_src, h_tgt, lo, hi - lo, args.icp_points, args.icp_mode,
                        steps=args.icp_steps, warmup=1, peaks=peaks, sm_count=sm_count)
        t_dev, t_e2e = max_over_ranks(r["device_total_ms"], r["e2e_total_ms"])
        parity_ok = r.get("parity", {}).get("ok", True)
        hist = r.get("iterations_hist", [])
        if dist is not None:
            allr = [None] * world
            dist.all_gather_object(allr, {"ok": parity_ok, "hist": hist, "pairs": r["pairs"], "ms": r.get("device_ms_per_batch", 0.0)})
            parity_ok = all(a["ok"] for a in allr)
            L = max(len(a["hist"]) for a in allr)
            hist = [sum(a["hist"][i] if i < len(a["hist"]) else 0 for a in allr) for i in range(L)]
            r["per_rank"] = [{"pairs": a["pairs"], "device_ms_per_batch": a["ms"]} for a in allr]
        icp = {"metric": "ICP pairs/s", "value": P * r["steps"] / (t_dev * 1e-3), "unit": "pairs/s",
               "ms_per_step": t_dev / r["steps"], "steps": r["steps"], "scaling": "strong",
               "e2e": {"value": P * r.get("e2e_steps", 1) / (t_e2e * 1e-3), "unit": "pairs/s",
                       "h2d_bytes_per_step": int(2 * P * args.icp_points * 12), "d2h_bytes_per_step": int(P * 264),
                       "timer": "host perf_counter around the synchronous C-ABI call (PCP_HOST, pinned clouds), max over ranks"},
               "iterations_hist_all_pairs": hist, "parity_all_ranks_ok": bool(parity_ok),
               "config": icp_config(P, args.icp_points, args.icp_mode),
               "rank0": r}
        if rank == 0 and not args.no_cpu_baseline:
            n_cpu = max(8, 2 * threads)
            v, per_pair, its = cpu_icp_port(n_cpu, args.icp_points, args.icp_mode, threads)
            icp["cpu_baseline"] = {"value": v, "unit": "pairs/s", "cores": min(threads, n_cpu), "kind": "port",
                                   "single_core_pairs_per_s": 1.0 / statistics.mean(per_pair),
                                   "sample": f"{n_cpu} pairs of the batch on a pool of {min(threads, n_cpu)} processes; float64 numpy + cKDTree "
                                             f"restatement of main.py:97-156, iterations {sorted(set(its))}"}
        h_src = h_tgt = None

    # ---- radius block (config 1, second half) and config 5
    radius = None
    if rank == 0 and not args.no_radius and args.workload == "kitti_knn8":
        radius = measure_radius(torch, pcp, ctx, stream, w["pts"], 1.0, 5, peaks)
    brute = None
    if rank == 0 and world == 1 and not args.no_radius and args.workload == "kitti_knn8":
        brute = measure_bruteforce(torch, pcp, ctx, stream, w["pts"], w["k"], 3, sm_count, float(peaks.get("sm_max_mhz", 1965.0)))
    c5 = None
    if world > 1 and not args.no_c5:
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import run_c5
        c5 = run_c5.run_c5(2_000_000 * world, 8_000_000 * world, 16, 2048)     # N = 8: config 5 at full size
        if c5 is not None:
            c5["scale_note"] = f"config 5 scaled to {world}/8 of its size (16 M reference, 64 M queries at N = 8)"

    # ---- the line
    common = {"n_gpus": world, "warmup": args.warmup, "higher_is_better": True, "vs_baseline": None,
              "clocks": knn["clocks"]}
    knn_block = {"metric": "kNN queries/s", "value": knn["value"], "unit": "queries/s", "steps": args.steps,
                 "ms_per_step": knn["ms_per_step"], "scaling": "weak", "dtype": "f32", "config": knn_cfg, "e2e": knn["e2e"],
                 "roofline": knn["roofline"], "latency": knn["latency"], "breakdown_ms": knn["breakdown_ms"],
                 "graph_nodes": knn["graph_nodes"], "stream_launches_per_step": knn["stream_launches_per_step"],
                 "per_rank": knn["per_rank"]}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        reps, tb, ts, single = cpu_knn_port(w["pts"], w["k"], threads, args.cpu_seconds)
        n = knn["n"]
        knn_block["cpu_baseline"] = {"value": n / (tb + ts), "unit": "queries/s", "cores": threads, "kind": "port",
                                     "sample": f"whole scan ({n} queries), build+search, mean of {reps} repetitions "
                                               f"(build {tb*1e3:.1f} ms single-threaded, search {ts*1e3:.1f} ms on {threads} threads)",
                                     "search_only_queries_per_s": n / ts, "single_core_search_queries_per_s": single}
    series = {"kNN queries/s": "`value` at N = 1, `also.knn.value` at N > 1 (replicas)",
              "ICP pairs/s": "`also.icp.value` at N = 1, `value` at N > 1 (config 4 split over the ranks)"}
    if world == 1 or icp is None:
        line = dict(common)
        line.update({k2: v for k2, v in knn_block.items() if k2 != "per_rank"})
        line.update({"data": "KITTI scan bundled with the reference (tests/golden copy)" if args.workload.startswith("kitti") else "synthetic",
                     "gpu_launches": knn["gpu_launches"], "series": series,
                     "also": {"icp": icp, "radius": radius, "bruteforce": brute}})
    else:
        line = dict(common)
        line.update({"metric": icp["metric"], "value": icp["value"], "unit": icp["unit"], "steps": icp["steps"],
                     "ms_per_step": icp["ms_per_step"], "scaling": "strong", "dtype": "f32 search + f64 poses", "data": "synthetic",
                     "config": icp["config"], "e2e": icp["e2e"], "roofline": icp["rank0"].get("roofline"),
                     "gpu_launches": int(icp["steps"] * (icp["rank0"].get("iteration_loop_launches_per_batch", 0) + 24)),
                     "iterations_hist_all_pairs": icp["iterations_hist_all_pairs"], "parity_all_ranks_ok": icp["parity_all_ranks_ok"],
                     "series": series, "rank0": icp["rank0"], "also": {"knn": knn_block, "c5": c5, "radius": radius}})
        if "cpu_baseline" in icp:
            line["cpu_baseline"] = icp["cpu_baseline"]
    if rank == 0:
        emit_line(line)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
