"""Device-resident timings of the BASELINE.json configs C1, C3, C4 on ONE GPU -> one JSON document.
python tools/run_configs.py [--pairs 4096]   (C5 and the sharded variants: tools/run_c5.py, bench.py --gpus N)"""
import argparse, json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import point_cloud_process_b200 as pcp
import bench
native = pcp.native


def timeit(stream, fn, reps=10, warm=2):
    for _ in range(warm):
        fn()
    stream.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for a, b in evs:
        a.record(stream); fn(); b.record(stream)
    stream.synchronize()
    return float(np.median([a.elapsed_time(b) for a, b in evs]))


def search_config(name, pts, k, r, ctx, stream):
    """Timed passes run with the work counters OFF; the evaluations come from ONE separate counted pass each."""
    n = len(pts)
    d = torch.from_numpy(pts).cuda()
    ix = native.Index(ctx, d.data_ptr(), n, 3)
    idx = torch.empty((n, k), dtype=torch.int32, device="cuda"); dist = torch.empty((n, k), dtype=torch.float32, device="cuda")
    cnt = torch.empty(n, dtype=torch.int32, device="cuda"); off = torch.empty(n + 1, dtype=torch.int64, device="cuda")
    out = {"config": name, "points": n, "k": k, "radius": r}
    out["build_ms"] = timeit(stream, lambda: ix.rebuild_device(d.data_ptr(), n, 3))
    out["knn_ms"] = timeit(stream, lambda: ix.knn_device(d.data_ptr(), n, 3, k, idx.data_ptr(), dist.data_ptr(), cnt.data_ptr()))
    out["knn_self_ms"] = timeit(stream, lambda: ix.knn_self_device(k, idx.data_ptr(), dist.data_ptr(), cnt.data_ptr()))
    ctx.enable_counters(True)
    ix.knn_device(d.data_ptr(), n, 3, k, idx.data_ptr(), dist.data_ptr(), cnt.data_ptr())
    c = ctx.read_counters()
    ctx.enable_counters(False)
    out["knn_evals_per_query"] = c[0] / n
    out["knn_fp32_frac_whole_search"] = 8.0 * c[0] / (out["knn_ms"] * 1e-3) / (148 * 128 * 1.965e9)
    out["radius_count_ms"] = timeit(stream, lambda: ix.radius_count_device(d.data_ptr(), n, 3, r, off.data_ptr()), reps=5)
    ctx.synchronize()
    m = int(off[-1].item())
    ridx = torch.empty(m, dtype=torch.int32, device="cuda"); rdist = torch.empty(m, dtype=torch.float32, device="cuda")
    out["radius_fill_ms"] = timeit(stream, lambda: ix.radius_fill_device(d.data_ptr(), n, 3, r, off.data_ptr(), ridx.data_ptr(), rdist.data_ptr(), False), reps=5)
    ctx.enable_counters(True)
    ix.radius_count_device(d.data_ptr(), n, 3, r, off.data_ptr())
    c = ctx.read_counters()
    ctx.enable_counters(False)
    out["radius_pairs"] = m
    out["radius_evals_per_query"] = c[0] / n
    out["radius_count_fp32_frac"] = 8.0 * c[0] / (out["radius_count_ms"] * 1e-3) / (148 * 128 * 1.965e9)
    out["radius_fill_out_gbs"] = m * 8 / (out["radius_fill_ms"] * 1e-3) / 1e9
    out["radius_fill_hbm_frac"] = (28 * n + 8 * (n + 1) + 8 * m) / (out["radius_fill_ms"] * 1e-3) / 1e9 / 6557.4
    out["knn_queries_per_s"] = n / (out["knn_ms"] * 1e-3)
    out["radius_queries_per_s"] = n / ((out["radius_count_ms"] + out["radius_fill_ms"]) * 1e-3)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pairs", type=int, default=4096)
    ap.add_argument("--skip-icp", action="store_true")
    a = ap.parse_args()
    ctx = native.Context(0); stream = torch.cuda.ExternalStream(ctx.stream())
    res = []
    kitti = np.fromfile(bench.KITTI, np.float32).reshape(-1, 3)
    res.append(search_config("C1 KITTI 000000.bin kNN k=8 + radius r=1.0, queries = all points", kitti, 8, 1.0, ctx, stream))
    uni = (np.random.default_rng(1234).random((1_000_000, 3)) * 25.4).astype(np.float32)
    res.append(search_config("C3 uniform 1M in [0,25.4)^3: build + radius r=0.5 + kNN k=16", uni, 16, 0.5, ctx, stream))
    if not a.skip_icp:
        for mode in ("A", "B"):
            h_src, h_tgt = bench.make_pairs(0, a.pairs, 20000, os.cpu_count() or 1) if mode == "A" else (h_src, h_tgt)
            r = bench.measure_icp(torch, pcp, ctx, stream, h_src, h_tgt, 0, a.pairs, 20000, mode, steps=2, warmup=1,
                                  peaks=bench.load_peaks()[0], sm_count=148)
            r["config"] = f"C4 {a.pairs} synthetic pairs x 20000 pts, mode {mode} {bench.ICP_MODES[mode]}"
            r["pairs_per_s"] = a.pairs / (r["device_ms_per_batch"] * 1e-3)
            r["pair_iterations_per_s"] = sum(i * c for i, c in enumerate(r["iterations_hist"])) / (r["device_ms_per_batch"] * 1e-3)
            res.append(r)
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
