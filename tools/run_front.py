"""Device-resident timings of the ICP front end on ONE GPU -> one JSON document:
voxel-grid down-sampling (pcp_voxel_filter: two host syncs inside -- bounding box, row count -- so it is timed on the
host clock around the call AND by CUDA events) and normals by k-NN PCA (pcp_normals = kNN of every point + PCA kernel).
The CPU restatement (oracle/front_oracle.py, numpy, 1 core) is timed beside it on a bounded sample.
python tools/run_front.py"""
import json, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import point_cloud_process_b200 as pcp
import bench
from oracle import front_oracle
native = pcp.native
PEAKS = json.load(open(os.path.join(os.path.dirname(bench.__file__), "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(os.path.dirname(bench.__file__), "MEASURED_PEAKS.json")) else {}


def hbm_peak():
    for k in ("hbm_gbs", "hbm_copy_gbs", "hbm_gbs_burst"):
        if k in PEAKS:
            return float(PEAKS[k]), k
    return 6533.5, "fallback"


def wall(ctx, fn, reps=20, warm=3):
    for _ in range(warm):
        fn()
    ctx.synchronize()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter(); fn(); ctx.synchronize(); ts.append((time.perf_counter() - t0) * 1e3)
    return float(np.median(ts))


def main():
    ctx = native.Context(0)
    res = []
    peak, peak_src = hbm_peak()
    kitti = np.fromfile(bench.KITTI, np.float32).reshape(-1, 3)
    uni = (np.random.default_rng(1234).random((1_000_000, 3)) * 25.4).astype(np.float32)
    big = (np.random.default_rng(99).random((16_000_000, 3)) * 64.0).astype(np.float32)
    for name, pts, leaf in (("KITTI 000000.bin", kitti, 0.5), ("uniform 1M in [0,25.4)^3", uni, 0.4), ("uniform 16M in [0,64)^3", big, 0.4)):
        n = len(pts)
        d = torch.from_numpy(pts).cuda()
        m = native.voxel_filter_device(ctx, d.data_ptr(), n, 3, leaf, 0, 0)
        ms = wall(ctx, lambda: native.voxel_filter_device(ctx, d.data_ptr(), n, 3, leaf, 0, 0))
        # algorithmic bytes: bbox 12 + keys 12+8 + sort passes (key+value both ways, 2-3 passes of 16 B) + heads/gather 8+12+4+16
        # + scan 8 + starts 8+8/voxel + reduce 16 + 24/voxel
        bits = max(1, int(np.ceil(np.log2(max(2.0, np.prod(np.ceil(np.ptp(pts, axis=0) / leaf) + 1))))))
        passes = (bits + 7) // 8
        bytes_ = n * (12 + 20 + passes * 16 + 40 + 8 + 8 + 16) + m * 32
        r = {"op": "voxel_filter centroid", "cloud": name, "points": n, "leaf": leaf, "rows_out": int(m), "ms": ms,
             "points_per_s": n / (ms * 1e-3), "sort_passes": passes, "algorithmic_bytes": int(bytes_),
             "achieved_gbs": bytes_ / (ms * 1e-3) / 1e9, "hbm_peak_gbs": peak, "hbm_peak_source": peak_src,
             "timer": "host perf_counter around the call + synchronize (the call syncs twice itself)"}
        r["hbm_frac"] = r["achieved_gbs"] / peak
        if n <= 1_000_000:
            t0 = time.perf_counter(); ref = front_oracle.voxel_filter(pts, leaf); cpu = time.perf_counter() - t0
            out = torch.empty((m, 3), dtype=torch.float64, device="cuda")
            native.voxel_filter_read_device(ctx, out.data_ptr()); ctx.synchronize()
            r["cpu_restatement_ms_1core"] = cpu * 1e3
            r["bit_exact_vs_restatement"] = bool(np.array_equal(out.cpu().numpy(), ref.reshape(-1, 3)))
        res.append(r)
        ms = wall(ctx, lambda: native.pca_device(ctx, d.data_ptr(), n, 3))
        res.append({"op": "whole-cloud PCA (pcp_pca)", "cloud": name, "points": n, "ms": ms, "algorithmic_bytes": 24 * n,
                    "achieved_gbs": 24 * n / (ms * 1e-3) / 1e9, "hbm_frac": 24 * n / (ms * 1e-3) / 1e9 / peak,
                    "timer": "host perf_counter around the call (4 launches + the 96-byte result read back)"})
        del d
    for name, pts, k in (("KITTI 000000.bin", kitti, 5), ("uniform 1M in [0,25.4)^3", uni, 5), ("uniform 1M in [0,25.4)^3", uni, 16)):
        n = len(pts)
        d = torch.from_numpy(pts).cuda()
        ix = native.Index(ctx, d.data_ptr(), n, 3)
        nrm = torch.empty((n, 3), dtype=torch.float64, device="cuda"); ev = torch.empty((n, 3), dtype=torch.float64, device="cuda")
        ms = wall(ctx, lambda: ix.normals_device(k, nrm.data_ptr(), ev.data_ptr()))
        idx = torch.empty((n, k), dtype=torch.int32, device="cuda"); dist = torch.empty((n, k), dtype=torch.float32, device="cuda"); cnt = torch.empty(n, dtype=torch.int32, device="cuda")
        ms_knn = wall(ctx, lambda: ix.knn_self_device(k, idx.data_ptr(), dist.data_ptr(), cnt.data_ptr()))
        sel = np.random.default_rng(0).choice(n, 2000, replace=False)
        t0 = time.perf_counter(); front_oracle.normals(pts, idx[torch.from_numpy(sel).cuda()].cpu().numpy()); cpu = time.perf_counter() - t0
        res.append({"op": "normals (kNN + PCA)", "cloud": name, "points": n, "k": k, "ms": ms, "knn_self_ms": ms_knn,
                    "pca_kernel_ms": ms - ms_knn, "points_per_s": n / (ms * 1e-3),
                    "pca_kernel_bytes": int(n * (4 * k + 12 * k + 48)), "pca_kernel_gbs": n * (16 * k + 48) / max(1e-9, (ms - ms_knn) * 1e-3) / 1e9,
                    "cpu_restatement_points_per_s_1core": 2000 / cpu, "cpu_sample": "PCA of 2000 sampled neighbourhoods (numpy), kNN not included"})
        del ix, d
    for name, pts, r in (("KITTI 000000.bin", kitti, 0.5), ("KITTI 000000.bin", kitti, 1.0), ("uniform 1M in [0,25.4)^3", uni, 0.5)):
        n = len(pts)
        d = torch.from_numpy(pts).cuda()
        ix = native.Index(ctx, d.data_ptr(), n, 3)
        kp, l3 = ix.iss_keypoints(r, 0.5, 0.5, r, 20)
        ms = wall(ctx, lambda: ix.iss_keypoints(r, 0.5, 0.5, r, 20), reps=5, warm=1)
        off = torch.empty(n + 1, dtype=torch.int64, device="cuda")
        ix.radius_count_device(d.data_ptr(), n, 3, r, off.data_ptr()); ctx.synchronize()
        m = int(off[-1].item())
        res.append({"op": "ISS keypoints (radius count + fill + saliency + NMS)", "cloud": name, "points": n, "radius": r,
                    "pairs": m, "candidates": int(np.count_nonzero(~np.isnan(l3))), "keypoints": len(kp), "ms": ms,
                    "points_per_s": n / (ms * 1e-3), "pairs_per_s": m / (ms * 1e-3),
                    "timer": "host perf_counter around the whole call incl. the lambda3 read back (n x 8 B)"})
        del ix, d
    for name, pts, r, mp in (("KITTI 000000.bin", kitti, 0.5, 10), ("KITTI 000000.bin", kitti, 1.0, 10), ("uniform 1M in [0,25.4)^3", uni, 0.5, 30)):
        n = len(pts)
        d = torch.from_numpy(pts).cuda()
        ix = native.Index(ctx, d.data_ptr(), n, 3)
        lab = torch.empty(n, dtype=torch.int32, device="cuda")
        l0 = ctx.launch_count()
        nc = ix.dbscan_device(r, mp, lab.data_ptr())
        launches = ctx.launch_count() - l0
        ms = wall(ctx, lambda: ix.dbscan_device(r, mp, lab.data_ptr()), reps=5, warm=1)
        res.append({"op": "DBSCAN (radius count + fill + components + seed order + labels)", "cloud": name, "points": n,
                    "radius": r, "min_pts": mp, "clusters": int(nc), "noise": int((lab < 0).sum().item()), "ms": ms,
                    "launches": int(launches), "points_per_s": n / (ms * 1e-3),
                    "timer": "host perf_counter around the whole call, labels left on the device"})
        del ix, d
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
