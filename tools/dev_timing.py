"""Developer timing probe (not the bench): device-resident kNN / radius / build timings on the KITTI scan."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.environ.get("PCP_ROOT") or os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import point_cloud_process_b200 as pcp   # noqa: E402

native = pcp.native
kitti = np.fromfile(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "kitti_000000_xyz.f32"),
                    np.float32).reshape(-1, 3)


def timeit(stream, fn, reps=20, warm=3):
    for _ in range(warm):
        fn()
    stream.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for a, b in evs:
        a.record(stream)
        fn()
        b.record(stream)
    stream.synchronize()
    ts = sorted(a.elapsed_time(b) for a, b in evs)
    return ts[len(ts) // 2], ts[0]


def run(name, pts, k, r):
    ctx = native.Context(0)
    stream = torch.cuda.ExternalStream(ctx.stream())
    d = torch.from_numpy(pts).cuda()
    n = len(pts)
    ix = native.Index(ctx, d.data_ptr(), n, 3)
    ctx.synchronize()
    print(name, "n", n, "buckets/levels", ix.stats())
    med, best = timeit(stream, lambda: ix.rebuild_device(d.data_ptr(), n, 3))
    print(f"  build           median {med*1e3:9.1f} us  best {best*1e3:9.1f} us")
    idx = torch.empty((n, k), dtype=torch.int32, device="cuda")
    dist = torch.empty((n, k), dtype=torch.float32, device="cuda")
    cnt = torch.empty(n, dtype=torch.int32, device="cuda")
    med, best = timeit(stream, lambda: ix.knn_self_device(k, idx.data_ptr(), dist.data_ptr(), cnt.data_ptr()))
    print(f"  knn_self k={k:<3d}  median {med*1e3:9.1f} us  best {best*1e3:9.1f} us  -> {n/med*1e3/1e6:8.1f} Mq/s")
    med, best = timeit(stream, lambda: ix.knn_device(d.data_ptr(), n, 3, k, idx.data_ptr(), dist.data_ptr(), cnt.data_ptr()))
    print(f"  knn      k={k:<3d}  median {med*1e3:9.1f} us  best {best*1e3:9.1f} us  -> {n/med*1e3/1e6:8.1f} Mq/s")
    off = torch.empty(n + 1, dtype=torch.int64, device="cuda")
    med, best = timeit(stream, lambda: ix.radius_count_device(d.data_ptr(), n, 3, r, off.data_ptr()), reps=5)
    ctx.synchronize()
    m = int(off[-1].item())
    print(f"  radius_count r={r} median {med*1e3:9.1f} us  best {best*1e3:9.1f} us  pairs {m}")
    ridx = torch.empty(m, dtype=torch.int32, device="cuda")
    rdist = torch.empty(m, dtype=torch.float32, device="cuda")
    med, best = timeit(stream, lambda: ix.radius_fill_device(d.data_ptr(), n, 3, r, off.data_ptr(), ridx.data_ptr(),
                                                               rdist.data_ptr(), False), reps=5)
    print(f"  radius_fill     median {med*1e3:9.1f} us  best {best*1e3:9.1f} us  -> {m*8/med/1e6:8.1f} GB/s out")
    # host API end to end
    six = pcp.SearchIndex(pts)
    six.knn(pts, k)
    t0 = time.perf_counter()
    for _ in range(5):
        six.knn(pts, k)
    t = (time.perf_counter() - t0) / 5
    print(f"  host knn e2e    {t*1e3:9.3f} ms -> {n/t/1e6:8.2f} Mq/s")


if __name__ == "__main__":
    run("kitti", kitti, 8, 1.0)
    rng = np.random.default_rng(1234)
    run("uniform1M", (rng.random((1_000_000, 3)) * 25.4).astype(np.float32), 16, 0.5)
