"""Developer probe: work counters of the kNN kernels (pcp_ctx_enable_counters) on the KITTI scan and the uniform 1M cloud."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.environ.get("PCP_ROOT") or os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import point_cloud_process_b200 as pcp   # noqa: E402

native = pcp.native
NAMES = ["evals", "box_tests", "opens", "rounds", "groups", "scans", "aborted", "r7"]


def run(name, pts, k, self_q):
    ctx = native.Context(0)
    stream = torch.cuda.ExternalStream(ctx.stream())
    d = torch.from_numpy(pts).cuda()
    n = len(pts)
    ix = native.Index(ctx, d.data_ptr(), n, 3)
    idx = torch.empty((n, k), dtype=torch.int32, device="cuda")
    dist = torch.empty((n, k), dtype=torch.float32, device="cuda")
    cnt = torch.empty(n, dtype=torch.int32, device="cuda")

    def go():
        if self_q:
            ix.knn_self_device(k, idx.data_ptr(), dist.data_ptr(), cnt.data_ptr())
        else:
            ix.knn_device(d.data_ptr(), n, 3, k, idx.data_ptr(), dist.data_ptr(), cnt.data_ptr())
    for _ in range(3):
        go()
    ctx.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(10)]
    for a, b in evs:
        a.record(stream); go(); b.record(stream)
    ctx.synchronize()
    ms = sorted(a.elapsed_time(b) for a, b in evs)[5]
    ctx.enable_counters(True)
    go()
    c = [int(x) for x in ctx.read_counters()]
    ctx.enable_counters(False)
    g = max(c[4], 1)
    print(f"{name} k={k} self={self_q}: {ms*1e3:.1f} us  " + " ".join(f"{nm}={v}" for nm, v in zip(NAMES, c)))
    print(f"    per group: scans {c[5]/g:.1f} tests {c[1]/g:.1f} opens {c[2]/g:.1f} rounds {c[3]/g:.1f}; evals/query {c[0]/n:.1f}; aborted groups {c[6]}")


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "both"
    kitti = np.fromfile(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "kitti_000000_xyz.f32"),
                        np.float32).reshape(-1, 3)
    if which in ("kitti", "both"):
        run("kitti", kitti, 8, True)
        run("kitti", kitti, 8, False)
    if which in ("uni", "both"):
        uni = (np.random.default_rng(1234).random((1_000_000, 3)) * 25.4).astype(np.float32)
        run("uniform1M", uni, 16, True)
        run("uniform1M", uni, 8, True)
