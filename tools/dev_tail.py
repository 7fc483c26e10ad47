"""Developer probe: self-kNN time on uniform 1M (and KITTI) for the current PCP_TILE_* environment."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import point_cloud_process_b200 as pcp   # noqa: E402
from tools.dev_timing import timeit, kitti   # noqa: E402

native = pcp.native


def run(name, pts, k):
    ctx = native.Context(0)
    stream = torch.cuda.ExternalStream(ctx.stream())
    d = torch.from_numpy(pts).cuda()
    n = len(pts)
    ix = native.Index(ctx, d.data_ptr(), n, 3)
    idx = torch.empty((n, k), dtype=torch.int32, device="cuda")
    dist = torch.empty((n, k), dtype=torch.float32, device="cuda")
    cnt = torch.empty(n, dtype=torch.int32, device="cuda")
    med, best = timeit(stream, lambda: ix.knn_self_device(k, idx.data_ptr(), dist.data_ptr(), cnt.data_ptr()))
    print(f"{name:10s} k={k:<3d} median {med*1e3:8.1f} us best {best*1e3:8.1f} us", flush=True)


if __name__ == "__main__":
    rng = np.random.default_rng(1234)
    uni = (rng.random((1_000_000, 3)) * 25.4).astype(np.float32)
    print({k: v for k, v in os.environ.items() if k.startswith("PCP_TILE")})
    run("kitti", kitti, 8)
    run("uniform1M", uni, 8)
    run("uniform1M", uni, 16)
