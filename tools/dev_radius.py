"""Developer probe: radius count / fill timings at several result densities (PCP_RADIUS_FILL=tile|unset)."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import point_cloud_process_b200 as pcp
from tools.dev_timing import timeit, kitti
native = pcp.native


def run(name, pts, radii):
    ctx = native.Context(0)
    stream = torch.cuda.ExternalStream(ctx.stream())
    d = torch.from_numpy(pts).cuda()
    n = len(pts)
    ix = native.Index(ctx, d.data_ptr(), n, 3)
    off = torch.empty(n + 1, dtype=torch.int64, device="cuda")
    for r in radii:
        medc, _ = timeit(stream, lambda: ix.radius_count_device(d.data_ptr(), n, 3, r, off.data_ptr()), reps=5)
        ctx.synchronize()
        m = int(off[-1].item())
        ridx = torch.empty(m, dtype=torch.int32, device="cuda")
        rdist = torch.empty(m, dtype=torch.float32, device="cuda")
        medf, _ = timeit(stream, lambda: ix.radius_fill_device(d.data_ptr(), n, 3, r, off.data_ptr(), ridx.data_ptr(),
                                                               rdist.data_ptr(), False), reps=5)
        print(f"{name:9s} r={r:<4} hits/query {m/n:7.1f}  count {medc*1e3:8.1f} us  fill {medf*1e3:8.1f} us "
              f"({m*8/medf/1e6:7.1f} GB/s out)  fill={os.environ.get('PCP_RADIUS_FILL', 'default')}", flush=True)
        del ridx, rdist


if __name__ == "__main__":
    uni = (np.random.default_rng(1234).random((1_000_000, 3)) * 25.4).astype(np.float32)
    run("kitti", kitti, [0.3, 0.5, 1.0])
    run("uniform1M", uni, [0.3, 0.5, 0.8, 1.0])
