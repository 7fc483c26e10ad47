"""Developer probe: batched ICP timing (device-resident), both modes."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import point_cloud_process_b200 as pcp
import bench
native = pcp.native
ctx = native.Context(0); stream = torch.cuda.ExternalStream(ctx.stream())
P = int(sys.argv[1]) if len(sys.argv) > 1 else 256
for mode in ("A", "B"):
    r = bench.measure_icp(torch, native, ctx, stream, 0, P, 20000, mode, steps=3, warmup=1, e2e=(mode == "A"))
    print(os.environ.get("PCP_ICP_KERNEL"), mode, {k: r[k] for k in ("pairs", "device_ms_per_batch", "launches_per_batch", "iterations_hist")},
          "pairs/s %.0f" % (P / r["device_ms_per_batch"] * 1e3), "pair-iterations/s %.0f" % (sum(i * c for i, c in enumerate(r["iterations_hist"])) / r["device_ms_per_batch"] * 1e3),
          "e2e ms", r.get("e2e_ms_per_batch"))
