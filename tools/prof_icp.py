"""Profiling target: one batched ICP call (mode A) on P synthetic pairs, device-resident."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import point_cloud_process_b200 as pcp
import bench
native = pcp.native
P = int(sys.argv[1]) if len(sys.argv) > 1 else 512
mode = sys.argv[2] if len(sys.argv) > 2 else "A"
ctx = native.Context(0); stream = torch.cuda.ExternalStream(ctx.stream())
r = bench.measure_icp(torch, native, ctx, stream, 0, P, 20000, mode, steps=2, warmup=1, e2e=False)
print({k: r[k] for k in ("pairs", "device_ms_per_batch", "launches_per_batch", "iterations_hist")})
