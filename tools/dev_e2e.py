"""Developer probe: where the end-to-end (host buffer) step goes."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import point_cloud_process_b200 as pcp
native = pcp.native
pts = np.fromfile(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "kitti_000000_xyz.f32"), np.float32).reshape(-1, 3)
n, k = len(pts), 8
ctx = native.Context(0)
h = torch.from_numpy(pts).pin_memory().numpy()
ix = native.Index(ctx, h)

def t(fn, reps=50):
    for _ in range(5): fn()
    t0 = time.perf_counter()
    for _ in range(reps): fn()
    return (time.perf_counter() - t0) / reps * 1e6

def both():
    ix.rebuild(h); return ix.knn(h, k)
for _ in range(3):
    print("rebuild + knn (host pinned) %.1f us" % t(both))
print("rebuild(host pinned)      %.1f us" % t(lambda: ix.rebuild(h)))
print("knn(host pinned)          %.1f us" % t(lambda: ix.knn(h, k)))
print("knn_self (D2H only)       %.1f us" % t(lambda: ix.knn_self(k)))
d = torch.from_numpy(pts).cuda()
idx = torch.empty((n, k), dtype=torch.int32, device="cuda"); dist = torch.empty((n, k), dtype=torch.float32, device="cuda"); cnt = torch.empty(n, dtype=torch.int32, device="cuda")
def dev():
    ix.rebuild_device(d.data_ptr(), n, 3); ix.knn_device(d.data_ptr(), n, 3, k, idx.data_ptr(), dist.data_ptr(), cnt.data_ptr()); ctx.synchronize()
print("device step + sync        %.1f us" % t(dev))
hp = torch.from_numpy(pts).pin_memory(); dd = torch.empty_like(d)
def h2d(): dd.copy_(hp, non_blocking=True); torch.cuda.synchronize()
print("torch H2D 1.5 MB + sync   %.1f us" % t(h2d))
ho = torch.empty((n, k), dtype=torch.int32).pin_memory()
def d2h(): ho.copy_(idx, non_blocking=True); torch.cuda.synchronize()
print("torch D2H 4 MB + sync     %.1f us" % t(d2h))
def empty_sync(): ctx.synchronize()
print("ctx.synchronize           %.1f us" % t(empty_sync))
