"""Generate golden ICP vectors by executing the REFERENCE's own icp_point2point.

Run in the build container only (needs /root/reference):

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden_icp.py

Registration/main.py imports open3d, which is not installable here.  The ICP loop touches exactly
three Open3D calls (main.py:105,110,117), so a minimal stand-in module is registered as `open3d`
before importing the reference file unmodified:
    geometry.PointCloud      .points, .transform(T) in place (returns self)
    utility.Vector3dVector   float64 (N,3) array
    geometry.KDTreeFlann     .search_knn_vector_3d(q, 1) -> (1, [argmin], [min squared distance])
Mode A runs icp_point2point exactly as shipped.  Mode B re-executes the same function source with
only its three constants edited (max_iteration 100->50, thresholds 0.5->1e-6) so that the
multi-iteration trajectory is pinned too.  Output: tests/golden/icp_golden.npz (inputs + outputs).
"""
import inspect
import os
import sys
import types

import numpy as np
from scipy.spatial import cKDTree

HERE = os.path.dirname(os.path.abspath(__file__))
sys.dont_write_bytecode = True
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle.icp_oracle import make_pair   # noqa: E402  (input generator only)


class _PointCloud:
    def __init__(self):
        self.points = np.zeros((0, 3))

    def transform(self, T):
        P = np.asarray(self.points)
        P[:] = P @ T[:3, :3].T + T[:3, 3]
        return self


class _KDTreeFlann:
    def __init__(self, pcd):
        self._t = cKDTree(np.asarray(pcd.points))

    def search_knn_vector_3d(self, q, k):
        d, j = self._t.query(np.asarray(q, np.float64).reshape(3), k=k)
        d, j = np.atleast_1d(d), np.atleast_1d(j)
        return k, [int(x) for x in j], [float(x * x) for x in d]


def install_stub():
    o3d = types.ModuleType("open3d")
    o3d.geometry = types.SimpleNamespace(PointCloud=_PointCloud, KDTreeFlann=_KDTreeFlann)
    o3d.utility = types.SimpleNamespace(Vector3dVector=lambda a: np.array(a, dtype=np.float64))
    sys.modules["open3d"] = o3d
    return o3d


def cloud(o3d, pts):
    pc = o3d.geometry.PointCloud()
    pc.points = o3d.utility.Vector3dVector(pts)
    return pc


def main():
    o3d = install_stub()
    sys.path.insert(0, "/root/reference/Registration")
    import main as ref_main                              # the reference file, unmodified
    src_text = inspect.getsource(ref_main.icp_point2point)
    edited = (src_text.replace("max_iteration = 100", "max_iteration = 50")
              .replace("R_diff_thres = 0.5", "R_diff_thres = 1e-6")
              .replace("t_diff_thres = 0.5", "t_diff_thres = 1e-6"))
    assert edited != src_text
    ns = dict(vars(ref_main))
    exec(edited, ns)                                      # Mode B variant of the same source
    icp_mode_b = ns["icp_point2point"]

    out = {}
    n_pts = 1500        # the reference builds a dense M x M centring matrix (main.py:134): keep M small
    specs = [
        # name, seed, n, box, rot, t, init
        ("a0", 1000, n_pts, (20.0, 20.0, 3.0), (2, 2, 5), 0.5, "identity"),
        ("a1", 1001, n_pts, (20.0, 20.0, 3.0), (2, 2, 5), 0.5, "identity"),
        ("a2", 1002, n_pts, (20.0, 20.0, 3.0), (2, 2, 5), 0.5, "identity"),
        ("a3", 1003, n_pts, (20.0, 20.0, 3.0), (2, 2, 5), 0.5, "identity"),
        ("a4", 1004, 800, (12.0, 12.0, 3.0), (1, 1, 3), 0.3, "perturbed"),
        ("a5", 1005, 800, (12.0, 12.0, 3.0), (1, 1, 3), 0.3, "perturbed"),
    ]
    for name, seed, n, box, rot, tmax, init in specs:
        tgt, src, T_gt = make_pair(seed, n=n, box=box, max_rot_deg=rot, max_t=tmax)
        T0 = np.eye(4)
        if init == "perturbed":
            T0 = T_gt.copy()
            T0[:3, 3] += np.random.default_rng(seed + 7).uniform(-0.1, 0.1, 3)
        out[f"{name}/target"] = tgt
        out[f"{name}/source"] = src
        out[f"{name}/T_gt"] = T_gt
        out[f"{name}/T_init"] = T0
        for mode, fn in (("A", ref_main.icp_point2point), ("B", icp_mode_b)):
            s, t = cloud(o3d, src), cloud(o3d, tgt)
            T = fn(s, t, T0.copy())
            out[f"{name}/{mode}_T_last"] = np.asarray(T, np.float64)
            # the reference transforms `source` in place (main.py:110): its final points pin the
            # composed pose and the iteration count
            out[f"{name}/{mode}_source_final"] = np.asarray(s.points, np.float64)
            print(name, mode, "t =", np.round(T[:3, 3], 4))
    # too-few-associations case (main.py:125-127): clouds 100 m apart
    tgt, src, _ = make_pair(1100, n=200, box=(5.0, 5.0, 2.0))
    src = src + np.float32(100.0)
    s, t = cloud(o3d, src), cloud(o3d, tgt)
    T = ref_main.icp_point2point(s, t, np.eye(4))
    out["fail/target"], out["fail/source"], out["fail/T_init"] = tgt, src, np.eye(4)
    out["fail/A_T_last"] = np.asarray(T, np.float64)
    out["fail/A_source_final"] = np.asarray(s.points, np.float64)

    # quaternion / writer helpers (main.py:158-180, 220-222)
    rng = np.random.default_rng(5)
    mats = []
    for _ in range(16):
        q, _r = np.linalg.qr(rng.normal(size=(3, 3)))
        if np.linalg.det(q) < 0:
            q[:, 0] *= -1
        M = np.eye(4)
        M[:3, :3] = q
        M[:3, 3] = rng.normal(size=3) * 3
        mats.append(M)
    mats = np.asarray(mats)
    out["quat/homo"] = mats
    out["quat/tq"] = np.asarray([ref_main.homo2tq(M) for M in mats], np.float64)
    np.savez_compressed(os.path.join(HERE, "icp_golden.npz"), **out)
    print("wrote icp_golden.npz")


if __name__ == "__main__":
    main()
