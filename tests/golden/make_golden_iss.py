"""Generates tests/golden/iss_golden.npz by EXECUTING the reference's own script Keypoint_detection_ISS/ISS.py
(its whole algorithm lives under `if __name__ == "__main__"`, ISS.py:17-86) on a seeded synthetic cloud written
where the script expects its ModelNet40 file (a relative path, ISS.py:30,35), from a scratch working directory,
with a MagicMock standing in for the absent open3d (only used for the visualisation after the result, :79-86).
The script's variables (iss_idx, cand_idx, lambda3_idx, hyper-parameters) are read from the globals runpy returns.

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden_iss.py
"""
import contextlib
import io
import os
import runpy
import sys
import tempfile
from unittest import mock

import numpy as np

REF = "/root/reference/Keypoint_detection_ISS/ISS.py"
HERE = os.path.dirname(os.path.abspath(__file__))


def cloud(seed, n):
    """A box surface with a few bumps: corners / edges give distinct eigenvalue ratios (float32-exact values)."""
    rng = np.random.default_rng(seed)
    face = rng.integers(0, 6, n)
    uv = rng.random((n, 2)) * 2 - 1
    pts = np.empty((n, 3))
    for a in range(3):
        sel = (face // 2) == a
        o = [i for i in range(3) if i != a]
        pts[sel, a] = np.where(face[sel] % 2 == 0, -1.0, 1.0) * (0.6 + 0.2 * a)
        pts[sel, o[0]] = uv[sel, 0] * (0.6 + 0.2 * o[0])
        pts[sel, o[1]] = uv[sel, 1] * (0.6 + 0.2 * o[1])
    pts += rng.normal(0, 0.01, pts.shape)
    return pts.astype(np.float32)


def main():
    sys.modules["open3d"] = mock.MagicMock()
    out = {}
    for name, seed, n in (("box1500", 5, 1500), ("box3000", 6, 3000)):
        pts = cloud(seed, n)
        with tempfile.TemporaryDirectory() as tmp:
            os.makedirs(os.path.join(tmp, "modelnet40_normal_resampled", "chair"))
            rows = np.c_[pts.astype(np.float64), np.zeros((n, 3))]
            with open(os.path.join(tmp, "modelnet40_normal_resampled", "chair", "chair_0001.txt"), "w") as f:
                for r in rows:
                    f.write(",".join(repr(float(x)) for x in r) + "\n")
            cwd = os.getcwd()
            os.chdir(tmp)
            try:
                with contextlib.redirect_stdout(io.StringIO()), contextlib.redirect_stderr(io.StringIO()):
                    g = runpy.run_path(REF, run_name="__main__")
            finally:
                os.chdir(cwd)
        assert np.array_equal(g["points"][:, :3], pts.astype(np.float64))
        out[f"{name}_pts"] = pts
        out[f"{name}_iss_idx"] = np.array(g["iss_idx"], np.int64)
        out[f"{name}_cand_idx"] = np.array(g["cand_idx"], np.int64)
        out[f"{name}_params"] = np.array([g["radius"], g["lambda21"], g["lambda32"], g["non_max_radius"], g["iss_count"]], np.float64)
        print(name, len(g["cand_idx"]), "candidates ->", g["iss_idx"])
    np.savez_compressed(os.path.join(HERE, "iss_golden.npz"), **out)


if __name__ == "__main__":
    main()
