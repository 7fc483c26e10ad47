"""Generates tests/golden/dbscan_golden.npz from the reference's OWN class (Cluster_dbscan/dbscan.py:4-41, imported
unmodified -- it only needs scipy and numpy) on seeded float32-exact clouds.

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden_dbscan.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def cloud(seed, n, dim3=True):
    """Gaussian blobs + a ring + uniform background noise, shuffled; float32 values."""
    rng = np.random.default_rng(seed)
    k = 4
    cen = rng.random((k, 3)) * np.array([6, 6, 1.0 if dim3 else 0.0])
    nb = n // 2
    blobs = cen[rng.integers(0, k, nb)] + rng.normal(0, 0.3, (nb, 3)) * np.array([1, 1, 0.3 if dim3 else 0.0])
    nr = n // 4
    a = rng.random(nr) * 2 * np.pi
    ring = np.c_[3 + 2.5 * np.cos(a), 3 + 2.5 * np.sin(a), np.zeros(nr)] + rng.normal(0, 0.05, (nr, 3)) * np.array([1, 1, float(dim3)])
    noise = rng.random((n - nb - nr, 3)) * np.array([6, 6, 1.0 if dim3 else 0.0])
    pts = np.concatenate([blobs, ring, noise])
    return pts[rng.permutation(len(pts))].astype(np.float32)


CASES = (("a", 11, 1500, 0.25, 10), ("b", 12, 3000, 0.2, 4), ("c", 13, 2500, 0.3, 3), ("d", 14, 2000, 0.15, 6), ("flat", 15, 2000, 0.2, 5))


def main():
    sys.path.insert(0, "/root/reference/Cluster_dbscan")
    from dbscan import DBSCAN
    out = {}
    for name, seed, n, r, mp in CASES:
        pts = cloud(seed, n, dim3=name != "flat")
        d = DBSCAN(radius=r, Min_Pts=mp)
        d.fit(pts.astype(np.float64))
        out[f"{name}_pts"] = pts
        out[f"{name}_params"] = np.array([r, mp], np.float64)
        out[f"{name}_labels"] = d.predict()
        print(name, n, r, mp, "clusters", out[f"{name}_labels"].max() + 1, "noise", int((out[f"{name}_labels"] < 0).sum()))
    np.savez_compressed(os.path.join(HERE, "dbscan_golden.npz"), **out)


if __name__ == "__main__":
    main()
