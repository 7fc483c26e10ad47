"""Generates tests/golden/front_golden.npz from the reference's OWN functions (run in the build container,
where /root/reference exists):

    voxel_filter(point_cloud, leaf_size, "centroid")   Pca_and_Voxel_filter/voxel_filter.py:10-66
    PCA(data)                                          Pca_and_Voxel_filter/pca_normal.py:10-36

Both modules import open3d / pyntcloud / matplotlib at the top but the two functions never use them, so empty
stand-in modules are registered before the import.  PYTHONDONTWRITEBYTECODE=1 (the reference tree is read-only).

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden_front.py
"""
import os
import sys
import types

import numpy as np

REF = "/root/reference/Pca_and_Voxel_filter"
HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    for name in ("open3d", "pyntcloud", "matplotlib", "matplotlib.pyplot"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["pyntcloud"].PyntCloud = object
    sys.path.insert(0, REF)
    import pca_normal as ref_pca
    import voxel_filter as ref_vf

    rng = np.random.default_rng(2024)
    out = {}
    # voxel filter cases: (points float32, leaf)
    kitti = np.fromfile(os.path.join(HERE, "kitti_000000_xyz.f32"), np.float32).reshape(-1, 3)
    cases = {
        "uniform": ((rng.random((3000, 3)) * np.float32([4, 3, 2])).astype(np.float32), 0.25),
        "dense": ((rng.random((4000, 3)) * np.float32([1, 1, 0.2])).astype(np.float32), 0.5),       # runs > 128 points
        "kitti_sub": (kitti[rng.choice(len(kitti), 5000, replace=False)].copy(), 2.0),
        "flat": (np.c_[rng.random((500, 2)), np.zeros(500)].astype(np.float32), 0.1),               # Dz = 0
        "one_voxel": ((rng.random((50, 3)) * 0.01).astype(np.float32), 1.0),                        # single run -> empty
    }
    for name, (pts, leaf) in cases.items():
        res = ref_vf.voxel_filter(pts.astype(np.float64), leaf, "centroid")
        out[f"vf_{name}_pts"] = pts
        out[f"vf_{name}_leaf"] = np.float64(leaf)
        out[f"vf_{name}_out"] = np.asarray(res, np.float64).reshape(-1, 3)
    # PCA cases: neighbourhoods of 5 / 12 / 30 points
    for k in (5, 12, 30):
        data = (rng.random((k, 3)) * np.float32([1, 1, 0.05]) + 10).astype(np.float32)
        w, v = ref_pca.PCA(data.astype(np.float64))
        out[f"pca_{k}_data"] = data
        out[f"pca_{k}_w"] = w
        out[f"pca_{k}_v"] = v
    # whole-cloud PCA (pca_normal.py:69): the voxel-filter clouds again
    for name in ("uniform", "kitti_sub", "flat"):
        w, v = ref_pca.PCA(cases[name][0].astype(np.float64))
        out[f"pca_cloud_{name}_w"] = w
        out[f"pca_cloud_{name}_v"] = v
    np.savez_compressed(os.path.join(HERE, "front_golden.npz"), **out)
    print({k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
