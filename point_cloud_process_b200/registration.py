"""Driver of BASELINE config 2: the `__main__` block of Registration/main.py:183-222 as a function.

    for every row (trg, src) of the pair list (registration_dataset/reg_result.txt, main.py:186-192):
        load point_clouds/<src>.bin and point_clouds/<trg>.bin          main.py:193-194, read_bin_velodyne :10-17
        [global initialisation: Open3D FPFH + RANSAC, main.py:196-209 -- an INPUT here: `inits`, identity by default]
        icp_point2point(source, target, init)                            main.py:211
        homo2tq -> [trg, src, tx, ty, tz, qw, qx, qy, qz]                main.py:212-218
    np.savetxt(reg_result.txt, fmt='%i,%i,%f,...', header=...)           main.py:220-222

All pairs go to the GPU in ONE pcp_icp_batch call (icp.register_pairs).  The Oxford clouds are not shipped with
the reference (Registration/.gitignore:1); `write_synthetic_dataset` writes seeded stand-ins in the same
6 x float32 layout, one pair of files per ROW (row j: seed 7000 + j).
"""
import os

import numpy as np

from . import icp


def cloud_path(cloud_dir, idx, row=None, role=None):
    """point_clouds/<idx>.bin (main.py:193-194); per-row stand-ins are <row>_<role>_<idx>.bin."""
    if row is not None:
        p = os.path.join(cloud_dir, "%03d_%s_%d.bin" % (row, role, idx))
        if os.path.exists(p):
            return p
    return os.path.join(cloud_dir, "%d.bin" % idx)


def write_synthetic_dataset(pair_ids, cloud_dir, make_pair, n_points=20000):
    """Row j -> make_pair(7000 + j, n_points) = (target, source, ...), stored as x,y,z,0,0,1 float32 rows."""
    os.makedirs(cloud_dir, exist_ok=True)
    for j, (trg, src) in enumerate(pair_ids):
        out = make_pair(7000 + j, n_points)
        for role, idx, pts in (("trg", trg, out[0]), ("src", src, out[1])):
            rows = np.zeros((len(pts), 6), np.float32)
            rows[:, :3] = pts
            rows[:, 5] = 1.0
            rows.tofile(os.path.join(cloud_dir, "%03d_%s_%d.bin" % (j, role, idx)))


def run(pair_list_path, cloud_dir, out_path, inits=None, init="identity", voxel_size=2.0, seed=0, **icp_kw):
    """-> (rows (P, 9), raw icp_batch output).  Writes out_path in the reference's reg_result.txt format.
    init = "ransac": the global initialisation of main.py:196-209 (voxel down-sample, normals, FPFH, feature-matching
    RANSAC; global_registration.py) is computed per pair on the GPU -- raw clouds in, reg_result.txt out, no
    caller-supplied transformation; "identity" (or explicit `inits`): ICP alone."""
    pair_ids = icp.read_pair_list(pair_list_path)
    sources = [icp.read_bin_velodyne(cloud_path(cloud_dir, s, j, "src")) for j, (t, s) in enumerate(pair_ids)]
    targets = [icp.read_bin_velodyne(cloud_path(cloud_dir, t, j, "trg")) for j, (t, s) in enumerate(pair_ids)]
    if inits is None and init == "ransac":
        from . import global_registration as gr
        inits = np.zeros((len(pair_ids), 4, 4))
        for j, (src, tgt) in enumerate(zip(sources, targets)):
            sd, sf = gr.preprocess_point_cloud(src, voxel_size)                       # main.py:63-64
            td, tf = gr.preprocess_point_cloud(tgt, voxel_size)
            inits[j] = gr.execute_global_registration(sd, td, sf, tf, voxel_size, seed=seed + j).transformation   # main.py:205-209
    elif inits is None:
        inits = np.tile(np.eye(4), (len(pair_ids), 1, 1))
    return icp.register_pairs(sources, targets, inits, pair_ids, out_path=out_path, **icp_kw)
