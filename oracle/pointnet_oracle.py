"""TEST INFRASTRUCTURE: numpy restatement of the reference's two PointNet++ search kernels,
checked on the GPU box against oracle/_ref (the reference's own sources compiled, oracle/build_oracle.py).

    ball_query   _ext-src/src/ball_query_gpu.cu:9-46   first nsample points in index order with d2 < radius^2;
                                                       the tail repeats the first hit; no hit: zeros
    three_nn     _ext-src/src/interpolate_gpu.cu:9-60  three smallest d2, strict '<' (lowest index wins)

d2 follows the compiled float32 arithmetic: fma(dz, dz, fma(dx, dx, dy * dy)); the fused steps are emulated in
float64 and rounded once to float32 (exact up to a double rounding that needs a 2^-29 coincidence).
"""
import ctypes
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def d2_matrix(a, p):
    """a (m, 3), p (n, 3) float32 -> (m, n) float32 squared distances in the kernels' arithmetic."""
    a, p = np.asarray(a, np.float32), np.asarray(p, np.float32)
    dx = (a[:, None, 0] - p[None, :, 0]).astype(np.float32)
    dy = (a[:, None, 1] - p[None, :, 1]).astype(np.float32)
    dz = (a[:, None, 2] - p[None, :, 2]).astype(np.float32)
    t = (dy * dy).astype(np.float32)
    t = (dx.astype(np.float64) * dx.astype(np.float64) + t.astype(np.float64)).astype(np.float32)
    return (dz.astype(np.float64) * dz.astype(np.float64) + t.astype(np.float64)).astype(np.float32)


def ball_query(radius, nsample, xyz, new_xyz):
    xyz, new_xyz = np.asarray(xyz, np.float32), np.asarray(new_xyz, np.float32)
    B, m = new_xyz.shape[0], new_xyz.shape[1]
    r2 = np.float32(radius) * np.float32(radius)
    idx = np.zeros((B, m, nsample), np.int32)
    for b in range(B):
        hit = d2_matrix(new_xyz[b], xyz[b]) < r2
        for j in range(m):
            ks = np.nonzero(hit[j])[0][:nsample]
            if len(ks):
                idx[b, j, :] = ks[0]
                idx[b, j, :len(ks)] = ks
    return idx


def three_nn(unknown, known):
    unknown, known = np.asarray(unknown, np.float32), np.asarray(known, np.float32)
    B, n = unknown.shape[0], unknown.shape[1]
    d2o = np.full((B, n, 3), np.inf, np.float32)
    idx = np.zeros((B, n, 3), np.int32)
    for b in range(B):
        d = d2_matrix(unknown[b], known[b])
        order = np.argsort(d, axis=1, kind="stable")[:, :3]          # stable: the lowest index wins a tie
        k = order.shape[1]
        idx[b, :, :k] = order
        d2o[b, :, :k] = np.take_along_axis(d, order, axis=1)
    return d2o, idx


def ref_lib():
    """oracle/_ref/libpointnet2_ref.so (the reference's own kernels), or None when it was not built."""
    p = os.path.join(HERE, "_ref", "libpointnet2_ref.so")
    return ctypes.CDLL(p) if os.path.exists(p) else None


def ref_ball_query(lib, radius, nsample, xyz, new_xyz):
    xyz, new_xyz = np.ascontiguousarray(xyz, np.float32), np.ascontiguousarray(new_xyz, np.float32)
    B, n, m = xyz.shape[0], xyz.shape[1], new_xyz.shape[1]
    idx = np.zeros((B, m, nsample), np.int32)
    fp, ip = ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_int)
    rc = lib.ref_ball_query(B, n, m, ctypes.c_float(radius), nsample, new_xyz.ctypes.data_as(fp), xyz.ctypes.data_as(fp),
                            idx.ctypes.data_as(ip))
    assert rc == 0, rc
    return idx


def ref_three_nn(lib, unknown, known):
    unknown, known = np.ascontiguousarray(unknown, np.float32), np.ascontiguousarray(known, np.float32)
    B, n, m = unknown.shape[0], unknown.shape[1], known.shape[1]
    d2 = np.zeros((B, n, 3), np.float32)
    idx = np.zeros((B, n, 3), np.int32)
    fp, ip = ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_int)
    rc = lib.ref_three_nn(B, n, m, unknown.ctypes.data_as(fp), known.ctypes.data_as(fp), d2.ctypes.data_as(fp), idx.ctypes.data_as(ip))
    assert rc == 0, rc
    return d2, idx
