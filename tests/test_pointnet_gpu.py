"""GPU parity of the PointNet++ neighbourhood layouts: pcp_ball_query / pcp_three_nn against
oracle/_ref -- the reference's OWN ball_query_gpu.cu / interpolate_gpu.cu compiled by oracle/build_oracle.py --
and against the numpy restatement.  Index-exact, including the "first nsample hits in index order" rule, rows
without a hit, ties and fewer than three known points."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pn():
    from point_cloud_process_b200 import pointnet2_utils
    return pointnet2_utils


@pytest.fixture(scope="module")
def ref():
    from oracle import pointnet_oracle
    lib = pointnet_oracle.ref_lib()
    if lib is None:
        pytest.skip("oracle/_ref/libpointnet2_ref.so was not built (needs /root/reference at build time)")
    return lib


def _cloud(rng, B, n, scale=1.0, dup=True):
    xyz = (rng.random((B, n, 3)) * scale).astype(np.float32)
    if dup and n > 40:
        xyz[:, 10:30] = xyz[:, 5:6]           # exact duplicates: ties in three_nn, bursts of hits in ball_query
    return xyz


@pytest.mark.parametrize("B,n,m,radius,nsample", [(1, 1, 1, 0.5, 4), (2, 100, 37, 0.2, 16), (3, 1500, 512, 0.1, 32),
                                                  (2, 5000, 1024, 0.05, 64), (1, 3000, 300, 1e-4, 8), (2, 2048, 600, 3.0, 5)])
def test_ball_query_vs_reference_kernel(pn, ref, B, n, m, radius, nsample):
    from oracle import pointnet_oracle
    rng = np.random.default_rng(B * 1000 + n + m)
    xyz = _cloud(rng, B, n)
    new_xyz = np.concatenate([xyz[:, rng.choice(n, min(n, m // 2 + 1))], _cloud(rng, B, m, 1.2, dup=False)], axis=1)[:, :m]
    got = pn.ball_query(radius, nsample, xyz, new_xyz)
    want = pointnet_oracle.ref_ball_query(ref, radius, nsample, xyz, new_xyz)
    assert got.shape == (B, m, nsample) and got.dtype == np.int32
    assert np.array_equal(got, want)
    if n * m <= 200_000:
        assert np.array_equal(pointnet_oracle.ball_query(radius, nsample, xyz, new_xyz), want)


@pytest.mark.parametrize("B,n,m", [(1, 5, 1), (2, 7, 2), (2, 300, 3), (3, 2000, 500), (2, 4096, 1024), (1, 1000, 5000)])
def test_three_nn_vs_reference_kernel(pn, ref, B, n, m):
    from oracle import pointnet_oracle
    rng = np.random.default_rng(B * 77 + n + m)
    known = _cloud(rng, B, m)
    unknown = np.concatenate([known[:, rng.choice(m, min(m, n // 3 + 1))], _cloud(rng, B, n, 1.1, dup=False)], axis=1)[:, :n]
    dist, idx = pn.three_nn(unknown, known)
    d2_ref, idx_ref = pointnet_oracle.ref_three_nn(ref, unknown, known)
    assert np.array_equal(idx, idx_ref)
    assert np.array_equal(dist, np.sqrt(d2_ref))                     # pointnet2_utils.py:124-125 returns sqrt(dist2)
    if n * m <= 200_000:
        d2_o, idx_o = pointnet_oracle.three_nn(unknown, known)
        assert np.array_equal(idx_o, idx_ref) and np.array_equal(d2_o, d2_ref)


def test_layout_errors(pn):
    with pytest.raises(ValueError):
        pn.ball_query(0.1, 4, np.zeros((2, 10, 2), np.float32), np.zeros((2, 3, 3), np.float32))
    with pytest.raises(ValueError):
        pn.three_nn(np.zeros((2, 10, 3), np.float32), np.zeros((3, 4, 3), np.float32))
    idx = pn.ball_query(0.1, 4, np.zeros((1, 0, 3), np.float32), np.zeros((1, 5, 3), np.float32))
    assert idx.shape == (1, 5, 4) and (idx == 0).all()
