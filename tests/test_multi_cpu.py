"""world_size-2 gloo tests (CPU) of the multi-GPU host logic: shard ranges, the all-gather of a replicated
reference cloud with ragged shards, and pair-sharded ICP result assembly (the per-rank solver is replaced by
the oracle so that no GPU is needed; the GPU path itself is covered by tests -m gpu and bench.py --gpus N)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from point_cloud_process_b200 import multi


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_shard_ranges_cover_everything():
    for n in (0, 1, 7, 8, 124668, 64_000_000):
        for w in (1, 2, 3, 8):
            r = [multi.shard_range(n, i, w) for i in range(w)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[i][1] == r[i + 1][0] for i in range(w - 1))
            sizes = [b - a for a, b in r]
            assert max(sizes) - min(sizes) <= 1
    b = multi.shard_pairs_by_points([10, 10, 10, 10, 1000, 10, 10], 2)
    assert b[0] == 0 and b[-1] == 7 and b == sorted(b)
    b = multi.shard_pairs_by_points([20000] * 4096, 8)
    assert b == [512 * i for i in range(9)]


def test_pair_range_matches_the_c_abi():
    """multi.shard_pairs_by_points (host numpy) and pcp_multi_pair_range (C ABI, host only) are the same rule."""
    from point_cloud_process_b200 import native
    rng = np.random.default_rng(9)
    for world in (1, 2, 3, 8):
        for P in (0, 1, 5, 64, 342):
            ns, nt = rng.integers(1, 30000, P), rng.integers(1, 30000, P)
            so, to = np.concatenate([[0], np.cumsum(ns)]).astype(np.int64), np.concatenate([[0], np.cumsum(nt)]).astype(np.int64)
            b = multi.shard_pairs_by_points(ns + nt, world)
            got = [native.multi_pair_range(so, to, r, world) for r in range(world)]
            assert got == [(b[r], b[r + 1]) for r in range(world)]


def test_product_package_does_not_import_torch():
    """north_star: no PyTorch in the product -- the multi-GPU collective lives behind the C ABI (pcp_multi_*)."""
    import glob
    root = os.path.dirname(os.path.abspath(multi.__file__))
    for f in glob.glob(os.path.join(root, "**", "*.py"), recursive=True):
        src = open(f).read()
        assert "import torch" not in src and "from torch" not in src, f


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # --- the rendezvous that carries the NCCL id of rank 0 (plain TCP, no torch)
        payload = bytes(range(128)) if rank == 0 else b""
        ok_id = multi.exchange_bytes(payload, rank, world) == bytes(range(128))
        # --- shards of a replicated reference cloud: ranges cover the cloud, ragged lengths
        rng = np.random.default_rng(5)
        full = rng.random((1001, 3)).astype(np.float32)
        lo, hi = multi.shard_range(len(full), rank, world)
        parts = [None] * world
        dist.all_gather_object(parts, full[lo:hi])
        ok_gather = bool(np.array_equal(np.concatenate(parts), full))
        # --- pair-sharded ICP with the oracle standing in for the GPU solver
        from oracle import icp_oracle
        pairs = [icp_oracle.make_pair(3000 + i, n=200 + 40 * i, box=(8.0, 8.0, 2.0)) for i in range(5)]
        srcs, tgts = [p[1] for p in pairs], [p[0] for p in pairs]

        def run(s, t, T_init=None, **kw):
            outs = [icp_oracle.icp_point2point(a, b, np.eye(4), return_info=True) for a, b in zip(s, t)]
            return {"T_last": np.asarray([o[0] for o in outs]).reshape(-1, 4, 4),
                    "iterations": np.asarray([o[1]["iterations"] for o in outs], np.int32)}
        lo, hi, local = multi.icp_batch_sharded(srcs, tgts, run=run)
        got = [None] * world
        dist.all_gather_object(got, (lo, hi, local))
        got.sort(key=lambda g: g[0])
        ok_cover = got[0][0] == 0 and got[-1][1] == len(srcs) and all(got[i][1] == got[i + 1][0] for i in range(world - 1))
        res = {k: np.concatenate([g[2][k] for g in got]) for k in ("T_last", "iterations")}
        want = run(srcs, tgts)
        ok_icp = bool(ok_cover and np.array_equal(res["T_last"], want["T_last"]) and np.array_equal(res["iterations"], want["iterations"]))
        q.put((rank, ok_id and ok_gather, ok_icp))
    finally:
        dist.destroy_process_group()


def test_gloo_world2_gather_and_pair_sharding():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    results = sorted(q.get(timeout=5) for _ in range(2))
    assert results == [(0, True, True), (1, True, True)]
