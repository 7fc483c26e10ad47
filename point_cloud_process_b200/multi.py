"""Multi-GPU plumbing: one process per GPU, work sharded where the path shards.

    registration pairs       independent -> contiguous ranges of pairs per rank (balanced by points), NO
                             collective on the data path; every rank keeps the rows of its own range
    queries of a large cloud independent given the whole reference cloud -> queries split per rank, the
                             reference cloud REPLICATED: each rank holds a shard, ONE ncclAllGather (NVLink)
                             rebuilds the full cloud on every GPU, each rank builds its own index (the build
                             is deterministic) and searches its own queries

Everything goes through the C ABI (pcp_multi_* of include/pcp.h: NCCL is bound inside libpcp_b200.so); there is
no torch in here.  The 128-byte NCCL id of rank 0 reaches the other ranks over a plain TCP socket
(`exchange_unique_id`, MASTER_ADDR / MASTER_PORT + 1 of the launcher by default) or by any means of the caller.
"""
import os
import socket
import struct
import time

import numpy as np


def shard_range(n, rank, world):
    """Balanced contiguous range [lo, hi) of n items for `rank` of `world`."""
    base, rem = divmod(int(n), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_pairs_by_points(sizes, world):
    """Contiguous pair ranges balanced by the number of points: sizes[i] = points of pair i.
    Returns world+1 boundaries (the rule of pcp_multi_pair_range)."""
    sizes = np.asarray(sizes, np.int64)
    total = int(sizes.sum())
    cum = np.concatenate([[0], np.cumsum(sizes)])
    bounds = [0]
    for r in range(1, world):
        target = total * r / world
        bounds.append(max(int(np.searchsorted(cum, target, side="left")), bounds[-1]))
    bounds.append(len(sizes))
    return bounds


def env_rank_world():
    """(rank, world, local_rank) from the launcher's environment (torchrun / mpirun style), 0/1/0 when absent."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


def _recv_exact(conn, n):
    buf = b""
    while len(buf) < n:
        part = conn.recv(n - len(buf))
        if not part:
            raise ConnectionError("peer closed the rendezvous socket")
        buf += part
    return buf


def exchange_bytes(payload, rank, world, addr=None, port=None, timeout=120.0):
    """Rank 0 hands `payload` (bytes) to every other rank over TCP; returns the payload on every rank.
    A one-shot star rendezvous: rank 0 listens on (addr, port), every other rank connects and reads."""
    if world == 1:
        return payload
    addr = addr or os.environ.get("MASTER_ADDR", "127.0.0.1")
    port = int(port or int(os.environ.get("MASTER_PORT", "29500")) + 1)
    if rank == 0:
        with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as srv:
            srv.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
            srv.bind((addr, port))
            srv.listen(world)
            srv.settimeout(timeout)
            for _ in range(world - 1):
                conn, _ = srv.accept()
                with conn:
                    conn.sendall(struct.pack("<I", len(payload)) + payload)
        return payload
    deadline = time.monotonic() + timeout
    while True:
        try:
            with socket.create_connection((addr, port), timeout=5.0) as conn:
                n = struct.unpack("<I", _recv_exact(conn, 4))[0]
                return _recv_exact(conn, n)
        except (ConnectionRefusedError, socket.timeout, OSError):
            if time.monotonic() > deadline:
                raise
            time.sleep(0.05)


def prefer_bundled_nccl():
    """Point the library's run-time NCCL binding (PCP_NCCL_LIB) at the NCCL wheel installed beside torch, when there
    is one and nothing was chosen: a process that later imports torch then finds ITS libnccl.so.2 already loaded
    instead of an older system copy under the same soname.  No torch import here."""
    if os.environ.get("PCP_NCCL_LIB"):
        return os.environ["PCP_NCCL_LIB"]
    try:
        import importlib.util
        spec = importlib.util.find_spec("nvidia.nccl")
        for d in (spec.submodule_search_locations if spec else []):
            p = os.path.join(d, "lib", "libnccl.so.2")
            if os.path.exists(p):
                os.environ["PCP_NCCL_LIB"] = p
                return p
    except Exception:   # noqa: BLE001
        pass
    return None


def exchange_unique_id(rank, world, addr=None, port=None):
    """The NCCL unique id of rank 0 (pcp_multi_unique_id) on every rank; None for a single rank."""
    if world == 1:
        return None
    from . import native
    prefer_bundled_nccl()
    uid = native.Multi.unique_id() if rank == 0 else b""
    return exchange_bytes(uid, rank, world, addr, port)


def create(rank=None, world=None, device=None, unique_id=None, ctx=None, addr=None, port=None):
    """pcp_multi_create for this process: rank / world default to the launcher's environment, the device to
    LOCAL_RANK, the unique id to a TCP exchange with rank 0."""
    from . import native
    from .api import default_context
    e_rank, e_world, e_local = env_rank_world()
    rank = e_rank if rank is None else rank
    world = e_world if world is None else world
    device = e_local if device is None else device
    ctx = ctx or default_context(device)
    if world > 1:
        prefer_bundled_nccl()
    if world > 1 and unique_id is None:
        unique_id = exchange_unique_id(rank, world, addr, port)
    return native.Multi(ctx, rank, world, unique_id)


class ShardedSearch:
    """Query-sharded exact search over a replicated reference cloud (BASELINE config 5)."""

    def __init__(self, reference_shard, multi=None, **kw):
        self.multi = multi or create(**kw)
        shard = np.ascontiguousarray(np.asarray(reference_shard, dtype=np.float32))
        self.n_total = self.multi.set_reference(shard)          # the one collective of this path + the index build
        self.allgather_ms, self.build_ms = self.multi.last_timing()

    def knn(self, queries, k):
        """queries: this rank's shard (Q, 3).  Returns (idx, dist, count); indices refer to rows of the gathered
        reference cloud (rank order)."""
        return self.multi.knn(np.ascontiguousarray(np.asarray(queries, dtype=np.float32)), int(k))


def icp_batch_sharded(sources, targets, T_init=None, multi=None, run=None, rank=None, world=None,
                      max_iteration=100, R_diff_thres=0.5, t_diff_thres=0.5, dist_thres=5.0, first_iter_broadcast=True):
    """Pair-sharded batched ICP: this rank registers its contiguous range of pairs (balanced by points) with no
    communication.  Returns (first_pair, last_pair, results of that range).  `run` replaces the GPU solver for
    CPU tests of the sharding logic: run(sources[lo:hi], targets[lo:hi], T_init=...) -> dict of arrays."""
    P = len(sources)
    if len(targets) != P:
        raise ValueError("sources and targets must have the same length")
    if run is not None:
        e_rank, e_world, _ = env_rank_world()
        rank = e_rank if rank is None else rank
        world = e_world if world is None else world
        bounds = shard_pairs_by_points([len(s) + len(t) for s, t in zip(sources, targets)], world)
        lo, hi = bounds[rank], bounds[rank + 1]
        T0 = None if T_init is None else np.asarray(T_init)[lo:hi]
        return lo, hi, run(sources[lo:hi], targets[lo:hi], T_init=T0)
    from .api import _as_cloud
    multi = multi or create(rank=rank, world=world)
    src = [_as_cloud(s, "source") for s in sources]
    tgt = [_as_cloud(t, "target") for t in targets]
    so = np.zeros(P + 1, np.int64)
    to = np.zeros(P + 1, np.int64)
    if P:
        so[1:] = np.cumsum([len(s) for s in src])
        to[1:] = np.cumsum([len(t) for t in tgt])
    sp = np.concatenate(src) if P else np.zeros((0, 3), np.float32)
    tp = np.concatenate(tgt) if P else np.zeros((0, 3), np.float32)
    T0 = np.tile(np.eye(4), (P, 1, 1)) if T_init is None else np.ascontiguousarray(np.asarray(T_init, np.float64).reshape(P, 4, 4))
    out = multi.icp_batch(sp, so, tp, to, T0, int(max_iteration), float(R_diff_thres), float(t_diff_thres),
                          float(dist_thres), bool(first_iter_broadcast))
    lo, hi = int(out.pop("first_pair")), int(out.pop("last_pair"))
    return lo, hi, {k: np.asarray(v)[lo:hi] for k, v in out.items()}
