"""Multi-GPU plumbing: one process per GPU (torch.distributed), work sharded where the path shards.

    registration pairs      independent -> contiguous ranges of pairs per rank, NO collective on the data
                            path; the (P,4,4) results are gathered on the host side afterwards
    queries of a large cloud independent given the whole reference cloud -> queries split per rank, the
                            reference cloud REPLICATED: each rank holds a shard, ONE all-gather (NCCL over
                            NVLink) rebuilds the full cloud on every GPU, each rank builds its own index
                            (the build is deterministic) and searches its own queries

torch is used for the process group and device buffers only; search / ICP go through the C ABI.
"""
import numpy as np


def shard_range(n, rank, world):
    """Balanced contiguous range [lo, hi) of n items for `rank` of `world`."""
    base, rem = divmod(int(n), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_pairs_by_points(sizes, world):
    """Contiguous pair ranges balanced by the number of points: sizes[i] = points of pair i.
    Returns world+1 boundaries."""
    sizes = np.asarray(sizes, np.int64)
    total = int(sizes.sum())
    cum = np.concatenate([[0], np.cumsum(sizes)])
    bounds = [0]
    for r in range(1, world):
        target = total * r / world
        bounds.append(max(int(np.searchsorted(cum, target, side="left")), bounds[-1]))
    bounds.append(len(sizes))
    return bounds


def all_gather_cloud(local, group=None):
    """All-gather row shards of a cloud (torch tensor (n_r, C), CPU for gloo / CUDA for nccl; shards may
    differ in length) -> the full (sum n_r, C) tensor on every rank, rank order."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    n_local = torch.tensor([local.shape[0]], dtype=torch.int64, device=local.device)
    sizes = [torch.zeros_like(n_local) for _ in range(world)]
    dist.all_gather(sizes, n_local, group=group)
    sizes = [int(s.item()) for s in sizes]
    n_max = max(sizes)
    padded = local
    if local.shape[0] != n_max:
        padded = torch.zeros((n_max,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        padded[:local.shape[0]] = local
    out = torch.empty((world * n_max,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, padded.contiguous(), group=group)
    if all(s == n_max for s in sizes):
        return out
    return torch.cat([out[r * n_max:r * n_max + sizes[r]] for r in range(world)], dim=0)


class ShardedSearch:
    """Query-sharded exact search over a replicated reference cloud (BASELINE config 5)."""

    def __init__(self, reference_shard, device=None, group=None, ctx=None):
        import torch

        from . import native
        from .api import default_context
        self.device = torch.cuda.current_device() if device is None else device
        self.ctx = ctx or default_context(self.device)
        shard = torch.as_tensor(reference_shard, dtype=torch.float32)
        if not shard.is_cuda:
            shard = shard.cuda(self.device)
        self.reference = all_gather_cloud(shard.contiguous(), group)        # the one collective of this path
        torch.cuda.current_stream().synchronize()
        self.index = native.Index(self.ctx, self.reference.data_ptr(), self.reference.shape[0], self.reference.shape[1])

    def knn(self, queries, k):
        """queries: this rank's shard, (Q, 3) CUDA tensor.  Returns CUDA tensors (idx, dist, count); indices
        refer to rows of the gathered reference cloud."""
        import torch
        q = torch.as_tensor(queries, dtype=torch.float32)
        if not q.is_cuda:
            q = q.cuda(self.device)
        q = q.contiguous()
        n = q.shape[0]
        idx = torch.empty((n, k), dtype=torch.int32, device=q.device)
        dist = torch.empty((n, k), dtype=torch.float32, device=q.device)
        cnt = torch.empty(n, dtype=torch.int32, device=q.device)
        torch.cuda.current_stream().synchronize()
        self.index.knn_device(q.data_ptr(), n, q.shape[1], k, idx.data_ptr(), dist.data_ptr(), cnt.data_ptr())
        self.ctx.synchronize()
        return idx, dist, cnt


def icp_batch_sharded(sources, targets, T_init=None, group=None, run=None, **kw):
    """Pair-sharded batched ICP: every rank registers its contiguous range of pairs (balanced by points) with
    no communication, then the small per-pair results are all-gathered as Python objects.  `run` defaults to
    the GPU icp_batch; tests on CPU inject a stand-in to exercise the sharding logic."""
    import torch.distributed as dist
    if run is None:
        from .api import icp_batch as run
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    bounds = shard_pairs_by_points([len(s) + len(t) for s, t in zip(sources, targets)], world)
    lo, hi = bounds[rank], bounds[rank + 1]
    T0 = None if T_init is None else np.asarray(T_init)[lo:hi]
    local = run(sources[lo:hi], targets[lo:hi], T_init=T0, **kw)
    parts = [None] * world
    dist.all_gather_object(parts, {k: np.asarray(v) for k, v in local.items()}, group=group)
    return {k: np.concatenate([p[k] for p in parts], axis=0) for k in parts[0]}
