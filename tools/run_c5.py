"""BASELINE config 5 (scaled by --scale): query-sharded kNN k=16 over a replicated reference cloud.
torchrun --nproc-per-node N tools/run_c5.py [--ref 16000000 --queries 64000000]
Each rank generates its 1/N of the reference (counter hash); the C ABI (pcp_multi_set_reference) all-gathers it
over NCCL and builds the index; each rank searches its own query shard.  torch here is only the test harness (device
buffers, barrier, max over ranks).  Parity: sampled queries vs the exhaustive device kernel."""
import argparse, json, os, sys, time
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from point_cloud_process_b200 import multi


def splitmix64(x):
    x = (x + np.uint64(0x9E3779B97F4A7C15)) & np.uint64(0xFFFFFFFFFFFFFFFF)
    z = x
    z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return z ^ (z >> np.uint64(31))


def gen(seed, lo, hi, L):
    """u = ((splitmix64(seed ^ (3 i + axis)) >> 40) * 2^-24) * L  """
    i = np.arange(lo, hi, dtype=np.uint64)
    out = np.empty((hi - lo, 3), np.float32)
    with np.errstate(over="ignore"):
        for a in range(3):
            h = splitmix64(np.uint64(seed) ^ (np.uint64(3) * i + np.uint64(a)))
            out[:, a] = ((h >> np.uint64(40)).astype(np.float64) * 2.0 ** -24 * L).astype(np.float32)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ref", type=int, default=16_000_000)
    ap.add_argument("--queries", type=int, default=64_000_000)
    ap.add_argument("--k", type=int, default=16)
    ap.add_argument("--check", type=int, default=4096)
    args = ap.parse_args()
    r = run_c5(args.ref, args.queries, args.k, args.check)
    if r is not None:
        print(json.dumps(r))
    if dist.is_initialized():
        dist.destroy_process_group()


def run_c5(n_ref, n_queries, k, check, use_dist=True):
    """Every rank: its shard of the reference -> pcp_multi_set_reference (ncclAllGather + build, timed by the
    library) -> kNN of its own query shard (CUDA events) -> parity of `check` sampled queries against the
    exhaustive DEVICE kernel (pcp_knn_bruteforce) over the gathered cloud.  Returns the report on rank 0."""
    import point_cloud_process_b200 as pcp
    rank, world, local = multi.env_rank_world()
    torch.cuda.set_device(local)
    if world > 1 and use_dist and not dist.is_initialized():
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    L = 64.0 * (n_ref / 16e6) ** (1 / 3)
    rlo, rhi = multi.shard_range(n_ref, rank, world)
    qlo, qhi = multi.shard_range(n_queries, rank, world)
    ref_shard = torch.from_numpy(gen(42, rlo, rhi, L)).cuda()
    q = torch.from_numpy(gen(43, qlo, qhi, L)).cuda()
    ctx = pcp.default_context(local)
    m = multi.create(rank=rank, world=world, device=local, ctx=ctx)
    st = torch.cuda.ExternalStream(ctx.stream())
    nq = qhi - qlo
    idx = torch.empty((nq, k), dtype=torch.int32, device="cuda")
    d = torch.empty((nq, k), dtype=torch.float32, device="cuda")
    cnt = torch.empty(nq, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    for timed in (False, True):      # first pass untimed: buffer allocation, NCCL channel set-up
        n_total = m.set_reference_device(ref_shard.data_ptr(), rhi - rlo, 3)    # the one collective + the build
        gather_ms, build_ms = m.last_timing()
        e = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        e[0].record(st)
        m.knn_device(q.data_ptr(), nq, 3, k, idx.data_ptr(), d.data_ptr(), cnt.data_ptr())
        e[1].record(st)
        ctx.synchronize()
        if world > 1 and not timed:
            dist.barrier()
    t = torch.tensor([gather_ms, build_ms, e[0].elapsed_time(e[1])], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ok = True
    if check:
        sel = torch.from_numpy(np.random.default_rng(rank).choice(nq, min(check, nq), replace=False)).cuda()
        qs = q[sel].contiguous()
        bi = torch.empty((len(sel), k), dtype=torch.int32, device="cuda")
        bd = torch.empty((len(sel), k), dtype=torch.float32, device="cuda")
        bc = torch.empty(len(sel), dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        m.knn_bruteforce_device(qs.data_ptr(), len(sel), 3, k, bi.data_ptr(), bd.data_ptr(), bc.data_ptr())
        ctx.synchronize()
        ok = bool(torch.equal(idx[sel], bi) and torch.equal(d[sel], bd))
    oks = [ok]
    if world > 1:
        oks = [None] * world
        dist.all_gather_object(oks, ok)
    info = m.info()
    gm, bm, sm = (float(x) for x in t.tolist())
    return {"config": "C5", "ref": n_ref, "queries": n_queries, "k": k, "n_gpus": world, "reference_rows_gathered": int(n_total),
            "allgather_ms": gm, "build_ms": bm, "search_ms": sm, "timing": "max over ranks, CUDA events",
            "queries_per_s_search": n_queries / (sm * 1e-3), "queries_per_s_incl_gather_and_build": n_queries / ((gm + bm + sm) * 1e-3),
            "allgather_bytes_received_per_gpu": int(n_total) * 12, "comm_nranks": info["comm_nranks"], "nccl_version": info["nccl_version"],
            "parity_sampled_vs_device_bruteforce": oks, "sampled_queries_per_rank": int(check)} if rank == 0 else None


if __name__ == "__main__":
    main()
