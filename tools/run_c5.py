"""BASELINE config 5 (scaled by --scale): query-sharded kNN k=16 over a replicated reference cloud.
torchrun --nproc-per-node N tools/run_c5.py [--ref 16000000 --queries 64000000]
Each rank generates its 1/N of the reference (counter hash), all-gathers it over NCCL, builds the
index locally and searches its own query shard.  Parity: sampled queries vs the brute-force oracle."""
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
    ap.add_argument("--check", type=int, default=512)
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    L = 64.0 * (args.ref / 16e6) ** (1 / 3)
    rlo, rhi = multi.shard_range(args.ref, rank, world)
    qlo, qhi = multi.shard_range(args.queries, rank, world)
    ref_shard = torch.from_numpy(gen(42, rlo, rhi, L)).cuda()
    q = torch.from_numpy(gen(43, qlo, qhi, L)).cuda()
    dist.barrier(); torch.cuda.synchronize()
    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    e[0].record()
    s = multi.ShardedSearch(ref_shard)                       # all-gather + build
    torch.cuda.synchronize(); t_build = time.perf_counter()
    st = torch.cuda.ExternalStream(s.ctx.stream())
    e[1].record(st)
    idx, d, cnt = s.knn(q, args.k)
    e[2].record(st)
    torch.cuda.synchronize()
    ms = torch.tensor([e[1].elapsed_time(e[2])], dtype=torch.float64, device="cuda")
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ok = True
    if args.check:
        from oracle import knn_oracle
        ref = s.reference.cpu().numpy()
        sel = np.random.default_rng(rank).choice(qhi - qlo, min(args.check, qhi - qlo), replace=False)
        bd, bi, _ = knn_oracle.bruteforce_knn(ref, q[sel].cpu().numpy(), args.k, threads=os.cpu_count() // world or 1)
        ok = bool(np.array_equal(idx[sel].cpu().numpy(), bi.astype(np.int32)) and np.array_equal(d[sel].cpu().numpy(), bd))
    oks = [None] * world
    dist.all_gather_object(oks, ok)
    if rank == 0:
        print(json.dumps({"config": "C5", "ref": args.ref, "queries": args.queries, "k": args.k, "n_gpus": world,
                          "search_ms_max_over_ranks": float(ms.item()), "queries_per_s": args.queries / (ms.item() * 1e-3),
                          "parity_sampled": oks}))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
