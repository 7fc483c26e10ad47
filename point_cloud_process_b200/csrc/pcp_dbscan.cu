// DBSCAN on the radius-search CSR: the labels of DBSCAN.fit of
// Cluster_dbscan/dbscan.py:10-38, which depend on the SEQUENTIAL order of that loop, computed in parallel.
//
// What the reference loop does (c_p = |ball(p, radius)|, self included):
//   * points are popped from the END of `unvisited` (:18): seeds are tried in DESCENDING index order;
//   * a popped point with c_p <  Min_Pts is noise for good (:20-22) -- it has left `unvisited`, so no later cluster
//     relabels it; with c_p >= Min_Pts it opens the next cluster label and claims its unvisited neighbours (:24-26);
//   * a claimed point expands the cluster only if c_p > Min_Pts (STRICT, :31-32).
// Hence: "expanders" E = {c_p > Min_Pts} form connected components that are always claimed as a whole, by the
// earliest of (a) their own highest index m(C) being popped, (b) a "weak seed" w (c_w == Min_Pts, still unvisited
// when popped) adjacent to the component.  Whether a weak seed is still unvisited at its turn depends only on
// events with HIGHER index, so the loop is a fixed point that rounds of a parallel rule reach:
//   w is taken   if an adjacent weak seed s > w is a seed, or an adjacent component has max(m(C), Vmax(C)) > w,
//   w is a seed  if no adjacent weak seed s > w and no weak seed > w adjacent to one of its components is undecided,
// with Vmax(C) / Umax(C) = the highest seed / undecided weak seed adjacent to C (atomicMax scatter every round).
// Every round decides at least the highest undecided weak seed.  Then: the claim time of a component is
// T(C) = max(m(C), Vmax(C)); a non-expander p belongs to the cluster with the highest claim time among its adjacent
// seeds / components if that time is > p, else it is noise (popped first) or, if it is a seed, its own cluster;
// cluster labels number the seeds in descending index order (:24).
#include <cub/device/device_scan.cuh>

#include "pcp_internal.cuh"

namespace pcp {
namespace {

enum : unsigned char { kNoise = 0, kWeak = 1, kExpander = 2 };
enum : unsigned char { kNone = 0, kUndecided = 1, kSeed = 2, kTaken = 3 };

__device__ __forceinline__ int warp_max_i(int v) {
#pragma unroll
    for (int m = 16; m > 0; m >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, m));
    return v;
}

__global__ void __launch_bounds__(256) db_classify_kernel(const long long *__restrict__ off, long long n, int min_pts,
                                                          unsigned char *__restrict__ kind,
                                                          unsigned char *__restrict__ state, int *__restrict__ comp,
                                                          int *__restrict__ vmax) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const long long c = off[i + 1] - off[i];
    const unsigned char k = c > min_pts ? kExpander : (c >= min_pts ? kWeak : kNoise);
    kind[i] = k;
    state[i] = k == kWeak ? kUndecided : kNone;
    comp[i] = k == kExpander ? (int)i : -1;
    vmax[i] = -1;
}

// max-index propagation over expander-expander edges; one warp per expander.  (Measured alternatives, both dropped:
// a single-pass lock-free union-find with path halving -- 58 ms on the KITTI scan, whose ground plane is ONE component
// of ~10^5 points, so 22 M edges contend for a handful of roots -- against 1.1 ms for these rounds; and reading the
// "changed" flag only every 2, 4, 8 rounds -- each extra round rescans the whole CSR, 1.57 -> 1.89 ms at r = 1.0.)
__global__ void __launch_bounds__(256) db_hook_kernel(const long long *__restrict__ off, const int32_t *__restrict__ nbr,
                                                      long long n, const unsigned char *__restrict__ kind,
                                                      int *__restrict__ comp, unsigned *__restrict__ changed) {
    const long long i = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (i >= n || kind[i] != kExpander) return;
    const int mine = comp[i];
    int m = mine;
    for (long long e = off[i] + lane; e < off[i + 1]; e += 32) {
        const int j = nbr[e];
        if (kind[j] == kExpander) m = max(m, comp[j]);
    }
    m = warp_max_i(m);
    if (lane == 0 && m > mine) {
        atomicMax(comp + i, m);
        atomicMax(comp + mine, m);      // hook the old representative as well
        *changed = 1u;
    }
}

// comp values are member indices with comp[c] >= c: follow them to the representative (comp[r] == r)
__global__ void __launch_bounds__(256) db_jump_kernel(long long n, const unsigned char *__restrict__ kind,
                                                      int *__restrict__ comp) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= n || kind[i] != kExpander) return;
    int c = comp[i];
    while (true) {
        const int d = comp[c];
        if (d == c) break;
        c = d;
    }
    comp[i] = c;
}

__global__ void __launch_bounds__(256) db_fill_kernel(int *__restrict__ a, long long n, int v) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i < n) a[i] = v;
}

// weak seeds publish themselves to the components they touch: Vmax (seeds), Umax (undecided)
__global__ void __launch_bounds__(256) db_scatter_kernel(const long long *__restrict__ off, const int32_t *__restrict__ nbr,
                                                         long long n, const unsigned char *__restrict__ kind,
                                                         const unsigned char *__restrict__ state,
                                                         const int *__restrict__ comp, int *__restrict__ vmax,
                                                         int *__restrict__ umax) {
    const long long w = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (w >= n) return;
    const unsigned char st = state[w];
    if (st != kUndecided && st != kSeed) return;
    int *dst = st == kSeed ? vmax : umax;
    for (long long e = off[w] + lane; e < off[w + 1]; e += 32) {
        const int j = nbr[e];
        if (kind[j] == kExpander) atomicMax(dst + comp[j], (int)w);
    }
}

__global__ void __launch_bounds__(256) db_decide_kernel(const long long *__restrict__ off, const int32_t *__restrict__ nbr,
                                                        long long n, const unsigned char *__restrict__ kind,
                                                        volatile unsigned char *__restrict__ state,
                                                        const int *__restrict__ comp, const int *__restrict__ vmax,
                                                        const int *__restrict__ umax, unsigned *__restrict__ undecided) {
    const long long w = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (w >= n || state[w] != kUndecided) return;
    bool taken = false, wait = false;
    for (long long e = off[w] + lane; e < off[w + 1]; e += 32) {
        const int j = nbr[e];
        const unsigned char k = kind[j];
        if (k == kWeak && j > w) {
            const unsigned char s = state[j];
            taken |= s == kSeed;
            wait |= s == kUndecided;
        } else if (k == kExpander) {
            const int c = comp[j];
            taken |= max(c, vmax[c]) > (int)w;
            wait |= umax[c] > (int)w;
        }
    }
    taken = __any_sync(0xffffffffu, taken);
    wait = __any_sync(0xffffffffu, wait);
    if (lane != 0) return;
    if (taken) state[w] = kTaken;
    else if (!wait) state[w] = kSeed;
    else atomicAdd(undecided, 1u);
}

__global__ void __launch_bounds__(256) db_seed_flag_kernel(long long n, const unsigned char *__restrict__ kind,
                                                           const unsigned char *__restrict__ state,
                                                           const int *__restrict__ comp, const int *__restrict__ vmax,
                                                           int *__restrict__ flag) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const bool weak_seed = kind[i] == kWeak && state[i] == kSeed;
    const bool comp_seed = kind[i] == kExpander && comp[i] == (int)i && vmax[i] < (int)i;   // T(C) == m(C)
    flag[i] = (weak_seed || comp_seed) ? 1 : 0;
}

__global__ void __launch_bounds__(256) db_label_kernel(const long long *__restrict__ off, const int32_t *__restrict__ nbr,
                                                       long long n, const unsigned char *__restrict__ kind,
                                                       const unsigned char *__restrict__ state,
                                                       const int *__restrict__ comp, const int *__restrict__ vmax,
                                                       const int *__restrict__ incl, int32_t *__restrict__ labels) {
    const long long p = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (p >= n) return;
    const int S = incl[n - 1];
    int t = -1;
    if (kind[p] == kExpander) {
        const int c = comp[p];
        t = max(c, vmax[c]);
    } else if (kind[p] == kWeak && state[p] == kSeed) {
        t = (int)p;
    } else {
        for (long long e = off[p] + lane; e < off[p + 1]; e += 32) {
            const int j = nbr[e];
            const unsigned char k = kind[j];
            if (k == kWeak && state[j] == kSeed) t = max(t, j);
            else if (k == kExpander) {
                const int c = comp[j];
                t = max(t, max(c, vmax[c]));
            }
        }
        t = warp_max_i(t);
        if (t < (int)p) t = -1;          // popped (and left `unvisited`) before any cluster could claim it: noise
    }
    if (lane == 0) labels[p] = t < 0 ? -1 : S - incl[t];
}

}  // namespace

int dbscan_device(pcp_index *ix, double radius, int min_pts, int32_t *d_labels, int32_t *h_n_clusters) {
    pcp_ctx *ctx = ix->ctx;
    cudaStream_t st = ctx->stream;
    const int64_t n = ix->n;
    const float *d_xyz = ix->built_from;
    const int64_t stride = ix->built_stride;
    DevBuf &b_off = ctx->iss_buf[0], &b_idx = ctx->iss_buf[1], &b_dist = ctx->iss_buf[2], &b_misc = ctx->iss_buf[3];
    PCP_TRY(b_off.ensure(sizeof(int64_t) * (size_t)(n + 1)));
    // neighbourhoods as scipy's float64 KDTree decides them (dbscan.py:13)
    PCP_TRY(radius_count_device(ix, d_xyz, n, stride, (float)radius, b_off.as<int64_t>(), true, radius));
    long long M = 0;
    PCP_CUDA(cudaMemcpyAsync(&M, b_off.as<int64_t>() + n, sizeof(M), cudaMemcpyDeviceToHost, st));
    PCP_CUDA(cudaStreamSynchronize(st));
    PCP_TRY(b_idx.ensure(sizeof(int32_t) * (size_t)M));
    PCP_TRY(b_dist.ensure(sizeof(float) * (size_t)M));
    PCP_TRY(radius_fill_device(ix, d_xyz, n, stride, (float)radius, b_off.as<int64_t>(), b_idx.as<int32_t>(), b_dist.as<float>(), 0, true, radius));
    const long long *off = b_off.as<long long>();
    const int32_t *nbr = b_idx.as<int32_t>();
    // misc: comp | vmax | umax | flag | incl (n int each) | counters (4 u32) | kind | state (n u8 each)
    PCP_TRY(b_misc.ensure((size_t)n * (5 * 4 + 2) + 64));
    int *comp = b_misc.as<int>(), *vmax = comp + n, *umax = vmax + n, *flag = umax + n, *incl = flag + n;
    unsigned *ctr = reinterpret_cast<unsigned *>(incl + n);
    unsigned char *kind = reinterpret_cast<unsigned char *>(ctr + 4), *state = kind + n;
    const unsigned blocks = (unsigned)((n + 255) / 256), wblocks = (unsigned)((n * 32 + 255) / 256);
    db_classify_kernel<<<blocks, 256, 0, st>>>(off, n, min_pts, kind, state, comp, vmax);
    ctx->launches += 1;
    unsigned h_ctr = 0;
    for (int round = 0; round < 1000000; ++round) {            // connected components of the expanders
        PCP_CUDA(cudaMemsetAsync(ctr, 0, 4, st));
        db_hook_kernel<<<wblocks, 256, 0, st>>>(off, nbr, n, kind, comp, ctr);
        db_jump_kernel<<<blocks, 256, 0, st>>>(n, kind, comp);
        ctx->launches += 2;
        PCP_CUDA(cudaMemcpyAsync(&h_ctr, ctr, 4, cudaMemcpyDeviceToHost, st));
        PCP_CUDA(cudaStreamSynchronize(st));
        if (h_ctr == 0) break;
    }
    for (int round = 0; round < 1000000; ++round) {            // which weak seeds are still unvisited at their turn
        db_fill_kernel<<<blocks, 256, 0, st>>>(umax, n, -1);
        PCP_CUDA(cudaMemsetAsync(ctr, 0, 4, st));
        db_scatter_kernel<<<wblocks, 256, 0, st>>>(off, nbr, n, kind, state, comp, vmax, umax);
        db_decide_kernel<<<wblocks, 256, 0, st>>>(off, nbr, n, kind, state, comp, vmax, umax, ctr);
        ctx->launches += 3;
        PCP_CUDA(cudaMemcpyAsync(&h_ctr, ctr, 4, cudaMemcpyDeviceToHost, st));
        PCP_CUDA(cudaStreamSynchronize(st));
        if (h_ctr == 0) break;
    }
    db_scatter_kernel<<<wblocks, 256, 0, st>>>(off, nbr, n, kind, state, comp, vmax, umax);   // Vmax with every seed
    db_seed_flag_kernel<<<blocks, 256, 0, st>>>(n, kind, state, comp, vmax, flag);
    size_t tmp = 0;
    PCP_CUDA(cub::DeviceScan::InclusiveSum(nullptr, tmp, flag, incl, n, st));
    PCP_TRY(ctx->cub_tmp.ensure(tmp));
    PCP_CUDA(cub::DeviceScan::InclusiveSum(ctx->cub_tmp.p, tmp, flag, incl, n, st));
    db_label_kernel<<<wblocks, 256, 0, st>>>(off, nbr, n, kind, state, comp, vmax, incl, d_labels);
    PCP_CUDA(cudaGetLastError());
    ctx->launches += 5;
    int S = 0;
    PCP_CUDA(cudaMemcpyAsync(&S, incl + (n - 1), sizeof(int), cudaMemcpyDeviceToHost, st));
    PCP_CUDA(cudaStreamSynchronize(st));
    *h_n_clusters = S;
    return PCP_OK;
}

}  // namespace pcp
