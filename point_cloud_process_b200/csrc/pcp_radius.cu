// Exact fixed-radius search over the bucket hierarchy, CSR output.
// Replaces kdtree_radius_search (kdtree.py:176-208), octree_radius_search(_fast) (octree.py:166-259)
// and RadiusNNResultSet.add_point (result_set.py:78-84: keep iff NOT dist > radius).
//
// One warp = one query.  Lane j tests child j of the current hierarchy node (coalesced AABB loads),
// surviving internal nodes go on a small per-warp stack, surviving buckets are scanned with one
// coalesced 512-byte load (lane j = point j).  Hits are compacted with ballot + popc, so the fill
// pass writes each query's CSR segment with contiguous, coalesced stores -- the C1 workload emits
// 80 M (index, distance) pairs = 642 MB and is output-bandwidth bound.
//
// The float32 test `sqrtf(d2) <= r` of the reference is evaluated as `d2 <= d2max` where d2max is the
// LARGEST float whose correctly rounded square root is <= r (computed on the host, exact because
// fl(sqrt(.)) is monotone).
#include <cub/device/device_scan.cuh>
#include <cub/device/device_segmented_sort.cuh>

#include <cmath>
#include <cstdlib>
#include <cstring>

#include "pcp_device.cuh"
#include "pcp_internal.cuh"

namespace pcp {

// Largest float x with sqrtf(x) <= r (r >= 0, finite).  Returns -1 for r < 0 (nothing matches).
float exact_d2_threshold(float r) {
    if (!(r >= 0.f)) return -1.f;
    if (std::isinf(r)) return r;
    float x = r * r;
    if (std::isinf(x)) x = 3.402823466e38f;
    while (sqrtf(x) > r) x = nextafterf(x, -1.f);
    while (true) {
        float y = nextafterf(x, INFINITY);
        if (std::isinf(y) || sqrtf(y) > r) break;
        x = y;
    }
    return x;
}

namespace {

// launch radius_kernel<FILL, LIST, EXACT> with EXACT chosen at run time
#define PCP_RADIUS_LAUNCH(FILL_, LIST_, rt_, grid_, ...)                                                           \
    do {                                                                                                          \
        if ((rt_).lo != (rt_).hi || (rt_).p64)                                                                    \
            radius_kernel<FILL_, LIST_, true><<<grid_, kWarpsPerCta * 32, 0, ctx->stream>>>(__VA_ARGS__);           \
        else                                                                                                      \
            radius_kernel<FILL_, LIST_, false><<<grid_, kWarpsPerCta * 32, 0, ctx->stream>>>(__VA_ARGS__);          \
    } while (0)

constexpr int kWarpsPerCta = 8;
constexpr int kStack = kMaxLevels * kFan;

// LIST = false: warp w of the grid takes query w.
// LIST = true:  persistent warps drain the hand-over list of the tiled kernel (pcp_radius_tile.cu): list[0] =
//               number of groups, list[1..] = group ids; item w -> regrouped position list[1 + w / 32] * 32 + w % 32.
template <bool FILL, bool LIST, bool EXACT>
__global__ void __launch_bounds__(kWarpsPerCta * 32)
radius_kernel(const IndexView ix, const float *__restrict__ queries, long long nq, long long stride, const RadiusTest rt,
              long long *__restrict__ counts, const long long *__restrict__ offsets, int *__restrict__ out_idx,
              float *__restrict__ out_dist, unsigned long long *__restrict__ ctr, const unsigned *__restrict__ list,
              const unsigned *__restrict__ perm) {
    __shared__ unsigned s_stack[kWarpsPerCta][kStack];   // entries: (level << 27) | node   (level >= 1)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long n_work = LIST ? (long long)list[0] * 32 : nq;
    for (long long item = blockIdx.x * (long long)kWarpsPerCta + warp; item < n_work;
         item += LIST ? gridDim.x * (long long)kWarpsPerCta : n_work) {
    long long qi = item;
    if (LIST) {
        const long long pos = (long long)list[1 + (item >> 5)] * 32 + (item & 31);
        if (pos >= nq) continue;
        qi = perm ? (long long)perm[pos] : pos;
    }
    const float *q = queries + qi * stride;
    const float qx = __ldg(q), qy = __ldg(q + 1), qz = __ldg(q + 2);
    unsigned *stack = s_stack[warp];
    long long total = 0;
    long long base = FILL ? offsets[qi] : 0;
    const unsigned lt_mask = (1u << lane) - 1u;
    unsigned n_scan = 0, n_open = 0;

    auto scan_bucket = [&](long long leaf) {
        ++n_scan;
        const float4 p = __ldg(ix.pts + leaf * kLeaf + lane);
        const float d2 = ref_d2(qx, qy, qz, p.x, p.y, p.z);
        // NaN pads never hit.  EXACT: borderline pairs are decided in float64 (RadiusTest, pcp_device.cuh)
        const bool hit = EXACT ? radius_hit(rt, d2, qx, qy, qz, p.x, p.y, p.z, qi, __float_as_uint(p.w)) : d2 <= rt.hi;
        const unsigned m = __ballot_sync(kFullMask, hit);
        if (FILL && hit) {
            const long long o = base + total + __popc(m & lt_mask);
            out_idx[o] = (int)__float_as_uint(p.w);
            out_dist[o] = __fsqrt_rn(d2);
        }
        total += __popc(m);
    };

    const int top = ix.n_levels - 1;
    if (top == 0) {
        scan_bucket(0);
    } else if (top > 0) {
        int sp = 0;
        if (lane == 0) stack[0] = ((unsigned)top << 27);
        sp = 1;
        __syncwarp();
        while (sp > 0) {
            const unsigned e = stack[sp - 1];
            --sp;
            __syncwarp();
            ++n_open;
            const int lv = (int)(e >> 27);
            const unsigned node = e & 0x07ffffffu;
            const long long c = (long long)node * kFan + lane;
            bool hit = false;
            if (c < ix.level_cnt[lv - 1]) {
                const long long b = ix.level_off[lv - 1] + c;
                hit = box_lb2(qx, qy, qz, __ldg(ix.box_lo + b), __ldg(ix.box_hi + b)) <= rt.hi;
            }
            unsigned m = __ballot_sync(kFullMask, hit);
            if (lv == 1) {
                while (m) {
                    const int j = __ffs(m) - 1;
                    m &= m - 1;
                    scan_bucket((long long)node * kFan + j);
                }
            } else {
                if (hit) stack[sp + __popc(m & lt_mask)] = ((unsigned)(lv - 1) << 27) | (unsigned)c;
                sp += __popc(m);
                __syncwarp();
            }
        }
    }
    if (!FILL && lane == 0) counts[qi] = total;
    if (ctr && lane == 0) {
        atomicAdd(ctr + 0, (unsigned long long)n_scan * kLeaf);
        atomicAdd(ctr + 1, (unsigned long long)n_open * kFan);
        atomicAdd(ctr + 2, (unsigned long long)n_open);
        atomicAdd(ctr + 4, 1ull);
        atomicAdd(ctr + 5, (unsigned long long)n_scan);
    }
    __syncwarp();
    }
}

// counts (nq) -> offsets (nq+1) with offsets[0] = 0
__global__ void set_zero_kernel(long long *p) { *p = 0; }

// Below this many queries the warp-per-query kernel wins (no query sort, 32x more warps).
constexpr int64_t kTileMinQueries = 8192;
constexpr int64_t kTileFillMinQueries = 262144;   // FILL pass: tiled kernel from here on (see radius_fill_device)

// 0 = by query count, 1 = warp-per-query, 2 = tiled (PCP_RADIUS_KERNEL=warp|tile, for tests)
int radius_kernel_choice() {
    static const int choice = [] {
        const char *e = getenv("PCP_RADIUS_KERNEL");
        if (!e) return 0;
        return strcmp(e, "warp") == 0 ? 1 : strcmp(e, "tile") == 0 ? 2 : 0;
    }();
    return choice;
}
bool use_tile(int64_t nq) {
    const int c = radius_kernel_choice();
    return c == 2 || (c == 0 && nq >= kTileMinQueries);
}

// Regrouped order of the query set (sort_queries), cached between the count and the fill pass.
int radius_query_perm(pcp_index *ix, const float *d_q, int64_t nq, int64_t stride, bool reuse, const uint32_t **out) {
    pcp_ctx *ctx = ix->ctx;
    auto &tag = ctx->radius_perm_tag;
    const bool valid = reuse && tag.q == d_q && tag.nq == nq && tag.stride == stride && tag.ix == ix && tag.serial == ix->serial;
    if (!valid) {
        PCP_TRY(ctx->radius_perm.ensure(sizeof(uint32_t) * (size_t)nq));
        PCP_TRY(sort_queries(ix, d_q, nq, stride, ctx->radius_perm.as<uint32_t>()));
        tag.q = d_q; tag.nq = nq; tag.stride = stride; tag.ix = ix; tag.serial = ix->serial;
    }
    *out = ctx->radius_perm.as<uint32_t>();
    return PCP_OK;
}

}  // namespace

// Groups the tiled kernel handed over (its stragglers), one warp per query, persistent grid.
int radius_warp_list_device(pcp_index *ix, const float *d_q, int64_t nq, int64_t stride, const uint32_t *d_perm,
                            const uint32_t *d_list, const RadiusTest &d2max, long long *d_counts, const long long *d_offsets,
                            int32_t *d_idx, float *d_dist) {
    pcp_ctx *ctx = ix->ctx;
    const IndexView v = make_view(ix);
    const unsigned grid = (unsigned)ctx->sm_count * 8;
    if (d_counts)
        PCP_RADIUS_LAUNCH(false, true, d2max, grid, v, d_q, nq, stride, d2max, d_counts, nullptr,
                                                                               nullptr, nullptr, ctx->ctr(), d_list, d_perm);
    else
        PCP_RADIUS_LAUNCH(true, true, d2max, grid, v, d_q, nq, stride, d2max, nullptr, d_offsets,
                                                                              d_idx, d_dist, ctx->ctr(), d_list, d_perm);
    ctx->launches += 1;
    PCP_CUDA(cudaGetLastError());
    return PCP_OK;
}

// exact64: decide borderline pairs in float64 (RadiusTest, pcp_device.cuh); the float64 rows attached to the index
// are used when the queries ARE the indexed cloud (the ISS / DBSCAN case).
RadiusTest make_radius_test(const pcp_index *ix, const float *d_q, int64_t nq, float radius, bool exact64, double radius64) {
    RadiusTest t;
    t.lo = t.hi = exact_d2_threshold(radius);
    t.r2 = (double)radius * (double)radius;
    t.q64 = t.p64 = nullptr;
    if (!exact64 || !(radius >= 0.f) || std::isinf(radius)) return t;
    double E = 0.0;                                     // float32 rounding of attached float64 coordinates
    const bool same = d_q == ix->built_from && nq == ix->n;
    if (ix->xyz64.p && ix->xyz64_n == ix->n) {
        t.p64 = ix->xyz64.as<double>();
        if (same) t.q64 = t.p64;
        E = std::sqrt(3.0) * ix->xyz64_max_abs * 1.2e-7 * (same ? 2.0 : 1.0);
    }
    const double r = radius64 >= 0.0 ? radius64 : (double)radius, rel = 1.0e-6;
    t.r2 = r * r;      // > 8 x 2^-24: the float32 evaluation's own rounding
    const double lo = r > E ? (r - E) * (r - E) * (1.0 - rel) : -1.0, hi = (r + E) * (r + E) * (1.0 + rel);
    t.lo = lo < 0.0 ? -1.f : nextafterf((float)lo, -1.f);
    t.hi = nextafterf((float)hi, INFINITY);
    if (!(t.hi < 3.0e38f)) t.hi = 3.0e38f;
    return t;
}

int radius_count_device(pcp_index *ix, const float *d_q, int64_t nq, int64_t stride, float radius,
                        int64_t *d_offsets, bool exact64, double radius64) {
    pcp_ctx *ctx = ix->ctx;
    set_zero_kernel<<<1, 1, 0, ctx->stream>>>(reinterpret_cast<long long *>(d_offsets));
    ctx->launches += 1;
    if (nq == 0) { PCP_CUDA(cudaGetLastError()); return PCP_OK; }
    PCP_TRY(ctx->scratch[3].ensure(sizeof(long long) * (size_t)nq));
    long long *counts = ctx->scratch[3].as<long long>();
    if (ix->n == 0) {
        PCP_CUDA(cudaMemsetAsync(d_offsets, 0, sizeof(int64_t) * (size_t)(nq + 1), ctx->stream));
        return PCP_OK;
    }
    const RadiusTest d2max = make_radius_test(ix, d_q, nq, radius, exact64, radius64);
    const IndexView v = make_view(ix);
    if (use_tile(nq)) {
        const uint32_t *perm = nullptr;
        PCP_TRY(radius_query_perm(ix, d_q, nq, stride, false, &perm));
        PCP_TRY(radius_tile_device(ix, d_q, nq, stride, perm, d2max, counts, nullptr, nullptr, nullptr));
    } else {
        const unsigned grid = (unsigned)((nq + kWarpsPerCta - 1) / kWarpsPerCta);
        PCP_RADIUS_LAUNCH(false, false, d2max, grid, v, d_q, nq, stride, d2max, counts, nullptr,
                                                                                nullptr, nullptr, ctx->ctr(), nullptr,
                                                                                nullptr);
        ctx->launches += 1;
    }
    size_t tmp = 0;
    PCP_CUDA(cub::DeviceScan::InclusiveSum(nullptr, tmp, counts, reinterpret_cast<long long *>(d_offsets) + 1, nq,
                                           ctx->stream));
    PCP_TRY(ctx->cub_tmp.ensure(tmp));
    PCP_CUDA(cub::DeviceScan::InclusiveSum(ctx->cub_tmp.p, tmp, counts, reinterpret_cast<long long *>(d_offsets) + 1,
                                           nq, ctx->stream));
    ctx->launches += 2;
    PCP_CUDA(cudaGetLastError());
    return PCP_OK;
}

namespace {

__global__ void __launch_bounds__(256) pack_pairs_kernel(const int *__restrict__ idx, const float *__restrict__ dist,
                                                         long long m, unsigned long long *__restrict__ keys) {
    long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i < m) keys[i] = pack_key(dist[i], (unsigned)idx[i]);
}

__global__ void __launch_bounds__(256) unpack_pairs_kernel(const unsigned long long *__restrict__ keys, long long m,
                                                           int *__restrict__ idx, float *__restrict__ dist) {
    long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i < m) { idx[i] = (int)key_index(keys[i]); dist[i] = key_dist(keys[i]); }
}

}  // namespace

int radius_fill_device(pcp_index *ix, const float *d_q, int64_t nq, int64_t stride, float radius,
                       const int64_t *d_offsets, int32_t *d_idx, float *d_dist, int sorted, bool exact64, double radius64) {
    pcp_ctx *ctx = ix->ctx;
    if (nq == 0 || ix->n == 0) return PCP_OK;
    const RadiusTest d2max = make_radius_test(ix, d_q, nq, radius, exact64, radius64);
    const IndexView v = make_view(ix);
    // FILL kernel choice.  The tiled FILL stages hits in shared memory and writes 64-byte runs; it walks the
    // hierarchy once per 32 queries and wins on large query sets at every result density (1 M uniform queries:
    // 0.56-... ms against 0.99-2.64 ms).  On a KITTI-sized set (3 896 groups: half the warp slots idle) the
    // warp-per-query kernel with its ballot-compacted rows is faster (r = 1.0: 80 M pairs at 1.9 TB/s).
    // PCP_RADIUS_FILL=tile|warp forces one (tests keep both paths alive).
    static const int fill_choice = [] {
        const char *e = getenv("PCP_RADIUS_FILL");
        return !e ? 0 : strcmp(e, "tile") == 0 ? 2 : strcmp(e, "warp") == 0 ? 1 : 0;
    }();
    const bool tile_fill = fill_choice == 2 || (fill_choice == 0 && nq >= kTileFillMinQueries);
    if (use_tile(nq) && tile_fill) {
        const uint32_t *perm = nullptr;
        PCP_TRY(radius_query_perm(ix, d_q, nq, stride, true, &perm));
        PCP_TRY(radius_tile_device(ix, d_q, nq, stride, perm, d2max, nullptr, reinterpret_cast<const long long *>(d_offsets),
                                   d_idx, d_dist));
    } else {
        const unsigned grid = (unsigned)((nq + kWarpsPerCta - 1) / kWarpsPerCta);
        PCP_RADIUS_LAUNCH(true, false, d2max, grid, v, d_q, nq, stride, d2max, nullptr, reinterpret_cast<const long long *>(d_offsets), d_idx, d_dist, ctx->ctr(),
            nullptr, nullptr);
        ctx->launches += 1;
    }
    PCP_CUDA(cudaGetLastError());
    if (!sorted) return PCP_OK;
    // ascending (distance, index) inside every CSR segment: what RadiusNNResultSet.__str__ prints
    long long m = 0;
    PCP_CUDA(cudaMemcpyAsync(&m, d_offsets + nq, sizeof(long long), cudaMemcpyDeviceToHost, ctx->stream));
    PCP_CUDA(cudaStreamSynchronize(ctx->stream));
    if (m == 0) return PCP_OK;
    PCP_TRY(ctx->scratch[4].ensure(sizeof(unsigned long long) * (size_t)m));
    PCP_TRY(ctx->scratch[5].ensure(sizeof(unsigned long long) * (size_t)m));
    unsigned long long *ka = ctx->scratch[4].as<unsigned long long>(), *kb = ctx->scratch[5].as<unsigned long long>();
    const unsigned g2 = (unsigned)((m + 255) / 256);
    pack_pairs_kernel<<<g2, 256, 0, ctx->stream>>>(d_idx, d_dist, m, ka);
    const long long *offs = reinterpret_cast<const long long *>(d_offsets);
    size_t tmp = 0;
    PCP_CUDA(cub::DeviceSegmentedSort::SortKeys(nullptr, tmp, ka, kb, m, nq, offs, offs + 1, ctx->stream));
    PCP_TRY(ctx->cub_tmp.ensure(tmp));
    PCP_CUDA(cub::DeviceSegmentedSort::SortKeys(ctx->cub_tmp.p, tmp, ka, kb, m, nq, offs, offs + 1, ctx->stream));
    unpack_pairs_kernel<<<g2, 256, 0, ctx->stream>>>(kb, m, d_idx, d_dist);
    ctx->launches += 5;
    PCP_CUDA(cudaGetLastError());
    return PCP_OK;
}

}  // namespace pcp
