// Exact kNN, tiled: one warp = one GROUP of 32 spatially adjacent queries, one query per lane.
// Replaces kdtree_knn_search (kdtree.py:141-172), octree_knn_search (octree.py:262-306) and
// KNNResultSet.add_point (result_set.py:35-53) for batched query sets.
//
// The group walks the 32-ary AABB hierarchy ONCE for all its queries, nearest-first, pruning with
// the exact box-to-box lower bound against the loosest k-th distance in the group.  Every surviving
// bucket (32 points, 512 B) is staged in shared memory with one coalesced float4 load per lane and
// then read back as 32 broadcast LDS.128: each lane evaluates its own query against all 32 points
// (8 unfused FP32 ops per pair, the reference's operation order) and only records a 32-bit pass
// mask against its own k-th-distance gate.  The few survivors are then inserted into the lane's
// sorted register list of K packed (distance, index) keys.  The inner loop is 11 instructions per
// distance evaluation, 8 of them FP32 pipe work -- this is the kernel the FP32 roofline is quoted on.
//
// Query groups: for pcp_knn_self the buckets themselves (Morton-sorted cloud); for arbitrary query
// sets the queries are first sorted by their Morton code in the index's frame (sort_queries).
#include <algorithm>
#include <cfloat>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "pcp_device.cuh"
#include "pcp_internal.cuh"

namespace pcp {

namespace {

constexpr int kTileWarps = 4;
constexpr unsigned kDone = 0xffffffffu;
constexpr unsigned kInfBits = 0x7f800000u;
// Work budget of one group, per result slot K, before it is handed to the warp-per-query kernel.  It only
// applies when the whole grid fits in about two waves of resident warps (e.g. the 124 668-point KITTI scan =
// 3 896 groups): there the kernel lasts as long as its slowest group, and the sparse outlier groups (a box far
// larger than its queries' neighbourhoods) are 8x the mean.  Larger inputs run unbounded.
constexpr unsigned kScansPerSlot = 4;
constexpr unsigned kTestsPerSlot = 12;

template <int K>
struct TopK {
    unsigned long long key[K];

    __device__ __forceinline__ void init() {
#pragma unroll
        for (int i = 0; i < K; ++i) key[i] = kEmptyKey;
    }
    // The gate is always taken from slot K-1: for a runtime k < K (k not a power of two) the list simply
    // keeps a few more candidates than needed; the first k are still the exact k nearest.
    __device__ __forceinline__ unsigned long long worst() const { return key[K - 1]; }
    // sorted insertion of nk (caller guarantees nk < key[K-1] or accepts a no-op otherwise)
    __device__ __forceinline__ void insert(unsigned long long nk) {
        bool g[K];
#pragma unroll
        for (int i = 0; i < K; ++i) g[i] = key[i] > nk;
#pragma unroll
        for (int i = K - 1; i > 0; --i) key[i] = g[i - 1] ? key[i - 1] : (g[i] ? nk : key[i]);
        key[0] = g[0] ? nk : key[0];
    }
};

// group-box lower bound of bucket c of the open level-1 node, from the boxes kept in shared memory
__device__ __forceinline__ unsigned lbs_bound(const float4 *boxes, const float3 &glo, const float3 &ghi, int c) {
    return __float_as_uint(boxbox_lb2(glo, ghi, boxes[2 * c], boxes[2 * c + 1]));
}

template <int K, bool SELF>
__global__ void __launch_bounds__(kTileWarps * 32)
knn_tile_kernel(const IndexView ix, const float *__restrict__ queries, const unsigned *__restrict__ perm, long long nq,
                long long stride, int k, int *__restrict__ out_idx, float *__restrict__ out_dist,
                int *__restrict__ out_cnt, unsigned long long *__restrict__ ctr, unsigned *__restrict__ dbg,
                unsigned *__restrict__ overflow, unsigned max_scans, unsigned max_tests) {
    __shared__ float4 s_tile[kTileWarps][32];
    __shared__ float4 s_box[kTileWarps][64];
    __shared__ unsigned s_lb[kTileWarps][kMaxLevels][32];
    __shared__ unsigned s_cd[kTileWarps][kMaxLevels][32];
    __shared__ unsigned s_node[kTileWarps][kMaxLevels];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long group = blockIdx.x * (long long)kTileWarps + warp;
    const long long n_groups = SELF ? ix.n_leaf : (nq + 31) / 32;
    if (group >= n_groups) return;

    // ---- this lane's query
    float qx = 0.f, qy = 0.f, qz = 0.f;
    long long qid = -1;
    if (SELF) {
        const float4 p = __ldg(ix.pts + group * kLeaf + lane);
        qx = p.x; qy = p.y; qz = p.z;
        qid = (long long)(int)__float_as_uint(p.w);            // -1 on pad lanes
    } else {
        const long long pos = group * 32 + lane;
        if (pos < nq) {
            qid = perm ? (long long)perm[pos] : pos;
            const float *q = queries + qid * stride;
            qx = __ldg(q); qy = __ldg(q + 1); qz = __ldg(q + 2);
        }
    }
    const bool active = qid >= 0;
    // group AABB over the active lanes
    float3 glo, ghi;
    glo.x = warp_min(active ? qx : FLT_MAX); glo.y = warp_min(active ? qy : FLT_MAX); glo.z = warp_min(active ? qz : FLT_MAX);
    ghi.x = warp_max(active ? qx : -FLT_MAX); ghi.y = warp_max(active ? qy : -FLT_MAX); ghi.z = warp_max(active ? qz : -FLT_MAX);
    const float gcx = 0.5f * (glo.x + ghi.x), gcy = 0.5f * (glo.y + ghi.y), gcz = 0.5f * (glo.z + ghi.z);

    TopK<K> top;
    top.init();
    float gate = active ? __uint_as_float(kInfBits) : -1.f;     // squared-distance gate of this lane (inactive: nothing passes)
    unsigned group_gate = kInfBits;                             // bits of the loosest gate in the group
    unsigned n_scan = 0, n_open = 0, n_test = 0, n_round = 0;
    bool aborted = false;
    const unsigned n_active = __popc(__ballot_sync(kFullMask, active));

    float4 *tile = s_tile[warp];
    float4 *boxes = s_box[warp];            // boxes of the 32 buckets below the open level-1 node: [2*c] = lo, [2*c+1] = hi
    // Exact per-lane lower bound against bucket c's box (shared-memory broadcast): can ANY lane still use it?
    auto bucket_needed = [&](int c) -> bool {
        ++n_test;
        return __ballot_sync(kFullMask, box_lb2(qx, qy, qz, boxes[2 * c], boxes[2 * c + 1]) <= gate) != 0;
    };
    // Scan one bucket whose points `pv` (lane j = point j) are already in registers.
    auto scan_bucket = [&](const float4 pv) {
        ++n_scan;
        __syncwarp();
        tile[lane] = pv;
        __syncwarp();
        unsigned mask = 0;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            const float4 p = tile[j];
            const float d2 = ref_d2(qx, qy, qz, p.x, p.y, p.z);
            if (d2 <= gate) mask |= 1u << j;                     // NaN pads never pass
        }
        while (__any_sync(kFullMask, mask != 0)) {
            ++n_round;
            if (mask) {
                const int j = __ffs(mask) - 1;
                mask &= mask - 1;
                const float4 p = tile[j];
                const float d2 = ref_d2(qx, qy, qz, p.x, p.y, p.z);
                if (d2 <= gate) {                                // the gate may have tightened since the mask was taken
                    const unsigned long long nk = pack_key(__fsqrt_rn(d2), __float_as_uint(p.w));
                    if (nk < top.worst()) {
                        top.insert(nk);
                        gate = d2_gate(key_dist(top.worst()));
                    }
                }
            }
        }
        group_gate = __reduce_max_sync(kFullMask, active ? __float_as_uint(gate) : 0u);
    };

    const int top_level = ix.n_levels - 1;      // >= 1 for every non-empty index (pcp_build.cu)
    if (top_level > 0) {
        unsigned(*lbs)[32] = s_lb[warp];
        unsigned(*cds)[32] = s_cd[warp];
        unsigned *nodes = s_node[warp];
        auto open_node = [&](int lv, unsigned node) {
            ++n_open;
            if (lane == 0) nodes[lv] = node;
            const long long c = (long long)node * kFan + lane;
            unsigned bits = kDone, cbits = kDone;
            if (c < ix.level_cnt[lv - 1]) {
                const long long b = ix.level_off[lv - 1] + c;
                const float4 lo = __ldg(ix.box_lo + b), hi = __ldg(ix.box_hi + b);
                bits = __float_as_uint(boxbox_lb2(glo, ghi, lo, hi));
                const float cx = 0.5f * (lo.x + hi.x) - gcx, cy = 0.5f * (lo.y + hi.y) - gcy, cz = 0.5f * (lo.z + hi.z) - gcz;
                cbits = __float_as_uint(cx * cx + cy * cy + cz * cz);   // ordering hint only, never used to prune
                if (lv == 1) { boxes[2 * lane] = lo; boxes[2 * lane + 1] = hi; }
            }
            lbs[lv][lane] = bits;
            cds[lv][lane] = cbits;
            __syncwarp();
        };
        // nearest unvisited child of the open node on `lv` that can still matter: -1 if none.  Marks it visited.
        auto pick = [&](int lv) -> int {
            const unsigned mine = lbs[lv][lane];
            const unsigned best = __reduce_min_sync(kFullMask, mine);
            // exact prune: every point below a child is at least `best` away from every query of the group
            if (best == kDone || best > group_gate) return -1;
            unsigned tie = __ballot_sync(kFullMask, mine == best);
            if (tie & (tie - 1)) {   // several children at the same bound (typically 0: they overlap the group box)
                const unsigned cd = (mine == best) ? cds[lv][lane] : kDone;
                const unsigned cbest = __reduce_min_sync(kFullMask, cd);
                tie = __ballot_sync(kFullMask, cd == cbest);
            }
            const int sel = __ffs(tie) - 1;
            if (lane == sel) lbs[lv][lane] = kDone;
            __syncwarp();
            return sel;
        };
        int level = top_level;
        open_node(top_level, 0u);
        while (true) {
            int sel = pick(level);
            if (sel < 0) {
                if (level == top_level) break;
                ++level;
                continue;
            }
            if (level > 1) {
                const unsigned child = nodes[level] * kFan + sel;
                --level;
                open_node(level, child);
                continue;
            }
            // level 1: buckets.  Software pipeline: the points of the NEXT needed bucket in visiting order are
            // loaded while the current one is scanned (the order depends only on the box bounds; the gates only
            // shrink, so a bucket rejected now stays rejected).
            const float4 *base = ix.pts + (long long)nodes[1] * kFan * kLeaf + lane;
            while (sel >= 0 && !bucket_needed(sel)) sel = pick(1);
            if (sel < 0) continue;
            float4 cur = __ldg(base + sel * kLeaf);
            while (true) {
                int nxt = pick(1);
                while (nxt >= 0 && !bucket_needed(nxt)) nxt = pick(1);
                float4 pf = cur;
                if (nxt >= 0) pf = __ldg(base + nxt * kLeaf);
                scan_bucket(cur);
                if (nxt < 0) break;
                sel = nxt;
                cur = pf;
                if (n_scan > max_scans || n_test > max_tests) break;
            }
            if (n_scan > max_scans || n_test > max_tests) { aborted = true; break; }
        }
    }
    if (aborted) {
        // work budget exceeded (a sparse group whose box is much larger than its neighbourhoods): hand the
        // group's queries to the warp-per-query kernel, which re-does them in parallel
        if (lane == 0) overflow[1 + atomicAdd(overflow, 1u)] = (unsigned)group;
        if (ctr && lane == 0) {
            atomicAdd(ctr + 0, (unsigned long long)n_scan * kLeaf * n_active);
            atomicAdd(ctr + 6, 1ull);
        }
        return;
    }

    if (ctr && lane == 0) {
        atomicAdd(ctr + 0, (unsigned long long)n_scan * kLeaf * n_active);
        atomicAdd(ctr + 1, (unsigned long long)n_open * kFan + n_test);
        atomicAdd(ctr + 2, (unsigned long long)n_open);
        atomicAdd(ctr + 3, (unsigned long long)n_round);
        atomicAdd(ctr + 4, 1ull);
        atomicAdd(ctr + 5, (unsigned long long)n_scan);
        if (dbg) { dbg[group * 4 + 0] = n_scan; dbg[group * 4 + 1] = n_test; dbg[group * 4 + 2] = n_round; dbg[group * 4 + 3] = n_open; }
    }
    if (!active) return;
    int cnt = 0;
#pragma unroll
    for (int i = 0; i < K; ++i) {
        if (i < k) {
            const unsigned id = key_index(top.key[i]);
            const bool filled = id != 0xffffffffu;
            cnt += filled ? 1 : 0;
            // result_set.py:19-22: unfilled slots read (1e10, 0)
            out_idx[qid * k + i] = filled ? (int)id : 0;
            out_dist[qid * k + i] = filled ? key_dist(top.key[i]) : PCP_KNN_EMPTY_DIST;
        }
    }
    if (out_cnt) out_cnt[qid] = cnt;
}

template <bool SELF>
int launch_tile(pcp_index *ix, const float *d_q, const unsigned *d_perm, long long nq, long long stride,
                int k, int32_t *d_idx, float *d_dist, int32_t *d_count) {
    pcp_ctx *ctx = ix->ctx;
    const IndexView v = make_view(ix);
    const long long n_groups = SELF ? v.n_leaf : (nq + 31) / 32;
    const unsigned grid = (unsigned)((n_groups + kTileWarps - 1) / kTileWarps);
    const unsigned block = kTileWarps * 32;
    unsigned long long *ctr = ctx->ctr();
    // overflow list: [0] = count, [1..] = group ids
    PCP_TRY(ctx->scratch[5].ensure(sizeof(unsigned) * (size_t)(n_groups + 1)));
    unsigned *overflow = ctx->scratch[5].as<unsigned>();
    PCP_CUDA(cudaMemsetAsync(overflow, 0, sizeof(unsigned), ctx->stream));
    const int K = k <= 2 ? k : k <= 4 ? 4 : k <= 8 ? 8 : k <= 16 ? 16 : 32;
    const bool small_grid = n_groups < 2LL * ctx->sm_count * 32;
    static const char *e_scans = getenv("PCP_TILE_MAX_SCANS"), *e_tests = getenv("PCP_TILE_MAX_TESTS");   // developer overrides
    const unsigned max_scans = e_scans ? (unsigned)atoi(e_scans) : small_grid ? kScansPerSlot * std::max(K, 4) : 0xffffffffu;
    const unsigned max_tests = e_tests ? (unsigned)atoi(e_tests) : small_grid ? kTestsPerSlot * std::max(K, 4) : 0xffffffffu;
    unsigned *dbg = nullptr;
    const char *dump = ctr ? getenv("PCP_DUMP_GROUPS") : nullptr;     // developer aid: per-group work counters
    if (dump) {
        PCP_TRY(ctx->scratch[4].ensure(sizeof(unsigned) * 4 * (size_t)n_groups));
        dbg = ctx->scratch[4].as<unsigned>();
    }
#define PCP_TILE(KK)                                                                                              \
    knn_tile_kernel<KK, SELF><<<grid, block, 0, ctx->stream>>>(v, d_q, d_perm, nq, stride, k, d_idx, d_dist, d_count, ctr, dbg, \
                                                               overflow, max_scans, max_tests)
    if (k == 1) PCP_TILE(1);
    else if (k == 2) PCP_TILE(2);
    else if (k <= 4) PCP_TILE(4);
    else if (k <= 8) PCP_TILE(8);
    else if (k <= 16) PCP_TILE(16);
    else PCP_TILE(32);
#undef PCP_TILE
    ctx->launches += 1;
    PCP_CUDA(cudaGetLastError());
    if (dump) {
        std::vector<unsigned> h(4 * (size_t)n_groups);
        PCP_CUDA(cudaMemcpyAsync(h.data(), dbg, sizeof(unsigned) * h.size(), cudaMemcpyDeviceToHost, ctx->stream));
        PCP_CUDA(cudaStreamSynchronize(ctx->stream));
        if (FILE *f = fopen(dump, "wb")) { fwrite(h.data(), sizeof(unsigned), h.size(), f); fclose(f); }
    }
    // groups that ran out of budget: one warp per query, persistent grid sized by the device-side count
    return knn_warp_list_device(ix, SELF, d_q, nq, stride, d_perm, overflow, k, d_idx, d_dist, d_count);
}

}  // namespace

int knn_tile_device(pcp_index *ix, const float *d_q, int64_t nq, int64_t stride, const uint32_t *d_perm, int k,
                    int32_t *d_idx, float *d_dist, int32_t *d_count) {
    return launch_tile<false>(ix, d_q, d_perm, nq, stride, k, d_idx, d_dist, d_count);
}

int knn_tile_self_device(pcp_index *ix, int k, int32_t *d_idx, float *d_dist, int32_t *d_count) {
    return launch_tile<true>(ix, nullptr, nullptr, ix->n, 3, k, d_idx, d_dist, d_count);
}

}  // namespace pcp
