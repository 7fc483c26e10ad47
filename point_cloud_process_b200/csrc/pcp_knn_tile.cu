// Exact kNN, tiled: one warp = one GROUP of 32 spatially adjacent queries, one query per lane.
// Replaces kdtree_knn_search (kdtree.py:141-172), octree_knn_search (octree.py:262-306) and
// KNNResultSet.add_point (result_set.py:35-53) for batched query sets.
//
// The group walks the 32-ary AABB hierarchy ONCE for all its queries, nearest-first, pruning with
// the exact box-to-box lower bound against the loosest k-th distance gate in the group.  Every
// surviving bucket (32 points, 512 B) is staged in shared memory with one coalesced float4 load per
// lane and read back as 32 broadcast LDS.128: each lane evaluates its own query against all 32
// points (8 unfused FP32 ops per pair, the reference's operation order).  A candidate inside the
// lane's gate is APPENDED to the lane's column of a shared-memory candidate buffer with one
// predicated store -- no sorted insertion, no divergence in the scan loop (12 instructions per
// distance evaluation, 8 of them FP32 pipe work: the kernel the FP32 roofline is quoted on).
// When a column could overflow on the next scan the whole warp COMPACTS: each lane selects the
// k-th smallest squared distance of its column with a register min/max network (2 FMNMX per slot),
// tightens its gate to it and drops the candidates beyond it.  The first scan (gate = +inf) fills the
// columns with 32 candidates and so bootstraps the gates.  After the walk one last compaction leaves
// k (+ distance ties) candidates per lane, which are ranked by the packed (float32 sqrt, index) key
// -- KNNResultSet order made canonical (distance, then index).
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

constexpr unsigned kDone = 0xffffffffu;
constexpr unsigned kInfBits = 0x7f800000u;
// Work budget of one group, per result slot K, before it is handed to the warp-per-query kernel.  It only
// applies when the whole grid fits in about two waves of resident warps (e.g. the 124 668-point KITTI scan =
// 3 896 groups): there the kernel lasts as long as its slowest group, and the sparse outlier groups (a box far
// larger than its queries' neighbourhoods) are 8x the mean.  Larger inputs run unbounded.
constexpr unsigned kScansPerSlot = 4;
constexpr unsigned kTestsPerSlot = 12;

// Candidate-buffer geometry per result capacity K: CAP entries per lane (a chunk of 8 evaluations appends at
// most 8, a compaction leaves k + ties), WARPS groups per CTA.  Shared memory per CTA = WARPS * CAP * 256 B.
constexpr int kChunk = 8;
template <int K> struct TileCfg { static constexpr int kCap = 29, kWarps = 4; };
template <> struct TileCfg<16> { static constexpr int kCap = 40, kWarps = 4; };
template <> struct TileCfg<32> { static constexpr int kCap = 72, kWarps = 2; };
// Children whose box overlaps the group box (lower bound 0) are visited in the order of their centre distance:
// the slot then holds the distance's float bits >> 9, i.e. a value below kMinNormalBits that never prunes.
constexpr unsigned kMinNormalBits = 0x00800000u;

template <int K>
struct TopK {
    unsigned long long key[K];

    __device__ __forceinline__ void init() {
#pragma unroll
        for (int i = 0; i < K; ++i) key[i] = kEmptyKey;
    }
    __device__ __forceinline__ unsigned long long worst() const { return key[K - 1]; }
    // sorted insertion of nk (a no-op when nk >= key[K-1])
    __device__ __forceinline__ void insert(unsigned long long nk) {
        bool g[K];
#pragma unroll
        for (int i = 0; i < K; ++i) g[i] = key[i] > nk;
#pragma unroll
        for (int i = K - 1; i > 0; --i) key[i] = g[i - 1] ? key[i - 1] : (g[i] ? nk : key[i]);
        key[0] = g[0] ? nk : key[0];
    }
};

template <int K, bool SELF>
__global__ void __launch_bounds__(TileCfg<K>::kWarps * 32)
knn_tile_kernel(const IndexView ix, const float *__restrict__ queries, const unsigned *__restrict__ perm, long long nq,
                long long stride, int k, int *__restrict__ out_idx, float *__restrict__ out_dist,
                int *__restrict__ out_cnt, unsigned long long *__restrict__ ctr, unsigned *__restrict__ dbg,
                unsigned *__restrict__ overflow, unsigned max_scans, unsigned max_tests, int trigger) {
    constexpr int kTileWarps = TileCfg<K>::kWarps;
    constexpr int kCap = TileCfg<K>::kCap;
    extern __shared__ uint2 s_cand[];                       // [kTileWarps][kCap][32]: (d2 bits, index) per lane column
    __shared__ float4 s_tile[kTileWarps][32];
    __shared__ float4 s_box[kTileWarps][128];
    __shared__ unsigned s_lb[kTileWarps][kMaxLevels][32];
    __shared__ unsigned s_node[kTileWarps][kMaxLevels];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long n_groups = SELF ? ix.n_leaf : (nq + 31) / 32;
    unsigned *sched = overflow + n_groups + 1;                  // next group to hand out (zeroed by the host)
    // Persistent warps: groups are handed out dynamically, so the kernel ends within one group's work of the
    // moment the queue runs dry instead of waiting for the unluckiest statically assigned warp.
    for (;;) {
    long long group;
    {
        unsigned g = 0;
        if (lane == 0) g = atomicAdd(sched, 1u);
        group = __shfl_sync(kFullMask, g, 0);
    }
    if (group >= n_groups) break;

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

    const float inf = __uint_as_float(kInfBits);
    float gate = active ? inf : -1.f;                           // squared-distance gate of this lane (inactive: nothing passes)
    unsigned group_gate = kInfBits;                             // bits of the loosest gate in the group
    unsigned n_scan = 0, n_half = 0, n_open = 0, n_test = 0, n_round = 0;
    bool aborted = false;
    const unsigned n_active = __popc(__ballot_sync(kFullMask, active));

    float4 *tile = s_tile[warp];
    float4 *boxes = s_box[warp];            // half-bucket boxes below the open level-1 node: [4*c + 2*h] = lo, [4*c + 2*h + 1] = hi
    uint2 *col = s_cand + warp * (kCap * 32) + lane;            // this lane's column: entry e at col[e * 32]
    const unsigned col_addr = (unsigned)__cvta_generic_to_shared(col);
    unsigned waddr = col_addr;                                  // append position (shared-memory byte address; 256 B per entry)

    // Running k smallest squared distances of this lane's column, ascending (k < K: the first K - k slots hold
    // a sentinel below every distance, so that l[K-1] is always the k-th smallest).
    float l[K];
#pragma unroll
    for (int i = 0; i < K; ++i) l[i] = i < K - k ? -1.f : inf;
    int n_old = 0;                                              // entries of the column already merged into l[]

    // Merge the candidates appended since the last compaction into l[] (register min/max network), tighten the
    // gates to the k-th smallest squared distance and drop the candidates beyond it.  Warp-uniform loops.
    auto compact = [&]() {
        const int cnt = (int)((waddr - col_addr) >> 8);
        const int n_new = __reduce_max_sync(kFullMask, cnt - n_old);
        const int nmax = __reduce_max_sync(kFullMask, cnt);
        const uint2 *src_new = col + n_old * 32;
#pragma unroll 2
        for (int e = 0; e < n_new; ++e) {
            float v = n_old + e < cnt ? __uint_as_float(src_new[e * 32].x) : inf;
            // the lists tighten as the merge proceeds: a round in which no lane's candidate still beats its
            // k-th distance changes nothing
            if (!__any_sync(kFullMask, v < l[K - 1])) continue;
            ++n_round;
#pragma unroll
            for (int i = 0; i < K; ++i) {
                const float lo = fminf(v, l[i]);
                v = fmaxf(v, l[i]);
                l[i] = lo;
            }
        }
        // candidates whose float32 sqrt can still tie the k-th distance stay (canonical (distance, index) order)
        if (l[K - 1] < inf) gate = fminf(gate, d2_gate(__fsqrt_rn(l[K - 1])));
        // In-place filter.  The write position never passes the read position, so loading 4 entries ahead of the
        // stores they may alias is safe (and lets the loads overlap); the loads may run up to 3 entries past the
        // longest column (the buffer is padded for it), the `< end` test discards them.
        unsigned w = col_addr;
        const unsigned end = waddr;
#pragma unroll 1
        for (unsigned r = col_addr; r < col_addr + (unsigned)nmax * 256u; r += 4 * 256u) {
            float d[4], x[4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
                asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=f"(d[i]), "=f"(x[i]) : "r"(r + i * 256u) : "memory");
#pragma unroll
            for (int i = 0; i < 4; ++i)
                if (r + i * 256u < end && d[i] <= gate) {
                    asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(w), "f"(d[i]), "f"(x[i]) : "memory");
                    w += 256u;
                }
        }
        waddr = w;
        n_old = (int)((w - col_addr) >> 8);
        group_gate = __reduce_max_sync(kFullMask, active ? __float_as_uint(gate) : 0u);
    };

    // Exact per-lane lower bounds against the boxes of bucket c's two 16-point halves (shared-memory broadcast):
    // bit h is set when ANY lane can still use half h.
    auto bucket_needed = [&](int c) -> unsigned {
        ++n_test;
        const bool a = box_lb2(qx, qy, qz, boxes[4 * c], boxes[4 * c + 1]) <= gate;
        const bool b = box_lb2(qx, qy, qz, boxes[4 * c + 2], boxes[4 * c + 3]) <= gate;
        return (__any_sync(kFullMask, a) ? 1u : 0u) | (__any_sync(kFullMask, b) ? 2u : 0u);
    };
    // Scan the needed halves of one bucket whose points `pv` (lane j = point j) are already in registers.
    auto scan_bucket = [&](const float4 pv, unsigned halves) {
        ++n_scan;
        n_half += __popc(halves);
        __syncwarp();
        tile[lane] = pv;
        __syncwarp();
        // 8 points at a time: the broadcast loads are issued ahead of the (possibly aliasing) candidate stores
#pragma unroll 1
        for (int j0 = 0; j0 < 32; j0 += kChunk) {
            if (!((halves >> (j0 >> 4)) & 1u)) continue;
            float4 p[kChunk];
#pragma unroll
            for (int j = 0; j < kChunk; ++j) p[j] = tile[j0 + j];
#pragma unroll
            for (int j = 0; j < kChunk; ++j) {
                const float d2 = ref_d2(qx, qy, qz, p[j].x, p[j].y, p[j].z);
                if (d2 <= gate) {                                // NaN pads never pass
                    asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(waddr), "f"(d2), "f"(p[j].w) : "memory");
                    waddr += 256;
                }
            }
            // a column that could overflow during the next chunk: compact all columns of the warp
            if (__any_sync(kFullMask, waddr > col_addr + (unsigned)trigger * 256)) {
                compact();
                // still too long = a mass of exactly tied distances: leave the group to the warp-per-query kernel
                if (__any_sync(kFullMask, waddr > col_addr + (unsigned)(kCap - kChunk) * 256)) { aborted = true; break; }
            }
        }
    };

    const int top_level = ix.n_levels - 1;      // >= 1 for every non-empty index (pcp_build.cu)
    if (top_level > 0) {
        unsigned(*lbs)[32] = s_lb[warp];
        unsigned *nodes = s_node[warp];
        auto open_node = [&](int lv, unsigned node) {
            ++n_open;
            if (lane == 0) nodes[lv] = node;
            const long long c = (long long)node * kFan + lane;
            unsigned bits = kDone;
            if (c < ix.level_cnt[lv - 1]) {
                float4 lo, hi;
                if (lv == 1) {
                    // buckets: the boxes of the two halves; the bucket's bound is the smaller of theirs
                    const float4 *h = ix.half + c * 4;
                    const float4 alo = __ldg(h), ahi = __ldg(h + 1), blo = __ldg(h + 2), bhi = __ldg(h + 3);
                    boxes[4 * lane] = alo; boxes[4 * lane + 1] = ahi; boxes[4 * lane + 2] = blo; boxes[4 * lane + 3] = bhi;
                    bits = min(__float_as_uint(boxbox_lb2(glo, ghi, alo, ahi)), __float_as_uint(boxbox_lb2(glo, ghi, blo, bhi)));
                    lo = make_float4(fminf(alo.x, blo.x), fminf(alo.y, blo.y), fminf(alo.z, blo.z), 0.f);
                    hi = make_float4(fmaxf(ahi.x, bhi.x), fmaxf(ahi.y, bhi.y), fmaxf(ahi.z, bhi.z), 0.f);
                } else {
                    const long long b = ix.level_off[lv - 1] + c;
                    lo = __ldg(ix.box_lo + b); hi = __ldg(ix.box_hi + b);
                    bits = __float_as_uint(boxbox_lb2(glo, ghi, lo, hi));
                }
                if (bits == 0u) {   // overlaps the group box: order by centre distance (a hint, never used to prune)
                    const float cx = 0.5f * (lo.x + hi.x) - gcx, cy = 0.5f * (lo.y + hi.y) - gcy, cz = 0.5f * (lo.z + hi.z) - gcz;
                    bits = __float_as_uint(cx * cx + cy * cy + cz * cz) >> 9;
                }
            }
            lbs[lv][lane] = bits;
            __syncwarp();
        };
        // nearest unvisited child of the open node on `lv` that can still matter: -1 if none.  Marks it visited.
        auto pick = [&](int lv) -> int {
            const unsigned mine = lbs[lv][lane];
            const unsigned best = __reduce_min_sync(kFullMask, mine);
            // exact prune: every point below a child is at least `best` away from every query of the group
            // (slots below kMinNormalBits are ordering hints of overlapping children or denormal bounds: kept)
            if (best == kDone || (best > group_gate && best >= kMinNormalBits)) return -1;
            const int sel = __ffs(__ballot_sync(kFullMask, mine == best)) - 1;
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
            unsigned halves = 0;
            while (sel >= 0 && !(halves = bucket_needed(sel))) sel = pick(1);
            if (sel < 0) continue;
            float4 cur = __ldg(base + sel * kLeaf);
            while (true) {
                int nxt = pick(1);
                unsigned nxt_halves = 0;
                while (nxt >= 0 && !(nxt_halves = bucket_needed(nxt))) nxt = pick(1);
                float4 pf = cur;
                if (nxt >= 0) pf = __ldg(base + nxt * kLeaf);
                scan_bucket(cur, halves);
                if (nxt < 0) break;
                sel = nxt;
                halves = nxt_halves;
                cur = pf;
                if (aborted || n_scan > max_scans || n_test > max_tests) break;
            }
            if (aborted || n_scan > max_scans || n_test > max_tests) { aborted = true; break; }
        }
    }
    if (aborted) {
        // work budget exceeded (a sparse group whose box is much larger than its neighbourhoods) or a mass of
        // tied distances: hand the group's queries to the warp-per-query kernel, which re-does them in parallel
        if (lane == 0) overflow[1 + atomicAdd(overflow, 1u)] = (unsigned)group;
        if (ctr && lane == 0) {
            atomicAdd(ctr + 0, (unsigned long long)n_half * (kLeaf / 2) * n_active);
            atomicAdd(ctr + 6, 1ull);
        }
        __syncwarp();
        continue;
    }

    // ---- final selection: k (+ ties) candidates per lane, ranked by (float32 sqrt, index)
    compact();
    TopK<K> top;
    top.init();
    {
        const int cnt = (int)((waddr - col_addr) >> 8);
        const int nmax = __reduce_max_sync(kFullMask, cnt);
        for (int e = 0; e < nmax; ++e) {
            if (e < cnt) {
                const uint2 ent = col[e * 32];
                top.insert(pack_key(__fsqrt_rn(__uint_as_float(ent.x)), ent.y));
            }
        }
    }

    if (ctr && lane == 0) {
        atomicAdd(ctr + 0, (unsigned long long)n_half * (kLeaf / 2) * n_active);
        atomicAdd(ctr + 1, (unsigned long long)n_open * kFan + n_test);
        atomicAdd(ctr + 2, (unsigned long long)n_open);
        atomicAdd(ctr + 3, (unsigned long long)n_round);
        atomicAdd(ctr + 4, 1ull);
        atomicAdd(ctr + 5, (unsigned long long)n_scan);
        if (dbg) { dbg[group * 4 + 0] = n_scan; dbg[group * 4 + 1] = n_test; dbg[group * 4 + 2] = n_round; dbg[group * 4 + 3] = n_open; }
    }
    if (active) {
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
    __syncwarp();
    }   // next group
}

template <bool SELF>
int launch_tile(pcp_index *ix, const float *d_q, const unsigned *d_perm, long long nq, long long stride,
                int k, int32_t *d_idx, float *d_dist, int32_t *d_count) {
    pcp_ctx *ctx = ix->ctx;
    const IndexView v = make_view(ix);
    const long long n_groups = SELF ? v.n_leaf : (nq + 31) / 32;
    unsigned long long *ctr = ctx->ctr();
    // overflow list: [0] = count, [1..] = group ids
    // [n_groups + 1] = next group of the persistent tile warps
    PCP_TRY(ctx->scratch[5].ensure(sizeof(unsigned) * (size_t)(n_groups + 2)));
    unsigned *overflow = ctx->scratch[5].as<unsigned>();
    PCP_CUDA(cudaMemsetAsync(overflow, 0, sizeof(unsigned), ctx->stream));
    PCP_CUDA(cudaMemsetAsync(overflow + n_groups + 1, 0, sizeof(unsigned), ctx->stream));
    const int K = k <= 2 ? k : k <= 4 ? 4 : k <= 8 ? 8 : k <= 16 ? 16 : 32;
    const bool small_grid = n_groups < 2LL * ctx->sm_count * 32;
    static const char *e_scans = getenv("PCP_TILE_MAX_SCANS"), *e_tests = getenv("PCP_TILE_MAX_TESTS"),
                      *e_trig = getenv("PCP_TILE_TRIGGER");   // developer overrides
    const unsigned max_scans = e_scans ? (unsigned)atoi(e_scans) : small_grid ? kScansPerSlot * std::max(K, 4) : 0xffffffffu;
    const unsigned max_tests = e_tests ? (unsigned)atoi(e_tests) : small_grid ? kTestsPerSlot * std::max(K, 4) : 0xffffffffu;
    unsigned *dbg = nullptr;
    const char *dump = ctr ? getenv("PCP_DUMP_GROUPS") : nullptr;     // developer aid: per-group work counters
    if (dump) {
        PCP_TRY(ctx->scratch[4].ensure(sizeof(unsigned) * 4 * (size_t)n_groups));
        dbg = ctx->scratch[4].as<unsigned>();
    }
#define PCP_TILE(KK)                                                                                                   \
    do {                                                                                                               \
        constexpr int W = TileCfg<KK>::kWarps, CAP = TileCfg<KK>::kCap;                                                \
        const size_t smem = (size_t)W * CAP * 32 * sizeof(uint2) + 3 * 32 * sizeof(uint2); /* + filter read-ahead */   \
        static int ctas_per_sm = 0;                                                                                    \
        if (!ctas_per_sm) {                                                                                            \
            PCP_CUDA(cudaFuncSetAttribute(knn_tile_kernel<KK, SELF>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
            PCP_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas_per_sm, knn_tile_kernel<KK, SELF>, W * 32, smem)); \
            ctas_per_sm = std::max(ctas_per_sm, 1);                                                                    \
        }                                                                                                              \
        /* compaction trigger: a column longer than this could overflow during the next chunk of evaluations */       \
        const int trig = e_trig ? std::min(atoi(e_trig), CAP - kChunk) : CAP - kChunk;                                 \
        /* persistent grid: every resident warp slot, or fewer when there are not that many groups */                  \
        const unsigned grid = (unsigned)std::min<long long>((n_groups + W - 1) / W, (long long)ctx->sm_count * ctas_per_sm); \
        knn_tile_kernel<KK, SELF><<<grid, W * 32, smem, ctx->stream>>>(v, d_q, d_perm, nq, stride, k, d_idx, d_dist,   \
                                                                       d_count, ctr, dbg, overflow, max_scans,         \
                                                                       max_tests, trig);                               \
    } while (0)
    cudaEvent_t ev_end = nullptr;
    if (ctx->time_kernels) {
        if (ctx->timing_used == ctx->timing_events.size()) {
            cudaEvent_t a, b;
            PCP_CUDA(cudaEventCreate(&a));
            PCP_CUDA(cudaEventCreate(&b));
            ctx->timing_events.emplace_back(a, b);
        }
        PCP_CUDA(cudaEventRecord(ctx->timing_events[ctx->timing_used].first, ctx->stream));
        ev_end = ctx->timing_events[ctx->timing_used].second;
        ++ctx->timing_used;
    }
    if (k == 1) PCP_TILE(1);
    else if (k == 2) PCP_TILE(2);
    else if (k <= 4) PCP_TILE(4);
    else if (k <= 8) PCP_TILE(8);
    else if (k <= 16) PCP_TILE(16);
    else PCP_TILE(32);
#undef PCP_TILE
    if (ev_end) PCP_CUDA(cudaEventRecord(ev_end, ctx->stream));
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
