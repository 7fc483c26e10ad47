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
// Tail hand-over: with the queue dry and fewer groups in flight than this share of the warp slots, the groups
// still running are handed over too (see the kernel).
constexpr double kTailShare = 0.05;
constexpr long long kHintGroups = 1 << 16;     // overflow groups whose gates are passed on (8 MB of scratch)

// Candidate-buffer geometry per result capacity K: CAP entries per lane (a half-chunk of 4 evaluations appends
// at most 4, a compaction leaves k + ties), WARPS groups per CTA.  The per-warp shared-memory state is
//   CAP * 256 B candidate columns + 512 B staged bucket + 1536 B half-bucket boxes + 128 B per tree level + 32 B,
// sized so that K <= 8 runs 28 warps per SM (7 CTAs; 72 registers) and K = 16 runs 20 (5 CTAs; 96 registers):
// the kernel is latency-bound, and at 28 warps/SM the 3 896 groups of a KITTI scan are all resident at once.
constexpr int kChunk = 8;                      // points loaded per step; the overflow check runs every kChunk / 2
template <int K> struct TileCfg { static constexpr int kCap = 21, kWarps = 4, kCtas = 7; };
template <> struct TileCfg<16> { static constexpr int kCap = 34, kWarps = 4, kCtas = 5; };
template <> struct TileCfg<32> { static constexpr int kCap = 68, kWarps = 2, kCtas = 8; };
constexpr int kBoxFloats = 64 * 6;             // 64 half boxes x (lo.xyz, hi.xyz)
__host__ __device__ constexpr size_t tile_warp_bytes(int cap, int levels) {
    return (size_t)cap * 256 + 512 + kBoxFloats * 4 + (size_t)levels * 128 + 32;
}
// lower bound against a box stored as 6 floats (same arithmetic as box_lb2)
__device__ __forceinline__ float box6_lb2(float qx, float qy, float qz, const float *b) {
    const float2 l01 = *reinterpret_cast<const float2 *>(b), l2h0 = *reinterpret_cast<const float2 *>(b + 2),
                 h12 = *reinterpret_cast<const float2 *>(b + 4);
    float dx = fmaxf(fmaxf(__fsub_rn(l01.x, qx), __fsub_rn(qx, l2h0.y)), 0.f);
    float dy = fmaxf(fmaxf(__fsub_rn(l01.y, qy), __fsub_rn(qy, h12.x)), 0.f);
    float dz = fmaxf(fmaxf(__fsub_rn(l2h0.x, qz), __fsub_rn(qz, h12.y)), 0.f);
    return __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
}
// free-running device timer in 32 ns units (wraps after 137 s; compared with wrap-safe differences)
__device__ __forceinline__ unsigned timer32() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return (unsigned)(t >> 5);
}
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
__global__ void __launch_bounds__(TileCfg<K>::kWarps * 32, TileCfg<K>::kCtas)
knn_tile_kernel(const IndexView ix, const float *__restrict__ queries, const unsigned *__restrict__ perm, long long nq,
                long long stride, int k, int *__restrict__ out_idx, float *__restrict__ out_dist,
                int *__restrict__ out_cnt, unsigned long long *__restrict__ ctr, unsigned *__restrict__ dbg,
                unsigned *__restrict__ overflow, unsigned max_scans, unsigned max_tests, int trigger,
                unsigned tail_slots, unsigned *__restrict__ sched, unsigned n_groups, float *__restrict__ hints,
                unsigned hint_groups, unsigned warp_stride) {
    constexpr int kTileWarps = TileCfg<K>::kWarps;
    constexpr int kCap = TileCfg<K>::kCap;
    // dynamic shared memory: [kTileWarps][kCap][32] candidate columns (d2 bits, index) first -- the filter's
    // read-ahead past the last column lands in the arrays behind them --, then per warp: staged bucket,
    // half-bucket boxes, open node per level, one row of child bounds per tree level
    extern __shared__ uint2 s_cand[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // (warp_stride = tile_warp_bytes(0, levels), a kernel parameter: the per-warp arrays sit at constant offsets
    // from one base that costs a single multiply-add to re-derive)
    unsigned char *s_rest = reinterpret_cast<unsigned char *>(s_cand + kTileWarps * kCap * 32) + warp * warp_stride;
    // sched (behind the overflow list): [0] next group to hand out, [1] groups finished (zeroed by the host);
    // it and n_groups are kernel parameters so that the tail poll below holds no registers across the hot loops
    // Persistent warps: groups are handed out dynamically, so the kernel ends within one group's work of the
    // moment the queue runs dry instead of waiting for the unluckiest statically assigned warp.
    // Tail hand-over (tail_slots != 0): once the queue is dry the kernel lasts as long as its slowest group
    // while most warp slots idle (a group whose 32 queries straddle a jump of the Morton curve tests thousands of
    // boxes and runs 10x the mean).  sched[1] counts finished groups; when fewer than tail_slots groups are still
    // in flight the stragglers hand their queries to the warp-per-query kernel (32 warps per group instead of 1).
    for (;;) {
    unsigned group;
    {
        unsigned g = 0;
        if (lane == 0) g = atomicAdd(sched, 1u);
        group = __shfl_sync(kFullMask, g, 0);
    }
    if (group >= n_groups) break;
    if (dbg && lane == 0) {                                     // developer timeline (PCP_DUMP_GROUPS)
        unsigned smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        dbg[(size_t)group * 8 + 4] = timer32();
        dbg[(size_t)group * 8 + 6] = smid;
        dbg[(size_t)group * 8 + 7] = 0u;
    }

    // ---- this lane's query
    float qx = 0.f, qy = 0.f, qz = 0.f;
    long long qid = -1;
    if (SELF) {
        const float4 p = __ldg(ix.pts + (long long)group * kLeaf + lane);
        qx = p.x; qy = p.y; qz = p.z;
        qid = (long long)(int)__float_as_uint(p.w);            // -1 on pad lanes
    } else {
        const long long pos = (long long)group * 32 + lane;
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

    float4 *tile = reinterpret_cast<float4 *>(s_rest);
    float *boxes = reinterpret_cast<float *>(s_rest + 512);   // half boxes below the open level-1 node: 6 floats at [(2*c + h) * 6]
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

    // Tail hand-over poll (every 4th scan / open): queue dry and fewer than tail_slots groups still in flight?
    auto poll_tail = [&]() {
        if (!tail_slots) return;
        uint2 w = make_uint2(0u, 0u);
        if (lane == 0) asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(w.x), "=r"(w.y) : "l"(sched));
        // (sched[0] runs past n_groups by at most one per warp; sched[1] <= n_groups)
        if (__shfl_sync(kFullMask, (w.x >= n_groups && n_groups - w.y < tail_slots) ? 1 : 0, 0)) aborted = true;
    };

    // Exact per-lane lower bounds against the boxes of bucket c's two 16-point halves (shared-memory broadcast):
    // bit h is set when ANY lane can still use half h.
    auto bucket_needed = [&](int c) -> unsigned {
        ++n_test;
        const bool a = box6_lb2(qx, qy, qz, boxes + c * 12) <= gate;
        const bool b = box6_lb2(qx, qy, qz, boxes + c * 12 + 6) <= gate;
        return (__any_sync(kFullMask, a) ? 1u : 0u) | (__any_sync(kFullMask, b) ? 2u : 0u);
    };
    // Scan the needed halves of one bucket whose points `pv` (lane j = point j) are already in registers.
    auto scan_bucket = [&](const float4 pv, unsigned halves) {
        ++n_scan;
        n_half += __popc(halves);
        __syncwarp();
        tile[lane] = pv;
        __syncwarp();
        // 8 points at a time: the broadcast loads are issued ahead of the (possibly aliasing) candidate stores; the
        // columns are checked after every 4 evaluations, so only 4 entries of headroom are reserved
        const int j_end = (halves & 2u) ? 32 : 16;             // (halves != 0: at least one of the two is scanned)
#pragma unroll 1
        for (int j0 = (halves & 1u) ? 0 : 16; j0 < j_end; j0 += kChunk) {
            float4 p[kChunk];
#pragma unroll
            for (int j = 0; j < kChunk; ++j) p[j] = tile[j0 + j];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
#pragma unroll
                for (int j = h * (kChunk / 2); j < (h + 1) * (kChunk / 2); ++j) {
                    const float d2 = ref_d2(qx, qy, qz, p[j].x, p[j].y, p[j].z);
                    if (d2 <= gate) {                            // NaN pads never pass
                        asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(waddr), "f"(d2), "f"(p[j].w) : "memory");
                        waddr += 256;
                    }
                }
                // a column that could overflow during the next 4 evaluations: compact all columns of the warp
                if (__any_sync(kFullMask, waddr > col_addr + (unsigned)trigger * 256)) {
                    compact();
                    // still too long = a mass of exactly tied distances: leave the group to the warp-per-query kernel
                    if (__any_sync(kFullMask, waddr > col_addr + (unsigned)(kCap - kChunk / 2) * 256)) aborted = true;
                }
                if (aborted) break;      // (before the second half appends past the end of the columns)
            }
            if (aborted) break;
        }
    };

    const int top_level = ix.n_levels - 1;      // >= 1 for every non-empty index (pcp_build.cu)
    if (top_level > 0) {
        unsigned *nodes = reinterpret_cast<unsigned *>(s_rest + 512 + kBoxFloats * 4) - 1;               // [lv - 1]
        unsigned(*lbs)[32] = reinterpret_cast<unsigned(*)[32]>(s_rest + 512 + kBoxFloats * 4 + 32) - 1;   // row lv - 1
        auto open_node = [&](int lv, unsigned node) {
            ++n_open;
            if (lv == 1 && (n_open & 3u) == 0u) poll_tail();
            if (lane == 0) nodes[lv] = node;
            const long long c = (long long)node * kFan + lane;
            unsigned bits = kDone;
            if (c < ix.level_cnt[lv - 1]) {
                float4 lo, hi;
                if (lv == 1) {
                    // buckets: the boxes of the two halves; the bucket's bound is the smaller of theirs
                    const float4 *h = ix.half + c * 4;
                    const float4 alo = __ldg(h), ahi = __ldg(h + 1), blo = __ldg(h + 2), bhi = __ldg(h + 3);
                    float2 *bw = reinterpret_cast<float2 *>(boxes + lane * 12);
                    bw[0] = make_float2(alo.x, alo.y); bw[1] = make_float2(alo.z, ahi.x); bw[2] = make_float2(ahi.y, ahi.z);
                    bw[3] = make_float2(blo.x, blo.y); bw[4] = make_float2(blo.z, bhi.x); bw[5] = make_float2(bhi.y, bhi.z);
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
        while (!aborted) {
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
                if ((n_scan & 3u) == 0u) poll_tail();
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
        unsigned pos = 0;
        if (lane == 0) {
            pos = atomicAdd(overflow, 1u);
            overflow[1 + pos] = group;
            atomicAdd(sched + 1, 1u);
        }
        // the gates reached so far seed the searches of the warp-per-query kernel
        pos = __shfl_sync(kFullMask, pos, 0);
        if (pos < hint_groups) hints[(size_t)pos * 32 + lane] = gate;
        if (ctr && lane == 0) {
            atomicAdd(ctr + 0, (unsigned long long)n_half * (kLeaf / 2) * n_active);
            atomicAdd(ctr + 6, 1ull);
            if (dbg) { dbg[(size_t)group * 8 + 0] = n_scan; dbg[(size_t)group * 8 + 1] = n_test; dbg[(size_t)group * 8 + 5] = timer32(); dbg[(size_t)group * 8 + 7] = 1u; }
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

    if (lane == 0) atomicAdd(sched + 1, 1u);
    if (ctr && lane == 0) {
        atomicAdd(ctr + 0, (unsigned long long)n_half * (kLeaf / 2) * n_active);
        atomicAdd(ctr + 1, (unsigned long long)n_open * kFan + n_test);
        atomicAdd(ctr + 2, (unsigned long long)n_open);
        atomicAdd(ctr + 3, (unsigned long long)n_round);
        atomicAdd(ctr + 4, 1ull);
        atomicAdd(ctr + 5, (unsigned long long)n_scan);
        if (dbg) {
            dbg[(size_t)group * 8 + 0] = n_scan; dbg[(size_t)group * 8 + 1] = n_test; dbg[(size_t)group * 8 + 2] = n_round; dbg[(size_t)group * 8 + 3] = n_open;
            dbg[(size_t)group * 8 + 5] = timer32();
        }
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
    // behind the list, 8-byte aligned: next group of the persistent tile warps, number of finished groups
    // and the gates of the first kHintGroups overflow groups (32 floats each)
    const long long sched_at = (n_groups + 2) & ~1LL, hints_at = (sched_at + 2 + 31) & ~31LL;
    const unsigned hint_groups = (unsigned)std::min<long long>(n_groups, kHintGroups);
    PCP_TRY(ctx->scratch[5].ensure(sizeof(unsigned) * (size_t)(hints_at + 32LL * hint_groups)));
    unsigned *overflow = ctx->scratch[5].as<unsigned>();
    float *hints = reinterpret_cast<float *>(overflow + hints_at);
    PCP_CUDA(cudaMemsetAsync(overflow, 0, sizeof(unsigned), ctx->stream));
    PCP_CUDA(cudaMemsetAsync(overflow + sched_at, 0, 2 * sizeof(unsigned), ctx->stream));
    const int K = k <= 2 ? k : k <= 4 ? 4 : k <= 8 ? 8 : k <= 16 ? 16 : 32;
    const bool small_grid = n_groups < 2LL * ctx->sm_count * 32;
    static const char *e_scans = getenv("PCP_TILE_MAX_SCANS"), *e_tests = getenv("PCP_TILE_MAX_TESTS"),
                      *e_trig = getenv("PCP_TILE_TRIGGER"), *e_tail = getenv("PCP_TILE_TAIL");   // developer overrides
    const double tail_share = e_tail ? atof(e_tail) : kTailShare;   // share of the warp slots; 0 = off
    const unsigned max_scans = e_scans ? (unsigned)atoi(e_scans) : small_grid ? kScansPerSlot * std::max(K, 4) : 0xffffffffu;
    const unsigned max_tests = e_tests ? (unsigned)atoi(e_tests) : small_grid ? kTestsPerSlot * std::max(K, 4) : 0xffffffffu;
    unsigned *dbg = nullptr;
    const char *dump = ctr ? getenv("PCP_DUMP_GROUPS") : nullptr;     // developer aid: per-group work counters
    if (dump) {
        PCP_TRY(ctx->scratch[4].ensure(sizeof(unsigned) * 8 * (size_t)n_groups));
        dbg = ctx->scratch[4].as<unsigned>();
    }
#define PCP_TILE(KK)                                                                                                   \
    do {                                                                                                               \
        constexpr int W = TileCfg<KK>::kWarps, CAP = TileCfg<KK>::kCap;                                                \
        const int n_lv = v.n_levels > 1 ? v.n_levels - 1 : 1;                                                          \
        const size_t smem = (size_t)W * tile_warp_bytes(CAP, n_lv);                                                   \
        static int occ_by_levels[kMaxLevels + 1] = {0};   /* resident CTAs per SM, per tree depth */                   \
        int &ctas_per_sm = occ_by_levels[n_lv];                                                                        \
        if (!ctas_per_sm) {                                                                                            \
            static size_t attr_smem = 0;                                                                               \
            if (smem > attr_smem) {                                                                                    \
                PCP_CUDA(cudaFuncSetAttribute(knn_tile_kernel<KK, SELF>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
                attr_smem = smem;                                                                                      \
            }                                                                                                          \
            PCP_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas_per_sm, knn_tile_kernel<KK, SELF>, W * 32, smem)); \
            ctas_per_sm = std::max(ctas_per_sm, 1);                                                                    \
            if (const char *e = getenv("PCP_TILE_CTAS")) ctas_per_sm = std::max(1, std::min(ctas_per_sm, atoi(e)));     \
            if (getenv("PCP_VERBOSE")) fprintf(stderr, "knn_tile_kernel<%d>: %zu B smem, %d CTAs/SM\n", KK, smem, ctas_per_sm); \
        }                                                                                                              \
        /* compaction trigger: a column longer than this could overflow during the next chunk of evaluations */       \
        const int trig = e_trig ? std::min(atoi(e_trig), CAP - kChunk / 2) : CAP - kChunk / 2;                         \
        /* persistent grid: every resident warp slot, or fewer when there are not that many groups */                  \
        const unsigned grid = (unsigned)std::min<long long>((n_groups + W - 1) / W, (long long)ctx->sm_count * ctas_per_sm); \
        const unsigned tail_slots = (unsigned)(tail_share * (double)grid * W);                                        \
        knn_tile_kernel<KK, SELF><<<grid, W * 32, smem, ctx->stream>>>(v, d_q, d_perm, nq, stride, k, d_idx, d_dist,   \
                                                                       d_count, ctr, dbg, overflow, max_scans,         \
                                                                       max_tests, trig, tail_slots,                    \
                                                                       overflow + sched_at, (unsigned)n_groups, hints, \
                                                                       hint_groups,                                    \
                                                                       (unsigned)tile_warp_bytes(0, n_lv));            \
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
        std::vector<unsigned> h(8 * (size_t)n_groups);
        PCP_CUDA(cudaMemcpyAsync(h.data(), dbg, sizeof(unsigned) * h.size(), cudaMemcpyDeviceToHost, ctx->stream));
        PCP_CUDA(cudaStreamSynchronize(ctx->stream));
        if (FILE *f = fopen(dump, "wb")) { fwrite(h.data(), sizeof(unsigned), h.size(), f); fclose(f); }
    }
    // groups that ran out of budget: one warp per query, persistent grid sized by the device-side count
    return knn_warp_list_device(ix, SELF, d_q, nq, stride, d_perm, overflow, hints, hint_groups, k, d_idx, d_dist, d_count);
}

}  // namespace

int knn_tile_device_v1(pcp_index *ix, const float *d_q, int64_t nq, int64_t stride, const uint32_t *d_perm, int k,
                    int32_t *d_idx, float *d_dist, int32_t *d_count) {
    return launch_tile<false>(ix, d_q, d_perm, nq, stride, k, d_idx, d_dist, d_count);
}

int knn_tile_self_device_v1(pcp_index *ix, int k, int32_t *d_idx, float *d_dist, int32_t *d_count) {
    return launch_tile<true>(ix, nullptr, nullptr, ix->n, 3, k, d_idx, d_dist, d_count);
}

}  // namespace pcp
