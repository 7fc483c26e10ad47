// Exact kNN, tiled: one warp = one GROUP of 32 spatially adjacent queries, one query per lane.
// Replaces kdtree_knn_search (kdtree.py:141-172), octree_knn_search (octree.py:262-306) and
// KNNResultSet.add_point (result_set.py:35-53) for batched query sets.
//
// The group walks the 32-ary AABB hierarchy ONCE for all its queries, nearest-first, pruning with the
// exact box-to-box lower bound against the loosest k-th distance gate in the group.  Every surviving
// bucket (32 points, 512 B in PAIR layout, IndexView::pts2) is staged in shared memory with one coalesced
// float4 load per lane and read back as broadcast LDS.128; each lane evaluates its own query against TWO
// points per step in packed FP32 (3 FADD2 + 3 FMUL2 + 4 FADD for two distances, every element rounded once
// per operation in the reference's order -- bit-equal to the scalar expression, 5 issue slots per evaluation
// instead of 8).  A candidate inside the lane's gate is APPENDED to the lane's column of a shared-memory
// candidate buffer with one predicated store -- no sorted insertion, no divergence in the scan loop.
//
// Gates.  The FIRST bucket a group scans (its own bucket for pcp_knn_self, the bucket nearest to the group
// otherwise) never touches the columns: its 32 distances stay in registers and a sorting network (batches of
// 8 sorted with 19 compare-exchanges, merged into the running list with a half-cleaner + bitonic merge) yields
// the k smallest, so every lane owns a tight gate before the walk proper starts and only the k (+ ties) survivors
// are appended.  Later, when a column could overflow, the whole warp COMPACTS: the candidates appended since
// the last compaction are merged into the register list with the same batch network, the gate tightens to the
// k-th smallest squared distance, and candidates beyond it are dropped in place.  After the walk the k (+ ties)
// survivors are ranked by the packed (float32 sqrt, index) key -- KNNResultSet order made canonical.
//
// Level-1 traversal state (one lower bound per child bucket) lives in a register per lane: the nearest unvisited
// child is ONE redux.min over (bound bits | lane); the per-lane "does any lane still need this half bucket" test
// evaluates both 16-point halves at once in packed FP32 against boxes staged pair-wise in shared memory.
//
// Query groups: for pcp_knn_self the buckets themselves; for arbitrary query sets the queries are first sorted by
// Morton code in the index's frame and regrouped by the kd rule (sort_queries).
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

// Candidate-buffer geometry per result capacity K: CAP entries per lane (4 evaluations append at most 4 between two
// overflow checks, a compaction leaves k + ties), WARPS groups per CTA.  Per-warp shared memory:
//   CAP * 256 B candidate columns + 512 B staged bucket + 1536 B half-bucket boxes + 32 B + 128 B per level >= 2.
template <int K> struct TileCfg { static constexpr int kCap = 21, kWarps = 4, kCtas = 7; };
template <> struct TileCfg<16> { static constexpr int kCap = 34, kWarps = 4, kCtas = 5; };
template <> struct TileCfg<32> { static constexpr int kCap = 68, kWarps = 2, kCtas = 8; };
constexpr int kBoxBytes = 32 * 48;             // per child bucket: 3 float4 = the two half boxes, pair-interleaved
__host__ __device__ constexpr size_t tile_warp_bytes(int cap, int upper_levels) {
    return (size_t)cap * 256 + 512 + kBoxBytes + 32 + (size_t)upper_levels * 128;
}
// Euclidean distances above KNNResultSet's initial worst_dist are never accepted (result_set.py:19,37): the
// initial gate is the squared image of 1e10, finite, so the +inf distances of pad lanes never pass it either.
constexpr float kWorstDist0 = 1e10f;

// free-running device timer in 32 ns units (wraps after 137 s; compared with wrap-safe differences)
__device__ __forceinline__ unsigned timer32() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return (unsigned)(t >> 5);
}
// Children whose box overlaps the group box (lower bound 0) are visited in the order of their centre distance:
// the slot then holds the distance's float bits >> 9, i.e. a value below kMinNormalBits that never prunes.
constexpr unsigned kMinNormalBits = 0x00800000u;

#define PCP_CE(a, b)                   \
    do {                               \
        const float _t = fminf(a, b);  \
        b = fmaxf(a, b);               \
        a = _t;                        \
    } while (0)

// optimal sorting networks: 5 compare-exchanges for 4 values, 19 for 8 (ascending)
__device__ __forceinline__ void sort_net(float (&s)[4]) {
    PCP_CE(s[0], s[1]); PCP_CE(s[2], s[3]); PCP_CE(s[0], s[2]); PCP_CE(s[1], s[3]); PCP_CE(s[1], s[2]);
}
__device__ __forceinline__ void sort_net(float (&s)[8]) {
    PCP_CE(s[0], s[1]); PCP_CE(s[2], s[3]); PCP_CE(s[4], s[5]); PCP_CE(s[6], s[7]);
    PCP_CE(s[0], s[2]); PCP_CE(s[1], s[3]); PCP_CE(s[4], s[6]); PCP_CE(s[5], s[7]);
    PCP_CE(s[1], s[2]); PCP_CE(s[5], s[6]); PCP_CE(s[0], s[4]); PCP_CE(s[3], s[7]);
    PCP_CE(s[1], s[5]); PCP_CE(s[2], s[6]);
    PCP_CE(s[1], s[4]); PCP_CE(s[3], s[6]);
    PCP_CE(s[2], s[4]); PCP_CE(s[3], s[5]);
    PCP_CE(s[3], s[4]);
}
// Merge an ascending batch s[B] into the ascending list l[K], keeping the K smallest: the upper min(B, K) slots of
// l against the batch reversed (lower half of a bitonic split: the result is a bitonic sequence holding the K
// smallest), then a bitonic merge of K.
template <int K, int B>
__device__ __forceinline__ void merge_batch(float (&l)[K], const float (&s)[B]) {
    constexpr int U = B < K ? B : K;
#pragma unroll
    for (int i = 0; i < U; ++i) l[K - U + i] = fminf(l[K - U + i], s[U - 1 - i]);
#pragma unroll
    for (int st = K / 2; st >= 1; st >>= 1) {
#pragma unroll
        for (int i = 0; i < K; ++i)
            if ((i & st) == 0) PCP_CE(l[i], l[i + st]);
    }
}
template <int B> __device__ __forceinline__ float batch_min(const float (&s)[B]);
template <> __device__ __forceinline__ float batch_min<4>(const float (&s)[4]) { return fminf(min3f(s[0], s[1], s[2]), s[3]); }
template <> __device__ __forceinline__ float batch_min<8>(const float (&s)[8]) {
    return min3f(min3f(s[0], s[1], s[2]), min3f(s[3], s[4], s[5]), fminf(s[6], s[7]));
}

template <int K>
struct TopK {
    unsigned long long key[K];

    __device__ __forceinline__ void init() {
#pragma unroll
        for (int i = 0; i < K; ++i) key[i] = kEmptyKey;
    }
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
                unsigned *__restrict__ overflow, unsigned max_scans, unsigned max_tests,
                unsigned tail_slots, unsigned *__restrict__ sched, unsigned n_groups, float *__restrict__ hints,
                unsigned hint_groups, unsigned warp_bytes) {
    constexpr int kCap = TileCfg<K>::kCap;
    constexpr int kB = K >= 8 ? 8 : 4;                           // batch of the selection network
    constexpr unsigned kTrigger = (unsigned)(kCap - 4) * 256u;   // a longer column could overflow during the next 4 evaluations
    // dynamic shared memory, per warp (warp_bytes apart): candidate columns [kCap][32] of (d2 bits, index), staged
    // bucket (pair layout), half-bucket boxes of the open level-1 node, open node per level, one row of child
    // bounds per level >= 2.  The compaction filter's read-ahead past the last column lands in the arrays behind.
    extern __shared__ __align__(16) unsigned char s_dyn[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // (all accesses go through 32-bit shared-space addresses off one opaque per-lane base: pcp_device.cuh)
    const unsigned col_addr = smem_opaque(s_dyn + (size_t)warp * warp_bytes + lane * 8);   // this lane's column: entry e at + e * 256
    const unsigned wb = smem_opaque(s_dyn + (size_t)warp * warp_bytes);
    constexpr unsigned kTileOff = kCap * 256, kBoxOff = kTileOff + 512, kNodeOff = kBoxOff + kBoxBytes, kLbsOff = kNodeOff + 32 - 2 * 128;
#define TILE_A(e) (wb + kTileOff + (unsigned)(e) * 16u)
#define BOX_A(c) (wb + kBoxOff + (unsigned)(c) * 48u)
#define NODE_A(lv) (wb + kNodeOff + (unsigned)(lv) * 4u)
#define LBS_A(lv, j) (wb + kLbsOff + (unsigned)(lv) * 128u + (unsigned)(j) * 4u)
    const float inf = __uint_as_float(kInfBits);
    // The overflow kernel is launched as a PROGRAMMATIC DEPENDENT of this grid (knn_warp_list_device): it may start
    // as soon as every CTA of this grid has passed this point -- i.e. is resident, the grid being one wave -- and
    // then takes whatever the finishing CTAs free, draining the overflow list while the stragglers still run.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

    // Persistent warps: groups are handed out dynamically (sched[0]); sched[1] counts finished groups.  Tail
    // hand-over (tail_slots != 0): once the queue is dry and fewer than tail_slots groups are still in flight, the
    // stragglers hand their queries to the warp-per-query kernel (32 warps per group instead of 1).
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

    // ---- this lane's query and the group's AABB
    float qx = 0.f, qy = 0.f, qz = 0.f;
    long long qid = -1;
    float3 glo, ghi;
    if (SELF) {
        const float4 p = __ldg(ix.pts + (long long)group * kLeaf + lane);
        qid = (long long)(int)__float_as_uint(p.w);            // -1 on pad lanes
        if (qid >= 0) { qx = p.x; qy = p.y; qz = p.z; }
        const float4 lo = __ldg(ix.box_lo + group), hi = __ldg(ix.box_hi + group);   // the bucket's own box (level 0)
        glo = make_float3(lo.x, lo.y, lo.z);
        ghi = make_float3(hi.x, hi.y, hi.z);
    } else {
        const long long pos = (long long)group * 32 + lane;
        if (pos < nq) {
            qid = perm ? (long long)perm[pos] : pos;
            const float *q = queries + qid * stride;
            qx = __ldg(q); qy = __ldg(q + 1); qz = __ldg(q + 2);
        }
        const bool a = qid >= 0;
        glo.x = warp_min(a ? qx : FLT_MAX); glo.y = warp_min(a ? qy : FLT_MAX); glo.z = warp_min(a ? qz : FLT_MAX);
        ghi.x = warp_max(a ? qx : -FLT_MAX); ghi.y = warp_max(a ? qy : -FLT_MAX); ghi.z = warp_max(a ? qz : -FLT_MAX);
    }
    const bool active = qid >= 0;
    const float gcx = 0.5f * (glo.x + ghi.x), gcy = 0.5f * (glo.y + ghi.y), gcz = 0.5f * (glo.z + ghi.z);
    const f32x2 QX = pk2(qx, qx), QY = pk2(qy, qy), QZ = pk2(qz, qz);

    const float gate0 = d2_gate(kWorstDist0);
    float gate = active ? gate0 : -1.f;                         // squared-distance gate of this lane (inactive: nothing passes)
    unsigned group_gate = __float_as_uint(gate0);               // bits of the loosest gate in the group
    unsigned n_scan = 0, n_half = 0, n_open = 0, n_test = 0, n_round = 0, gate_gen = 0;
    bool aborted = false, booted = false;
    const unsigned n_active = __popc(__ballot_sync(kFullMask, active));
    unsigned waddr = col_addr;                                  // append position (shared-memory byte address; 256 B per entry)

    // Running k smallest squared distances of this lane, ascending (k < K: the first K - k slots hold a sentinel
    // below every distance, so that l[K-1] is always the k-th smallest).
    float l[K];
#pragma unroll
    for (int i = 0; i < K; ++i) l[i] = i < K - k ? -1.f : inf;
    int n_old = 0;                                              // entries of the column already merged into l[]

    auto append = [&](float d2, float w) {
        if (d2 <= gate) {
            asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(waddr), "f"(d2), "f"(w) : "memory");
            waddr += 256;
        }
    };
    // In-place filter: gate <- k-th smallest squared distance (candidates whose float32 sqrt can still tie it
    // stay: canonical order), candidates beyond it are dropped.  The write position never passes the read
    // position, so loading 4 entries ahead of the stores they may alias is safe (and lets the loads overlap); the
    // loads may run up to 3 entries past the longest column (they land in the arrays behind the columns), the
    // `< end` test discards them.
    auto filter = [&](int nmax) {
        if (l[K - 1] < inf) gate = fminf(gate, d2_gate(__fsqrt_rn(l[K - 1])));
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
        ++gate_gen;
    };

    // Merge the candidates appended since the last compaction into l[] (batch network), tighten the gates and
    // drop the candidates beyond them.  Warp-uniform loops.
    auto compact = [&]() {
        const int cnt = (int)((waddr - col_addr) >> 8);
        const int n_new = __reduce_max_sync(kFullMask, cnt - n_old);
        const int nmax = __reduce_max_sync(kFullMask, cnt);
        unsigned src_new = col_addr + (unsigned)n_old * 256u;
        int rem = cnt - n_old;                                  // this lane's candidates not merged yet
#pragma unroll 1
        for (int b = 0; b < n_new; b += kB, rem -= kB, src_new += kB * 256u) {
            float s[kB];
#pragma unroll
            for (int i = 0; i < kB; ++i) {
                s[i] = inf;
                if (i < rem) s[i] = lds_f1(src_new + i * 256u); // (predicated: never reads past this lane's entries)
            }
            // the lists tighten as the merge proceeds: a batch in which no lane's candidate beats its k-th
            // distance changes nothing
            if (!__any_sync(kFullMask, batch_min<kB>(s) < l[K - 1])) continue;
            ++n_round;
            sort_net(s);
            merge_batch<K, kB>(l, s);
        }
        filter(nmax);
    };

    // Tail hand-over poll: queue dry and fewer than tail_slots groups still in flight?
    auto poll_tail = [&]() {
        if (!tail_slots) return;
        uint2 w = make_uint2(0u, 0u);
        if (lane == 0) asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(w.x), "=r"(w.y) : "l"(sched));
        // (sched[0] runs past n_groups by at most one per warp; sched[1] <= n_groups)
        if (__shfl_sync(kFullMask, (w.x >= n_groups && n_groups - w.y < tail_slots) ? 1 : 0, 0)) aborted = true;
    };

    // Exact per-lane lower bounds against the boxes of bucket c's two 16-point halves, both at once in packed
    // FP32 (box_lb2's arithmetic, element by element): bit h is set when ANY lane can still use half h.
    auto bucket_needed = [&](int c) -> unsigned {
        ++n_test;
        // b0 = (loA.x, loB.x, loA.y, loB.y), b1 = (loA.z, loB.z, hiA.x, hiB.x), b2 = (hiA.y, hiB.y, hiA.z, hiB.z)
        const float4 b0 = lds_f4(BOX_A(c)), b1 = lds_f4(BOX_A(c) + 16u), b2 = lds_f4(BOX_A(c) + 32u);
        float ax, bx, ay, by, az, bz, t0, t1, u0, u1;
        upk2(sub2(pk2(b0.x, b0.y), QX), t0, t1);
        upk2(sub2(QX, pk2(b1.z, b1.w)), u0, u1);
        ax = max3f(t0, u0, 0.f); bx = max3f(t1, u1, 0.f);
        upk2(sub2(pk2(b0.z, b0.w), QY), t0, t1);
        upk2(sub2(QY, pk2(b2.x, b2.y)), u0, u1);
        ay = max3f(t0, u0, 0.f); by = max3f(t1, u1, 0.f);
        upk2(sub2(pk2(b1.x, b1.y), QZ), t0, t1);
        upk2(sub2(QZ, pk2(b2.z, b2.w)), u0, u1);
        az = max3f(t0, u0, 0.f); bz = max3f(t1, u1, 0.f);
        const f32x2 X = pk2(ax, bx), Y = pk2(ay, by), Z = pk2(az, bz);
        float x0, x1, y0, y1, z0, z1;
        upk2(mul2(X, X), x0, x1);
        upk2(mul2(Y, Y), y0, y1);
        upk2(mul2(Z, Z), z0, z1);
        const bool a = __fadd_rn(__fadd_rn(x0, y0), z0) <= gate;
        const bool b = __fadd_rn(__fadd_rn(x1, y1), z1) <= gate;
        return (__any_sync(kFullMask, a) ? 1u : 0u) | (__any_sync(kFullMask, b) ? 2u : 0u);
    };

    // First bucket of the group: its distances go through the batch network straight from registers (kB at a
    // time: evaluate, sort, merge into l[], tighten the gate) and a batch's candidates are appended against the gate
    // it has just produced -- about 2k entries reach the columns instead of 32, and every lane owns its k-th
    // distance among 32 real points before the walk proper starts.
    auto bootstrap = [&]() {
#pragma unroll 1
        for (int g = 0; g < 32; g += kB) {
            float d[kB];
#pragma unroll
            for (int m = 0; m < kB / 2; ++m) {
                const float4 a = lds_f4(TILE_A(g + 2 * m)), b = lds_f4(TILE_A(g + 2 * m + 1));
                ref_d2_pair(QX, QY, QZ, pk2(a.x, a.y), pk2(a.z, a.w), pk2(b.x, b.y), d[2 * m], d[2 * m + 1]);
            }
            {
                float s[kB];
#pragma unroll
                for (int i = 0; i < kB; ++i) s[i] = d[i];
                sort_net(s);
                merge_batch<K, kB>(l, s);
            }
            if (l[K - 1] < inf) gate = fminf(gate, d2_gate(__fsqrt_rn(l[K - 1])));
#pragma unroll
            for (int m = 0; m < kB / 2; ++m) {
                const float2 w = lds_f2(TILE_A(g + 2 * m + 1) + 8u);
                append(d[2 * m], w.x);
                append(d[2 * m + 1], w.y);
            }
            // columns that could overflow during the next batch: drop what the tightened gates no longer admit; still
            // too long = a mass of exactly tied distances: leave the group to the warp-per-query kernel
            if (g + kB < 32 && __any_sync(kFullMask, waddr > col_addr + (unsigned)(kCap - kB) * 256u)) {
                filter(__reduce_max_sync(kFullMask, (int)((waddr - col_addr) >> 8)));
                if (__any_sync(kFullMask, waddr > col_addr + (unsigned)(kCap - kB) * 256u)) { aborted = true; break; }
            }
        }
        n_round += 32 / kB;
        filter(__reduce_max_sync(kFullMask, (int)((waddr - col_addr) >> 8)));
        if (__any_sync(kFullMask, waddr > col_addr + kTrigger)) aborted = true;
    };

    // Scan the needed halves of one bucket whose pair-layout row `pv` (one float4 per lane) is already in registers.
    auto scan_bucket = [&](const float4 pv, unsigned halves) {
        ++n_scan;
        __syncwarp();
        sts_f4(TILE_A(lane), pv);
        __syncwarp();
        if (!booted) {
            booted = true;
            n_half += 2;
            bootstrap();
            return;
        }
        n_half += __popc(halves);
        // 8 points (= 8 float4 of the pair layout) at a time: the broadcast loads are issued ahead of the (possibly
        // aliasing) candidate stores; the columns are checked after every 4 evaluations
        const int e_end = (halves & 2u) ? 32 : 16;             // (halves != 0: at least one of the two is scanned)
#pragma unroll 1
        for (int e0 = (halves & 1u) ? 0 : 16; e0 < e_end; e0 += 8) {
            float4 p[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) p[j] = lds_f4(TILE_A(e0 + j));
#pragma unroll
            for (int h = 0; h < 2; ++h) {
#pragma unroll
                for (int m = 2 * h; m < 2 * h + 2; ++m) {
                    float d0, d1;
                    ref_d2_pair(QX, QY, QZ, pk2(p[2 * m].x, p[2 * m].y), pk2(p[2 * m].z, p[2 * m].w),
                                pk2(p[2 * m + 1].x, p[2 * m + 1].y), d0, d1);
                    append(d0, p[2 * m + 1].z);
                    append(d1, p[2 * m + 1].w);
                }
                // a column that could overflow during the next 4 evaluations: compact all columns of the warp
                if (__any_sync(kFullMask, waddr > col_addr + kTrigger)) {
                    compact();
                    // still too long = a mass of exactly tied distances: leave the group to the warp-per-query kernel
                    if (__any_sync(kFullMask, waddr > col_addr + kTrigger)) aborted = true;
                }
                if (aborted) break;      // (before the second half appends past the end of the columns)
            }
            if (aborted) break;
        }
    };

    // ---- upper levels (>= 2): child bounds of the open node per level in shared memory
    const int top_level = ix.n_levels - 1;      // >= 1 for every non-empty index (pcp_build.cu)
    auto open_upper = [&](int lv, unsigned node) {
        ++n_open;
        if (lane == 0) sts_u1(NODE_A(lv), node);
        const long long c = (long long)node * kFan + lane;
        unsigned bits = kDone;
        if (c < ix.level_cnt[lv - 1]) {
            const long long b = ix.level_off[lv - 1] + c;
            const float4 lo = __ldg(ix.box_lo + b), hi = __ldg(ix.box_hi + b);
            bits = __float_as_uint(boxbox_lb2(glo, ghi, lo, hi));
            if (bits == 0u) {   // overlaps the group box: order by centre distance (a hint, never used to prune)
                const float cx = 0.5f * (lo.x + hi.x) - gcx, cy = 0.5f * (lo.y + hi.y) - gcy, cz = 0.5f * (lo.z + hi.z) - gcz;
                bits = __float_as_uint(cx * cx + cy * cy + cz * cz) >> 9;
            }
        }
        sts_u1(LBS_A(lv, lane), bits);
        __syncwarp();
    };
    // nearest unvisited child of the open node on `lv` that can still matter: -1 if none.  Marks it visited.
    auto pick_upper = [&](int lv) -> int {
        const unsigned mine = lds_u1(LBS_A(lv, lane));
        const unsigned best = __reduce_min_sync(kFullMask, mine);
        // exact prune: every point below a child is at least `best` away from every query of the group
        // (slots below kMinNormalBits are ordering hints of overlapping children or denormal bounds: kept)
        if (best == kDone || (best > group_gate && best >= kMinNormalBits)) return -1;
        const int sel = __ffs(__ballot_sync(kFullMask, mine == best)) - 1;
        if (lane == sel) sts_u1(LBS_A(lv, lane), kDone);
        __syncwarp();
        return sel;
    };

    // ---- one level-1 node: its 32 buckets.  Lane j keeps (lower bound of bucket j, low 5 bits = j) in a register.
    auto process_l1 = [&](unsigned node) {
        ++n_open;
        if ((n_open & 3u) == 0u) poll_tail();
        unsigned key = kDone;
        {
            const long long c = (long long)node * kFan + lane;
            if (c < ix.level_cnt[0]) {
                // the boxes of the bucket's two halves; the bucket's bound is the smaller of theirs
                const float4 *h = ix.half + c * 4;
                const float4 alo = __ldg(h), ahi = __ldg(h + 1), blo = __ldg(h + 2), bhi = __ldg(h + 3);
                sts_f4(BOX_A(lane), make_float4(alo.x, blo.x, alo.y, blo.y));
                sts_f4(BOX_A(lane) + 16u, make_float4(alo.z, blo.z, ahi.x, bhi.x));
                sts_f4(BOX_A(lane) + 32u, make_float4(ahi.y, bhi.y, ahi.z, bhi.z));
                unsigned bits = min(__float_as_uint(boxbox_lb2(glo, ghi, alo, ahi)), __float_as_uint(boxbox_lb2(glo, ghi, blo, bhi)));
                if (bits == 0u) {   // overlaps the group box: order by centre distance (a hint, never used to prune)
                    const float cx = 0.5f * (fminf(alo.x, blo.x) + fmaxf(ahi.x, bhi.x)) - gcx;
                    const float cy = 0.5f * (fminf(alo.y, blo.y) + fmaxf(ahi.y, bhi.y)) - gcy;
                    const float cz = 0.5f * (fminf(alo.z, blo.z) + fmaxf(ahi.z, bhi.z)) - gcz;
                    bits = __float_as_uint(cx * cx + cy * cy + cz * cz) >> 9;
                }
                // (clearing the low mantissa bits only lowers the bound: still exact to prune with)
                key = (bits & ~31u) | (unsigned)lane;
            }
        }
        __syncwarp();
        // nearest unvisited bucket that some lane still needs: -1 if none
        auto next_needed = [&](unsigned &halves) -> int {
            for (;;) {
                const unsigned best = __reduce_min_sync(kFullMask, key);
                const unsigned bound = best & ~31u;
                if (best == kDone || (bound > group_gate && bound >= kMinNormalBits)) return -1;
                const int sel = (int)(best & 31u);
                if (lane == sel) key = kDone;
                if ((halves = bucket_needed(sel)) != 0u) return sel;
            }
        };
        // Software pipeline: the points of the NEXT needed bucket in visiting order are loaded while the current
        // one is scanned (the order depends only on the box bounds; the gates only shrink, so a bucket rejected
        // now stays rejected; one accepted before the gates tightened is tested again).
        const float4 *base = ix.pts2 + (long long)node * kFan * kLeaf + lane;
        unsigned halves = 0;
        int sel = next_needed(halves);
        if (sel < 0) return;
        float4 cur = __ldg(base + sel * kLeaf);
        for (;;) {
            unsigned nxt_halves = 0;
            int nxt = booted ? next_needed(nxt_halves) : -2;    // (before the bootstrap every bucket passes: decide after it)
            float4 pf = cur;
            if (nxt >= 0) pf = __ldg(base + nxt * kLeaf);
            const unsigned gen = gate_gen;
            scan_bucket(cur, halves);
            if ((n_scan & 3u) == 0u) poll_tail();
            if (aborted || n_scan > max_scans || n_test > max_tests) { aborted = true; return; }
            if (nxt == -2 || (nxt >= 0 && gate_gen != gen)) {
                // the gates tightened during the scan: test the prefetched bucket again, move on if it dropped out
                if (nxt >= 0) nxt_halves = bucket_needed(nxt);
                if (nxt < 0 || nxt_halves == 0u) {
                    nxt = next_needed(nxt_halves);
                    if (nxt >= 0) pf = __ldg(base + nxt * kLeaf);
                }
            }
            if (nxt < 0) return;
            sel = nxt;
            halves = nxt_halves;
            cur = pf;
        }
    };

    {
        int level = top_level;
        bool have_l1 = top_level == 1;
        unsigned l1node = 0;
        if (!have_l1) open_upper(top_level, 0u);
        for (;;) {
            if (!have_l1) {
                const int sel = pick_upper(level);
                if (sel < 0) {
                    if (level == top_level) break;
                    ++level;
                    continue;
                }
                const unsigned child = lds_u1(NODE_A(level)) * kFan + sel;
                if (level > 2) {
                    --level;
                    open_upper(level, child);
                    continue;
                }
                l1node = child;
            }
            process_l1(l1node);
            have_l1 = false;
            if (aborted || top_level == 1) break;
        }
    }
    if (aborted) {
        // work budget exceeded (a sparse group whose box is much larger than its neighbourhoods) or a mass of
        // tied distances: hand the group's queries to the warp-per-query kernel, which re-does them in parallel
        unsigned pos = 0;
        if (lane == 0) pos = atomicAdd(overflow, 1u);
        // the gates reached so far seed the searches of the warp-per-query kernel
        pos = __shfl_sync(kFullMask, pos, 0);
        if (pos < hint_groups) hints[(size_t)pos * 32 + lane] = active ? gate : inf;
        // publish: the consumer may be running already (it polls the entry, then the finished count)
        __threadfence();
        __syncwarp();
        if (lane == 0) {
            atomicExch(overflow + 1 + pos, group + 1u);
            __threadfence();
            atomicAdd(sched + 1, 1u);
        }
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
                const float2 ent = lds_f2(col_addr + (unsigned)e * 256u);
                top.insert(pack_key(__fsqrt_rn(ent.x), __float_as_uint(ent.y)));
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
#undef TILE_A
#undef BOX_A
#undef NODE_A
#undef LBS_A
}

template <bool SELF>
int launch_tile(pcp_index *ix, const float *d_q, const unsigned *d_perm, long long nq, long long stride,
                int k, int32_t *d_idx, float *d_dist, int32_t *d_count) {
    pcp_ctx *ctx = ix->ctx;
    const IndexView v = make_view(ix);
    const long long n_groups = SELF ? v.n_leaf : (nq + 31) / 32;
    unsigned long long *ctr = ctx->ctr();
    // overflow list: [0] = count, [1..] = group ids + 1
    // behind the list, 8-byte aligned: next group of the persistent tile warps, number of finished groups
    // and the gates of the first kHintGroups overflow groups (32 floats each)
    const long long sched_at = (n_groups + 2) & ~1LL, hints_at = (sched_at + 2 + 31) & ~31LL;
    const unsigned hint_groups = (unsigned)std::min<long long>(n_groups, kHintGroups);
    PCP_TRY(ctx->scratch[5].ensure(sizeof(unsigned) * (size_t)(hints_at + 32LL * hint_groups)));
    unsigned *overflow = ctx->scratch[5].as<unsigned>();
    float *hints = reinterpret_cast<float *>(overflow + hints_at);
    // count, entries (0 = not published yet; group + 1 otherwise) and the two scheduler words in one clear
    PCP_CUDA(cudaMemsetAsync(overflow, 0, sizeof(unsigned) * (size_t)(sched_at + 2), ctx->stream));
    const int K = k <= 2 ? k : k <= 4 ? 4 : k <= 8 ? 8 : k <= 16 ? 16 : 32;
    const bool small_grid = n_groups < 2LL * ctx->sm_count * 32;
    static const char *e_scans = getenv("PCP_TILE_MAX_SCANS"), *e_tests = getenv("PCP_TILE_MAX_TESTS"),
                      *e_tail = getenv("PCP_TILE_TAIL");   // developer overrides
    const double tail_share = e_tail ? atof(e_tail) : kTailShare;   // share of the warp slots; 0 = off
    const unsigned max_scans = e_scans ? (unsigned)atoi(e_scans) : small_grid ? kScansPerSlot * std::max(K, 4) : 0xffffffffu;
    const unsigned max_tests = e_tests ? (unsigned)atoi(e_tests) : small_grid ? kTestsPerSlot * std::max(K, 4) : 0xffffffffu;
    unsigned *dbg = nullptr;
    const char *dump = ctr && !ctx->capturing ? getenv("PCP_DUMP_GROUPS") : nullptr;     // developer aid: per-group work counters
    if (dump) {
        PCP_TRY(ctx->scratch[4].ensure(sizeof(unsigned) * 8 * (size_t)n_groups));
        dbg = ctx->scratch[4].as<unsigned>();
    }
#define PCP_TILE(KK)                                                                                                   \
    do {                                                                                                               \
        constexpr int W = TileCfg<KK>::kWarps, CAP = TileCfg<KK>::kCap;                                                \
        const int n_up = v.n_levels > 2 ? v.n_levels - 2 : 0;                                                          \
        const size_t wbytes = tile_warp_bytes(CAP, n_up);                                                              \
        const size_t smem = (size_t)W * wbytes;                                                                        \
        static int occ_by_levels[kMaxDevices][kMaxLevels + 1] = {{0}};   /* resident CTAs per SM, per device and tree depth */ \
        int &ctas_per_sm = occ_by_levels[ctx->device & (kMaxDevices - 1)][n_up];                                       \
        if (!ctas_per_sm) {                                                                                            \
            static size_t attr_smem_dev[kMaxDevices] = {0};                                                            \
            size_t &attr_smem = attr_smem_dev[ctx->device & (kMaxDevices - 1)];                                        \
            if (smem > attr_smem) {                                                                                    \
                PCP_CUDA(cudaFuncSetAttribute(knn_tile_kernel<KK, SELF>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
                attr_smem = smem;                                                                                      \
            }                                                                                                          \
            PCP_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas_per_sm, knn_tile_kernel<KK, SELF>, W * 32, smem)); \
            ctas_per_sm = std::max(ctas_per_sm, 1);                                                                    \
            if (const char *e = getenv("PCP_TILE_CTAS")) ctas_per_sm = std::max(1, std::min(ctas_per_sm, atoi(e)));     \
            if (getenv("PCP_VERBOSE")) fprintf(stderr, "knn_tile_kernel<%d>: %zu B smem, %d CTAs/SM\n", KK, smem, ctas_per_sm); \
        }                                                                                                              \
        /* persistent grid: every resident warp slot, or fewer when there are not that many groups */                  \
        const unsigned grid = (unsigned)std::min<long long>((n_groups + W - 1) / W, (long long)ctx->sm_count * ctas_per_sm); \
        const unsigned tail_slots = (unsigned)(tail_share * (double)grid * W);                                        \
        knn_tile_kernel<KK, SELF><<<grid, W * 32, smem, ctx->stream>>>(v, d_q, d_perm, nq, stride, k, d_idx, d_dist,   \
                                                                       d_count, ctr, dbg, overflow, max_scans,         \
                                                                       max_tests, tail_slots,                          \
                                                                       overflow + sched_at, (unsigned)n_groups, hints, \
                                                                       hint_groups, (unsigned)wbytes);                 \
    } while (0)
    cudaEvent_t ev_end = nullptr;
    if (ctx->time_kernels && !ctx->capturing) {
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
    // groups that ran out of budget or were handed over in the tail: one warp per query, persistent grid that runs
    // BESIDE the tail of the tiled kernel (not when its launches are being timed: the event sits in between)
    return knn_warp_list_device(ix, SELF, d_q, nq, stride, d_perm, overflow, overflow + sched_at, (uint32_t)n_groups, hints,
                                hint_groups, k, d_idx, d_dist, d_count, /*overlap=*/ev_end == nullptr && !dump);
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
