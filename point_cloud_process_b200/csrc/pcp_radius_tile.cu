// Exact fixed-radius search, tiled: one warp = one GROUP of 32 spatially adjacent queries, one query per lane.
// Replaces kdtree_radius_search (kdtree.py:176-208), octree_radius_search(_fast) (octree.py:166-259) and
// RadiusNNResultSet.add_point (result_set.py:78-84: keep iff NOT dist > radius) for batched query sets.
//
// The warp-per-query kernel (pcp_radius.cu) re-reads every bucket it needs from L2 for every query: at 1 M
// queries x ~40 buckets x 512 B per pass it runs at the L2 bandwidth.  Here a bucket is fetched ONCE per group
// of 32 queries: the group walks the hierarchy with its own AABB (depth-first, the radius never changes, so no
// ordering is needed), lane j tests child j of the open node, and every bucket half some lane still needs is
// staged in shared memory and read back as broadcast LDS.128 -- each lane evaluates its own query against the
// 16 points (the reference's unfused float32 expression, 8 FP32 ops; count pass: 11 instructions per evaluation).
// COUNT pass: per-lane hit counters -> counts[query].  FILL pass: every hit is stored at the lane's own cursor
// inside its CSR segment (offsets from the scanned counts); the order inside a segment is the visiting order
// (RadiusNNResultSet keeps visiting order too, result_set.py:84; `sorted` is applied afterwards).
#include <algorithm>
#include <cfloat>
#include <cstdlib>

#include "pcp_device.cuh"
#include "pcp_internal.cuh"

namespace pcp {

namespace {

constexpr int kRtStack = kMaxLevels * kFan;
constexpr int kRtChunk = 8;
// FILL: hits are staged per lane in shared memory (kRtCap entries, rows padded to kRtCap + 1 so that the 64-bit
// appends of a half-warp fall into distinct banks) and flushed by the whole warp, half a warp per lane's run:
// coalesced runs of up to 64 B instead of one scattered 4-byte store per hit.
constexpr int kRtCap = 16;
template <bool FILL> struct RtCfg { static constexpr int kWarps = FILL ? 4 : 8; };

// EXACT: borderline pairs are decided in float64 (RadiusTest); else the plain float32 compare d2 <= rt.hi (== rt.lo)
template <bool FILL, bool EXACT>
__global__ void __launch_bounds__(RtCfg<FILL>::kWarps * 32)
radius_tile_kernel(const IndexView ix, const float *__restrict__ queries, const unsigned *__restrict__ perm,
                   long long nq, long long stride, const RadiusTest rt, long long *__restrict__ counts,
                   const long long *__restrict__ offsets, int *__restrict__ out_idx, float *__restrict__ out_dist,
                   unsigned long long *__restrict__ ctr, unsigned *__restrict__ sched, unsigned *__restrict__ handover,
                   unsigned n_groups, unsigned tail_slots) {
    constexpr int kRtWarps = RtCfg<FILL>::kWarps;
    __shared__ float4 s_tile[kRtWarps][32];
    __shared__ float4 s_box[kRtWarps][128];            // half-bucket boxes of the open level-1 node
    __shared__ unsigned s_stack[kRtWarps][kRtStack];   // entries: (level << 27) | node   (level >= 1)
    __shared__ uint2 s_hit[FILL ? kRtWarps : 1][FILL ? 32 * (kRtCap + 1) : 1];   // staged (index, d2 bits)
    __shared__ long long s_cur[FILL ? kRtWarps : 1][32];
    __shared__ int s_cnt[FILL ? kRtWarps : 1][32];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned lt_mask = (1u << lane) - 1u;
    float4 *tile = s_tile[warp];
    float4 *boxes = s_box[warp];
    unsigned *stack = s_stack[warp];

    // sched[0] = next group to hand out, sched[1] = groups finished.  Tail hand-over as in knn_tile_kernel: with
    // the queue dry and fewer than tail_slots groups in flight, the stragglers (a group whose queries straddle a
    // jump of the Morton curve opens thousands of nodes) go on the `handover` list for the warp-per-query kernel.
    for (;;) {
        unsigned group;
        {
            unsigned g = 0;
            if (lane == 0) g = atomicAdd(sched, 1u);
            group = __shfl_sync(kFullMask, g, 0);
        }
        if (group >= n_groups) break;
        const long long pos = (long long)group * 32 + lane;
        long long qid = -1;
        float qx = 0.f, qy = 0.f, qz = 0.f;
        if (pos < nq) {
            qid = perm ? (long long)perm[pos] : pos;
            const float *q = queries + qid * stride;
            qx = __ldg(q); qy = __ldg(q + 1); qz = __ldg(q + 2);
        }
        const bool active = qid >= 0;
        const float gate = active ? rt.hi : -1.f;       // inactive lanes: nothing passes (box tests; points: radius_hit)
        float3 glo, ghi;
        glo.x = warp_min(active ? qx : FLT_MAX); glo.y = warp_min(active ? qy : FLT_MAX); glo.z = warp_min(active ? qz : FLT_MAX);
        ghi.x = warp_max(active ? qx : -FLT_MAX); ghi.y = warp_max(active ? qy : -FLT_MAX); ghi.z = warp_max(active ? qz : -FLT_MAX);
        long long cursor = (FILL && active) ? offsets[qid] : 0;   // COUNT: hits so far; FILL: next output slot
        unsigned n_half = 0, n_open = 0;
        // FILL: this lane's staging row and its append position (shared-memory byte address), entries pending
        const unsigned row_addr = (unsigned)__cvta_generic_to_shared(&s_hit[FILL ? warp : 0][FILL ? lane * (kRtCap + 1) : 0]);
        unsigned waddr = row_addr;
        auto flush = [&]() {
            const int n_pend = (int)((waddr - row_addr) >> 3);
            s_cnt[FILL ? warp : 0][lane] = n_pend;
            s_cur[FILL ? warp : 0][lane] = cursor;
            __syncwarp();
            const int j = lane & 15;
#pragma unroll 4
            for (int src = lane >> 4; src < 32; src += 2) {       // half a warp writes one lane's run
                if (j < s_cnt[FILL ? warp : 0][src]) {
                    const uint2 e = s_hit[FILL ? warp : 0][src * (kRtCap + 1) + j];
                    const long long o = s_cur[FILL ? warp : 0][src] + j;
                    out_idx[o] = (int)e.x;
                    out_dist[o] = __fsqrt_rn(__uint_as_float(e.y));
                }
            }
            __syncwarp();
            cursor += n_pend;
            waddr = row_addr;
        };

        const int top = ix.n_levels - 1;                 // >= 1 for every non-empty index
        int sp = 1;
        if (lane == 0) stack[0] = ((unsigned)top << 27);
        __syncwarp();
        bool aborted = false;
        while (sp > 0) {
            const unsigned e = stack[sp - 1];
            --sp;
            __syncwarp();
            ++n_open;
            if (tail_slots && (n_open & 7u) == 0u) {
                uint2 w = make_uint2(0u, 0u);
                if (lane == 0) asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(w.x), "=r"(w.y) : "l"(sched));
                if (__shfl_sync(kFullMask, (w.x >= n_groups && n_groups - w.y < tail_slots) ? 1 : 0, 0)) {
                    aborted = true;
                    break;
                }
            }
            const int lv = (int)(e >> 27);
            const unsigned node = e & 0x07ffffffu;
            const long long c = (long long)node * kFan + lane;
            const bool present = c < ix.level_cnt[lv - 1];
            if (lv > 1) {
                bool hit = false;
                if (present) {
                    const long long b = ix.level_off[lv - 1] + c;
                    hit = boxbox_lb2(glo, ghi, __ldg(ix.box_lo + b), __ldg(ix.box_hi + b)) <= rt.hi;
                }
                const unsigned m = __ballot_sync(kFullMask, hit);
                if (hit) stack[sp + __popc(m & lt_mask)] = ((unsigned)(lv - 1) << 27) | (unsigned)c;
                sp += __popc(m);
                __syncwarp();
                continue;
            }
            // level 1: the children are buckets; lane j owns bucket j's two half boxes
            bool ha = false, hb = false;
            if (present) {
                const float4 *h = ix.half + c * 4;
                const float4 alo = __ldg(h), ahi = __ldg(h + 1), blo = __ldg(h + 2), bhi = __ldg(h + 3);
                boxes[4 * lane] = alo; boxes[4 * lane + 1] = ahi; boxes[4 * lane + 2] = blo; boxes[4 * lane + 3] = bhi;
                ha = boxbox_lb2(glo, ghi, alo, ahi) <= rt.hi;
                hb = boxbox_lb2(glo, ghi, blo, bhi) <= rt.hi;
            }
            const unsigned ma = __ballot_sync(kFullMask, ha), mb = __ballot_sync(kFullMask, hb);
            unsigned m = ma | mb;
            const float4 *base = ix.pts + (long long)node * kFan * kLeaf + lane;
            while (m) {
                const int j = __ffs(m) - 1;
                m &= m - 1;
                // exact per-lane tests: does ANY lane reach into the half?
                unsigned halves = 0;
                if ((ma >> j) & 1u)
                    halves |= __any_sync(kFullMask, box_lb2(qx, qy, qz, boxes[4 * j], boxes[4 * j + 1]) <= gate) ? 1u : 0u;
                if ((mb >> j) & 1u)
                    halves |= __any_sync(kFullMask, box_lb2(qx, qy, qz, boxes[4 * j + 2], boxes[4 * j + 3]) <= gate) ? 2u : 0u;
                if (!halves) continue;
                n_half += __popc(halves);
                const float4 pv = __ldg(base + j * kLeaf);
                __syncwarp();
                tile[lane] = pv;
                __syncwarp();
#pragma unroll 1
                for (int j0 = 0; j0 < 32; j0 += kRtChunk) {
                    if (!((halves >> (j0 >> 4)) & 1u)) continue;
                    // a row that could overflow during the next chunk: the warp flushes all rows
                    if (FILL && __any_sync(kFullMask, waddr > row_addr + (kRtCap - kRtChunk) * 8)) flush();
                    float4 p[kRtChunk];
#pragma unroll
                    for (int i = 0; i < kRtChunk; ++i) p[i] = tile[j0 + i];
#pragma unroll
                    for (int i = 0; i < kRtChunk; ++i) {
                        const float d2 = ref_d2(qx, qy, qz, p[i].x, p[i].y, p[i].z);
                        const bool hit = EXACT ? (active && radius_hit(rt, d2, qx, qy, qz, p[i].x, p[i].y, p[i].z, qid, __float_as_uint(p[i].w)))
                                               : d2 <= gate;                 // NaN pads never hit
                        if (hit) {
                            if (FILL) {
                                asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(waddr), "f"(p[i].w), "f"(d2) : "memory");
                                waddr += 8;
                            } else {
                                ++cursor;
                            }
                        }
                    }
                }
            }
            __syncwarp();     // the boxes of this node are dead: the next open may overwrite them
        }
        if (FILL && !aborted) flush();
        if (lane == 0) {
            if (aborted) handover[1 + atomicAdd(handover, 1u)] = group;
            atomicAdd(sched + 1, 1u);
        }
        if (!FILL && active && !aborted) counts[qid] = cursor;
        const unsigned n_active = __popc(__ballot_sync(kFullMask, active));   // (all lanes: a full-mask vote)
        if (ctr && lane == 0) {
            atomicAdd(ctr + 0, (unsigned long long)n_half * (kLeaf / 2) * n_active);
            atomicAdd(ctr + 1, (unsigned long long)n_open * kFan);
            atomicAdd(ctr + 2, (unsigned long long)n_open);
            atomicAdd(ctr + 4, 1ull);
            atomicAdd(ctr + 5, (unsigned long long)n_half);
        }
        __syncwarp();
    }
}

}  // namespace

// COUNT (counts != nullptr) or FILL (offsets/out_* != nullptr) over queries regrouped by d_perm.
int radius_tile_device(pcp_index *ix, const float *d_q, int64_t nq, int64_t stride, const uint32_t *d_perm, const RadiusTest &d2max,
                       long long *d_counts, const long long *d_offsets, int32_t *d_idx, float *d_dist) {
    pcp_ctx *ctx = ix->ctx;
    const IndexView v = make_view(ix);
    const long long n_groups = (nq + 31) / 32;
    // scratch: [0] next group, [1] groups finished, [2] hand-over count, [3..] hand-over list
    PCP_TRY(ctx->scratch[5].ensure(sizeof(unsigned) * (size_t)(n_groups + 4)));
    unsigned *sched = ctx->scratch[5].as<unsigned>();
    unsigned *handover = sched + 2;
    PCP_CUDA(cudaMemsetAsync(sched, 0, 3 * sizeof(unsigned), ctx->stream));
    const int warps = d_counts ? RtCfg<false>::kWarps : RtCfg<true>::kWarps;
    const unsigned grid = (unsigned)std::min<long long>((n_groups + warps - 1) / warps, (long long)ctx->sm_count * 6);
    static const char *e_tail = getenv("PCP_TILE_TAIL");   // developer override (share of the warp slots; 0 = off)
    const unsigned tail_slots = (unsigned)((e_tail ? atof(e_tail) : 0.05) * (double)grid * warps);
    const bool exact = d2max.lo != d2max.hi || d2max.p64 != nullptr;
#define PCP_RT(FILL_, EXACT_, counts_, offs_, idx_, dist_)                                                              \
    radius_tile_kernel<FILL_, EXACT_><<<grid, warps * 32, 0, ctx->stream>>>(v, d_q, d_perm, nq, stride, d2max, counts_, \
                                                                            offs_, idx_, dist_, ctx->ctr(), sched,       \
                                                                            handover, (unsigned)n_groups, tail_slots)
    if (d_counts) {
        if (exact) PCP_RT(false, true, d_counts, nullptr, nullptr, nullptr);
        else PCP_RT(false, false, d_counts, nullptr, nullptr, nullptr);
    } else {
        if (exact) PCP_RT(true, true, nullptr, d_offsets, d_idx, d_dist);
        else PCP_RT(true, false, nullptr, d_offsets, d_idx, d_dist);
    }
#undef PCP_RT
    ctx->launches += 1;
    PCP_CUDA(cudaGetLastError());
    return radius_warp_list_device(ix, d_q, nq, stride, d_perm, handover, d2max, d_counts, d_offsets, d_idx, d_dist);
}

}  // namespace pcp
