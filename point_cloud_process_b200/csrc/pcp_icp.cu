// Point-to-point ICP, batched over independent registration pairs, every iteration on the device.
// Replaces icp_point2point (Registration/main.py:97-156) and the ICP / find_associations /
// procrustes_transformation contract of icp_template.py:43-54,113-126,128-200.
//
// Per iteration two kernels run over ALL still-active pairs (no host round trip on the data path;
// the host only reads one "pairs still active" word at geometrically spaced iterations to stop
// enqueueing launches):
//
//   icp_assoc_kernel   grid (chunks, pairs).  Fuses, in one pass over a chunk of source points:
//       main.py:110      source.transform(T)        -> p = applied * p0 in float64 (applied = composed pose)
//       main.py:116-117  1-NN in the target          -> warp-per-query walk of the target's bucket hierarchy,
//                                                       float32 box/point filter with a proven error margin,
//                                                       float64 distance on the survivors (Open3D computes in doubles)
//       main.py:119      keep iff d^2 < dist_thres
//       main.py:131-137  centroids + cross-covariance -> 16 float64 sums (M, sum a, sum b, sum b a^T), warp
//                                                       shuffle-free lane-0 accumulation + block reduction,
//                                                       one partial record per CTA (deterministic, no atomics)
//   icp_solve_kernel   one warp per pair:
//       main.py:125-127  fewer than 3 associations -> stop
//       main.py:134-140  H = sum (b - bm)(a - am)^T;  U S V^T = H by one-sided Jacobi (float64);  R = U V^T (no det fix);
//                        t = bm - R am
//       main.py:149-154  R_diff = ||R - R_last||_F, t_diff = ||t - t_last|| with the first-iteration (3,1)-(3,)
//                        broadcast of main.py:150; stop when both <= thresholds
//       main.py:156      T_last = last increment;   icp_template.py:198 homo_mat_total = composed pose
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "pcp_device.cuh"
#include "pcp_internal.cuh"
#include "pcp_linalg.cuh"

namespace pcp {

namespace {

constexpr int kAssocWarps = 8;
constexpr int kChunk = 512;               // source points per CTA
constexpr int kSums = 16;                 // M, a(3), b(3), b a^T (9)
constexpr unsigned kDone = 0xffffffffu;

struct IcpState {
    double applied[12];     // 3x4 row-major: pose applied to the ORIGINAL source in the current iteration
    double cur[12];         // last increment (main.py:143-146); the initial transformation before iteration 1
    double total[12];       // composed pose incl. the last increment (icp_template.py:198)
    double R_last[9];
    double t_last[3];
    int active;
    int iters;
    int status;
    int pad;
};

__device__ __forceinline__ unsigned long long shfl_xor64(unsigned long long v, int m) {
    return __shfl_xor_sync(kFullMask, v, m);
}

// Exact 1-NN of the float64 point q in one segment, restricted to d^2 < limit.  Warp-cooperative:
// every lane returns the same (d2, idx); idx == kDone when no point qualifies.
// Float32 filter: q is rounded to float (|e_i| <= 2^-24 |q_i|) and distances / box bounds are
// evaluated in the library's float32 arithmetic; |d2_f - d2| <= 2 d E + E^2 + 4 eps d2 with
// E = sqrt(3) 2^-24 max|q_i|, so every candidate that can beat the current float64 best passes
// "d2_f <= gate" with gate = best (1 + 2^-20) + 2 sqrt(best) E + E^2 (rounded up).
struct Nn1 {
    double d2;
    unsigned idx;
    float bx, by, bz;       // coordinates of the winner (exact float32 target point)
};

__device__ __forceinline__ float nn1_gate(double best, double E) {
    if (!(best < 1e300)) return __int_as_float(0x7f800000);
    return __double2float_ru(best * (1.0 + 9.6e-7) + 2.0 * sqrt(best) * E + E * E);
}

__device__ Nn1 nn1_search(const IndexView &ix, double qxd, double qyd, double qzd, double limit, unsigned (*lbs)[32],
                          unsigned *nodes, int lane) {
    Nn1 r;
    r.d2 = limit;
    r.idx = kDone;
    r.bx = r.by = r.bz = 0.f;
    if (ix.n_levels == 0) return r;
    const float qx = (float)qxd, qy = (float)qyd, qz = (float)qzd;
    const double E = 1.04e-7 * (fabs(qxd) + fabs(qyd) + fabs(qzd));
    float gate = nn1_gate(r.d2, E);

    auto scan_bucket = [&](long long leaf) {
        const float4 p = __ldg(ix.pts + leaf * kLeaf + lane);
        const float d2f = ref_d2(qx, qy, qz, p.x, p.y, p.z);
        const bool pass = d2f <= gate;                 // NaN pads never pass
        if (__ballot_sync(kFullMask, pass) == 0) return;
        double d2 = 1e308;
        unsigned id = kDone;
        if (pass) {
            const double dx = qxd - (double)p.x, dy = qyd - (double)p.y, dz = qzd - (double)p.z;
            d2 = dx * dx + dy * dy + dz * dz;
            id = __float_as_uint(p.w);
        }
        // warp arg-min over (d2, index): non-negative doubles order like their bit patterns
        unsigned long long key = __double_as_longlong(d2);
        unsigned wid = id;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const unsigned long long ok = shfl_xor64(key, o);
            const unsigned oi = __shfl_xor_sync(kFullMask, wid, o);
            if (ok < key || (ok == key && oi < wid)) { key = ok; wid = oi; }
        }
        const double best = __longlong_as_double(key);
        if (best < r.d2 || (best == r.d2 && wid < r.idx)) {      // warp-uniform branch
            const int src = __ffs(__ballot_sync(kFullMask, pass && id == wid)) - 1;
            r.d2 = best;
            r.idx = wid;
            r.bx = __shfl_sync(kFullMask, p.x, src);
            r.by = __shfl_sync(kFullMask, p.y, src);
            r.bz = __shfl_sync(kFullMask, p.z, src);
            gate = nn1_gate(r.d2, E);
        }
    };

    const int top = ix.n_levels - 1;
    if (top == 0) {
        scan_bucket(0);
        return r;
    }
    auto open_node = [&](int lv, unsigned node) {
        if (lane == 0) nodes[lv] = node;
        const long long c = (long long)node * kFan + lane;
        unsigned bits = kDone;
        if (c < ix.level_cnt[lv - 1]) {
            const long long b = ix.level_off[lv - 1] + c;
            bits = __float_as_uint(box_lb2(qx, qy, qz, __ldg(ix.box_lo + b), __ldg(ix.box_hi + b)));
        }
        lbs[lv][lane] = bits;
        __syncwarp();
    };
    int level = top;
    open_node(top, 0u);
    while (true) {
        const unsigned mine = lbs[level][lane];
        const unsigned best = __reduce_min_sync(kFullMask, mine);
        if (best == kDone || !(__uint_as_float(best) <= gate)) {
            if (level == top) break;
            ++level;
            continue;
        }
        const int sel = __ffs(__ballot_sync(kFullMask, mine == best)) - 1;
        if (lane == sel) lbs[level][lane] = kDone;
        const unsigned child = nodes[level] * kFan + sel;
        __syncwarp();
        if (level == 1) {
            scan_bucket(child);
        } else {
            --level;
            open_node(level, child);
        }
    }
    return r;
}

__global__ void __launch_bounds__(kAssocWarps * 32)
icp_assoc_kernel(const IndexView *__restrict__ views, const float *__restrict__ src, const long long *__restrict__ src_off,
                 long long stride, const IcpState *__restrict__ states, double *__restrict__ partials,
                 int chunks_per_pair, double dist_thres) {
    __shared__ unsigned s_lb[kAssocWarps][kMaxLevels][32];
    __shared__ unsigned s_node[kAssocWarps][kMaxLevels];
    __shared__ double s_sum[kAssocWarps][kSums];
    const int pair = blockIdx.y, chunk = blockIdx.x;
    const IcpState &st = states[pair];
    if (!st.active) return;
    const long long s0 = src_off[pair], ns = src_off[pair + 1] - s0;
    const long long first = (long long)chunk * kChunk;
    if (first >= ns) return;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const IndexView &ix = views[pair];
    double T[12];
#pragma unroll
    for (int i = 0; i < 12; ++i) T[i] = st.applied[i];
    double acc[kSums];
#pragma unroll
    for (int i = 0; i < kSums; ++i) acc[i] = 0.0;
    const long long last = min(first + (long long)kChunk, ns);
    for (long long i = first + warp; i < last; i += kAssocWarps) {
        const float *p = src + (s0 + i) * stride;
        const double x = (double)__ldg(p), y = (double)__ldg(p + 1), z = (double)__ldg(p + 2);
        // main.py:110 -- the reference transforms its float64 copy cumulatively; composing the pose and
        // transforming the original differs by O(1e-16) relative
        const double ax = T[0] * x + T[1] * y + T[2] * z + T[3];
        const double ay = T[4] * x + T[5] * y + T[6] * z + T[7];
        const double az = T[8] * x + T[9] * y + T[10] * z + T[11];
        const Nn1 nn = nn1_search(ix, ax, ay, az, dist_thres, s_lb[warp], s_node[warp], lane);
        if (nn.idx != kDone && nn.d2 < dist_thres) {       // main.py:119 (squared distance vs dist_thres)
            const double bx = (double)nn.bx, by = (double)nn.by, bz = (double)nn.bz;
            acc[0] += 1.0;
            acc[1] += ax; acc[2] += ay; acc[3] += az;
            acc[4] += bx; acc[5] += by; acc[6] += bz;
            acc[7] += bx * ax; acc[8] += bx * ay; acc[9] += bx * az;
            acc[10] += by * ax; acc[11] += by * ay; acc[12] += by * az;
            acc[13] += bz * ax; acc[14] += bz * ay; acc[15] += bz * az;
        }
    }
    // every lane of a warp holds identical sums; fixed-order block reduction -> one record per CTA
    if (lane == 0) {
#pragma unroll
        for (int i = 0; i < kSums; ++i) s_sum[warp][i] = acc[i];
    }
    __syncthreads();
    if (threadIdx.x < kSums) {
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < kAssocWarps; ++w) s += s_sum[w][threadIdx.x];
        partials[((long long)pair * chunks_per_pair + chunk) * kSums + threadIdx.x] = s;
    }
}

// ---- tiled association: one warp = one bucket of 32 Morton-adjacent SOURCE points, one point per lane --------
// Same structure as knn_tile_kernel (pcp_knn_tile.cu) with K = 1: the group walks the pair's TARGET hierarchy once,
// buckets are staged in shared memory and every lane evaluates its own (transformed) point against the 32 staged
// points in float32; survivors of the proven-margin filter are re-evaluated in float64 (nn1 semantics above).
constexpr int kTileAssocWarps = 8;

__device__ __forceinline__ float nn1_gate_fast(double best, float E) {
    // upper bound of best (1 + 2^-20) + 2 sqrt(best) E + E^2 in float, every step rounded up
    if (!(best < 1e30)) return __int_as_float(0x7f800000);
    const float b = __double2float_ru(best * (1.0 + 9.6e-7));
    const float r = __fsqrt_ru(b);
    return __fadd_ru(__fadd_ru(b, __fmul_ru(__fmul_ru(2.f, r), E)), __fmul_ru(E, E));
}

__global__ void __launch_bounds__(kTileAssocWarps * 32, 32 / kTileAssocWarps)
icp_assoc_tile_kernel(const IndexView *__restrict__ tgt_views, const IndexView *__restrict__ src_views,
                      const IcpState *__restrict__ states, double *__restrict__ partials, int chunks_per_pair,
                      double dist_thres, unsigned long long *__restrict__ ctr) {
    __shared__ float4 s_tile[kTileAssocWarps][32];
    __shared__ float4 s_box[kTileAssocWarps][128];   // half-bucket boxes of the open level-1 node
    __shared__ unsigned s_lb[kTileAssocWarps][kMaxLevels][32];
    __shared__ unsigned s_cd[kTileAssocWarps][kMaxLevels][32];
    __shared__ unsigned s_node[kTileAssocWarps][kMaxLevels];
    __shared__ double s_sum[kTileAssocWarps][kSums];
    const int pair = blockIdx.y, chunk = blockIdx.x;
    const IcpState &st = states[pair];
    if (!st.active) return;
    const IndexView &sx = src_views[pair];
    if ((long long)chunk * kTileAssocWarps >= sx.n_leaf) return;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long group = (long long)chunk * kTileAssocWarps + warp;
    const IndexView &ix = tgt_views[pair];
    // the matched pair of this lane (set by the search below); the 16 sums are formed only afterwards so that
    // they do not occupy registers during the search
    double ax = 0.0, ay = 0.0, az = 0.0;
    float bx = 0.f, by = 0.f, bz = 0.f;
    bool matched = false;

    if (group < sx.n_leaf) {                                   // warp-uniform
        const float4 sp = __ldg(sx.pts + group * kLeaf + lane);
        const bool active = (int)__float_as_uint(sp.w) >= 0;   // pad lanes carry index -1
        // main.py:110 in float64: composed pose applied to the original point
        const double x = (double)sp.x, y = (double)sp.y, z = (double)sp.z;
        ax = st.applied[0] * x + st.applied[1] * y + st.applied[2] * z + st.applied[3];
        ay = st.applied[4] * x + st.applied[5] * y + st.applied[6] * z + st.applied[7];
        az = st.applied[8] * x + st.applied[9] * y + st.applied[10] * z + st.applied[11];
        const float qx = (float)ax, qy = (float)ay, qz = (float)az;
        const float E = __double2float_ru(1.04e-7 * (fabs(ax) + fabs(ay) + fabs(az)));
        float3 glo, ghi;
        glo.x = warp_min(active ? qx : 3.0e38f); glo.y = warp_min(active ? qy : 3.0e38f); glo.z = warp_min(active ? qz : 3.0e38f);
        ghi.x = warp_max(active ? qx : -3.0e38f); ghi.y = warp_max(active ? qy : -3.0e38f); ghi.z = warp_max(active ? qz : -3.0e38f);
        const float gcx = 0.5f * (glo.x + ghi.x), gcy = 0.5f * (glo.y + ghi.y), gcz = 0.5f * (glo.z + ghi.z);

        double best = dist_thres;                              // only d^2 < dist_thres matters (main.py:119)
        unsigned best_id = kDone;
        float gate = active ? nn1_gate_fast(best, E) : -1.f;
        unsigned group_gate = __reduce_max_sync(kFullMask, active ? __float_as_uint(gate) : 0u);

        float4 *tile = s_tile[warp];
        float4 *boxes = s_box[warp];
        // bit h: some lane can still use the 16-point half h of bucket c (exact per-lane bounds)
        auto bucket_needed = [&](int c) -> unsigned {
            const bool a = box_lb2(qx, qy, qz, boxes[4 * c], boxes[4 * c + 1]) <= gate;
            const bool b = box_lb2(qx, qy, qz, boxes[4 * c + 2], boxes[4 * c + 3]) <= gate;
            return (__any_sync(kFullMask, a) ? 1u : 0u) | (__any_sync(kFullMask, b) ? 2u : 0u);
        };
        float m32 = __int_as_float(0x7f800000);                // smallest float32 d^2 this lane has seen
        unsigned n_pts = 0;                                    // work counter: points evaluated per lane
        const f32x2 qx2 = pk2(qx, qx), qy2 = pk2(qy, qy), qz2 = pk2(qz, qz);
        // pv: this lane's float4 of the bucket in the PAIR layout (IndexView::pts2): entries 2m / 2m+1 hold
        // (x0, x1, y0, y1) / (z0, z1, w0, w1) of points 2m, 2m+1; pad points have x = +inf (d^2 = +inf)
        auto scan_bucket = [&](const float4 pv, unsigned halves) {
            n_pts += 16u * __popc(halves);
            __syncwarp();
            tile[lane] = pv;
            __syncwarp();
            unsigned mask = 0;
            // 8 points at a time, two per packed FP32 instruction (bit-equal to ref_d2).  The gate is tightened INSIDE
            // the scan from the running float32 minimum: a point whose float32 d^2 is m has a true d^2 <=
            // (sqrt(m) + 2.5 E)^2 (1 + 6 eps), and whatever beats or ties it has a float32 d^2 <=
            // (sqrt(m) + 3.5 E)^2 (1 + 12 eps) -- bounded here by (sqrt(m)(1 + 2e-6) + 4 E)^2 rounded up.  Only "record
            // setters" of the float32 scan reach the float64 evaluation below (a handful per bucket instead of every
            // point inside the initial dist_thres ball); a chunk whose minimum misses the gate costs no mask work.
#pragma unroll
            for (int j0 = 0; j0 < 32; j0 += 8) {
                if (!((halves >> (j0 >> 4)) & 1u)) continue;
                float d2[8];
#pragma unroll
                for (int m = 0; m < 4; ++m) {
                    const float4 a = tile[j0 + 2 * m];
                    const float2 b = *reinterpret_cast<const float2 *>(tile + j0 + 2 * m + 1);
                    ref_d2_pair(qx2, qy2, qz2, pk2(a.x, a.y), pk2(a.z, a.w), pk2(b.x, b.y), d2[2 * m], d2[2 * m + 1]);
                }
                const float cm = min3f(min3f(d2[0], d2[1], d2[2]), min3f(d2[3], d2[4], d2[5]), fminf(d2[6], d2[7]));
                if (cm <= gate) {
                    if (cm < m32) {
                        m32 = cm;
                        const float r = __fadd_ru(__fmul_ru(__fsqrt_ru(cm), 1.000002f), __fmul_ru(4.f, E));
                        gate = fminf(gate, __fmul_ru(r, r));
                    }
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        if (d2[j] <= gate) mask |= 1u << (j0 + j);
                }
            }
            bool changed = false;
            while (__any_sync(kFullMask, mask != 0)) {
                if (mask) {
                    const int j = __ffs(mask) - 1;
                    mask &= mask - 1;
                    const float *e = reinterpret_cast<const float *>(tile) + (j >> 1) * 8 + (j & 1);
                    const float4 p = make_float4(e[0], e[2], e[4], e[6]);
                    const double dx = ax - (double)p.x, dy = ay - (double)p.y, dz = az - (double)p.z;
                    const double d2 = dx * dx + dy * dy + dz * dz;
                    const unsigned id = __float_as_uint(p.w);
                    if (d2 < best || (d2 == best && id < best_id)) {
                        best = d2; best_id = id; bx = p.x; by = p.y; bz = p.z;
                        changed = true;
                    }
                }
            }
            if (changed) gate = fminf(gate, nn1_gate_fast(best, E));
            group_gate = __reduce_max_sync(kFullMask, active ? __float_as_uint(gate) : 0u);
        };

        const int top_level = ix.n_levels - 1;
        if (top_level > 0) {
            unsigned(*lbs)[32] = s_lb[warp];
            unsigned(*cds)[32] = s_cd[warp];
            unsigned *nodes = s_node[warp];
            auto open_node = [&](int lv, unsigned node) {
                if (lane == 0) nodes[lv] = node;
                const long long c = (long long)node * kFan + lane;
                unsigned bits = kDone, cbits = kDone;
                if (c < ix.level_cnt[lv - 1]) {
                    const long long b = ix.level_off[lv - 1] + c;
                    float4 lo, hi;
                    if (lv == 1) {
                        // buckets: the boxes of the two 16-point halves; the bucket's bound is the smaller of theirs
                        const float4 *h = ix.half + c * 4;
                        const float4 alo = __ldg(h), ahi = __ldg(h + 1), blo = __ldg(h + 2), bhi = __ldg(h + 3);
                        boxes[4 * lane] = alo; boxes[4 * lane + 1] = ahi; boxes[4 * lane + 2] = blo; boxes[4 * lane + 3] = bhi;
                        bits = min(__float_as_uint(boxbox_lb2(glo, ghi, alo, ahi)), __float_as_uint(boxbox_lb2(glo, ghi, blo, bhi)));
                        lo = make_float4(fminf(alo.x, blo.x), fminf(alo.y, blo.y), fminf(alo.z, blo.z), 0.f);
                        hi = make_float4(fmaxf(ahi.x, bhi.x), fmaxf(ahi.y, bhi.y), fmaxf(ahi.z, bhi.z), 0.f);
                    } else {
                        lo = __ldg(ix.box_lo + b); hi = __ldg(ix.box_hi + b);
                        bits = __float_as_uint(boxbox_lb2(glo, ghi, lo, hi));
                    }
                    const float cx = 0.5f * (lo.x + hi.x) - gcx, cy = 0.5f * (lo.y + hi.y) - gcy, cz = 0.5f * (lo.z + hi.z) - gcz;
                    cbits = __float_as_uint(cx * cx + cy * cy + cz * cz);
                }
                lbs[lv][lane] = bits;
                cds[lv][lane] = cbits;
                __syncwarp();
            };
            auto pick = [&](int lv) -> int {
                const unsigned mine = lbs[lv][lane];
                const unsigned bestb = __reduce_min_sync(kFullMask, mine);
                if (bestb == kDone || bestb > group_gate) return -1;
                unsigned tie = __ballot_sync(kFullMask, mine == bestb);
                if (tie & (tie - 1)) {
                    const unsigned cd = (mine == bestb) ? cds[lv][lane] : kDone;
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
                const float4 *base = ix.pts2 + (long long)nodes[1] * kFan * kLeaf + lane;
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
                }
            }
        }
        matched = active && best_id != kDone && best < dist_thres;   // main.py:119
        if (ctr) {   // pcp_ctx_enable_counters: [0] distance evaluations, [4] source groups
            const unsigned n_act = __popc(__ballot_sync(kFullMask, active));
            if (lane == 0) { atomicAdd(ctr + 0, (unsigned long long)n_pts * n_act); atomicAdd(ctr + 4, 1ull); }
        }
    }
    double acc[kSums];
#pragma unroll
    for (int i = 0; i < kSums; ++i) acc[i] = 0.0;
    if (matched) {
        const double tx = (double)bx, ty = (double)by, tz = (double)bz;
        acc[0] = 1.0;
        acc[1] = ax; acc[2] = ay; acc[3] = az;
        acc[4] = tx; acc[5] = ty; acc[6] = tz;
        acc[7] = tx * ax; acc[8] = tx * ay; acc[9] = tx * az;
        acc[10] = ty * ax; acc[11] = ty * ay; acc[12] = ty * az;
        acc[13] = tz * ax; acc[14] = tz * ay; acc[15] = tz * az;
    }
    // fixed-shape reductions -> deterministic.  Inside the warp a halving butterfly: at distance 16 a lane keeps the
    // 8 sums of its half and hands the other 8 over, at distance 8 it keeps 4, ... -- 16 double shuffles instead of
    // the 80 of sixteen separate trees; sum i ends up in lanes 2 i and 2 i + 1.  Then the warps in order.
    {
        static_assert(kSums == 16, "butterfly below is written for 16 sums");
        const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4, b1 = lane & 2;
        double v8[8], v4[4], v2[2], v1;
#pragma unroll
        for (int i = 0; i < 8; ++i)
            v8[i] = (b4 ? acc[8 + i] : acc[i]) + __shfl_xor_sync(kFullMask, b4 ? acc[i] : acc[8 + i], 16);
#pragma unroll
        for (int i = 0; i < 4; ++i)
            v4[i] = (b3 ? v8[4 + i] : v8[i]) + __shfl_xor_sync(kFullMask, b3 ? v8[i] : v8[4 + i], 8);
#pragma unroll
        for (int i = 0; i < 2; ++i)
            v2[i] = (b2 ? v4[2 + i] : v4[i]) + __shfl_xor_sync(kFullMask, b2 ? v4[i] : v4[2 + i], 4);
        v1 = (b1 ? v2[1] : v2[0]) + __shfl_xor_sync(kFullMask, b1 ? v2[0] : v2[1], 2);
        v1 += __shfl_xor_sync(kFullMask, v1, 1);
        if ((lane & 1) == 0) s_sum[warp][lane >> 1] = v1;
    }
    __syncthreads();
    if (threadIdx.x < kSums) {
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < kTileAssocWarps; ++w) s += s_sum[w][threadIdx.x];
        partials[((long long)pair * chunks_per_pair + chunk) * kSums + threadIdx.x] = s;
    }
}

// out = inc o prev for 3x4 row-major rigid transforms
__device__ __forceinline__ void compose(const double inc[12], const double prev[12], double out[12]) {
#pragma unroll
    for (int i = 0; i < 3; ++i) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            double s = inc[i * 4 + 0] * prev[0 * 4 + j] + inc[i * 4 + 1] * prev[1 * 4 + j] + inc[i * 4 + 2] * prev[2 * 4 + j];
            if (j == 3) s += inc[i * 4 + 3];
            out[i * 4 + j] = s;
        }
    }
}

__global__ void __launch_bounds__(128)
icp_solve_kernel(IcpState *__restrict__ states, const double *__restrict__ partials, const long long *__restrict__ src_off,
                 long long n_pairs, int chunks_per_pair, int iteration, int max_iteration, double R_thr, double t_thr,
                 int first_iter_broadcast, double *__restrict__ diff_log, int *__restrict__ n_active,
                 const IndexView *__restrict__ src_views) {
    const long long pair = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (pair >= n_pairs) return;
    IcpState &st = states[pair];
    if (!st.active) return;
    const long long ns = src_off[pair + 1] - src_off[pair];
    // chunk records written for this pair: tiled association -> one per 8 source buckets, else one per 512 points
    const int chunks = src_views ? (int)((src_views[pair].n_leaf + kTileAssocWarps - 1) / kTileAssocWarps)
                                 : (int)((ns + kChunk - 1) / kChunk);
    // lane l < 16 sums component l over the pair's chunk records in chunk order (deterministic)
    double s = 0.0;
    if (lane < kSums)
        for (int c = 0; c < chunks; ++c) s += partials[((long long)pair * chunks_per_pair + c) * kSums + lane];
    double S[kSums];
#pragma unroll
    for (int i = 0; i < kSums; ++i) S[i] = __shfl_sync(kFullMask, s, i);
    if (lane != 0) return;

    st.iters = iteration + 1;
    const double M = S[0];
    if (M < 3.0) {                                  // main.py:125-127 "ICP failed, cannot find enough associations!"
        st.status = PCP_ICP_FEW_PAIRS;
        st.active = 0;
#pragma unroll
        for (int i = 0; i < 12; ++i) st.total[i] = st.applied[i];
        atomicSub(n_active, 1);
        return;
    }
    const double inv = 1.0 / M;
    const double am[3] = {S[1] * inv, S[2] * inv, S[3] * inv}, bm[3] = {S[4] * inv, S[5] * inv, S[6] * inv};
    double H[9], R[9];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) H[i * 3 + j] = S[7 + i * 3 + j] - M * bm[i] * am[j];   // Bp @ Ap.T, main.py:137
    polar_uvt(H, R);                                // main.py:138-139
    double t[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) t[i] = bm[i] - (R[i * 3] * am[0] + R[i * 3 + 1] * am[1] + R[i * 3 + 2] * am[2]);   // main.py:140
    double rd = 0.0, td = 0.0;
#pragma unroll
    for (int i = 0; i < 9; ++i) rd += (R[i] - st.R_last[i]) * (R[i] - st.R_last[i]);
    if (iteration == 0 && first_iter_broadcast) {
        // main.py:150 on the first pass: t is (3,1), t_last is (3,) -> the difference broadcasts to 3x3
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) td += (t[i] - st.t_last[j]) * (t[i] - st.t_last[j]);
    } else {
#pragma unroll
        for (int i = 0; i < 3; ++i) td += (t[i] - st.t_last[i]) * (t[i] - st.t_last[i]);
    }
    rd = sqrt(rd);
    td = sqrt(td);
    if (diff_log) {
        diff_log[((long long)pair * max_iteration + iteration) * 2 + 0] = rd;
        diff_log[((long long)pair * max_iteration + iteration) * 2 + 1] = td;
    }
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        st.cur[i * 4 + 0] = R[i * 3 + 0]; st.cur[i * 4 + 1] = R[i * 3 + 1]; st.cur[i * 4 + 2] = R[i * 3 + 2];
        st.cur[i * 4 + 3] = t[i];
        st.t_last[i] = t[i];
    }
#pragma unroll
    for (int i = 0; i < 9; ++i) st.R_last[i] = R[i];
    double next[12];
    compose(st.cur, st.applied, next);
#pragma unroll
    for (int i = 0; i < 12; ++i) st.total[i] = next[i];
    const bool converged = rd <= R_thr && td <= t_thr;        // main.py:153-154
    if (converged || iteration + 1 >= max_iteration) {
        st.status = converged ? PCP_ICP_CONVERGED : PCP_ICP_MAX_ITER;
        st.active = 0;
        atomicSub(n_active, 1);
    } else {
#pragma unroll
        for (int i = 0; i < 12; ++i) st.applied[i] = next[i];     // main.py:110 of the next pass
    }
}

// ---- procrustes_transformation(A, B) of main.py:131-141 / icp_template.py:43-54 on given correspondences.
// A, B: 3 x n float64, row-major (one COLUMN per point, the reference's layout).
constexpr int kProcThreads = 256;

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFullMask, v, o);
    return v;
}

__global__ void __launch_bounds__(kProcThreads)
procrustes_sum_kernel(const double *__restrict__ A, const double *__restrict__ B, long long n, double *__restrict__ partials) {
    __shared__ double s_sum[kProcThreads / 32][kSums];
    double acc[kSums];
#pragma unroll
    for (int i = 0; i < kSums; ++i) acc[i] = 0.0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const double ax = A[i], ay = A[n + i], az = A[2 * n + i];
        const double bx = B[i], by = B[n + i], bz = B[2 * n + i];
        acc[0] += 1.0;
        acc[1] += ax; acc[2] += ay; acc[3] += az;
        acc[4] += bx; acc[5] += by; acc[6] += bz;
        acc[7] += bx * ax; acc[8] += bx * ay; acc[9] += bx * az;
        acc[10] += by * ax; acc[11] += by * ay; acc[12] += by * az;
        acc[13] += bz * ax; acc[14] += bz * ay; acc[15] += bz * az;
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int i = 0; i < kSums; ++i) {
        const double v = warp_sum(acc[i]);
        if (lane == 0) s_sum[warp][i] = v;
    }
    __syncthreads();
    if (threadIdx.x < kSums) {
        double v = 0.0;
#pragma unroll
        for (int w = 0; w < kProcThreads / 32; ++w) v += s_sum[w][threadIdx.x];
        partials[(long long)blockIdx.x * kSums + threadIdx.x] = v;
    }
}

// out: R (9), t (3)
__global__ void procrustes_solve_kernel(const double *__restrict__ partials, int n_blocks, double *__restrict__ out) {
    const int lane = threadIdx.x & 31;
    double s = 0.0;
    if (lane < kSums)
        for (int b = 0; b < n_blocks; ++b) s += partials[(long long)b * kSums + lane];
    double S[kSums];
#pragma unroll
    for (int i = 0; i < kSums; ++i) S[i] = __shfl_sync(kFullMask, s, i);
    if (lane != 0) return;
    const double M = S[0], inv = M > 0.0 ? 1.0 / M : 0.0;
    const double am[3] = {S[1] * inv, S[2] * inv, S[3] * inv}, bm[3] = {S[4] * inv, S[5] * inv, S[6] * inv};
    double H[9], R[9];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) H[i * 3 + j] = S[7 + i * 3 + j] - M * bm[i] * am[j];
    polar_uvt(H, R);
    for (int i = 0; i < 9; ++i) out[i] = R[i];
    for (int i = 0; i < 3; ++i) out[9 + i] = bm[i] - (R[i * 3] * am[0] + R[i * 3 + 1] * am[1] + R[i * 3 + 2] * am[2]);
}

// sum over points of |b - (R a + t)|^2 -> partials[block]
__global__ void __launch_bounds__(kProcThreads)
procrustes_cost_kernel(const double *__restrict__ A, const double *__restrict__ B, long long n, const double *__restrict__ Rt,
                       double *__restrict__ partials) {
    __shared__ double s_sum[kProcThreads / 32];
    double R[12];
#pragma unroll
    for (int i = 0; i < 12; ++i) R[i] = Rt[i];
    double acc = 0.0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const double ax = A[i], ay = A[n + i], az = A[2 * n + i];
        const double rx = B[i] - (R[0] * ax + R[1] * ay + R[2] * az + R[9]);
        const double ry = B[n + i] - (R[3] * ax + R[4] * ay + R[5] * az + R[10]);
        const double rz = B[2 * n + i] - (R[6] * ax + R[7] * ay + R[8] * az + R[11]);
        acc += rx * rx + ry * ry + rz * rz;
    }
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) s_sum[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double v = 0.0;
        for (int w = 0; w < kProcThreads / 32; ++w) v += s_sum[w];
        partials[blockIdx.x] = v;
    }
}

}  // namespace

int procrustes_device(pcp_ctx *ctx, const double *d_A, const double *d_B, int64_t n, double *h_R, double *h_t, double *h_cost) {
    cudaStream_t st = ctx->stream;
    const int blocks = (int)std::max<int64_t>(1, std::min<int64_t>((n + kProcThreads - 1) / kProcThreads, ctx->sm_count * 4));
    DevBuf &b_part = ctx->scratch[0], &b_out = ctx->scratch[1];
    PCP_TRY(b_part.ensure(sizeof(double) * kSums * (size_t)blocks));
    PCP_TRY(b_out.ensure(sizeof(double) * 12));
    procrustes_sum_kernel<<<blocks, kProcThreads, 0, st>>>(d_A, d_B, n, b_part.as<double>());
    procrustes_solve_kernel<<<1, 32, 0, st>>>(b_part.as<double>(), blocks, b_out.as<double>());
    procrustes_cost_kernel<<<blocks, kProcThreads, 0, st>>>(d_A, d_B, n, b_out.as<double>(), b_part.as<double>());
    ctx->launches += 3;
    PCP_CUDA(cudaGetLastError());
    std::vector<double> h_part((size_t)blocks), h_out(12);
    PCP_CUDA(cudaMemcpyAsync(h_out.data(), b_out.p, sizeof(double) * 12, cudaMemcpyDeviceToHost, st));
    PCP_CUDA(cudaMemcpyAsync(h_part.data(), b_part.p, sizeof(double) * (size_t)blocks, cudaMemcpyDeviceToHost, st));
    PCP_CUDA(cudaStreamSynchronize(st));
    if (h_R) memcpy(h_R, h_out.data(), sizeof(double) * 9);
    if (h_t) memcpy(h_t, h_out.data() + 9, sizeof(double) * 3);
    if (h_cost) {
        double c = 0.0;
        for (double v : h_part) c += v;
        *h_cost = std::sqrt(c);                       // np.linalg.norm(B - (R A + t 1^T)), main.py:141
    }
    return PCP_OK;
}

int icp_batch_device(pcp_ctx *ctx, const pcp_icp_params &prm, int64_t P, const float *d_src, const int64_t *h_src_off,
                     const float *d_tgt, const int64_t *h_tgt_off, int64_t stride, const double *h_T_init,
                     double *h_T_last, double *h_T_total, int32_t *h_iters, int32_t *h_status, double *h_log,
                     cudaEvent_t src_ready) {
    if (P == 0) return PCP_OK;
    if (P > 0x7fffffffLL) return fail(PCP_ERR_UNSUPPORTED, "pcp_icp_batch: too many pairs");
    cudaStream_t st = ctx->stream;
    const int max_it = prm.max_iteration;

    // target hierarchies of all pairs in one segmented build (kept in the context between calls)
    BatchIndex &tb = ctx->icp_targets;
    tb.ctx = ctx;
    PCP_TRY(batch_build_device(&tb, d_tgt, stride, P, h_tgt_off));
    if (src_ready) PCP_CUDA(cudaStreamWaitEvent(st, src_ready, 0));   // sources still in flight on the side stream

    int64_t max_ns = 0;
    for (int64_t p = 0; p < P; ++p) {
        if (h_src_off[p + 1] < h_src_off[p]) return fail(PCP_ERR_INVALID, "pcp_icp_batch: src_offsets must be non-decreasing");
        max_ns = std::max<int64_t>(max_ns, h_src_off[p + 1] - h_src_off[p]);
    }
    // Tiled association (default): the sources are Morton-bucketed once per call with the same segmented build,
    // so that a warp's 32 source points are spatial neighbours.  PCP_ICP_KERNEL=warp keeps the warp-per-point walk.
    static const bool use_tile = !(getenv("PCP_ICP_KERNEL") && strcmp(getenv("PCP_ICP_KERNEL"), "warp") == 0);
    BatchIndex &sb = ctx->icp_sources;
    if (use_tile) {
        sb.ctx = ctx;
        PCP_TRY(batch_build_device(&sb, d_src, stride, P, h_src_off));
    }
    const int chunks = use_tile ? (int)std::max<int64_t>(1, (sb.max_leaf + kTileAssocWarps - 1) / kTileAssocWarps)
                                : (int)std::max<int64_t>(1, (max_ns + kChunk - 1) / kChunk);

    std::vector<IcpState> h_states((size_t)P);
    for (int64_t p = 0; p < P; ++p) {
        IcpState &s = h_states[(size_t)p];
        memset(&s, 0, sizeof s);
        double T[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
        if (h_T_init) memcpy(T, h_T_init + p * 16, sizeof T);
        for (int i = 0; i < 12; ++i) s.applied[i] = s.cur[i] = s.total[i] = T[i];
        for (int i = 0; i < 3; ++i) {                 // main.py:99-100
            for (int j = 0; j < 3; ++j) s.R_last[i * 3 + j] = T[i * 4 + j];
            s.t_last[i] = T[i * 4 + 3];
        }
        s.active = max_it > 0 ? 1 : 0;
        s.status = PCP_ICP_MAX_ITER;
    }
    DevBuf &b_states = ctx->scratch[0], &b_part = ctx->scratch[1], &b_off = ctx->scratch[2], &b_log = ctx->scratch[3],
           &b_act = ctx->scratch[4];
    PCP_TRY(b_states.ensure(sizeof(IcpState) * (size_t)P));
    PCP_TRY(b_part.ensure(sizeof(double) * kSums * (size_t)P * (size_t)chunks));
    PCP_TRY(b_off.ensure(sizeof(long long) * (size_t)(P + 1)));
    PCP_TRY(b_act.ensure(sizeof(int)));
    double *d_log = nullptr;
    if (h_log && max_it > 0) {
        PCP_TRY(b_log.ensure(sizeof(double) * 2 * (size_t)P * (size_t)max_it));
        d_log = b_log.as<double>();
        PCP_CUDA(cudaMemsetAsync(d_log, 0xff, sizeof(double) * 2 * (size_t)P * (size_t)max_it, st));   // NaN fill
    }
    PCP_TRY(small_h2d(ctx, b_states.p, h_states.data(), sizeof(IcpState) * (size_t)P, st));
    PCP_TRY(small_h2d(ctx, b_off.p, h_src_off, sizeof(long long) * (size_t)(P + 1), st));
    int h_active = max_it > 0 ? (int)P : 0;
    PCP_TRY(small_h2d(ctx, b_act.p, &h_active, sizeof(int), st));

    // No host round trip inside an iteration, and none between the iterations of a block: blocks of 4, 8, 16, ...
    // iterations are enqueued back to back (a pair that has stopped makes its CTAs return at once, so the
    // iterations enqueued past convergence cost a launch each), and the copy that ends a block brings back the
    // final states together with the "pairs still active" word.  The reference's own setting (main.py:98-102:
    // 1-2 iterations) therefore finishes with ONE synchronisation -- the result copy itself.
    const int64_t launches0 = ctx->launches;
    int it = 0, block = 4, syncs = 0;
    while (it < max_it && h_active > 0) {
        const int upto = std::min(max_it, it + block);
        for (; it < upto; ++it) {
            for (int64_t p0 = 0; p0 < P; p0 += 65535) {
                const unsigned gy = (unsigned)std::min<int64_t>(65535, P - p0);
                if (use_tile)
                    icp_assoc_tile_kernel<<<dim3((unsigned)chunks, gy), kTileAssocWarps * 32, 0, st>>>(
                        tb.d_views() + p0, sb.d_views() + p0, b_states.as<IcpState>() + p0,
                        b_part.as<double>() + (size_t)p0 * chunks * kSums, chunks, prm.dist_thres, ctx->ctr());
                else
                    icp_assoc_kernel<<<dim3((unsigned)chunks, gy), kAssocWarps * 32, 0, st>>>(
                        tb.d_views() + p0, d_src, b_off.as<long long>() + p0, stride, b_states.as<IcpState>() + p0,
                        b_part.as<double>() + (size_t)p0 * chunks * kSums, chunks, prm.dist_thres);
                ctx->launches += 1;
            }
            icp_solve_kernel<<<(unsigned)((P * 32 + 127) / 128), 128, 0, st>>>(
                b_states.as<IcpState>(), b_part.as<double>(), b_off.as<long long>(), P, chunks, it, max_it, prm.R_diff_thres,
                prm.t_diff_thres, prm.first_iter_broadcast, d_log, b_act.as<int>(), use_tile ? sb.d_views() : nullptr);
            ctx->launches += 1;
        }
        PCP_CUDA(cudaGetLastError());
        PCP_CUDA(cudaMemcpyAsync(h_states.data(), b_states.p, sizeof(IcpState) * (size_t)P, cudaMemcpyDeviceToHost, st));
        PCP_CUDA(cudaMemcpyAsync(&h_active, b_act.p, sizeof(int), cudaMemcpyDeviceToHost, st));
        PCP_CUDA(cudaStreamSynchronize(st));
        ++syncs;
        block *= 2;
    }
    if (syncs == 0) {   // max_iteration == 0: the states as initialised
        PCP_CUDA(cudaMemcpyAsync(h_states.data(), b_states.p, sizeof(IcpState) * (size_t)P, cudaMemcpyDeviceToHost, st));
        PCP_CUDA(cudaStreamSynchronize(st));
        ++syncs;
    }
    if (d_log) {
        PCP_CUDA(cudaMemcpyAsync(h_log, d_log, sizeof(double) * 2 * (size_t)P * (size_t)max_it, cudaMemcpyDeviceToHost, st));
        PCP_CUDA(cudaStreamSynchronize(st));
        ++syncs;
    }
    ctx->icp_last_launches = ctx->launches - launches0;
    ctx->icp_last_syncs = syncs;
    ctx->icp_last_enqueued = it;
    for (int64_t p = 0; p < P; ++p) {
        const IcpState &s = h_states[(size_t)p];
        const double bottom[4] = {0, 0, 0, 1};
        if (h_T_last) { memcpy(h_T_last + p * 16, s.cur, sizeof s.cur); memcpy(h_T_last + p * 16 + 12, bottom, sizeof bottom); }
        if (h_T_total) { memcpy(h_T_total + p * 16, s.total, sizeof s.total); memcpy(h_T_total + p * 16 + 12, bottom, sizeof bottom); }
        if (h_iters) h_iters[p] = s.iters;
        if (h_status) h_status[p] = s.status;
    }
    return PCP_OK;
}

}  // namespace pcp
