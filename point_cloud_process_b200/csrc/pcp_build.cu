// Index build: Morton-sorted 32-point buckets + implicit 32-ary AABB hierarchy,
// for P independent clouds ("segments") in one pass.  P = 1 is the pcp_index case; the ICP batch
// builds the hierarchies of all P target clouds with ONE segmented sort.
// Replaces kdtree_recursive_build (kdtree.py:59-97, np.argsort per node) and
// octree_recursive_build (octree.py:51-103, per-point Python loop computing the 3-bit child code).
//
// Kernels (all HBM-bound, one pass each; blockIdx.y = segment):
//   bbox_kernel        float3[n] -> 6 ordered-uint min/max words per segment  (np.amin/np.amax, octree.py:319-320)
//   key_kernel         float3[n] -> (segment | 30-bit Morton key, row)        (child-code loop, octree.py:75-85)
//   cub::DeviceRadixSort::SortPairs                                            (np.argsort, kdtree.py:46,76)
//   refine_leaf_kernel sorted rows -> per 1024-point node: 5 levels of widest-axis median splits in shared
//                      memory (the kd-tree's split rule, kdtree.py:74-97, applied below the Morton order)
//                      -> float4 bucket rows + bucket AABBs                    (Node / Octant objects)
//   node_box_kernel    32 child AABBs -> parent AABB, once per upper level
#include <cub/block/block_radix_sort.cuh>
#include <cub/device/device_radix_sort.cuh>

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "pcp_device.cuh"
#include "pcp_internal.cuh"

namespace pcp {

namespace {

constexpr int kMaxGridY = 65535;

// per segment: 8 floats (lo.xyz, scale, hi.xyz, extent) followed by 8 ordered-uint accumulators
constexpr int kBboxWords = 16;

// Bounding box of every segment in ONE launch.  Accumulators: words 8..13 of the segment's record, merged with
// atomicMax only -- the minima are kept complemented -- so that all-zero words are the identity; word 14 counts
// the CTAs that are done.  The last CTA of a segment turns the accumulators into the record morton30() reads
// {lo.xyz, cells-per-unit, hi.xyz, extent} and zeroes them again for the next build (the buffer is zeroed once
// when it is allocated).
__global__ void __launch_bounds__(256) bbox_kernel(const float *__restrict__ xyz, long long stride,
                                                   const IndexView *__restrict__ views) {
    const IndexView &v = views[blockIdx.y];
    const long long n = v.n;
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    const long long first = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    for (long long i = first; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float *p = xyz + (v.src_off + i) * stride;
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            float x = p[a];
            lo[a] = fminf(lo[a], x);
            hi[a] = fmaxf(hi[a], x);
        }
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        lo[a] = warp_min(lo[a]);
        hi[a] = warp_max(hi[a]);
    }
    // warp results -> shared memory -> one set of 6 atomics per CTA (thousands of same-address atomics
    // from every warp serialise in L2)
    __shared__ float s_lo[8][3], s_hi[8][3];
    __shared__ bool s_last;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) {
#pragma unroll
        for (int a = 0; a < 3; ++a) { s_lo[warp][a] = lo[a]; s_hi[warp][a] = hi[a]; }
    }
    __syncthreads();
    float *bbox = const_cast<float *>(v.bbox);
    unsigned *acc = reinterpret_cast<unsigned *>(bbox + 8);
    if (threadIdx.x < 6 && blockIdx.x * (long long)blockDim.x < n) {
        const int a = threadIdx.x % 3;
        const bool is_hi = threadIdx.x >= 3;
        float r = is_hi ? s_hi[0][a] : s_lo[0][a];
        for (int w = 1; w < 8; ++w) r = is_hi ? fmaxf(r, s_hi[w][a]) : fminf(r, s_lo[w][a]);
        atomicMax(&acc[threadIdx.x], is_hi ? float_to_ordered(r) : ~float_to_ordered(r));
        __threadfence();
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        s_last = atomicAdd(&acc[6], 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!s_last || threadIdx.x != 0) return;
    __threadfence();
    float l[3], h[3];
    for (int a = 0; a < 3; ++a) {
        l[a] = ordered_to_float(~atomicExch(&acc[a], 0u));         // (empty segment: ~0 and 0, as before)
        h[a] = ordered_to_float(atomicExch(&acc[3 + a], 0u));
    }
    acc[6] = 0u;
    float ext = fmaxf(fmaxf(h[0] - l[0], h[1] - l[1]), h[2] - l[2]);
    float scale = (ext > 0.f && isfinite(ext)) ? 1024.f / (ext * 1.0001f) : 0.f;
    bbox[0] = l[0]; bbox[1] = l[1]; bbox[2] = l[2]; bbox[3] = scale;
    bbox[4] = h[0]; bbox[5] = h[1]; bbox[6] = h[2]; bbox[7] = ext;
}

// key = (segment << seg_shift) | (morton30 >> skip): one index sorts plain 30-bit codes (skip = seg_shift = 0, the
// radix passes start at the first significant bit); a batch packs segment id and the significant Morton bits
// into as few bits as they need, so that 1 024 clouds of 20 000 points sort as 24-bit keys in 3 passes.
template <class KeyT>
__global__ void __launch_bounds__(256) key_kernel(const float *__restrict__ xyz, long long stride,
                                                  const IndexView *__restrict__ views, long long seg0, int skip,
                                                  int seg_shift, KeyT *__restrict__ keys, unsigned *__restrict__ vals) {
    const IndexView &v = views[blockIdx.y];
    long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= v.n) return;
    const float *p = xyz + (v.src_off + i) * stride;
    const KeyT seg = static_cast<KeyT>(seg0 + blockIdx.y);
    keys[v.src_off + i] = static_cast<KeyT>(seg << seg_shift) | static_cast<KeyT>(morton30(p[0], p[1], p[2], v.bbox) >> skip);
    vals[v.src_off + i] = static_cast<unsigned>(v.src_off + i);
}

// ---- kd refinement of one block of 1024 Morton-consecutive points ------------------------------
// A run of 32 points in Morton order is a poor bucket: it straddles cell boundaries and its AABB is
// 2-3x larger than the points' share of space.  Each block of 32 buckets (one level-1 node) is
// therefore re-partitioned in shared memory with the reference kd-tree's rule -- split at the
// median of the coordinate of largest extent, round-robin replaced by the widest axis
// (kdtree.py:74-97) -- 5 levels deep: 1024 -> 2 x 512 -> ... -> 32 buckets of 32.  Every level is a
// bitonic sort of (coordinate, slot) keys inside aligned power-of-two segments: strides below 32
// exchange by warp shuffle, larger ones through shared memory.  Pad slots (beyond the cloud) carry
// the largest key on every axis (no finite float maps to it), so real points always stay at the front of their segment and
// bucket b of the node holds sorted positions [32b, 32b+32) exactly as the hierarchy expects.
constexpr int kRefine = 1024;                 // points per block (one level-1 node)
constexpr int kRefineE = 8;                   // points per thread: slot s of the block lives in thread s / 8, register s % 8
constexpr int kRefineT = kRefine / kRefineE;  // threads per CTA
constexpr unsigned kPadKey = 0xffffffffu;

struct RefineSmem {
    unsigned ox[kRefine], oy[kRefine], oz[kRefine];   // coordinates as ordered integers, addressed by element id
    unsigned key[2][kRefine];                  // double-buffered exchange: one barrier per cross-warp stage
    unsigned box[2][6];                        // segments of 512 / 1024 slots span warps: ordered-uint min xyz, max xyz
};

// One compare-exchange stage of the direction-free bitonic network: slot s against slot s ^ MASK, the slot whose
// BIT is clear keeps the smaller key.  With eight consecutive slots per thread the strides 1, 2, 4 (and the mirror
// masks 1, 3, 7) stay inside the thread -- one min and one max per pair, no select, no exchange: 117 of the 200
// stages of a block.  Strides up to 128 exchange registers by warp shuffle, 256 and 512 through shared memory
// (4 stages of a block).
template <unsigned MASK, unsigned BIT>
__device__ __forceinline__ void refine_stage(RefineSmem &sm, unsigned (&k)[kRefineE], unsigned t, unsigned &phase) {
    if (MASK < kRefineE) {
#pragma unroll
        for (int r = 0; r < kRefineE; ++r) {
            const int p = r ^ (int)MASK;
            if (p > r) {
                const unsigned lo = min(k[r], k[p]), hi = max(k[r], k[p]);
                const bool first_low = (r & (int)BIT) == 0;
                k[r] = first_low ? lo : hi;
                k[p] = first_low ? hi : lo;
            }
        }
    } else {
        constexpr unsigned TM = MASK / kRefineE, RM = MASK % kRefineE;   // partner thread, partner register
        unsigned o[kRefineE];
        if (TM >= 32) {
            uint4 *mine = reinterpret_cast<uint4 *>(sm.key[phase] + t * kRefineE);
            mine[0] = make_uint4(k[0], k[1], k[2], k[3]);
            mine[1] = make_uint4(k[4], k[5], k[6], k[7]);
            __syncthreads();
            const uint4 *theirs = reinterpret_cast<const uint4 *>(sm.key[phase] + (t ^ TM) * kRefineE);
            const uint4 a = theirs[0], b = theirs[1];
            const unsigned v[kRefineE] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
            for (int r = 0; r < kRefineE; ++r) o[r] = v[r ^ (int)RM];
            phase ^= 1u;                       // the next exchange writes the other buffer: no second barrier
        } else {
#pragma unroll
            for (int r = 0; r < kRefineE; ++r) o[r] = __shfl_xor_sync(kFullMask, k[r ^ (int)RM], TM);
        }
        const bool keep_low = (t & (BIT / kRefineE)) == 0;
#pragma unroll
        for (int r = 0; r < kRefineE; ++r) k[r] = keep_low ? min(k[r], o[r]) : max(k[r], o[r]);
    }
}

// merge step K of the network (blocks of K become sorted): mirror stage, then strides K/4 ... 1
template <unsigned K, unsigned J>
struct RefineMerge {
    static __device__ __forceinline__ void run(RefineSmem &sm, unsigned (&k)[kRefineE], unsigned t, unsigned &phase) {
        refine_stage<J, J>(sm, k, t, phase);
        RefineMerge<K, J / 2>::run(sm, k, t, phase);
    }
};
template <unsigned K>
struct RefineMerge<K, 0> {
    static __device__ __forceinline__ void run(RefineSmem &, unsigned (&)[kRefineE], unsigned, unsigned &) {}
};
// ascending sort of every aligned segment of S slots, fully unrolled: K = 2, 4, ..., S
template <unsigned S, unsigned K = 2>
struct RefineSort {
    static __device__ __forceinline__ void run(RefineSmem &sm, unsigned (&k)[kRefineE], unsigned t, unsigned &phase) {
        refine_stage<K - 1, K / 2>(sm, k, t, phase);            // slot i of a block against slot K-1-i
        RefineMerge<K, K / 4>::run(sm, k, t, phase);
        RefineSort<S, K * 2>::run(sm, k, t, phase);
    }
};
template <unsigned S>
struct RefineSort<S, 2 * S> {
    static __device__ __forceinline__ void run(RefineSmem &, unsigned (&)[kRefineE], unsigned, unsigned &) {}
};

// One level of the kd refinement: every aligned segment of S slots is sorted along its widest axis, which splits
// it at the median into two segments of S/2.  id[r]: the element in slot 8 t + r; bit r of realm: it is a real point
// (pad elements carry the coordinates of a real one, so they never widen a box).
template <unsigned S>
__device__ __forceinline__ void refine_level(RefineSmem &sm, unsigned (&id)[kRefineE], unsigned &realm, unsigned &phase) {
    const unsigned t = threadIdx.x;
    constexpr unsigned TS = S / kRefineE;      // threads per segment: 4 .. 128
    // ---- widest axis of every segment
    unsigned mn[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu}, mx[3] = {0u, 0u, 0u};
#pragma unroll
    for (int r = 0; r < kRefineE; ++r) {
        const unsigned ox = sm.ox[id[r]], oy = sm.oy[id[r]], oz = sm.oz[id[r]];
        mn[0] = min(mn[0], ox); mn[1] = min(mn[1], oy); mn[2] = min(mn[2], oz);
        mx[0] = max(mx[0], ox); mx[1] = max(mx[1], oy); mx[2] = max(mx[2], oz);
    }
    if (TS == 32) {
#pragma unroll
        for (int a = 0; a < 3; ++a) { mn[a] = __reduce_min_sync(kFullMask, mn[a]); mx[a] = __reduce_max_sync(kFullMask, mx[a]); }
    } else if (TS < 32) {                      // the segment's threads are an aligned group of lanes: butterfly
#pragma unroll
        for (unsigned d = 1; d < TS; d *= 2) {
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                mn[a] = min(mn[a], __shfl_xor_sync(kFullMask, mn[a], d));
                mx[a] = max(mx[a], __shfl_xor_sync(kFullMask, mx[a], d));
            }
        }
    } else {
        const unsigned seg = t / TS;
        if (t < 12) sm.box[t / 6][t % 6] = (t % 6) < 3 ? 0xffffffffu : 0u;
        __syncthreads();
#pragma unroll
        for (int a = 0; a < 3; ++a) { mn[a] = __reduce_min_sync(kFullMask, mn[a]); mx[a] = __reduce_max_sync(kFullMask, mx[a]); }
        if ((t & 31u) == 0) {
#pragma unroll
            for (int a = 0; a < 3; ++a) { atomicMin(&sm.box[seg][a], mn[a]); atomicMax(&sm.box[seg][3 + a], mx[a]); }
        }
        __syncthreads();
#pragma unroll
        for (int a = 0; a < 3; ++a) { mn[a] = sm.box[seg][a]; mx[a] = sm.box[seg][3 + a]; }
    }
    const float wx = ordered_to_float(mx[0]) - ordered_to_float(mn[0]);
    const float wy = ordered_to_float(mx[1]) - ordered_to_float(mn[1]);
    const float wz = ordered_to_float(mx[2]) - ordered_to_float(mn[2]);
    const unsigned *c = sm.ox + ((wx >= wy && wx >= wz) ? 0 : (wy >= wz ? kRefine : 2 * kRefine));   // ox, oy, oz are contiguous
    // 22 coordinate bits + 10 element bits in one word: the split is a partition heuristic, the element id makes keys
    // unique and the order deterministic; pads carry the largest key
    unsigned k[kRefineE];
#pragma unroll
    for (int r = 0; r < kRefineE; ++r)
        k[r] = ((((realm >> r) & 1u) ? c[id[r]] : kPadKey) & ~(kRefine - 1u)) | id[r];
    RefineSort<S>::run(sm, k, t, phase);
    realm = 0;
#pragma unroll
    for (int r = 0; r < kRefineE; ++r) {
        id[r] = k[r] & (kRefine - 1u);
        realm |= ((k[r] | (kRefine - 1u)) != kPadKey ? 1u : 0u) << r;
    }
}

// Loads done (sm.ox/oy/oz and the caller's per-element payload filled for all 1024 slots, barrier NOT yet passed):
// gives pad elements the coordinates of the block's first point (always real), then runs the levels.
// Out: id[r] = the element (0..1023, the slot that loaded it) that ends up in slot 8 t + r.
// HALVES: a sixth level splits the 32-slot segments (buckets) into halves of 16.
template <bool HALVES>
__device__ __forceinline__ void refine_block(RefineSmem &sm, unsigned (&id)[kRefineE], unsigned &realm) {
    const unsigned t = threadIdx.x;
    __syncthreads();
    if (realm != 0xffu) {
        const unsigned x0 = sm.ox[0], y0 = sm.oy[0], z0 = sm.oz[0];
#pragma unroll
        for (int r = 0; r < kRefineE; ++r)
            if (!((realm >> r) & 1u)) { sm.ox[t * kRefineE + r] = x0; sm.oy[t * kRefineE + r] = y0; sm.oz[t * kRefineE + r] = z0; }
    }
    __syncthreads();
    unsigned phase = 0;
#pragma unroll
    for (int r = 0; r < kRefineE; ++r) id[r] = t * kRefineE + r;
    refine_level<1024>(sm, id, realm, phase);
    refine_level<512>(sm, id, realm, phase);
    refine_level<256>(sm, id, realm, phase);
    refine_level<128>(sm, id, realm, phase);
    refine_level<64>(sm, id, realm, phase);
    if (HALVES) refine_level<32>(sm, id, realm, phase);
}

__device__ __forceinline__ void refine_store8(unsigned *dst, const unsigned (&v)[kRefineE]) {
    reinterpret_cast<uint4 *>(dst)[0] = make_uint4(v[0], v[1], v[2], v[3]);
    reinterpret_cast<uint4 *>(dst)[1] = make_uint4(v[4], v[5], v[6], v[7]);
}

// One CTA per level-1 node: gather the node's points through the Morton order, refine, write the float4
// bucket rows (pad lanes NaN / index -1, so that every distance against them is NaN and fails every "<="
// gate), their pair-layout copy, and the bucket / half-bucket AABBs.
__global__ void __launch_bounds__(kRefineT) refine_leaf_kernel(const float *__restrict__ xyz, long long stride,
                                                                const IndexView *__restrict__ views,
                                                                const unsigned *__restrict__ sorted_rows, int fuse_upper) {
    __shared__ __align__(16) RefineSmem sm;
    __shared__ __align__(16) unsigned s_idx[kRefine];
    __shared__ unsigned s_node[6];                               // AABB of the whole block = the level-1 node
    const IndexView &v = views[blockIdx.y];
    const long long node = blockIdx.x;
    if (node * kFan >= v.n_leaf) return;
    const unsigned t = threadIdx.x;
    if (t < 6) s_node[t] = t < 3 ? 0xffffffffu : 0u;             // (ordered by the barriers inside refine_block)
    const long long pos0 = node * kRefine + t * kRefineE;
    unsigned id[kRefineE], realm = 0;
    {
        unsigned ox[kRefineE], oy[kRefineE], oz[kRefineE], idx[kRefineE];
#pragma unroll
        for (int r = 0; r < kRefineE; ++r) {
            ox[r] = oy[r] = oz[r] = 0u;
            idx[r] = 0xffffffffu;
            if (pos0 + r < v.n) {
                const unsigned row = sorted_rows[v.src_off + pos0 + r];   // sorted by (segment, morton): rows of this segment
                const float *p = xyz + (long long)row * stride;
                ox[r] = float_to_ordered(p[0]); oy[r] = float_to_ordered(p[1]); oz[r] = float_to_ordered(p[2]);
                idx[r] = row - static_cast<unsigned>(v.src_off);
                realm |= 1u << r;
            }
        }
        refine_store8(sm.ox + t * kRefineE, ox);
        refine_store8(sm.oy + t * kRefineE, oy);
        refine_store8(sm.oz + t * kRefineE, oz);
        refine_store8(s_idx + t * kRefineE, idx);
    }
    refine_block<true>(sm, id, realm);

    // slots 8 t .. 8 t + 7: a quarter of bucket t / 4, half of its 16-point half (t / 2) & 1
    const long long leaf = pos0 >> 5;
    const bool store = leaf < v.n_leaf;                          // else: bucket beyond the cloud
    const float nan = __uint_as_float(0x7fc00000u), inf = __uint_as_float(0x7f800000u);
    unsigned mn[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu}, mx[3] = {0u, 0u, 0u};
    float4 P[kRefineE];
#pragma unroll
    for (int r = 0; r < kRefineE; ++r) {
        const bool real = (realm >> r) & 1u;
        const unsigned ox = sm.ox[id[r]], oy = sm.oy[id[r]], oz = sm.oz[id[r]];
        mn[0] = min(mn[0], real ? ox : 0xffffffffu); mn[1] = min(mn[1], real ? oy : 0xffffffffu); mn[2] = min(mn[2], real ? oz : 0xffffffffu);
        mx[0] = max(mx[0], real ? ox : 0u); mx[1] = max(mx[1], real ? oy : 0u); mx[2] = max(mx[2], real ? oz : 0u);
        P[r] = real ? make_float4(ordered_to_float(ox), ordered_to_float(oy), ordered_to_float(oz), __uint_as_float(s_idx[id[r]]))
                    : make_float4(nan, nan, nan, __uint_as_float(0xffffffffu));
    }
    if (store) {
        float4 *dst = const_cast<float4 *>(v.pts) + pos0;
#pragma unroll
        for (int r = 0; r < kRefineE; ++r) dst[r] = P[r];
        // pair layout (IndexView::pts2): entries 2m / 2m+1 = (x0, x1, y0, y1) / (z0, z1, w0, w1) of points 2m, 2m+1
        float4 *dst2 = const_cast<float4 *>(v.pts2) + pos0;
#pragma unroll
        for (int m = 0; m < kRefineE / 2; ++m) {
            const bool ra = (realm >> (2 * m)) & 1u, rb = (realm >> (2 * m + 1)) & 1u;
            const float4 a = P[2 * m], b = P[2 * m + 1];
            dst2[2 * m] = make_float4(ra ? a.x : inf, rb ? b.x : inf, ra ? a.y : 0.f, rb ? b.y : 0.f);
            dst2[2 * m + 1] = make_float4(ra ? a.z : 0.f, rb ? b.z : 0.f, a.w, b.w);
        }
    }
    // AABBs of the two 16-point halves (the sixth split): an empty half gets (+inf, -inf), infinitely far from everything
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        mn[a] = min(mn[a], __shfl_xor_sync(kFullMask, mn[a], 1));
        mx[a] = max(mx[a], __shfl_xor_sync(kFullMask, mx[a], 1));
    }
    if (store && (t & 1u) == 0) {
        const bool any = mn[0] != 0xffffffffu;
        float4 *h = const_cast<float4 *>(v.half) + leaf * 4 + ((t & 2u) ? 2 : 0);
        h[0] = any ? make_float4(ordered_to_float(mn[0]), ordered_to_float(mn[1]), ordered_to_float(mn[2]), 0.f) : make_float4(inf, inf, inf, 0.f);
        h[1] = any ? make_float4(ordered_to_float(mx[0]), ordered_to_float(mx[1]), ordered_to_float(mx[2]), 0.f) : make_float4(-inf, -inf, -inf, 0.f);
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        mn[a] = min(mn[a], __shfl_xor_sync(kFullMask, mn[a], 2));
        mx[a] = max(mx[a], __shfl_xor_sync(kFullMask, mx[a], 2));
    }
    if (store && (t & 3u) == 0) {
        const_cast<float4 *>(v.box_lo)[leaf] = make_float4(ordered_to_float(mn[0]), ordered_to_float(mn[1]), ordered_to_float(mn[2]), 0.f);
        const_cast<float4 *>(v.box_hi)[leaf] = make_float4(ordered_to_float(mx[0]), ordered_to_float(mx[1]), ordered_to_float(mx[2]), 0.f);
    }
    // the block is one level-1 node: its box is the union of the bucket boxes (same values node_box_kernel
    // would compute: min / max of the same floats)
#pragma unroll
    for (int a = 0; a < 3; ++a) { mn[a] = __reduce_min_sync(kFullMask, mn[a]); mx[a] = __reduce_max_sync(kFullMask, mx[a]); }
    if ((t & 31u) == 0) {
#pragma unroll
        for (int a = 0; a < 3; ++a) { atomicMin(&s_node[a], mn[a]); atomicMax(&s_node[3 + a], mx[a]); }
    }
    __syncthreads();
    if (t == 0 && v.n_levels > 1) {
        const long long at = v.level_off[1] + node;
        const_cast<float4 *>(v.box_lo)[at] = make_float4(ordered_to_float(s_node[0]), ordered_to_float(s_node[1]), ordered_to_float(s_node[2]), 0.f);
        const_cast<float4 *>(v.box_hi)[at] = make_float4(ordered_to_float(s_node[3]), ordered_to_float(s_node[4]), ordered_to_float(s_node[5]), 0.f);
    }
    // fuse_upper (small hierarchies: <= 1024 level-1 nodes per cloud): the LAST CTA of a cloud to get here also
    // computes the few levels above (KITTI: 122 -> 4 -> 1 nodes), instead of a separate launch behind this kernel.
    // Ticket = word 15 of the cloud's bounding-box record (zero between builds).
    if (fuse_upper && v.n_levels > 2) {                          // (uniform over the CTA)
        __shared__ bool s_last;
        unsigned *ticket = reinterpret_cast<unsigned *>(const_cast<float *>(v.bbox)) + 15;
        if (t == 0) {
            __threadfence();
            s_last = atomicAdd(ticket, 1u) == (unsigned)v.level_cnt[1] - 1u;
        }
        __syncthreads();
        if (s_last) {
            __threadfence();
            float4 *box_lo = const_cast<float4 *>(v.box_lo), *box_hi = const_cast<float4 *>(v.box_hi);
            const unsigned lane = t & 31u;
            for (int level = 2; level < v.n_levels; ++level) {
                for (long long nd = t >> 5; nd < v.level_cnt[level]; nd += kRefineT / 32) {
                    const long long c = nd * kFan + lane;
                    float4 lo = make_float4(FLT_MAX, FLT_MAX, FLT_MAX, 0.f), hi = make_float4(-FLT_MAX, -FLT_MAX, -FLT_MAX, 0.f);
                    if (c < v.level_cnt[level - 1]) {            // (written by other CTAs / the previous round: past L1)
                        lo = __ldcg(box_lo + v.level_off[level - 1] + c);
                        hi = __ldcg(box_hi + v.level_off[level - 1] + c);
                    }
                    const float lx = warp_min(lo.x), ly = warp_min(lo.y), lz = warp_min(lo.z);
                    const float hx = warp_max(hi.x), hy = warp_max(hi.y), hz = warp_max(hi.z);
                    if (lane == 0) {
                        box_lo[v.level_off[level] + nd] = make_float4(lx, ly, lz, 0.f);
                        box_hi[v.level_off[level] + nd] = make_float4(hx, hy, hz, 0.f);
                    }
                }
                __threadfence();
                __syncthreads();
            }
            if (t == 0) *ticket = 0u;
        }
    }
}

// Same refinement for a sorted QUERY set: blocks of 1024 Morton-consecutive queries are re-grouped into 32
// tight groups of 32 (one group = one warp of the tiled kNN kernel).  In place on the permutation.
__global__ void __launch_bounds__(kRefineT) refine_queries_kernel(const float *__restrict__ q, long long nq, long long stride,
                                                                   unsigned *__restrict__ perm) {
    __shared__ __align__(16) RefineSmem sm;
    __shared__ __align__(16) unsigned s_idx[kRefine];
    const unsigned t = threadIdx.x;
    const long long pos0 = blockIdx.x * (long long)kRefine + t * kRefineE;
    unsigned id[kRefineE], realm = 0;
    {
        unsigned ox[kRefineE], oy[kRefineE], oz[kRefineE], idx[kRefineE];
#pragma unroll
        for (int r = 0; r < kRefineE; ++r) {
            ox[r] = oy[r] = oz[r] = 0u;
            idx[r] = 0xffffffffu;
            if (pos0 + r < nq) {
                idx[r] = perm[pos0 + r];
                const float *p = q + (long long)idx[r] * stride;
                ox[r] = float_to_ordered(p[0]); oy[r] = float_to_ordered(p[1]); oz[r] = float_to_ordered(p[2]);
                realm |= 1u << r;
            }
        }
        refine_store8(sm.ox + t * kRefineE, ox);
        refine_store8(sm.oy + t * kRefineE, oy);
        refine_store8(sm.oz + t * kRefineE, oz);
        refine_store8(s_idx + t * kRefineE, idx);
    }
    refine_block<false>(sm, id, realm);
    // real queries stay at the front of the block: slot s < nq - base is real
#pragma unroll
    for (int r = 0; r < kRefineE; ++r)
        if (pos0 + r < nq) perm[pos0 + r] = s_idx[id[r]];
}

// ---- small clouds (<= 20 480 points each, e.g. the 20 000-point registration clouds): ONE CTA per cloud does
// what bbox_kernel + key_kernel + the device-wide radix sort do for a batch -- bounding box, Morton keys, stable
// sort by the significant Morton bits -- with the keys in registers / shared memory instead of three radix passes
// over 16 B per point in HBM.  The word that is sorted is (key << 15 | index in the cloud); only the key bits take
// part and cub's block sort is stable on the blocked arrangement (= index order), so the permutation is the one
// the device-wide sort of (segment | key) produces and the index that comes out is bit-identical.
constexpr int kCsThreads = 512, kCsItems = 40, kCsMaxPoints = kCsThreads * kCsItems, kCsIdxBits = 15;
using CloudSort = cub::BlockRadixSort<unsigned, kCsThreads, kCsItems, cub::NullType, 5>;

__global__ void __launch_bounds__(kCsThreads) cloud_sort_kernel(const float *__restrict__ xyz, long long stride,
                                                                 const IndexView *__restrict__ views, int skip,
                                                                 int key_bits, unsigned *__restrict__ sorted_rows) {
    extern __shared__ __align__(16) unsigned char s_cs[];
    CloudSort::TempStorage &temp = *reinterpret_cast<CloudSort::TempStorage *>(s_cs);
    __shared__ float s_lo[kCsThreads / 32][3], s_hi[kCsThreads / 32][3];
    __shared__ float s_bbox[4];
    const IndexView &v = views[blockIdx.x];
    const int n = (int)v.n;
    const int t = threadIdx.x;
    const float *base = xyz + v.src_off * stride;
    // bounding box (the record bbox_kernel writes)
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int i = t; i < n; i += kCsThreads) {
        const float *p = base + (long long)i * stride;
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const float x = p[a];
            lo[a] = fminf(lo[a], x);
            hi[a] = fmaxf(hi[a], x);
        }
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        lo[a] = warp_min(lo[a]);
        hi[a] = warp_max(hi[a]);
    }
    if ((t & 31) == 0) {
#pragma unroll
        for (int a = 0; a < 3; ++a) { s_lo[t >> 5][a] = lo[a]; s_hi[t >> 5][a] = hi[a]; }
    }
    __syncthreads();
    if (t == 0) {
        float l[3], h[3];
        for (int a = 0; a < 3; ++a) {
            l[a] = s_lo[0][a]; h[a] = s_hi[0][a];
            for (int w = 1; w < kCsThreads / 32; ++w) { l[a] = fminf(l[a], s_lo[w][a]); h[a] = fmaxf(h[a], s_hi[w][a]); }
            if (n == 0) { l[a] = ordered_to_float(0xffffffffu); h[a] = ordered_to_float(0u); }   // (what bbox_kernel leaves)
            else { l[a] = ordered_to_float(float_to_ordered(l[a])); h[a] = ordered_to_float(float_to_ordered(h[a])); }
        }
        const float ext = fmaxf(fmaxf(h[0] - l[0], h[1] - l[1]), h[2] - l[2]);
        const float scale = (ext > 0.f && isfinite(ext)) ? 1024.f / (ext * 1.0001f) : 0.f;
        float *bbox = const_cast<float *>(v.bbox);
        bbox[0] = l[0]; bbox[1] = l[1]; bbox[2] = l[2]; bbox[3] = scale;
        bbox[4] = h[0]; bbox[5] = h[1]; bbox[6] = h[2]; bbox[7] = ext;
        s_bbox[0] = l[0]; s_bbox[1] = l[1]; s_bbox[2] = l[2]; s_bbox[3] = scale;
    }
    __syncthreads();
    // keys, blocked arrangement: thread t holds points 40 t .. 40 t + 39 (pads: the largest word, they stay last)
    unsigned key[kCsItems];
#pragma unroll
    for (int j = 0; j < kCsItems; ++j) {
        const int i = t * kCsItems + j;
        key[j] = 0xffffffffu;
        if (i < n) {
            const float *p = base + (long long)i * stride;
            key[j] = ((morton30(p[0], p[1], p[2], s_bbox) >> skip) << kCsIdxBits) | (unsigned)i;
        }
    }
    CloudSort(temp).SortBlockedToStriped(key, kCsIdxBits, kCsIdxBits + key_bits);
    const unsigned off = (unsigned)v.src_off;
#pragma unroll
    for (int j = 0; j < kCsItems; ++j) {
        const int r = j * kCsThreads + t;
        if (r < n) sorted_rows[v.src_off + r] = off + (key[j] & ((1u << kCsIdxBits) - 1u));
    }
}

__global__ void __launch_bounds__(256) node_box_kernel(const IndexView *__restrict__ views, int level) {
    const IndexView &v = views[blockIdx.y];
    if (level >= v.n_levels) return;
    long long node = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (node >= v.level_cnt[level]) return;
    float4 *box_lo = const_cast<float4 *>(v.box_lo), *box_hi = const_cast<float4 *>(v.box_hi);
    long long c = node * kFan + lane;
    float4 lo = make_float4(FLT_MAX, FLT_MAX, FLT_MAX, 0.f), hi = make_float4(-FLT_MAX, -FLT_MAX, -FLT_MAX, 0.f);
    if (c < v.level_cnt[level - 1]) {
        lo = box_lo[v.level_off[level - 1] + c];
        hi = box_hi[v.level_off[level - 1] + c];
    }
    float lx = warp_min(lo.x), ly = warp_min(lo.y), lz = warp_min(lo.z);
    float hx = warp_max(hi.x), hy = warp_max(hi.y), hz = warp_max(hi.z);
    if (lane == 0) {
        box_lo[v.level_off[level] + node] = make_float4(lx, ly, lz, 0.f);
        box_hi[v.level_off[level] + node] = make_float4(hx, hy, hz, 0.f);
    }
}

// Levels first_level .. top of every segment in one launch (one CTA per segment): used once a level is small
// enough (<= kFusedLevelNodes nodes) that a launch per level costs more than the serial walk.
constexpr long long kFusedLevelNodes = 256;
__global__ void __launch_bounds__(256) upper_boxes_kernel(const IndexView *__restrict__ views, int first_level) {
    const IndexView &v = views[blockIdx.y];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float4 *box_lo = const_cast<float4 *>(v.box_lo), *box_hi = const_cast<float4 *>(v.box_hi);
    for (int level = first_level; level < v.n_levels; ++level) {
        for (long long node = warp; node < v.level_cnt[level]; node += 8) {
            const long long c = node * kFan + lane;
            float4 lo = make_float4(FLT_MAX, FLT_MAX, FLT_MAX, 0.f), hi = make_float4(-FLT_MAX, -FLT_MAX, -FLT_MAX, 0.f);
            if (c < v.level_cnt[level - 1]) {
                lo = box_lo[v.level_off[level - 1] + c];
                hi = box_hi[v.level_off[level - 1] + c];
            }
            const float lx = warp_min(lo.x), ly = warp_min(lo.y), lz = warp_min(lo.z);
            const float hx = warp_max(hi.x), hy = warp_max(hi.y), hz = warp_max(hi.z);
            if (lane == 0) {
                box_lo[v.level_off[level] + node] = make_float4(lx, ly, lz, 0.f);
                box_hi[v.level_off[level] + node] = make_float4(hx, hy, hz, 0.f);
            }
        }
        __syncthreads();                       // the next level reads what this one wrote (same CTA)
    }
}

__global__ void __launch_bounds__(256) query_key_kernel(const float *__restrict__ q, long long nq, long long stride,
                                                        const float *__restrict__ bbox, unsigned *__restrict__ keys,
                                                        unsigned *__restrict__ vals) {
    long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= nq) return;
    const float *p = q + i * stride;
    keys[i] = morton30(p[0], p[1], p[2], bbox);
    vals[i] = static_cast<unsigned>(i);
}

// ---- one cloud / one query set of up to 2^18 points: Morton keys and the stable two-pass radix sort in ONE kernel.
// cub's device sort is four launches here (histogram, scan, two onesweep passes of ~11 us each: 22 CTAs of 5 888 keys
// with a serial look-back chain) on the critical path of the KITTI step.  This kernel spreads the cloud over up to
// one CTA per SM (256 threads, <= 8 keys per thread, kept in registers across the passes) and synchronises the grid
// itself.  Per pass (8-bit digit, LSD):
//   rank    every warp owns a contiguous run of its CTA's slice, lane-striped; it walks its <= 8 rows in order:
//           peers = match_any(digit), rank inside the warp = the warp's running count of the digit (shared memory,
//           private to the warp) + lower peers.  No shared-memory atomics; the running counts end as the warp's
//           digit histogram.
//   count   thread d adds the 8 warp histograms -> hist[cta][d] in global memory, and turns the warp histograms
//           into exclusive warp offsets.                                                     -- grid barrier --
//   scatter thread d: digit base = exclusive scan over d of the column totals + the column's CTAs before this one;
//           destination = base + warp offset + rank inside the warp: the order of equal digits is (CTA, warp, row,
//           lane) = input order, so each pass is stable and the permutation is the one cub produces.
// Grid barriers: one counter each, G arrivals (the grid is at most one CTA per SM of 256 threads, so it is
// co-resident beside anything else the library runs, and nothing it runs beside waits for it); counters are zero on
// entry and the last CTA to leave zeroes them again.  Data written by other CTAs is read with ld.cg (past L1).
constexpr int kSsThreads = 256, kSsItems = 8, kSsWarps = kSsThreads / 32;
constexpr long long kSsMaxPoints = 1LL << 18;

__device__ __forceinline__ void ss_grid_barrier(unsigned *counter, unsigned arrivals) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(counter, 1u);
        unsigned seen;
        do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(counter) : "memory");
            if (seen < arrivals) __nanosleep(64);
        } while (seen < arrivals);
        __threadfence();
    }
    __syncthreads();
}

__global__ void __launch_bounds__(kSsThreads)
small_sort_kernel(const float *__restrict__ xyz, long long n, long long stride, const float *__restrict__ bbox, int skip,
                  int items, unsigned *tmp_key, unsigned *tmp_val, unsigned *__restrict__ out_val, unsigned *hist,
                  unsigned *sync) {
    __shared__ unsigned s_cw[kSsWarps][256];   // per warp: running count of a digit, then its histogram, then its offset
    __shared__ unsigned s_base[256];
    __shared__ unsigned s_wsum[kSsWarps];
    const unsigned t = threadIdx.x, lane = t & 31u, warp = t >> 5, G = gridDim.x, cta = blockIdx.x;
    const unsigned lt_mask = (1u << lane) - 1u;
    // this warp's run of rows: [first, first + items * 32), row (it, lane) = first + it * 32 + lane
    const long long first = ((long long)cta * kSsWarps + warp) * items * 32;
    unsigned key[kSsItems], val[kSsItems];
#pragma unroll
    for (int it = 0; it < kSsItems; ++it) {
        const long long i = first + it * 32 + lane;
        key[it] = 0xffffffffu;                 // (beyond the run / the cloud: no digit, sorts nowhere)
        val[it] = 0u;
        if (it < items && i < n) {
            const float *p = xyz + i * stride;
            key[it] = morton30(__ldg(p), __ldg(p + 1), __ldg(p + 2), bbox) >> skip;     // 16 significant bits
            val[it] = (unsigned)i;
        }
    }
#pragma unroll 1
    for (int pass = 0; pass < 2; ++pass) {
        const int shift = pass * 8;
        unsigned *h = hist + (size_t)pass * G * 256;
        for (unsigned d = lane; d < 256; d += 32) s_cw[warp][d] = 0u;
        __syncwarp();
        unsigned rank[kSsItems];
#pragma unroll
        for (int it = 0; it < kSsItems; ++it) {
            rank[it] = 0;
            if (it < items) {                                        // (uniform)
                const bool live = key[it] != 0xffffffffu;
                const unsigned dg = live ? (key[it] >> shift) & 255u : 256u + lane;   // dead rows match nobody
                const unsigned peers = __match_any_sync(kFullMask, dg);
                unsigned prior = 0;
                if (live) prior = s_cw[warp][dg];
                __syncwarp();
                rank[it] = prior + __popc(peers & lt_mask);
                if (live && (peers & lt_mask) == 0u) s_cw[warp][dg] = prior + __popc(peers);
                __syncwarp();
            }
        }
        __syncthreads();
        {   // thread d: the CTA's count of digit d; warp histograms -> exclusive warp offsets
            unsigned run = 0;
#pragma unroll
            for (int w = 0; w < kSsWarps; ++w) {
                const unsigned c = s_cw[w][t];
                s_cw[w][t] = run;
                run += c;
            }
            h[(size_t)cta * 256 + t] = run;
        }
        ss_grid_barrier(sync + 2 * pass, G);
        {   // digit bases
            unsigned total = 0, before = 0;
            for (unsigned c0 = 0; c0 < G; c0 += 16) {            // 16 independent loads in flight (L2 latency each)
                unsigned v[16];
#pragma unroll
                for (unsigned j = 0; j < 16; ++j) v[j] = c0 + j < G ? __ldcg(h + (size_t)(c0 + j) * 256 + t) : 0u;
#pragma unroll
                for (unsigned j = 0; j < 16; ++j) {
                    total += v[j];
                    before += c0 + j < cta ? v[j] : 0u;
                }
            }
            unsigned incl = total;
#pragma unroll
            for (int d = 1; d < 32; d *= 2) {
                const unsigned o = __shfl_up_sync(kFullMask, incl, d);
                if ((int)lane >= d) incl += o;
            }
            if (lane == 31u) s_wsum[warp] = incl;
            __syncthreads();
            unsigned wbase = 0;
            for (unsigned w = 0; w < warp; ++w) wbase += s_wsum[w];
            s_base[t] = wbase + incl - total + before;
        }
        __syncthreads();
        if (pass == 0) {
#pragma unroll
            for (int it = 0; it < kSsItems; ++it) {
                if (it < items && key[it] != 0xffffffffu) {
                    const unsigned dg = key[it] & 255u;
                    const unsigned dst = s_base[dg] + s_cw[warp][dg] + rank[it];
                    tmp_key[dst] = key[it];
                    tmp_val[dst] = val[it];
                }
            }
            ss_grid_barrier(sync + 1, G);
#pragma unroll
            for (int it = 0; it < kSsItems; ++it) {
                const long long i = first + it * 32 + lane;
                key[it] = 0xffffffffu;
                if (it < items && i < n) { key[it] = __ldcg(tmp_key + i); val[it] = __ldcg(tmp_val + i); }
            }
        } else {
#pragma unroll
            for (int it = 0; it < kSsItems; ++it) {
                if (it < items && key[it] != 0xffffffffu) {
                    const unsigned dg = (key[it] >> 8) & 255u;
                    out_val[s_base[dg] + s_cw[warp][dg] + rank[it]] = val[it];
                }
            }
        }
        __syncthreads();                       // (s_cw / s_base are rewritten by the next pass)
    }
    // the last CTA to leave resets the barrier counters (every CTA has passed all three by then)
    if (t == 0) {
        __threadfence();
        if (atomicAdd(sync + 3, 1u) == G - 1) { sync[0] = 0u; sync[1] = 0u; sync[2] = 0u; sync[3] = 0u; __threadfence(); }
    }
}

// Launch helper: false if the set is outside the kernel's range (the caller falls back to cub).
bool small_sort_applies(pcp_ctx *ctx, long long n, int skip) {
    static const bool off = getenv("PCP_SMALL_SORT") && atoi(getenv("PCP_SMALL_SORT")) == 0;   // developer A/B
    return !off && n > 0 && n <= kSsMaxPoints && skip == 14 && (long long)ctx->sm_count * kSsThreads * kSsItems >= n;
}

int small_sort(pcp_ctx *ctx, const float *d_xyz, long long n, long long stride, const float *d_bbox, int skip, unsigned *tmp_key,
               unsigned *tmp_val, unsigned *out_val, DevBuf &state, cudaStream_t st) {
    static const long long per_cta = getenv("PCP_SS_PER_CTA") ? atoll(getenv("PCP_SS_PER_CTA")) : 2048;   // developer sweep (KITTI: 24.4 us at 1024, 23.1 us at 2048)
    const long long g_min = (n + (long long)kSsThreads * kSsItems - 1) / ((long long)kSsThreads * kSsItems);
    const unsigned G = (unsigned)std::min<long long>(ctx->sm_count, std::max<long long>(g_min, (n + per_cta - 1) / per_cta));
    const int items = (int)((n + (long long)G * kSsThreads - 1) / ((long long)G * kSsThreads));
    // state: 4 barrier counters (zero between launches) + 2 x G x 256 histogram words
    const size_t need = sizeof(unsigned) * (64 + 2 * (size_t)ctx->sm_count * 256);
    if (state.cap < need) {
        PCP_TRY(state.ensure(need));
        PCP_CUDA(cudaMemsetAsync(state.p, 0, state.cap, st));
    }
    unsigned *sync = state.as<unsigned>(), *hist = sync + 64;
    small_sort_kernel<<<G, kSsThreads, 0, st>>>(d_xyz, n, stride, d_bbox, skip, items, tmp_key, tmp_val, out_val, hist, sync);
    ctx->launches += 1;
    PCP_CUDA(cudaGetLastError());
    return PCP_OK;
}

template <class KeyT>
int radix_sort_pairs(pcp_ctx *ctx, const KeyT *k_in, KeyT *k_out, const unsigned *v_in, unsigned *v_out, long long n,
                     int begin_bit, int end_bit, cudaStream_t st = nullptr, DevBuf *tmp_buf = nullptr) {
    if (!st) st = ctx->stream;
    DevBuf &cub_tmp = tmp_buf ? *tmp_buf : ctx->cub_tmp;
    size_t tmp_bytes = 0;
    PCP_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, k_in, k_out, v_in, v_out, n, begin_bit, end_bit, st));
    PCP_TRY(cub_tmp.ensure(tmp_bytes));
    PCP_CUDA(cub::DeviceRadixSort::SortPairs(cub_tmp.p, tmp_bytes, k_in, k_out, v_in, v_out, n, begin_bit, end_bit, st));
    ctx->launches += 2 + (end_bit - begin_bit + 7) / 8;   // onesweep: histogram + scan + one pass per 8-bit digit (CUB internal)
    return PCP_OK;
}

// Low Morton bits that the sort may ignore: the order only has to make every run of 1024 points spatially
// coherent (the kd refinement re-partitions inside the run), so cells holding ~4 points on average are fine
// enough -- 16 significant bits up to 256 k points, 24 up to 64 M, all 30 beyond.  One radix pass per 8 bits.
inline int morton_skip_bits(long long n) {
    static const char *e = getenv("PCP_MORTON_SKIP");   // developer override
    if (e) return atoi(e);
    return n <= (1LL << 18) ? 14 : n <= (1LL << 26) ? 6 : 0;
}

inline unsigned grid_for(long long n, int block) { return static_cast<unsigned>(std::max<long long>(1, (n + block - 1) / block)); }

}  // namespace

namespace {
__global__ void small_copy_kernel(unsigned *__restrict__ dst, const unsigned *__restrict__ src, size_t words) {
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i < words) dst[i] = src[i];
}
}  // namespace

int small_h2d(pcp_ctx *ctx, void *dst, const void *src, size_t bytes, cudaStream_t st) {
    if (bytes == 0) return PCP_OK;
    const size_t at = (ctx->small_used + 15) & ~(size_t)15;
    if (!ctx->h2d_by_kernel || (bytes & 3) || at + bytes > ctx->small_stage.cap) {
        PCP_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, st));
        return PCP_OK;
    }
    unsigned char *stage = static_cast<unsigned char *>(ctx->small_stage.p) + at;
    memcpy(stage, src, bytes);
    ctx->small_used = at + bytes;
    const size_t words = bytes / 4;
    small_copy_kernel<<<(unsigned)((words + 255) / 256), 256, 0, st>>>(static_cast<unsigned *>(dst),
                                                                      reinterpret_cast<const unsigned *>(stage), words);
    ctx->launches += 1;
    PCP_CUDA(cudaGetLastError());
    return PCP_OK;
}

void BatchIndex::release() {
    if (ev_bbox) { cudaEventDestroy(ev_bbox); ev_bbox = nullptr; }
    for (DevBuf *d : {&pts, &pts2, &box_lo, &box_hi, &half, &bbox, &views, &keys_a, &keys_b, &vals_a, &vals_b}) d->release();
    h_views.clear();
    P = 0;
}

int batch_build_device(BatchIndex *b, const float *d_xyz, int64_t stride, int64_t P, const int64_t *h_offsets) {
    pcp_ctx *ctx = b->ctx;
    cudaStream_t st = ctx->stream;
    b->P = P;
    b->h_views.assign((size_t)P, IndexView());
    b->total_n = P ? h_offsets[P] - h_offsets[0] : 0;
    b->total_leaf = b->total_nodes = b->max_n = b->max_leaf = 0;
    b->max_levels = 0;
    for (int l = 0; l < kMaxLevels; ++l) b->level_max_cnt[l] = 0;
    if (P == 0) return PCP_OK;
    if (h_offsets[0] != 0) return fail(PCP_ERR_INVALID, "index build: offsets[0] must be 0");
    if (b->total_n > 0xfffffffeLL) return fail(PCP_ERR_UNSUPPORTED, "index build: more than 2^32-2 points in one batch");

    // hierarchy shapes (host): per segment, levels of ceil(cnt/32) nodes down to a single root
    for (int64_t p = 0; p < P; ++p) {
        IndexView &v = b->h_views[(size_t)p];
        const int64_t n = h_offsets[p + 1] - h_offsets[p];
        if (n < 0) return fail(PCP_ERR_INVALID, "index build: offsets must be non-decreasing");
        if (n > 0x7fffffffLL) return fail(PCP_ERR_UNSUPPORTED, "index build: more than 2^31-1 points in one cloud");
        v.src_off = h_offsets[p];
        v.n = n;
        v.n_leaf = (n + kLeaf - 1) / kLeaf;
        v.n_levels = 0;
        // a non-empty segment always has a root ABOVE the bucket level (n_levels >= 2), so the
        // search kernels have a single traversal loop and a single bucket-scan site
        long long cnt = v.n_leaf, off = 0;
        while (cnt > 0) {
            if (v.n_levels >= kMaxLevels) return fail(PCP_ERR_UNSUPPORTED, "index build: hierarchy too deep");
            v.level_off[v.n_levels] = off;
            v.level_cnt[v.n_levels] = cnt;
            b->level_max_cnt[v.n_levels] = std::max<int64_t>(b->level_max_cnt[v.n_levels], cnt);
            ++v.n_levels;
            off += cnt;
            if (cnt == 1 && v.n_levels >= 2) break;
            cnt = (cnt + kFan - 1) / kFan;
        }
        // pointers are patched below once the arrays are sized: keep element offsets for now
        v.pts = reinterpret_cast<const float4 *>((uintptr_t)(b->total_leaf * kLeaf));
        v.box_lo = reinterpret_cast<const float4 *>((uintptr_t)b->total_nodes);
        b->total_leaf += v.n_leaf;
        b->total_nodes += off;
        b->max_n = std::max<int64_t>(b->max_n, n);
        b->max_leaf = std::max<int64_t>(b->max_leaf, v.n_leaf);
        b->max_levels = std::max(b->max_levels, v.n_levels);
    }

    PCP_TRY(b->pts.ensure(sizeof(float4) * (size_t)std::max<int64_t>(1, b->total_leaf) * kLeaf));
    PCP_TRY(b->pts2.ensure(sizeof(float4) * (size_t)std::max<int64_t>(1, b->total_leaf) * kLeaf));
    PCP_TRY(b->box_lo.ensure(sizeof(float4) * (size_t)std::max<int64_t>(1, b->total_nodes)));
    PCP_TRY(b->box_hi.ensure(sizeof(float4) * (size_t)std::max<int64_t>(1, b->total_nodes)));
    PCP_TRY(b->half.ensure(sizeof(float4) * 4 * (size_t)std::max<int64_t>(1, b->total_leaf)));
    {
        const size_t before = b->bbox.cap;
        PCP_TRY(b->bbox.ensure(sizeof(float) * kBboxWords * (size_t)P));
        // fresh allocation: zero = the identity of bbox_kernel's accumulators (which clean up after themselves)
        if (b->bbox.cap != before) PCP_CUDA(cudaMemsetAsync(b->bbox.p, 0, b->bbox.cap, st));
    }
    PCP_TRY(b->views.ensure(sizeof(IndexView) * (size_t)P));
    for (int64_t p = 0; p < P; ++p) {
        IndexView &v = b->h_views[(size_t)p];
        const uintptr_t pt_off = reinterpret_cast<uintptr_t>(v.pts), node_off = reinterpret_cast<uintptr_t>(v.box_lo);
        v.pts = b->pts.as<float4>() + pt_off;
        v.pts2 = b->pts2.as<float4>() + pt_off;
        v.box_lo = b->box_lo.as<float4>() + node_off;
        v.box_hi = b->box_hi.as<float4>() + node_off;
        v.half = b->half.as<float4>() + pt_off / kLeaf * 4;
        v.bbox = b->bbox.as<float>() + p * kBboxWords;
    }
    // views are small (184 B per segment): pageable copy, staged by the driver before it returns
    PCP_TRY(small_h2d(ctx, b->views.p, b->h_views.data(), sizeof(IndexView) * (size_t)P, st));
    if (b->total_n == 0) return PCP_OK;

    // Key layout.  Batch: segment id above the significant Morton bits -- per cloud one bit fewer than
    // ceil(log2 n) (~2 points per cell, what 16 bits are to the KITTI scan; never more than a single index of that
    // size would sort) -- in 32-bit keys when both fit, else 64-bit keys.
    const bool wide = P > 1;
    int seg_bits = 0, mbits = 30 - morton_skip_bits(wide ? b->max_n : b->total_n);
    if (wide) {
        seg_bits = 1;
        while ((1LL << seg_bits) < P) ++seg_bits;
        int lg = 0;
        while ((1LL << lg) < b->max_n) ++lg;
        mbits = std::max(8, std::min(mbits, lg - 1));
    }
    static const bool force64 = getenv("PCP_BUILD_KEYS") && strcmp(getenv("PCP_BUILD_KEYS"), "64") == 0;   // tests
    const bool wide64 = wide && (seg_bits + mbits > 32 || force64);
    const size_t key_bytes = wide64 ? 8 : 4;
    PCP_TRY(b->keys_a.ensure(key_bytes * (size_t)b->total_n));
    PCP_TRY(b->keys_b.ensure(key_bytes * (size_t)b->total_n));
    PCP_TRY(b->vals_a.ensure(sizeof(unsigned) * (size_t)b->total_n));
    PCP_TRY(b->vals_b.ensure(sizeof(unsigned) * (size_t)b->total_n));
    const IndexView *dv = b->d_views();

    static const bool no_cloud_sort = getenv("PCP_BUILD_CLOUD_SORT") && atoi(getenv("PCP_BUILD_CLOUD_SORT")) == 0;   // developer A/B
    const bool cloud_sort = wide && !force64 && !no_cloud_sort && b->max_n <= kCsMaxPoints && mbits + kCsIdxBits <= 32;
    if (cloud_sort) {
        static bool attr_dev[kMaxDevices] = {false};   // function attributes are per device
        bool &attr = attr_dev[ctx->device & (kMaxDevices - 1)];
        if (!attr) {
            PCP_CUDA(cudaFuncSetAttribute(cloud_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)sizeof(CloudSort::TempStorage)));
            attr = true;
        }
        cloud_sort_kernel<<<(unsigned)P, kCsThreads, sizeof(CloudSort::TempStorage), st>>>(d_xyz, stride, dv, 30 - mbits, mbits,
                                                                                          b->vals_b.as<unsigned>());
        ctx->launches += 1;
        if (b->ev_bbox_record) PCP_CUDA(cudaEventRecord(b->ev_bbox_record, st));
    } else {
        const unsigned bx = (unsigned)std::min<long long>(grid_for(b->max_n, 256),
                                                          P == 1 ? (long long)ctx->sm_count * 2 : 16);
        for (int64_t p0 = 0; p0 < P; p0 += kMaxGridY) {
            const unsigned gy = (unsigned)std::min<int64_t>(kMaxGridY, P - p0);
            bbox_kernel<<<dim3(bx, gy), 256, 0, st>>>(d_xyz, stride, dv + p0);
            ctx->launches += 1;
        }
        if (b->ev_bbox_record) PCP_CUDA(cudaEventRecord(b->ev_bbox_record, st));   // query keys only need the boxes, not the sorted cloud
        const bool one_kernel_sort = !wide && small_sort_applies(ctx, b->total_n, 30 - mbits);
        if (one_kernel_sort)
            PCP_TRY(small_sort(ctx, d_xyz, b->total_n, stride, b->bbox.as<float>(), 30 - mbits, b->keys_a.as<unsigned>(),
                               b->vals_a.as<unsigned>(), b->vals_b.as<unsigned>(), ctx->ss_state[0], st));
        for (int64_t p0 = 0; p0 < P && !one_kernel_sort; p0 += kMaxGridY) {
            const unsigned gy = (unsigned)std::min<int64_t>(kMaxGridY, P - p0);
            if (wide64)
                key_kernel<unsigned long long><<<dim3(grid_for(b->max_n, 256), gy), 256, 0, st>>>(
                    d_xyz, stride, dv + p0, p0, 30 - mbits, mbits, b->keys_a.as<unsigned long long>(), b->vals_a.as<unsigned>());
            else
                key_kernel<unsigned><<<dim3(grid_for(b->max_n, 256), gy), 256, 0, st>>>(
                    d_xyz, stride, dv + p0, p0, wide ? 30 - mbits : 0, wide ? mbits : 0, b->keys_a.as<unsigned>(),
                    b->vals_a.as<unsigned>());
            ctx->launches += 1;
        }
        if (one_kernel_sort) {
        } else if (wide64) {
            PCP_TRY(radix_sort_pairs<unsigned long long>(ctx, b->keys_a.as<unsigned long long>(),
                                                         b->keys_b.as<unsigned long long>(), b->vals_a.as<unsigned>(),
                                                         b->vals_b.as<unsigned>(), b->total_n, 0, mbits + seg_bits));
        } else if (wide) {
            PCP_TRY(radix_sort_pairs<unsigned>(ctx, b->keys_a.as<unsigned>(), b->keys_b.as<unsigned>(),
                                               b->vals_a.as<unsigned>(), b->vals_b.as<unsigned>(), b->total_n, 0,
                                               mbits + seg_bits));
        } else {
            PCP_TRY(radix_sort_pairs<unsigned>(ctx, b->keys_a.as<unsigned>(), b->keys_b.as<unsigned>(),
                                               b->vals_a.as<unsigned>(), b->vals_b.as<unsigned>(), b->total_n,
                                               30 - mbits, 30));
        }
    }
    for (int64_t p0 = 0; p0 < P; p0 += kMaxGridY) {
        const unsigned gy = (unsigned)std::min<int64_t>(kMaxGridY, P - p0);
        const bool fuse_upper = b->level_max_cnt[1] <= 1024;     // the last CTA of a cloud finishes its hierarchy
        refine_leaf_kernel<<<dim3((unsigned)((b->max_leaf + kFan - 1) / kFan), gy), kRefineT, 0, st>>>(
            d_xyz, stride, dv + p0, b->vals_b.as<unsigned>(), fuse_upper ? 1 : 0);
        ctx->launches += 1;
        // level 1 came out of refine_leaf_kernel; wide levels get a launch each, the small ones above share one
        for (int l = 2; l < b->max_levels && !fuse_upper; ++l) {
            if (b->level_max_cnt[l] <= kFusedLevelNodes) {
                upper_boxes_kernel<<<dim3(1, gy), 256, 0, st>>>(dv + p0, l);
                ctx->launches += 1;
                break;
            }
            node_box_kernel<<<dim3(grid_for(b->level_max_cnt[l] * 32, 256), gy), 256, 0, st>>>(dv + p0, l);
            ctx->launches += 1;
        }
    }
    PCP_CUDA(cudaGetLastError());
    return PCP_OK;
}

int index_build_device(pcp_index *ix, const float *d_xyz, int64_t n, int64_t stride) {
    if (n > 0x7fffffffLL) return fail(PCP_ERR_UNSUPPORTED, "pcp_index_build: more than 2^31-1 points");
    ix->n = n;
    ++ix->serial;
    ix->built_from = d_xyz;
    ix->built_stride = stride;
    if (!ix->b.ev_bbox) PCP_CUDA(cudaEventCreateWithFlags(&ix->b.ev_bbox, cudaEventDisableTiming));
    ix->b.ctx = ix->ctx;
    ix->b.ev_bbox_record = ix->ctx->capturing ? ix->ctx->ev_cap_bbox : ix->b.ev_bbox;
    if (ix->ctx->capturing) ix->ctx->cap_has_bbox = true;
    const int64_t offs[2] = {0, n};
    return batch_build_device(&ix->b, d_xyz, stride, 1, offs);
}

int sort_queries_on(pcp_index *ix, const float *d_q, int64_t nq, int64_t stride, uint32_t *d_perm, cudaStream_t st,
                    DevBuf *keys, DevBuf &cub_tmp) {
    pcp_ctx *ctx = ix->ctx;
    if (nq == 0) return PCP_OK;
    for (int i = 0; i < 3; ++i) PCP_TRY(keys[i].ensure(sizeof(unsigned) * (size_t)nq));
    unsigned *ka = keys[0].as<unsigned>(), *kb = keys[1].as<unsigned>(), *va = keys[2].as<unsigned>();
    if (small_sort_applies(ctx, nq, morton_skip_bits(nq))) {
        PCP_TRY(small_sort(ctx, d_q, nq, stride, make_view(ix).bbox, morton_skip_bits(nq), ka, va, d_perm,
                           ctx->ss_state[st == ctx->stream ? 0 : 1], st));
    } else {
        query_key_kernel<<<grid_for(nq, 256), 256, 0, st>>>(d_q, nq, stride, make_view(ix).bbox, ka, va);
        ctx->launches += 1;
        PCP_TRY(radix_sort_pairs<unsigned>(ctx, ka, kb, va, d_perm, nq, morton_skip_bits(nq), 30, st, &cub_tmp));
    }
    refine_queries_kernel<<<grid_for(nq, kRefine), kRefineT, 0, st>>>(d_q, nq, stride, d_perm);
    ctx->launches += 1;
    PCP_CUDA(cudaGetLastError());
    return PCP_OK;
}

int sort_queries(pcp_index *ix, const float *d_q, int64_t nq, int64_t stride, uint32_t *d_perm) {
    pcp_ctx *ctx = ix->ctx;
    return sort_queries_on(ix, d_q, nq, stride, d_perm, ctx->stream, ctx->scratch, ctx->cub_tmp);
}

}  // namespace pcp
