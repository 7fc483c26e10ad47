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
constexpr int kRefine = 1024;                 // points per block == threads per CTA
constexpr unsigned kPadKey = 0xffffffffu;

struct RefineSmem {
    float x[kRefine], y[kRefine], z[kRefine];
    unsigned key[2][kRefine];                  // double-buffered exchange: one barrier per cross-warp stage
    unsigned box[32][6];                       // per segment: ordered-uint min xyz, max xyz
};

// One compare-exchange stage of the direction-free bitonic network: slot t against slot t ^ MASK, the slot
// whose BIT is clear keeps the smaller key.  Masks below 32 exchange by shuffle, larger ones through shared memory.
template <unsigned MASK, unsigned BIT>
__device__ __forceinline__ void refine_stage(RefineSmem &sm, unsigned &key, unsigned t, unsigned &phase) {
    unsigned other;
    if (MASK >= 32) {
        sm.key[phase][t] = key;
        __syncthreads();
        other = sm.key[phase][t ^ MASK];
        phase ^= 1u;                           // the next exchange writes the other buffer: no second barrier
    } else {
        other = __shfl_xor_sync(kFullMask, key, MASK);
    }
    key = (t & BIT) == 0 ? min(key, other) : max(key, other);
}

// merge step K of the network (blocks of K become sorted): mirror stage, then strides K/4 ... 1
template <unsigned K, unsigned J>
struct RefineMerge {
    static __device__ __forceinline__ void run(RefineSmem &sm, unsigned &key, unsigned t, unsigned &phase) {
        refine_stage<J, J>(sm, key, t, phase);
        RefineMerge<K, J / 2>::run(sm, key, t, phase);
    }
};
template <unsigned K>
struct RefineMerge<K, 0> {
    static __device__ __forceinline__ void run(RefineSmem &, unsigned &, unsigned, unsigned &) {}
};
// ascending sort of every aligned segment of S slots, fully unrolled: K = 2, 4, ..., S
template <unsigned S, unsigned K = 2>
struct RefineSort {
    static __device__ __forceinline__ void run(RefineSmem &sm, unsigned &key, unsigned t, unsigned &phase) {
        refine_stage<K - 1, K / 2>(sm, key, t, phase);          // slot i of a block against slot K-1-i
        RefineMerge<K, K / 4>::run(sm, key, t, phase);
        RefineSort<S, K * 2>::run(sm, key, t, phase);
    }
};
template <unsigned S>
struct RefineSort<S, 2 * S> {
    static __device__ __forceinline__ void run(RefineSmem &, unsigned &, unsigned, unsigned &) {}
};

// One level of the kd refinement: every aligned segment of S slots is sorted along its widest axis, which splits
// it at the median into two segments of S/2.
template <unsigned S>
__device__ __forceinline__ void refine_level(RefineSmem &sm, unsigned &id, bool &real, unsigned &phase) {
    const unsigned t = threadIdx.x;
    const unsigned seg = t / S;
    // ---- widest axis of every segment
    if (t < 32 * 6) sm.box[t / 6][t % 6] = (t % 6) < 3 ? 0xffffffffu : 0u;
    __syncthreads();                           // also orders the sm.x/y/z stores before the first gather
    const unsigned ox = float_to_ordered(sm.x[id]), oy = float_to_ordered(sm.y[id]), oz = float_to_ordered(sm.z[id]);
    {
        // a warp never straddles segments (S >= 32)
        const unsigned mnx = __reduce_min_sync(kFullMask, real ? ox : 0xffffffffu), mxx = __reduce_max_sync(kFullMask, real ? ox : 0u);
        const unsigned mny = __reduce_min_sync(kFullMask, real ? oy : 0xffffffffu), mxy = __reduce_max_sync(kFullMask, real ? oy : 0u);
        const unsigned mnz = __reduce_min_sync(kFullMask, real ? oz : 0xffffffffu), mxz = __reduce_max_sync(kFullMask, real ? oz : 0u);
        if ((t & 31) == 0) {
            if (S > 32) {
                atomicMin(&sm.box[seg][0], mnx); atomicMin(&sm.box[seg][1], mny); atomicMin(&sm.box[seg][2], mnz);
                atomicMax(&sm.box[seg][3], mxx); atomicMax(&sm.box[seg][4], mxy); atomicMax(&sm.box[seg][5], mxz);
            } else {
                sm.box[seg][0] = mnx; sm.box[seg][1] = mny; sm.box[seg][2] = mnz;
                sm.box[seg][3] = mxx; sm.box[seg][4] = mxy; sm.box[seg][5] = mxz;
            }
        }
    }
    __syncthreads();
    const float wx = ordered_to_float(sm.box[seg][3]) - ordered_to_float(sm.box[seg][0]);
    const float wy = ordered_to_float(sm.box[seg][4]) - ordered_to_float(sm.box[seg][1]);
    const float wz = ordered_to_float(sm.box[seg][5]) - ordered_to_float(sm.box[seg][2]);
    // empty segments give NaN / negative widths: any axis will do
    const unsigned okey = (wx >= wy && wx >= wz) ? ox : (wy >= wz ? oy : oz);
    // 22 coordinate bits + 10 slot bits in one word: the split is a partition heuristic, the slot makes keys
    // unique and the order deterministic; pads carry the largest key
    unsigned key = ((real ? okey : kPadKey) & ~(kRefine - 1u)) | id;
    RefineSort<S>::run(sm, key, t, phase);
    id = key & (kRefine - 1u);
    real = (key | (kRefine - 1u)) != kPadKey;
}

// Thread t owns slot t.  In: its element's coordinates (valid = real point).  Out: the element id (0..1023,
// the slot that loaded it) that ends up in slot t; coordinates stay addressable by id in sm.x/y/z.
// HALVES: a sixth level splits the 32-slot segments (buckets) into halves of 16.
template <bool HALVES>
__device__ __forceinline__ unsigned refine_block(RefineSmem &sm, float x, float y, float z, bool valid) {
    const unsigned t = threadIdx.x;
    sm.x[t] = x; sm.y[t] = y; sm.z[t] = z;
    unsigned id = t, phase = 0;
    bool real = valid;
    refine_level<1024>(sm, id, real, phase);
    refine_level<512>(sm, id, real, phase);
    refine_level<256>(sm, id, real, phase);
    refine_level<128>(sm, id, real, phase);
    refine_level<64>(sm, id, real, phase);
    if (HALVES) refine_level<32>(sm, id, real, phase);
    return id;
}

// One CTA per level-1 node: gather the node's points through the Morton order, refine, write the float4
// bucket rows (pad lanes NaN / index -1, so that every distance against them is NaN and fails every "<="
// gate) and the bucket AABBs.
__global__ void __launch_bounds__(kRefine) refine_leaf_kernel(const float *__restrict__ xyz, long long stride,
                                                               const IndexView *__restrict__ views,
                                                               const unsigned *__restrict__ sorted_rows) {
    __shared__ RefineSmem sm;
    __shared__ unsigned s_idx[kRefine];
    __shared__ unsigned s_node[6];                               // AABB of the whole block = the level-1 node
    const IndexView &v = views[blockIdx.y];
    const long long node = blockIdx.x;
    if (node * kFan >= v.n_leaf) return;
    const unsigned t = threadIdx.x;
    if (t < 6) s_node[t] = t < 3 ? 0xffffffffu : 0u;             // (ordered by the barriers inside refine_block)
    const long long pos = node * kRefine + t;
    float x = nanf(""), y = x, z = x;
    unsigned idx = 0xffffffffu;
    const bool valid = pos < v.n;
    if (valid) {
        const unsigned row = sorted_rows[v.src_off + pos];       // sorted by (segment, morton): rows of this segment
        const float *p = xyz + (long long)row * stride;
        x = p[0]; y = p[1]; z = p[2];
        idx = row - static_cast<unsigned>(v.src_off);
    }
    s_idx[t] = idx;
    const unsigned id = refine_block<true>(sm, x, y, z, valid);
    const long long leaf = pos >> 5;
    if (leaf < v.n_leaf) {                                       // (whole warp) else: bucket beyond the cloud
    x = sm.x[id]; y = sm.y[id]; z = sm.z[id];
    idx = s_idx[id];
    const_cast<float4 *>(v.pts)[pos] = make_float4(x, y, z, __uint_as_float(idx));
    const bool real = idx != 0xffffffffu;
    {
        // pair layout (IndexView::pts2): the even lane of a pair writes (x0, x1, y0, y1), the odd one (z0, z1, w0, w1)
        const bool odd = (t & 1u) != 0u;
        const float px = real ? x : __uint_as_float(0x7f800000u), py = real ? y : 0.f, pz = real ? z : 0.f;
        const float pw = __uint_as_float(idx);
        const float a = __shfl_xor_sync(kFullMask, odd ? px : pz, 1);
        const float b = __shfl_xor_sync(kFullMask, odd ? py : pw, 1);
        const_cast<float4 *>(v.pts2)[pos] = odd ? make_float4(a, pz, b, pw) : make_float4(px, a, py, b);
    }
    const unsigned ox = float_to_ordered(x), oy = float_to_ordered(y), oz = float_to_ordered(z);
    const unsigned lx = __reduce_min_sync(kFullMask, real ? ox : 0xffffffffu), hx = __reduce_max_sync(kFullMask, real ? ox : 0u);
    const unsigned ly = __reduce_min_sync(kFullMask, real ? oy : 0xffffffffu), hy = __reduce_max_sync(kFullMask, real ? oy : 0u);
    const unsigned lz = __reduce_min_sync(kFullMask, real ? oz : 0xffffffffu), hz = __reduce_max_sync(kFullMask, real ? oz : 0u);
    if ((t & 31) == 0) {
        const_cast<float4 *>(v.box_lo)[leaf] = make_float4(ordered_to_float(lx), ordered_to_float(ly), ordered_to_float(lz), 0.f);
        const_cast<float4 *>(v.box_hi)[leaf] = make_float4(ordered_to_float(hx), ordered_to_float(hy), ordered_to_float(hz), 0.f);
        atomicMin(&s_node[0], lx); atomicMin(&s_node[1], ly); atomicMin(&s_node[2], lz);
        atomicMax(&s_node[3], hx); atomicMax(&s_node[4], hy); atomicMax(&s_node[5], hz);
    }
    // AABBs of the two 16-point halves (the sixth split): an empty half gets (+inf, -inf), infinitely far from everything
    const unsigned hm = (t & 16) ? 0xffff0000u : 0x0000ffffu;
    const unsigned alx = __reduce_min_sync(hm, real ? ox : 0xffffffffu), ahx = __reduce_max_sync(hm, real ? ox : 0u);
    const unsigned aly = __reduce_min_sync(hm, real ? oy : 0xffffffffu), ahy = __reduce_max_sync(hm, real ? oy : 0u);
    const unsigned alz = __reduce_min_sync(hm, real ? oz : 0xffffffffu), ahz = __reduce_max_sync(hm, real ? oz : 0u);
    if ((t & 15) == 0) {
        const bool any = alx != 0xffffffffu;
        const float inf = __uint_as_float(0x7f800000u);
        float4 *h = const_cast<float4 *>(v.half) + leaf * 4 + ((t & 16) ? 2 : 0);
        h[0] = any ? make_float4(ordered_to_float(alx), ordered_to_float(aly), ordered_to_float(alz), 0.f) : make_float4(inf, inf, inf, 0.f);
        h[1] = any ? make_float4(ordered_to_float(ahx), ordered_to_float(ahy), ordered_to_float(ahz), 0.f) : make_float4(-inf, -inf, -inf, 0.f);
    }
    }
    // the block is one level-1 node: its box is the union of the bucket boxes (same values node_box_kernel
    // would compute: min / max of the same floats)
    __syncthreads();
    if (t == 0 && v.n_levels > 1) {
        const long long at = v.level_off[1] + node;
        const_cast<float4 *>(v.box_lo)[at] = make_float4(ordered_to_float(s_node[0]), ordered_to_float(s_node[1]), ordered_to_float(s_node[2]), 0.f);
        const_cast<float4 *>(v.box_hi)[at] = make_float4(ordered_to_float(s_node[3]), ordered_to_float(s_node[4]), ordered_to_float(s_node[5]), 0.f);
    }
}

// Same refinement for a sorted QUERY set: blocks of 1024 Morton-consecutive queries are re-grouped into 32
// tight groups of 32 (one group = one warp of the tiled kNN kernel).  In place on the permutation.
__global__ void __launch_bounds__(kRefine) refine_queries_kernel(const float *__restrict__ q, long long nq, long long stride,
                                                                  unsigned *__restrict__ perm) {
    __shared__ RefineSmem sm;
    __shared__ unsigned s_idx[kRefine];
    const unsigned t = threadIdx.x;
    const long long pos = blockIdx.x * (long long)kRefine + t;
    float x = 0.f, y = 0.f, z = 0.f;
    unsigned row = 0xffffffffu;
    const bool valid = pos < nq;
    if (valid) {
        row = perm[pos];
        const float *p = q + (long long)row * stride;
        x = p[0]; y = p[1]; z = p[2];
    }
    s_idx[t] = row;
    const unsigned id = refine_block<false>(sm, x, y, z, valid);
    if (valid) perm[pos] = s_idx[id];          // real queries stay at the front of the block: slot t < nq - base is real
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
    PCP_CUDA(cudaMemcpyAsync(b->views.p, b->h_views.data(), sizeof(IndexView) * (size_t)P, cudaMemcpyHostToDevice, st));
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

    const unsigned bx = (unsigned)std::min<long long>(grid_for(b->max_n, 256),
                                                      P == 1 ? (long long)ctx->sm_count * 2 : 16);
    for (int64_t p0 = 0; p0 < P; p0 += kMaxGridY) {
        const unsigned gy = (unsigned)std::min<int64_t>(kMaxGridY, P - p0);
        bbox_kernel<<<dim3(bx, gy), 256, 0, st>>>(d_xyz, stride, dv + p0);
        ctx->launches += 1;
    }
    if (b->ev_bbox_record) PCP_CUDA(cudaEventRecord(b->ev_bbox_record, st));   // query keys only need the boxes, not the sorted cloud
    for (int64_t p0 = 0; p0 < P; p0 += kMaxGridY) {
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
    if (wide64) {
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
    for (int64_t p0 = 0; p0 < P; p0 += kMaxGridY) {
        const unsigned gy = (unsigned)std::min<int64_t>(kMaxGridY, P - p0);
        refine_leaf_kernel<<<dim3((unsigned)((b->max_leaf + kFan - 1) / kFan), gy), kRefine, 0, st>>>(
            d_xyz, stride, dv + p0, b->vals_b.as<unsigned>());
        ctx->launches += 1;
        // level 1 came out of refine_leaf_kernel; wide levels get a launch each, the small ones above share one
        for (int l = 2; l < b->max_levels; ++l) {
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
    query_key_kernel<<<grid_for(nq, 256), 256, 0, st>>>(d_q, nq, stride, make_view(ix).bbox, ka, va);
    ctx->launches += 1;
    PCP_TRY(radix_sort_pairs<unsigned>(ctx, ka, kb, va, d_perm, nq, morton_skip_bits(nq), 30, st, &cub_tmp));
    refine_queries_kernel<<<grid_for(nq, kRefine), kRefine, 0, st>>>(d_q, nq, stride, d_perm);
    ctx->launches += 1;
    PCP_CUDA(cudaGetLastError());
    return PCP_OK;
}

int sort_queries(pcp_index *ix, const float *d_q, int64_t nq, int64_t stride, uint32_t *d_perm) {
    pcp_ctx *ctx = ix->ctx;
    return sort_queries_on(ix, d_q, nq, stride, d_perm, ctx->stream, ctx->scratch, ctx->cub_tmp);
}

}  // namespace pcp
