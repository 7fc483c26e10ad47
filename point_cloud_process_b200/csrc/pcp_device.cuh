// Device-side arithmetic shared by the search and ICP kernels.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace pcp {

constexpr unsigned kFullMask = 0xffffffffu;

// Squared float32 distance with the reference's exact operation order and NO fused multiply-add:
// np.linalg.norm(query - p) of kdtree.py:157-158 evaluates sqrt((dx*dx + dy*dy) + dz*dz) with
// dx = q - p and one rounding per operation.  The __f*_rn intrinsics are
// never contracted into FFMA by nvcc.
__device__ __forceinline__ float ref_d2(float qx, float qy, float qz, float px, float py, float pz) {
    float dx = __fsub_rn(qx, px);
    float dy = __fsub_rn(qy, py);
    float dz = __fsub_rn(qz, pz);
    return __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
}

// Lower bound, in the same arithmetic, of ref_d2(q, p) over every p inside the box [lo, hi]:
// per axis the clamped difference is <= |fl(q - p)| because rounding is monotone, and squaring and
// the two additions are monotone too, so lb <= ref_d2(q, p) holds EXACTLY in float arithmetic --
// pruning with "lb > threshold" can never drop a candidate.
__device__ __forceinline__ float box_lb2(float qx, float qy, float qz, const float4 &lo, const float4 &hi) {
    float dx = fmaxf(fmaxf(__fsub_rn(lo.x, qx), __fsub_rn(qx, hi.x)), 0.f);
    float dy = fmaxf(fmaxf(__fsub_rn(lo.y, qy), __fsub_rn(qy, hi.y)), 0.f);
    float dz = fmaxf(fmaxf(__fsub_rn(lo.z, qz), __fsub_rn(qz, hi.z)), 0.f);
    return __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
}

// Same bound between two boxes (group of queries [glo, ghi] vs node [lo, hi]).
__device__ __forceinline__ float boxbox_lb2(const float3 &glo, const float3 &ghi, const float4 &lo, const float4 &hi) {
    float dx = fmaxf(fmaxf(__fsub_rn(lo.x, ghi.x), __fsub_rn(glo.x, hi.x)), 0.f);
    float dy = fmaxf(fmaxf(__fsub_rn(lo.y, ghi.y), __fsub_rn(glo.y, hi.y)), 0.f);
    float dz = fmaxf(fmaxf(__fsub_rn(lo.z, ghi.z), __fsub_rn(glo.z, hi.z)), 0.f);
    return __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
}

// Conservative squared-distance gate for a Euclidean threshold s:  fl(sqrt(d2)) <= s  ==>  d2 <= gate(s).
// (d2 > s^2 (1 + 2^-22) pushes sqrt(d2) past the midpoint to the next float above s.)
__device__ __forceinline__ float d2_gate(float s) {
    return __fmul_ru(__fmul_ru(s, s), 1.00000024f);
}

// (distance, index) packed so that one unsigned 64-bit compare is the lexicographic order:
// distances are >= +0, whose IEEE bit patterns are monotone as unsigned integers.
__device__ __forceinline__ unsigned long long pack_key(float s, unsigned idx) {
    return (static_cast<unsigned long long>(__float_as_uint(s)) << 32) | idx;
}
__device__ __forceinline__ float key_dist(unsigned long long k) { return __uint_as_float(static_cast<unsigned>(k >> 32)); }
__device__ __forceinline__ unsigned key_index(unsigned long long k) { return static_cast<unsigned>(k); }

constexpr unsigned long long kEmptyKey = 0x7f800000ffffffffull;   // (+inf, 0xffffffff)

// 10 bits per axis -> 30-bit Morton code
__host__ __device__ __forceinline__ unsigned part1by2(unsigned v) {
    v &= 0x3ffu;
    v = (v | (v << 16)) & 0x030000ffu;
    v = (v | (v << 8)) & 0x0300f00fu;
    v = (v | (v << 4)) & 0x030c30c3u;
    v = (v | (v << 2)) & 0x09249249u;
    return v;
}

__device__ __forceinline__ unsigned morton30(float x, float y, float z, const float *bbox) {
    // bbox: lo.x lo.y lo.z, scale (cells per unit, isotropic)
    float s = bbox[3];
    float fx = (x - bbox[0]) * s, fy = (y - bbox[1]) * s, fz = (z - bbox[2]) * s;
    // clamp (queries may lie outside the indexed cloud's box; NaN maps to 0)
    unsigned ix = static_cast<unsigned>(fminf(fmaxf(fx, 0.f), 1023.f));
    unsigned iy = static_cast<unsigned>(fminf(fmaxf(fy, 0.f), 1023.f));
    unsigned iz = static_cast<unsigned>(fminf(fmaxf(fz, 0.f), 1023.f));
    return part1by2(ix) | (part1by2(iy) << 1) | (part1by2(iz) << 2);
}

// order-preserving float <-> uint mapping for atomicMin/atomicMax on floats of either sign
__device__ __forceinline__ unsigned float_to_ordered(float f) {
    unsigned u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ __forceinline__ float ordered_to_float(unsigned u) {
    unsigned v = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
#ifdef __CUDA_ARCH__
    return __uint_as_float(v);
#else
    float f;
    memcpy(&f, &v, 4);
    return f;
#endif
}

__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(kFullMask, v, o));
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(kFullMask, v, o));
    return v;
}

// min / max over the warp of NON-NEGATIVE floats (or +inf) through the integer redux unit.
__device__ __forceinline__ float warp_min_nonneg(float v) {
    return __uint_as_float(__reduce_min_sync(kFullMask, __float_as_uint(v)));
}
__device__ __forceinline__ float warp_max_nonneg(float v) {
    return __uint_as_float(__reduce_max_sync(kFullMask, __float_as_uint(v)));
}

}  // namespace pcp
