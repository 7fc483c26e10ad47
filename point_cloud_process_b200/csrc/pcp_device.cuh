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

// ---- packed FP32 (sm_100a FADD2 / FMUL2): two lanes of IEEE float32 per 64-bit register pair, every element
// rounded once per operation exactly like the scalar __f*_rn intrinsics.  A packed multiply is never fed into
// a packed add here (ptxas would contract that pair into FFMA2): products are unpacked and summed with scalar
// __fadd_rn, so a pair of distances costs 3 FADD2 + 3 FMUL2 + 4 FADD issue slots instead of 16.
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk2(float lo, float hi) {
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void upk2(f32x2 v, float &lo, float &hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 sub2(f32x2 a, f32x2 b) {
    f32x2 r;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
    f32x2 r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ float max3f(float a, float b, float c) {
    float r;
    asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
    return r;
}
__device__ __forceinline__ float min3f(float a, float b, float c) {
    float r;
    asm("min.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
    return r;
}
// ref_d2 of one query against a PAIR of points (x0, x1), (y0, y1), (z0, z1): bit-equal to two scalar ref_d2
__device__ __forceinline__ void ref_d2_pair(f32x2 qx, f32x2 qy, f32x2 qz, f32x2 px, f32x2 py, f32x2 pz, float &d0, float &d1) {
    const f32x2 dx = sub2(qx, px), dy = sub2(qy, py), dz = sub2(qz, pz);
    float x0, x1, y0, y1, z0, z1;
    upk2(mul2(dx, dx), x0, x1);
    upk2(mul2(dy, dy), y0, y1);
    upk2(mul2(dz, dz), z0, z1);
    d0 = __fadd_rn(__fadd_rn(x0, y0), z0);
    d1 = __fadd_rn(__fadd_rn(x1, y1), z1);
}

// ---- shared memory through explicit 32-bit shared-space addresses.  The hot kernels keep ONE opaque base address
// in a register (smem_opaque): the compiler can neither re-derive it from %tid / kernel parameters inside the loops
// (which it otherwise does, a dozen instructions per loop trip) nor fall back to generic addressing.
__device__ __forceinline__ unsigned smem_opaque(const void *p) {
    unsigned a = (unsigned)__cvta_generic_to_shared(p), r;
    asm volatile("mov.u32 %0, %1;" : "=r"(r) : "r"(a));
    return r;
}
__device__ __forceinline__ float4 lds_f4(unsigned a) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ float2 lds_f2(unsigned a) {
    float2 v;
    asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ float lds_f1(unsigned a) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ unsigned lds_u1(unsigned a) {
    unsigned v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ void sts_f4(unsigned a, float4 v) {
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(a), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void sts_f2(unsigned a, float x, float y) {
    asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(a), "f"(x), "f"(y) : "memory");
}
__device__ __forceinline__ void sts_u1(unsigned a, unsigned v) {
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory");
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

// Radius predicate.  Default (lesson2 contract, result_set.py:80 on float32 distances): d2 <= lo with lo = hi = the
// largest float32 d2 whose rounded sqrt is <= r.  exact64 (the consumers whose reference searches with scipy's
// float64 cKDTree: ISS.py:40, dbscan.py:13): a float32 d2 at or below `lo` is certainly inside and one above `hi`
// certainly outside the float64 ball; in the band between them the pair is decided in float64 -- differences and
// squares in cKDTree's order (x, y, z accumulated left to right, unfused) against r * r -- on the float64 originals
// of the rows when they are attached (pcp_index_attach_f64), else on the float32 rows widened exactly.
struct RadiusTest {
    float lo, hi;
    double r2;
    const double *q64, *p64;     // float64 rows of the queries / the indexed cloud (3 doubles per row), or nullptr
};
__device__ __forceinline__ bool radius_hit(const RadiusTest &t, float d2, float qx, float qy, float qz, float px, float py,
                                           float pz, long long qrow, unsigned prow) {
    if (d2 <= t.lo) return true;                 // NaN pads never hit
    if (!(d2 <= t.hi)) return false;
    double ax = qx, ay = qy, az = qz, bx = px, by = py, bz = pz;
    if (t.q64) { ax = t.q64[qrow * 3]; ay = t.q64[qrow * 3 + 1]; az = t.q64[qrow * 3 + 2]; }
    if (t.p64) { bx = t.p64[(long long)prow * 3]; by = t.p64[(long long)prow * 3 + 1]; bz = t.p64[(long long)prow * 3 + 2]; }
    const double dx = __dsub_rn(ax, bx), dy = __dsub_rn(ay, by), dz = __dsub_rn(az, bz);
    const double d = __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
    return d <= t.r2;
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
