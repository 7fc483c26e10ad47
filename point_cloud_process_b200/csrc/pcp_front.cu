// ICP front end on the device: voxel-grid down-sampling and per-point normals by
// k-NN PCA.  Both reuse what the search path already has (radix sort, exact kNN) and restate the arithmetic of
//   voxel_filter(point_cloud, leaf_size, type)                Pca_and_Voxel_filter/voxel_filter.py:10-66
//   PCA(data) on the k nearest neighbours of every point      Pca_and_Voxel_filter/pca_normal.py:10-36,85-90
// in float64 (the reference hands Open3D float64 points to numpy).
//
// Voxel filter, what the reference computes (and what is kept, quirks included):
//   D = (max - min) // leaf_size per axis          voxel_filter.py:22-24  (numpy floor_divide, NOT +1)
//   h = hx + hy*Dx + hz*Dx*Dy, hx = floor((x - min_x) / leaf_size)      :29-33  (cells of the last column collide
//                                                   with the first column of the next row: grouping is by h)
//   stable sort by h                                :36   -> points of a voxel keep their input order
//   one output point per run of equal h, in ascending h; the LAST run is never emitted            :40-50
//   centroid = np.mean(np.hstack(group), axis=1)    :48   -> numpy's pairwise summation order, restated below
// HBM-bound: bounding box (12 B/pt read), keys (12 B read + 8-12 B written), radix sort (2-3 passes of 8 B/pt each
// way on 32-bit keys), gather (4 + 12 B read, 16 B written), centroids (16 B read).
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include <cmath>
#include <cstring>
#include <vector>

#include "pcp_internal.cuh"
#include "pcp_linalg.cuh"

namespace pcp {
namespace {

// ---------------------------------------------------------------------------------------------- bounding box

__device__ __forceinline__ unsigned f32_ordered(float f) {
    const unsigned u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
inline float f32_from_ordered(unsigned u) {
    u = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
    float f;
    memcpy(&f, &u, 4);
    return f;
}

// acc[0..2] = min x,y,z   acc[3..5] = max x,y,z   (order-preserving unsigned images of the floats).
// Four independent rows in flight per thread, one atomic set per CTA.
__global__ void __launch_bounds__(256) voxel_bbox_kernel(const float *__restrict__ xyz, long long n, long long stride,
                                                         unsigned *__restrict__ acc) {
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    const long long step = (long long)gridDim.x * blockDim.x;
    long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    for (; i + 3 * step < n; i += 4 * step) {
        float v[4][3];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const float *p = xyz + (i + u * step) * stride;
            v[u][0] = p[0]; v[u][1] = p[1]; v[u][2] = p[2];
        }
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                lo[a] = fminf(lo[a], v[u][a]);
                hi[a] = fmaxf(hi[a], v[u][a]);
            }
    }
    for (; i < n; i += step) {
        const float *p = xyz + i * stride;
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            lo[a] = fminf(lo[a], p[a]);
            hi[a] = fmaxf(hi[a], p[a]);
        }
    }
    __shared__ float s_lo[8][3], s_hi[8][3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
#pragma unroll
        for (int m = 16; m > 0; m >>= 1) {
            lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], m));
            hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], m));
        }
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            s_lo[threadIdx.x >> 5][a] = lo[a];
            s_hi[threadIdx.x >> 5][a] = hi[a];
        }
    }
    __syncthreads();
    if (threadIdx.x < 3) {
        const int a = threadIdx.x;
        float l = s_lo[0][a], h = s_hi[0][a];
        for (int w = 1; w < 8; ++w) {
            l = fminf(l, s_lo[w][a]);
            h = fmaxf(h, s_hi[w][a]);
        }
        atomicMin(acc + a, f32_ordered(l));
        atomicMax(acc + 3 + a, f32_ordered(h));
    }
}

__global__ void voxel_bbox_init_kernel(unsigned *acc) {
    if (threadIdx.x < 3) acc[threadIdx.x] = 0xffffffffu;
    else if (threadIdx.x < 6) acc[threadIdx.x] = 0u;
}

// ---------------------------------------------------------------------------------------------- keys

struct VoxelGrid {
    double min[3];
    double leaf;
    double Dx, Dy;            // h = hx + hy*Dx + (hz*Dx)*Dy
};

__device__ __forceinline__ double voxel_hash(const float *p, const VoxelGrid &g) {
    // voxel_filter.py:29-32 in float64; the products are integer-valued, __dmul_rn / __dadd_rn pin the
    // evaluation order ((hx + hy*Dx) + (hz*Dx)*Dy) against contraction
    const double hx = floor(((double)p[0] - g.min[0]) / g.leaf);
    const double hy = floor(((double)p[1] - g.min[1]) / g.leaf);
    const double hz = floor(((double)p[2] - g.min[2]) / g.leaf);
    return __dadd_rn(__dadd_rn(hx, __dmul_rn(hy, g.Dx)), __dmul_rn(__dmul_rn(hz, g.Dx), g.Dy));
}

template <class KeyT>
__global__ void __launch_bounds__(256) voxel_key_kernel(const float *__restrict__ xyz, long long n, long long stride,
                                                        VoxelGrid g, KeyT *__restrict__ keys,
                                                        unsigned *__restrict__ vals) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double h = voxel_hash(xyz + i * stride, g);
    if (sizeof(KeyT) == 4) keys[i] = (KeyT)(unsigned)h;                       // h < 2^32 checked on the host
    else keys[i] = (KeyT)(unsigned long long)__double_as_longlong(h);         // h >= 0: IEEE bits order like the values
    vals[i] = (unsigned)i;
}

// ---------------------------------------------------------------------------------------------- runs

template <class KeyT>
__global__ void __launch_bounds__(256) voxel_heads_kernel(const KeyT *__restrict__ keys,
                                                          const unsigned *__restrict__ vals,
                                                          const float *__restrict__ xyz, long long n, long long stride,
                                                          int *__restrict__ flags, float4 *__restrict__ sorted_pts) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= n) return;
    flags[i] = (i == 0 || keys[i] != keys[i - 1]) ? 1 : 0;
    const float *p = xyz + (long long)vals[i] * stride;
    sorted_pts[i] = make_float4(p[0], p[1], p[2], 0.f);
}

// starts[g] = first sorted position of run g; starts[G] = n; meta[0] = G
__global__ void __launch_bounds__(256) voxel_starts_kernel(const int *__restrict__ incl, long long n,
                                                           long long *__restrict__ starts,
                                                           long long *__restrict__ meta) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (i == 0 || incl[i] != incl[i - 1]) starts[incl[i] - 1] = i;   // the prefix sum steps exactly at the run heads
    if (i == n - 1) {
        starts[incl[i]] = n;
        meta[0] = incl[i];
    }
}

// ---------------------------------------------------------------------------------------------- centroids

struct D3 {
    double x, y, z;
};
__device__ __forceinline__ D3 d3_add(D3 a, D3 b) { return {__dadd_rn(a.x, b.x), __dadd_rn(a.y, b.y), __dadd_rn(a.z, b.z)}; }
__device__ __forceinline__ D3 d3_of(float4 p) { return {(double)p.x, (double)p.y, (double)p.z}; }

// numpy's DOUBLE_pairwise_sum on a contiguous run of at most 128 values (numpy/_core/src/umath/loops_utils.h.src):
// n < 8: left to right from 0; else 8 interleaved accumulators, combined ((r0+r1)+(r2+r3))+((r4+r5)+(r6+r7)),
// then the n % 8 tail left to right
__device__ D3 pairwise_leaf(const float4 *__restrict__ a, long long n) {
    if (n < 8) {
        D3 r = {0.0, 0.0, 0.0};
        for (long long i = 0; i < n; ++i) r = d3_add(r, d3_of(a[i]));
        return r;
    }
    D3 r[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) r[j] = d3_of(a[j]);
    long long i = 8;
    for (; i < n - (n % 8); i += 8) {
#pragma unroll
        for (int j = 0; j < 8; ++j) r[j] = d3_add(r[j], d3_of(a[i + j]));
    }
    D3 res = d3_add(d3_add(d3_add(r[0], r[1]), d3_add(r[2], r[3])), d3_add(d3_add(r[4], r[5]), d3_add(r[6], r[7])));
    for (; i < n; ++i) res = d3_add(res, d3_of(a[i]));
    return res;
}

// the recursive split above 128 values (n2 = n/2 rounded down to a multiple of 8), as an explicit stack
__device__ D3 pairwise_sum(const float4 *__restrict__ a, long long n) {
    if (n <= 128) return pairwise_leaf(a, n);
    struct Frame {
        long long start, n;
        int stage;
        D3 left;
    };
    Frame st[48];
    int sp = 0;
    st[sp++] = {0, n, 0, {0, 0, 0}};
    D3 ret = {0, 0, 0};
    while (sp > 0) {
        Frame &f = st[sp - 1];
        if (f.stage == 0) {
            if (f.n <= 128) {
                ret = pairwise_leaf(a + f.start, f.n);
                --sp;
                continue;
            }
            long long n2 = f.n / 2;
            n2 -= n2 % 8;
            f.stage = 1;
            st[sp++] = {f.start, n2, 0, {0, 0, 0}};
        } else if (f.stage == 1) {
            long long n2 = f.n / 2;
            n2 -= n2 % 8;
            f.left = ret;
            f.stage = 2;
            st[sp++] = {f.start + n2, f.n - n2, 0, {0, 0, 0}};
        } else {
            ret = d3_add(f.left, ret);
            --sp;
        }
    }
    return ret;
}

__device__ __forceinline__ unsigned long long splitmix64(unsigned long long x) {
    x += 0x9e3779b97f4a7c15ull;
    x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ull;
    x = (x ^ (x >> 27)) * 0x94d049bb133111ebull;
    return x ^ (x >> 31);
}

// one thread per run; the last run is skipped (voxel_filter.py:40-50 never appends it)
__global__ void __launch_bounds__(128) voxel_reduce_kernel(const float4 *__restrict__ sorted_pts,
                                                           const long long *__restrict__ starts,
                                                           const long long *__restrict__ meta, long long n, int mode,
                                                           unsigned long long seed, double *__restrict__ out) {
    const long long g = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (g >= n) return;
    const long long G = meta[0];
    if (g >= G - 1) return;
    const long long s = starts[g], cnt = starts[g + 1] - s;
    D3 r;
    if (mode == PCP_VOXEL_CENTROID) {
        r = d3_add({0.0, 0.0, 0.0}, pairwise_sum(sorted_pts + s, cnt));   // add.reduce starts from the identity 0
        const double c = (double)cnt;
        r = {r.x / c, r.y / c, r.z / c};                                  // np.mean: true_divide(sum, count)
    } else {
        r = d3_of(sorted_pts[s + (long long)(splitmix64(seed ^ (unsigned long long)g) % (unsigned long long)cnt)]);
    }
    out[g * 3 + 0] = r.x;
    out[g * 3 + 1] = r.y;
    out[g * 3 + 2] = r.z;
}

// numpy's floor_divide for float64 (npy_divmod, numpy/_core/src/npymath/npy_math_internal.h.src)
double npy_floor_divide(double a, double b) {
    if (b == 0.0) return a / b;
    double mod = fmod(a, b);
    double div = (a - mod) / b;
    if (mod != 0.0 && ((b < 0) != (mod < 0))) div -= 1.0;
    if (div != 0.0) {
        double fl = floor(div);
        if (div - fl > 0.5) fl += 1.0;
        return fl;
    }
    return copysign(0.0, a / b);
}

template <class KeyT>
int voxel_sort_and_reduce(pcp_ctx *ctx, const float *d_xyz, int64_t n, int64_t stride, const VoxelGrid &grid,
                          int end_bit, int mode, unsigned long long seed) {
    cudaStream_t st = ctx->stream;
    const unsigned blocks = (unsigned)((n + 255) / 256);
    PCP_TRY(ctx->scratch[0].ensure(sizeof(KeyT) * (size_t)n));
    PCP_TRY(ctx->scratch[1].ensure(sizeof(KeyT) * (size_t)n));
    PCP_TRY(ctx->scratch[2].ensure(sizeof(unsigned) * (size_t)n));
    PCP_TRY(ctx->scratch[3].ensure(sizeof(unsigned) * (size_t)n));
    KeyT *k_a = ctx->scratch[0].as<KeyT>(), *k_b = ctx->scratch[1].as<KeyT>();
    unsigned *v_a = ctx->scratch[2].as<unsigned>(), *v_b = ctx->scratch[3].as<unsigned>();
    voxel_key_kernel<KeyT><<<blocks, 256, 0, st>>>(d_xyz, n, stride, grid, k_a, v_a);
    size_t tmp_bytes = 0, scan_bytes = 0;
    PCP_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, k_a, k_b, v_a, v_b, n, 0, end_bit, st));
    PCP_CUDA(cub::DeviceScan::InclusiveSum(nullptr, scan_bytes, (int *)nullptr, (int *)nullptr, n, st));
    PCP_TRY(ctx->cub_tmp.ensure(tmp_bytes > scan_bytes ? tmp_bytes : scan_bytes));
    PCP_CUDA(cub::DeviceRadixSort::SortPairs(ctx->cub_tmp.p, tmp_bytes, k_a, k_b, v_a, v_b, n, 0, end_bit, st));
    // runs of equal keys; the key / value input buffers are free again and hold flags / their prefix sum
    int *flags = reinterpret_cast<int *>(v_a);
    PCP_TRY(ctx->scratch[4].ensure(sizeof(float4) * (size_t)n));
    PCP_TRY(ctx->scratch[5].ensure(sizeof(long long) * (size_t)(n + 2) + sizeof(int) * (size_t)n));
    float4 *sorted_pts = ctx->scratch[4].as<float4>();
    long long *starts = ctx->scratch[5].as<long long>();
    int *incl = reinterpret_cast<int *>(starts + n + 2);
    long long *meta = ctx->voxel_meta.as<long long>();
    voxel_heads_kernel<KeyT><<<blocks, 256, 0, st>>>(k_b, v_b, d_xyz, n, stride, flags, sorted_pts);
    PCP_CUDA(cub::DeviceScan::InclusiveSum(ctx->cub_tmp.p, scan_bytes, flags, incl, n, st));
    voxel_starts_kernel<<<blocks, 256, 0, st>>>(incl, n, starts, meta);
    PCP_TRY(ctx->voxel_out.ensure(sizeof(double) * 3 * (size_t)n));
    voxel_reduce_kernel<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(sorted_pts, starts, meta, n, mode, seed,
                                                                    ctx->voxel_out.as<double>());
    PCP_CUDA(cudaGetLastError());
    ctx->launches += 4 + 2 + (end_bit + 7) / 8 + 2;
    return PCP_OK;
}

}  // namespace

int voxel_filter_device(pcp_ctx *ctx, const float *d_xyz, int64_t n, int64_t stride, double leaf, int mode,
                        unsigned long long seed, int64_t *n_out) {
    cudaStream_t st = ctx->stream;
    ctx->voxel_n_out = -1;
    PCP_TRY(ctx->voxel_meta.ensure(64));
    unsigned *acc = ctx->voxel_meta.as<unsigned>() + 4;   // meta[0] (8 bytes) = number of runs; acc behind it
    voxel_bbox_init_kernel<<<1, 32, 0, st>>>(acc);
    int blocks = (int)((n + 255) / 256);
    if (blocks > ctx->sm_count * 16) blocks = ctx->sm_count * 16;
    voxel_bbox_kernel<<<blocks, 256, 0, st>>>(d_xyz, n, stride, acc);
    unsigned h_acc[6];
    PCP_CUDA(cudaMemcpyAsync(h_acc, acc, sizeof(h_acc), cudaMemcpyDeviceToHost, st));
    PCP_CUDA(cudaStreamSynchronize(st));
    ctx->launches += 2;
    VoxelGrid grid;
    double D[3], hmax[3];
    for (int a = 0; a < 3; ++a) {
        const double lo = (double)f32_from_ordered(h_acc[a]), hi = (double)f32_from_ordered(h_acc[3 + a]);
        if (!(std::isfinite(lo) && std::isfinite(hi)))
            return fail(PCP_ERR_INVALID, "pcp_voxel_filter: the cloud holds NaN / infinite coordinates");
        grid.min[a] = lo;
        D[a] = npy_floor_divide(hi - lo, leaf);          // voxel_filter.py:22-24
        hmax[a] = floor((hi - lo) / leaf);
    }
    grid.leaf = leaf;
    grid.Dx = D[0];
    grid.Dy = D[1];
    // largest hash any point can take; beyond 2^53 the float64 sums round exactly as numpy's do (same IEEE operations in
    // the same order), so only the key width depends on it
    const double bound = hmax[0] + hmax[1] * D[0] + hmax[2] * D[0] * D[1];
    if (!std::isfinite(bound))
        return fail(PCP_ERR_UNSUPPORTED, "pcp_voxel_filter: leaf_size too small for this extent (voxel hash overflows float64)");
    if (bound < 4294967296.0) {
        int bits = 1;
        while (bits < 32 && (double)(1ull << bits) <= bound) ++bits;
        PCP_TRY(voxel_sort_and_reduce<unsigned>(ctx, d_xyz, n, stride, grid, bits, mode, seed));
    } else {
        PCP_TRY(voxel_sort_and_reduce<unsigned long long>(ctx, d_xyz, n, stride, grid, 64, mode, seed));
    }
    long long G = 0;
    PCP_CUDA(cudaMemcpyAsync(&G, ctx->voxel_meta.p, sizeof(G), cudaMemcpyDeviceToHost, st));
    PCP_CUDA(cudaStreamSynchronize(st));
    ctx->voxel_n_out = G - 1;
    *n_out = G - 1;
    return PCP_OK;
}

// ================================================================================================ normals

namespace {

// One thread per point: PCA(pc_array[idx, :]) of pca_normal.py:10-36 on the point's k nearest neighbours
// (self included, pca_normal.py:88) -- mean (np.mean: left-to-right below 8 values, pairwise above: k <= 32 uses
// the 8-accumulator leaf), centring, np.cov (centres again by its own average, divides by k-1), symmetric
// eigen-decomposition, eigenvalues descending (:30-33); the normal is the last column (:89).
__global__ void __launch_bounds__(128) normals_kernel(const float *__restrict__ xyz, long long stride, long long n,
                                                      const int32_t *__restrict__ nn_idx,
                                                      const int32_t *__restrict__ nn_cnt, int k,
                                                      double *__restrict__ normals, double *__restrict__ evals) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int m = nn_cnt ? min(nn_cnt[i], k) : k;
    const int32_t *row = nn_idx + i * k;
    double X[PCP_MAX_K][3];
    for (int j = 0; j < m; ++j) {
        const float *p = xyz + (long long)row[j] * stride;
        X[j][0] = p[0];
        X[j][1] = p[1];
        X[j][2] = p[2];
    }
    // two centring passes: pca_normal.py:24-25, then np.cov's own `X -= avg[:, None]`
    for (int pass = 0; pass < 2; ++pass) {
        for (int a = 0; a < 3; ++a) {
            double s;
            if (m < 8) {
                s = 0.0;
                for (int j = 0; j < m; ++j) s = __dadd_rn(s, X[j][a]);
            } else {
                double r[8];
                for (int j = 0; j < 8; ++j) r[j] = X[j][a];
                int j = 8;
                for (; j < m - (m % 8); j += 8)
                    for (int u = 0; u < 8; ++u) r[u] = __dadd_rn(r[u], X[j + u][a]);
                s = __dadd_rn(__dadd_rn(__dadd_rn(r[0], r[1]), __dadd_rn(r[2], r[3])),
                              __dadd_rn(__dadd_rn(r[4], r[5]), __dadd_rn(r[6], r[7])));
                for (; j < m; ++j) s = __dadd_rn(s, X[j][a]);
            }
            const double mean = s / (double)m;
            for (int j = 0; j < m; ++j) X[j][a] = __dsub_rn(X[j][a], mean);
        }
    }
    double c[6] = {0, 0, 0, 0, 0, 0};
    for (int j = 0; j < m; ++j) {
        c[0] += X[j][0] * X[j][0];
        c[1] += X[j][0] * X[j][1];
        c[2] += X[j][0] * X[j][2];
        c[3] += X[j][1] * X[j][1];
        c[4] += X[j][1] * X[j][2];
        c[5] += X[j][2] * X[j][2];
    }
    const double f = 1.0 / (double)(m - 1);   // np.cov: c *= np.true_divide(1, fact); m == 1 gives inf/nan as numpy does
    for (int u = 0; u < 6; ++u) c[u] *= f;
    double w[3], V[9];
    sym_eig3(c[0], c[1], c[2], c[3], c[4], c[5], w, V);
    // descending eigenvalues (pca_normal.py:30-33)
    int o0 = 0, o1 = 1, o2 = 2, t;
    if (w[o0] < w[o1]) { t = o0; o0 = o1; o1 = t; }
    if (w[o1] < w[o2]) { t = o1; o1 = o2; o2 = t; }
    if (w[o0] < w[o1]) { t = o0; o0 = o1; o1 = t; }
    double nx = V[0 * 3 + o2], ny = V[1 * 3 + o2], nz = V[2 * 3 + o2];
    // eigh leaves the sign to LAPACK; here: the component of largest magnitude is positive
    const double ax = fabs(nx), ay = fabs(ny), az = fabs(nz);
    const double lead = ax >= ay && ax >= az ? nx : (ay >= az ? ny : nz);
    if (lead < 0.0) { nx = -nx; ny = -ny; nz = -nz; }
    normals[i * 3 + 0] = nx;
    normals[i * 3 + 1] = ny;
    normals[i * 3 + 2] = nz;
    if (evals) {
        evals[i * 3 + 0] = w[o0];
        evals[i * 3 + 1] = w[o1];
        evals[i * 3 + 2] = w[o2];
    }
}

}  // namespace

int normals_device(pcp_index *ix, int k, double *d_normals, double *d_evals) {
    pcp_ctx *ctx = ix->ctx;
    const int64_t n = ix->n;
    const size_t nres = (size_t)n * (size_t)k;
    PCP_TRY(ctx->normals_nn.ensure(nres * 8 + sizeof(int32_t) * (size_t)n));
    int32_t *d_idx = ctx->normals_nn.as<int32_t>();
    float *d_dist = reinterpret_cast<float *>(d_idx + nres);
    int32_t *d_cnt = reinterpret_cast<int32_t *>(d_dist + nres);
    PCP_TRY(knn_self_device(ix, k, d_idx, d_dist, d_cnt));
    normals_kernel<<<(unsigned)((n + 127) / 128), 128, 0, ctx->stream>>>(ix->built_from, ix->built_stride, n, d_idx,
                                                                         d_cnt, k, d_normals, d_evals);
    PCP_CUDA(cudaGetLastError());
    ctx->launches += 1;
    return PCP_OK;
}

// ================================================================================================ whole-cloud PCA
// PCA(data, correlation=False, sort=True) of pca_normal.py:10-36 on a whole cloud: mean, centred covariance with
// 1/(n-1), symmetric eigen-decomposition -- two deterministic reduction passes in float64 (per-thread partial sums
// over a grid-stride loop, warp shuffle, shared memory, fixed-order final sum in one CTA), 12 B/point each.

namespace {

constexpr int kPcaThreads = 256;

template <int NV>
__device__ __forceinline__ void block_sum_store(double (&v)[NV], double *__restrict__ partial) {
    __shared__ double sh[kPcaThreads / 32][NV];
#pragma unroll
    for (int a = 0; a < NV; ++a)
#pragma unroll
        for (int m = 16; m > 0; m >>= 1) v[a] += __shfl_xor_sync(0xffffffffu, v[a], m);
    if ((threadIdx.x & 31) == 0)
#pragma unroll
        for (int a = 0; a < NV; ++a) sh[threadIdx.x >> 5][a] = v[a];
    __syncthreads();
    if (threadIdx.x < NV) {
        double t = 0.0;
        for (int w = 0; w < kPcaThreads / 32; ++w) t += sh[w][threadIdx.x];
        partial[blockIdx.x * NV + threadIdx.x] = t;
    }
}

__global__ void __launch_bounds__(kPcaThreads) pca_sum_kernel(const float *__restrict__ xyz, long long n, long long stride,
                                                              double *__restrict__ partial) {
    double v[3] = {0.0, 0.0, 0.0};
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float *p = xyz + i * stride;
        v[0] += (double)p[0];
        v[1] += (double)p[1];
        v[2] += (double)p[2];
    }
    block_sum_store<3>(v, partial);
}

// fixed-order sum of the per-CTA partials by one CTA: thread t takes CTAs t, t+256, ...; shared-memory tree
template <int NV>
__device__ __forceinline__ void final_sum(const double *__restrict__ partial, int blocks, double (&tot)[NV]) {
    __shared__ double sh[kPcaThreads][NV];
    double v[NV];
#pragma unroll
    for (int a = 0; a < NV; ++a) v[a] = 0.0;
    for (int b = threadIdx.x; b < blocks; b += kPcaThreads)
#pragma unroll
        for (int a = 0; a < NV; ++a) v[a] += partial[b * NV + a];
#pragma unroll
    for (int a = 0; a < NV; ++a) sh[threadIdx.x][a] = v[a];
    __syncthreads();
    for (int h = kPcaThreads / 2; h > 0; h >>= 1) {
        if (threadIdx.x < h)
#pragma unroll
            for (int a = 0; a < NV; ++a) sh[threadIdx.x][a] += sh[threadIdx.x + h][a];
        __syncthreads();
    }
#pragma unroll
    for (int a = 0; a < NV; ++a) tot[a] = sh[0][a];
}

__global__ void __launch_bounds__(kPcaThreads) pca_mean_kernel(const double *__restrict__ partial, int blocks, long long n,
                                                               double *__restrict__ mean) {
    double tot[3];
    final_sum<3>(partial, blocks, tot);
    if (threadIdx.x < 3) mean[threadIdx.x] = tot[threadIdx.x] / (double)n;
}

__global__ void __launch_bounds__(kPcaThreads) pca_cov_kernel(const float *__restrict__ xyz, long long n, long long stride,
                                                              const double *__restrict__ mean, double *__restrict__ partial) {
    const double mx = mean[0], my = mean[1], mz = mean[2];
    double v[6] = {0, 0, 0, 0, 0, 0};
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float *p = xyz + i * stride;
        const double x = (double)p[0] - mx, y = (double)p[1] - my, z = (double)p[2] - mz;
        v[0] += x * x; v[1] += x * y; v[2] += x * z;
        v[3] += y * y; v[4] += y * z; v[5] += z * z;
    }
    block_sum_store<6>(v, partial);
}

// out[0..2] eigenvalues descending, out[3..11] eigenvectors as columns (row-major 3x3) in the same order, each
// with its component of largest magnitude positive
__global__ void __launch_bounds__(kPcaThreads) pca_solve_kernel(const double *__restrict__ partial, int blocks, long long n,
                                                                double *__restrict__ out) {
    double c[6];
    final_sum<6>(partial, blocks, c);
    if (threadIdx.x != 0) return;
    for (int a = 0; a < 6; ++a) c[a] *= 1.0 / (double)(n - 1);   // np.cov: c *= np.true_divide(1, n - 1)
    double w[3], V[9];
    sym_eig3(c[0], c[1], c[2], c[3], c[4], c[5], w, V);
    int o[3] = {0, 1, 2}, t;
    if (w[o[0]] < w[o[1]]) { t = o[0]; o[0] = o[1]; o[1] = t; }
    if (w[o[1]] < w[o[2]]) { t = o[1]; o[1] = o[2]; o[2] = t; }
    if (w[o[0]] < w[o[1]]) { t = o[0]; o[0] = o[1]; o[1] = t; }
    for (int j = 0; j < 3; ++j) {
        out[j] = w[o[j]];
        double x = V[0 * 3 + o[j]], y = V[1 * 3 + o[j]], z = V[2 * 3 + o[j]];
        const double ax = fabs(x), ay = fabs(y), az = fabs(z);
        const double lead = ax >= ay && ax >= az ? x : (ay >= az ? y : z);
        const double sgn = lead < 0.0 ? -1.0 : 1.0;
        out[3 + 0 * 3 + j] = sgn * x;
        out[3 + 1 * 3 + j] = sgn * y;
        out[3 + 2 * 3 + j] = sgn * z;
    }
}

}  // namespace

int pca_device(pcp_ctx *ctx, const float *d_xyz, int64_t n, int64_t stride, double *h_w, double *h_V) {
    cudaStream_t st = ctx->stream;
    int blocks = (int)((n + kPcaThreads - 1) / kPcaThreads);
    if (blocks > ctx->sm_count * 8) blocks = ctx->sm_count * 8;
    PCP_TRY(ctx->scratch[0].ensure(sizeof(double) * ((size_t)blocks * 6 + 16)));
    double *partial = ctx->scratch[0].as<double>();
    double *mean = partial + (size_t)blocks * 6, *out = mean + 3;
    pca_sum_kernel<<<blocks, kPcaThreads, 0, st>>>(d_xyz, n, stride, partial);
    pca_mean_kernel<<<1, kPcaThreads, 0, st>>>(partial, blocks, n, mean);
    pca_cov_kernel<<<blocks, kPcaThreads, 0, st>>>(d_xyz, n, stride, mean, partial);
    pca_solve_kernel<<<1, kPcaThreads, 0, st>>>(partial, blocks, n, out);
    PCP_CUDA(cudaGetLastError());
    ctx->launches += 4;
    double h[12];
    PCP_CUDA(cudaMemcpyAsync(h, out, sizeof(h), cudaMemcpyDeviceToHost, st));
    PCP_CUDA(cudaStreamSynchronize(st));
    for (int i = 0; i < 3; ++i) h_w[i] = h[i];
    for (int i = 0; i < 9; ++i) h_V[i] = h[3 + i];
    return PCP_OK;
}

// ================================================================================================ ISS keypoints
// The detection loop of Keypoint_detection_ISS/ISS.py:40-73 on top of the radius search:
//   c_j  = |ball(p_j, radius)|                                              ISS.py:49   (pcp_radius_count)
//   M_i  = sum_j (1/c_j) (p_j - p_i)(p_j - p_i)^T / sum_j (1/c_j), j in ball(p_i)   :43-51 (CSR of pcp_radius_fill,
//          one warp per point, float64)
//   candidates: l2/l1 < lambda21 and l3/l2 < lambda32 (eigenvalues descending; NaN ratios fail as in numpy)  :52-57
//   non-maximum suppression in descending l3 (ties: ascending index -- sorted(..., reverse=True) is stable) :59-73,
//   stopping once MORE than iss_count keypoints are kept (:72) -- as rounds of a parallel rule that reaches the
//   same fixed point as the sequential loop: a candidate is kept once every earlier-ranked neighbour is suppressed,
//   suppressed once an earlier-ranked neighbour is kept.

namespace {

__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int m = 16; m > 0; m >>= 1) v += __shfl_xor_sync(0xffffffffu, v, m);
    return v;
}

// x, y, z and the neighbour count (exact in float below 2^24) of every point in ONE aligned 16-byte record: the
// saliency pass then pays one scattered sector per pair instead of a 12-byte row plus two CSR offsets
__global__ void __launch_bounds__(256) iss_pack_kernel(const float *__restrict__ xyz, long long stride, long long n,
                                                       const long long *__restrict__ off, float4 *__restrict__ packed) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float *p = xyz + i * stride;
    packed[i] = make_float4(p[0], p[1], p[2], (float)(off[i + 1] - off[i]));
}

__global__ void __launch_bounds__(256) iss_saliency_kernel(const float4 *__restrict__ packed, long long n,
                                                           const long long *__restrict__ off,
                                                           const int32_t *__restrict__ nbr, double lambda21,
                                                           double lambda32, double *__restrict__ lambda3,
                                                           unsigned long long *__restrict__ keys,
                                                           unsigned *__restrict__ vals) {
    const long long i = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (i >= n) return;
    const float4 pi = packed[i];
    const double px = pi.x, py = pi.y, pz = pi.z;
    double a[7] = {0, 0, 0, 0, 0, 0, 0};
    for (long long e = off[i] + lane; e < off[i + 1]; e += 32) {
        const float4 pj = packed[nbr[e]];
        const double w = 1.0 / (double)pj.w;                           // ISS.py:49
        const double dx = (double)pj.x - px, dy = (double)pj.y - py, dz = (double)pj.z - pz;
        a[0] += w * dx * dx; a[1] += w * dx * dy; a[2] += w * dx * dz;
        a[3] += w * dy * dy; a[4] += w * dy * dz; a[5] += w * dz * dz;
        a[6] += w;
    }
#pragma unroll
    for (int u = 0; u < 7; ++u) a[u] = warp_sum_d(a[u]);
    if (lane != 0) return;
    double w[3], V[9];
    sym_eig3(a[0] / a[6], a[1] / a[6], a[2] / a[6], a[3] / a[6], a[4] / a[6], a[5] / a[6], w, V);
    const double l1 = fmax(w[0], fmax(w[1], w[2])), l3 = fmin(w[0], fmin(w[1], w[2]));
    const double l2 = fmax(fmin(w[0], w[1]), fmin(fmax(w[0], w[1]), w[2]));          // median of the three
    const bool cand = (l2 / l1 < lambda21) && (l3 / l2 < lambda32);    // NaN (isolated point: 0/0) compares false
    lambda3[i] = cand ? l3 : nan("");
    // sort key: candidates first in DESCENDING l3 (order-preserving image of the double, inverted), others last
    unsigned long long b = (unsigned long long)__double_as_longlong(l3);
    b = (b >> 63) ? ~b : (b | 0x8000000000000000ull);
    keys[i] = cand ? ~b : 0xffffffffffffffffull;
    vals[i] = (unsigned)i;
}

// state: 0 = not a candidate, 1 = undecided, 2 = kept, 3 = suppressed.  rank[i] = position in the sorted order.
__global__ void __launch_bounds__(256) iss_rank_kernel(const unsigned long long *__restrict__ keys_sorted,
                                                       const unsigned *__restrict__ vals_sorted, long long n,
                                                       unsigned *__restrict__ rank, unsigned char *__restrict__ state,
                                                       unsigned *__restrict__ meta) {
    const long long s = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (s >= n) return;
    const bool cand = keys_sorted[s] != 0xffffffffffffffffull;
    const unsigned i = vals_sorted[s];
    rank[i] = (unsigned)s;
    state[i] = cand ? 1 : 0;
    if (cand && (s + 1 == n || keys_sorted[s + 1] == 0xffffffffffffffffull)) meta[0] = (unsigned)(s + 1);   // candidates
}

// one warp per sorted candidate position; meta[1] counts the undecided left after this round
__global__ void __launch_bounds__(256) iss_nms_round_kernel(const unsigned *__restrict__ vals_sorted, unsigned n_cand,
                                                            const long long *__restrict__ off,
                                                            const int32_t *__restrict__ nbr,
                                                            const unsigned *__restrict__ rank,
                                                            volatile unsigned char *__restrict__ state,
                                                            unsigned *__restrict__ meta) {
    const unsigned s = (unsigned)((blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5);
    const int lane = threadIdx.x & 31;
    if (s >= n_cand) return;
    const unsigned i = vals_sorted[s];
    if (state[i] != 1) return;
    bool kept_before = false, open_before = false;
    for (long long e = off[i] + lane; e < off[i + 1]; e += 32) {
        const unsigned j = (unsigned)nbr[e];
        if (j == i || rank[j] >= s) continue;
        const unsigned char st = state[j];
        kept_before |= st == 2;
        open_before |= st == 1;
    }
    kept_before = __any_sync(0xffffffffu, kept_before);
    open_before = __any_sync(0xffffffffu, open_before);
    if (lane != 0) return;
    if (kept_before) state[i] = 3;
    else if (!open_before) state[i] = 2;
    else atomicAdd(meta + 1, 1u);
}

}  // namespace

int iss_device(pcp_index *ix, double radius, double lambda21, double lambda32, double nms_radius, int iss_count,
               int32_t *h_keypoints, int32_t *h_n_keypoints, double *d_lambda3_out) {
    pcp_ctx *ctx = ix->ctx;
    cudaStream_t st = ctx->stream;
    const int64_t n = ix->n;
    const float *d_xyz = ix->built_from;
    const int64_t stride = ix->built_stride;
    DevBuf &b_off = ctx->iss_buf[0], &b_idx = ctx->iss_buf[1], &b_dist = ctx->iss_buf[2], &b_misc = ctx->iss_buf[3];
    auto csr = [&](double r, long long *M) -> int {   // neighbourhoods as scipy's float64 cKDTree decides them (ISS.py:40)
        PCP_TRY(b_off.ensure(sizeof(int64_t) * (size_t)(n + 1)));
        PCP_TRY(radius_count_device(ix, d_xyz, n, stride, (float)r, b_off.as<int64_t>(), true, r));
        PCP_CUDA(cudaMemcpyAsync(M, b_off.as<int64_t>() + n, sizeof(long long), cudaMemcpyDeviceToHost, st));
        PCP_CUDA(cudaStreamSynchronize(st));
        PCP_TRY(b_idx.ensure(sizeof(int32_t) * (size_t)*M));
        PCP_TRY(b_dist.ensure(sizeof(float) * (size_t)*M));
        return radius_fill_device(ix, d_xyz, n, stride, (float)r, b_off.as<int64_t>(), b_idx.as<int32_t>(), b_dist.as<float>(), 0, true, r);
    };
    long long M = 0;
    PCP_TRY(csr(radius, &M));
    // misc: packed (n float4) | lambda3 (n f64) | keys a/b (n u64 each) | vals a/b (n u32 each) | rank (n u32) | meta (4 u32) | state (n u8)
    if (n >= (1 << 24)) return fail(PCP_ERR_UNSUPPORTED, "pcp_iss_keypoints: more than 2^24 points");   // counts ride in a float
    PCP_TRY(b_misc.ensure((size_t)n * (16 + 8 + 16 + 8 + 4 + 1) + 64));
    float4 *packed = b_misc.as<float4>();
    double *lambda3 = reinterpret_cast<double *>(packed + n);
    unsigned long long *k_a = reinterpret_cast<unsigned long long *>(lambda3 + n), *k_b = k_a + n;
    unsigned *v_a = reinterpret_cast<unsigned *>(k_b + n), *v_b = v_a + n, *rank = v_b + n, *meta = rank + n;
    unsigned char *state = reinterpret_cast<unsigned char *>(meta + 4);
    const unsigned warp_blocks = (unsigned)((n * 32 + 255) / 256), blocks = (unsigned)((n + 255) / 256);
    iss_pack_kernel<<<blocks, 256, 0, st>>>(d_xyz, stride, n, b_off.as<long long>(), packed);
    iss_saliency_kernel<<<warp_blocks, 256, 0, st>>>(packed, n, b_off.as<long long>(), b_idx.as<int32_t>(), lambda21,
                                                    lambda32, lambda3, k_a, v_a);
    size_t tmp = 0;
    PCP_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp, k_a, k_b, v_a, v_b, n, 0, 64, st));
    PCP_TRY(ctx->cub_tmp.ensure(tmp));
    PCP_CUDA(cub::DeviceRadixSort::SortPairs(ctx->cub_tmp.p, tmp, k_a, k_b, v_a, v_b, n, 0, 64, st));
    PCP_CUDA(cudaMemsetAsync(meta, 0, 16, st));
    iss_rank_kernel<<<blocks, 256, 0, st>>>(k_b, v_b, n, rank, state, meta);
    ctx->launches += 3 + 10;
    if (d_lambda3_out)
        PCP_CUDA(cudaMemcpyAsync(d_lambda3_out, lambda3, sizeof(double) * (size_t)n, cudaMemcpyDeviceToDevice, st));
    if (nms_radius != radius) PCP_TRY(csr(nms_radius, &M));          // ISS.py:67 searches with non_max_radius
    unsigned h_meta[2] = {0, 0};
    PCP_CUDA(cudaMemcpyAsync(h_meta, meta, sizeof(h_meta), cudaMemcpyDeviceToHost, st));
    PCP_CUDA(cudaStreamSynchronize(st));
    const unsigned n_cand = h_meta[0];
    *h_n_keypoints = 0;
    if (n_cand == 0) return PCP_OK;
    const unsigned cand_blocks = (unsigned)(((long long)n_cand * 32 + 255) / 256);
    for (int batch = 1;; batch = batch < 8 ? batch * 2 : 8) {   // counter read back after 1, 2, 4, 8, 8, ... rounds
        for (int r = 0; r < batch; ++r) {
            PCP_CUDA(cudaMemsetAsync(meta + 1, 0, 4, st));
            iss_nms_round_kernel<<<cand_blocks, 256, 0, st>>>(v_b, n_cand, b_off.as<long long>(), b_idx.as<int32_t>(), rank, state, meta);
            ctx->launches += 1;
        }
        PCP_CUDA(cudaMemcpyAsync(h_meta, meta, sizeof(h_meta), cudaMemcpyDeviceToHost, st));
        PCP_CUDA(cudaStreamSynchronize(st));
        if (h_meta[1] == 0) break;
    }
    // kept candidates in rank order; the loop of ISS.py:61-73 stops once more than iss_count are kept
    std::vector<unsigned> order(n_cand);
    std::vector<unsigned char> h_state((size_t)n);
    PCP_CUDA(cudaMemcpyAsync(order.data(), v_b, sizeof(unsigned) * n_cand, cudaMemcpyDeviceToHost, st));
    PCP_CUDA(cudaMemcpyAsync(h_state.data(), state, (size_t)n, cudaMemcpyDeviceToHost, st));
    PCP_CUDA(cudaStreamSynchronize(st));
    int kept = 0;
    for (unsigned s = 0; s < n_cand && kept <= iss_count; ++s)
        if (h_state[order[s]] == 2) h_keypoints[kept++] = (int32_t)order[s];
    *h_n_keypoints = kept;
    return PCP_OK;
}

}  // namespace pcp
