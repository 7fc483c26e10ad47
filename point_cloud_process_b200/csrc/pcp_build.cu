// Index build: Morton-sorted 32-point buckets + implicit 32-ary AABB hierarchy.
// Replaces kdtree_recursive_build (kdtree.py:59-97, np.argsort per node) and
// octree_recursive_build (octree.py:51-103, per-point Python loop computing the 3-bit child code).
//
// Kernels (all HBM-bound, one pass each):
//   bbox_kernel        float3[N] -> 6 ordered-uint min/max words         (np.amin/np.amax, octree.py:319-320)
//   morton_kernel      float3[N] -> (30-bit Morton key, original index)  (child-code loop, octree.py:75-85)
//   cub::DeviceRadixSort::SortPairs                                       (np.argsort, kdtree.py:46,76)
//   gather_leaf_kernel sorted index -> float4 bucket rows + bucket AABBs  (Node / Octant objects)
//   node_box_kernel    32 child AABBs -> parent AABB, once per upper level
#include <cub/device/device_radix_sort.cuh>

#include <cfloat>
#include <cmath>
#include <cstring>

#include "pcp_device.cuh"
#include "pcp_internal.cuh"

namespace pcp {

namespace {

__global__ void bbox_init_kernel(unsigned *acc) {
    int t = threadIdx.x;
    if (t < 3) acc[t] = 0xffffffffu;        // min slots
    else if (t < 6) acc[t] = 0u;            // max slots
}

__global__ void __launch_bounds__(256) bbox_kernel(const float *__restrict__ xyz, long long n, long long stride,
                                                   unsigned *__restrict__ acc) {
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float *p = xyz + i * stride;
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            float v = p[a];
            lo[a] = fminf(lo[a], v);
            hi[a] = fmaxf(hi[a], v);
        }
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        lo[a] = warp_min(lo[a]);
        hi[a] = warp_max(hi[a]);
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            atomicMin(&acc[a], float_to_ordered(lo[a]));
            atomicMax(&acc[3 + a], float_to_ordered(hi[a]));
        }
    }
}

// bbox words -> {lo.x, lo.y, lo.z, cells-per-unit} used by morton30()
__global__ void bbox_finish_kernel(const unsigned *acc, float *bbox) {
    float lo[3], hi[3];
    for (int a = 0; a < 3; ++a) {
        lo[a] = ordered_to_float(acc[a]);
        hi[a] = ordered_to_float(acc[3 + a]);
    }
    float ext = fmaxf(fmaxf(hi[0] - lo[0], hi[1] - lo[1]), hi[2] - lo[2]);
    float scale = (ext > 0.f && isfinite(ext)) ? 1024.f / (ext * 1.0001f) : 0.f;
    bbox[0] = lo[0]; bbox[1] = lo[1]; bbox[2] = lo[2]; bbox[3] = scale;
    bbox[4] = hi[0]; bbox[5] = hi[1]; bbox[6] = hi[2]; bbox[7] = ext;
}

__global__ void __launch_bounds__(256) morton_kernel(const float *__restrict__ xyz, long long n, long long stride,
                                                     const float *__restrict__ bbox, unsigned *__restrict__ keys,
                                                     unsigned *__restrict__ vals) {
    long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float *p = xyz + i * stride;
    keys[i] = morton30(p[0], p[1], p[2], bbox);
    vals[i] = static_cast<unsigned>(i);
}

// One warp per bucket: lane j takes sorted position leaf*32+j, gathers its point and the warp
// reduces the bucket AABB.  Pad lanes of the last bucket are written as NaN / index -1 so that every
// distance against them is NaN and fails every "<=" gate.
__global__ void __launch_bounds__(256) gather_leaf_kernel(const float *__restrict__ xyz, long long n, long long stride,
                                                          const unsigned *__restrict__ sorted_idx,
                                                          float4 *__restrict__ pts, float4 *__restrict__ box_lo,
                                                          float4 *__restrict__ box_hi, long long n_leaf) {
    long long leaf = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (leaf >= n_leaf) return;
    long long pos = leaf * kLeaf + lane;
    float x = nanf(""), y = x, z = x;
    unsigned idx = 0xffffffffu;
    if (pos < n) {
        idx = sorted_idx[pos];
        const float *p = xyz + (long long)idx * stride;
        x = p[0]; y = p[1]; z = p[2];
    }
    pts[pos] = make_float4(x, y, z, __uint_as_float(idx));
    // fminf/fmaxf return the non-NaN operand, so pad lanes do not disturb the box
    float lx = warp_min(x), ly = warp_min(y), lz = warp_min(z);
    float hx = warp_max(x), hy = warp_max(y), hz = warp_max(z);
    if (lane == 0) {
        box_lo[leaf] = make_float4(lx, ly, lz, 0.f);
        box_hi[leaf] = make_float4(hx, hy, hz, 0.f);
    }
}

__global__ void __launch_bounds__(256) node_box_kernel(float4 *__restrict__ box_lo, float4 *__restrict__ box_hi,
                                                       long long child_off, long long child_cnt,
                                                       long long node_off, long long node_cnt) {
    long long node = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (node >= node_cnt) return;
    long long c = node * kFan + lane;
    float4 lo = make_float4(FLT_MAX, FLT_MAX, FLT_MAX, 0.f), hi = make_float4(-FLT_MAX, -FLT_MAX, -FLT_MAX, 0.f);
    if (c < child_cnt) {
        lo = box_lo[child_off + c];
        hi = box_hi[child_off + c];
    }
    float lx = warp_min(lo.x), ly = warp_min(lo.y), lz = warp_min(lo.z);
    float hx = warp_max(hi.x), hy = warp_max(hi.y), hz = warp_max(hi.z);
    if (lane == 0) {
        box_lo[node_off + node] = make_float4(lx, ly, lz, 0.f);
        box_hi[node_off + node] = make_float4(hx, hy, hz, 0.f);
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

int radix_sort_pairs(pcp_ctx *ctx, const unsigned *k_in, unsigned *k_out, const unsigned *v_in, unsigned *v_out,
                     long long n) {
    size_t tmp_bytes = 0;
    PCP_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, k_in, k_out, v_in, v_out, n, 0, 30, ctx->stream));
    PCP_TRY(ctx->cub_tmp.ensure(tmp_bytes));
    PCP_CUDA(cub::DeviceRadixSort::SortPairs(ctx->cub_tmp.p, tmp_bytes, k_in, k_out, v_in, v_out, n, 0, 30, ctx->stream));
    ctx->launches += 5;   // onesweep: histogram + 4 digit passes (approximate, CUB internal)
    return PCP_OK;
}

inline unsigned grid_for(long long n, int block) { return static_cast<unsigned>((n + block - 1) / block); }

}  // namespace

IndexView make_view(const pcp_index *ix) {
    IndexView v;
    v.pts = ix->pts.as<float4>();
    v.box_lo = ix->box_lo.as<float4>();
    v.box_hi = ix->box_hi.as<float4>();
    v.bbox = ix->bbox.as<float>();
    v.n = ix->n;
    v.n_leaf = ix->n_leaf;
    v.n_levels = ix->n_levels;
    for (int l = 0; l < kMaxLevels; ++l) {
        v.level_off[l] = ix->level_off[l];
        v.level_cnt[l] = ix->level_cnt[l];
    }
    return v;
}

int index_build_device(pcp_index *ix, const float *d_xyz, int64_t n, int64_t stride) {
    pcp_ctx *ctx = ix->ctx;
    cudaStream_t st = ctx->stream;
    ix->n = n;
    ix->n_leaf = (n + kLeaf - 1) / kLeaf;
    ix->n_levels = 0;
    if (n == 0) return PCP_OK;
    if (n > 0x7fffffffLL) return fail(PCP_ERR_UNSUPPORTED, "pcp_index_build: more than 2^31-1 points");

    // hierarchy shape
    long long cnt = ix->n_leaf, off = 0;
    int levels = 0;
    while (true) {
        if (levels >= kMaxLevels) return fail(PCP_ERR_UNSUPPORTED, "pcp_index_build: hierarchy too deep");
        ix->level_off[levels] = off;
        ix->level_cnt[levels] = cnt;
        ++levels;
        off += cnt;
        if (cnt == 1) break;
        cnt = (cnt + kFan - 1) / kFan;
    }
    ix->n_levels = levels;
    long long total_nodes = off;

    PCP_TRY(ix->pts.ensure(sizeof(float4) * (size_t)ix->n_leaf * kLeaf));
    PCP_TRY(ix->box_lo.ensure(sizeof(float4) * (size_t)total_nodes));
    PCP_TRY(ix->box_hi.ensure(sizeof(float4) * (size_t)total_nodes));
    PCP_TRY(ix->bbox.ensure(sizeof(float) * 8 + sizeof(unsigned) * 8));
    PCP_TRY(ix->keys_a.ensure(sizeof(unsigned) * (size_t)n));
    PCP_TRY(ix->keys_b.ensure(sizeof(unsigned) * (size_t)n));
    PCP_TRY(ix->vals_a.ensure(sizeof(unsigned) * (size_t)n));
    PCP_TRY(ix->vals_b.ensure(sizeof(unsigned) * (size_t)n));

    float *bbox = ix->bbox.as<float>();
    unsigned *acc = reinterpret_cast<unsigned *>(bbox + 8);

    bbox_init_kernel<<<1, 32, 0, st>>>(acc);
    int bgrid = (int)std::min<long long>(grid_for(n, 256), (long long)ctx->sm_count * 8);
    bbox_kernel<<<bgrid, 256, 0, st>>>(d_xyz, n, stride, acc);
    bbox_finish_kernel<<<1, 1, 0, st>>>(acc, bbox);
    morton_kernel<<<grid_for(n, 256), 256, 0, st>>>(d_xyz, n, stride, bbox, ix->keys_a.as<unsigned>(),
                                                    ix->vals_a.as<unsigned>());
    ctx->launches += 4;
    PCP_TRY(radix_sort_pairs(ctx, ix->keys_a.as<unsigned>(), ix->keys_b.as<unsigned>(), ix->vals_a.as<unsigned>(),
                             ix->vals_b.as<unsigned>(), n));
    gather_leaf_kernel<<<grid_for(ix->n_leaf * 32, 256), 256, 0, st>>>(
        d_xyz, n, stride, ix->vals_b.as<unsigned>(), ix->pts.as<float4>(), ix->box_lo.as<float4>(),
        ix->box_hi.as<float4>(), ix->n_leaf);
    ctx->launches += 1;
    for (int l = 1; l < levels; ++l) {
        node_box_kernel<<<grid_for(ix->level_cnt[l] * 32, 256), 256, 0, st>>>(
            ix->box_lo.as<float4>(), ix->box_hi.as<float4>(), ix->level_off[l - 1], ix->level_cnt[l - 1],
            ix->level_off[l], ix->level_cnt[l]);
        ctx->launches += 1;
    }
    PCP_CUDA(cudaGetLastError());
    return PCP_OK;
}

int sort_queries(pcp_index *ix, const float *d_q, int64_t nq, int64_t stride, uint32_t *d_perm) {
    pcp_ctx *ctx = ix->ctx;
    if (nq == 0) return PCP_OK;
    PCP_TRY(ctx->scratch[0].ensure(sizeof(unsigned) * (size_t)nq));
    PCP_TRY(ctx->scratch[1].ensure(sizeof(unsigned) * (size_t)nq));
    PCP_TRY(ctx->scratch[2].ensure(sizeof(unsigned) * (size_t)nq));
    unsigned *ka = ctx->scratch[0].as<unsigned>(), *kb = ctx->scratch[1].as<unsigned>();
    unsigned *va = ctx->scratch[2].as<unsigned>();
    query_key_kernel<<<grid_for(nq, 256), 256, 0, ctx->stream>>>(d_q, nq, stride, ix->bbox.as<float>(), ka, va);
    ctx->launches += 1;
    PCP_TRY(radix_sort_pairs(ctx, ka, kb, va, d_perm, nq));
    PCP_CUDA(cudaGetLastError());
    return PCP_OK;
}

}  // namespace pcp
