// Brute-force exact kNN / radius count over the WHOLE indexed cloud: the on-device exhaustive cross-check
// (SURVEY section 7 step 2 / kernel K5).  Replaces the check the reference prints next to its tree
// searches -- np.argsort(np.linalg.norm(db - query, axis=1))[:k], kdtree.py:234-238, octree.py:352-356 --
// for full query sets: every (query, point) pair is evaluated, no spatial pruning, so it shares NO traversal
// logic with the tiled kernels it is used to check (tests: full-set parity at C3, per-rank shards of C5).
//
// One query per thread, 128 queries per CTA.  The reference points stream through shared memory in tiles of
// 1024 points (16 KB of the pair layout, IndexView::pts2), fetched by the TMA unit as ONE 1-D bulk copy per
// tile (cp.async.bulk.shared::cluster.global + mbarrier complete_tx, double buffered: the copy of tile t+1 is
// in flight while tile t is evaluated); every thread reads the tile as broadcast LDS.128 and evaluates two
// points per step in packed FP32 (3 FADD2 + 3 FMUL2 + 4 FADD, the reference's unfused operation order).  The
// top-k is a register list of packed (float32 sqrt bits, index) keys -- canonical (distance, index) order --
// guarded by a squared-distance gate, so after the first few tiles almost every pair costs the 5 arithmetic
// issue slots, its share of the two loads, and a compare.
#include <cfloat>
#include <cstdlib>
#include <mutex>
#include <string>

#include "pcp_device.cuh"
#include "pcp_internal.cuh"

namespace pcp {

namespace {

constexpr int kBruteThreads = 128;
constexpr int kTilePts = 1024;                       // points per tile
constexpr unsigned kTileBytes = kTilePts * 16;       // pair layout: 32 B per pair of points

__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@!p bra WAIT_%=;\n"
        "}\n" ::"r"(bar), "r"(parity)
        : "memory");
}
// 1-D bulk copy global -> shared by the TMA unit; completion is signalled on the mbarrier as transferred bytes
__device__ __forceinline__ void tma_load_1d(unsigned dst, const void *src, unsigned bytes, unsigned bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}

template <int K>
struct KeyList {
    unsigned long long key[K];
    // k < K: the first K - k slots hold a sentinel below every key, so that key[K-1] is always the k-th best
    // (a select chain over a run-time slot number would turn the list into a local-memory array)
    __device__ __forceinline__ void init(int k) {
#pragma unroll
        for (int i = 0; i < K; ++i) key[i] = i < K - k ? 0ull : kEmptyKey;
    }
    __device__ __forceinline__ void insert(unsigned long long nk) {
        bool g[K];
#pragma unroll
        for (int i = 0; i < K; ++i) g[i] = key[i] > nk;
#pragma unroll
        for (int i = K - 1; i > 0; --i) key[i] = g[i - 1] ? key[i - 1] : (g[i] ? nk : key[i]);
        key[0] = g[0] ? nk : key[0];
    }
};

// MODE 0: kNN (K = list width >= k).  MODE 1: radius count (d2 <= d2max), K unused.
template <int K, int MODE>
__global__ void __launch_bounds__(kBruteThreads)
brute_kernel(const float4 *__restrict__ pts2, long long n_slots, const float *__restrict__ queries, long long nq,
             long long stride, int k, float d2max, int *__restrict__ out_idx, float *__restrict__ out_dist,
             int *__restrict__ out_cnt, long long *__restrict__ out_count64) {
    extern __shared__ __align__(128) unsigned char s_raw[];
    const unsigned s_base = (unsigned)__cvta_generic_to_shared(s_raw);
    const unsigned bar0 = s_base + 2 * kTileBytes;            // two mbarriers behind the two tile buffers
    const int tid = threadIdx.x;
    const long long qid = blockIdx.x * (long long)kBruteThreads + tid;
    const bool active = qid < nq;
    float qx = 0.f, qy = 0.f, qz = 0.f;
    if (active) {
        const float *q = queries + qid * stride;
        qx = __ldg(q); qy = __ldg(q + 1); qz = __ldg(q + 2);
    }
    const f32x2 QX = pk2(qx, qx), QY = pk2(qy, qy), QZ = pk2(qz, qz);
    KeyList<K> top;
    top.init(k);
    // result_set.py:19,37: nothing beyond the initial worst_dist 1e10 is ever accepted; finite, so pad slots
    // (x = +inf -> d2 = +inf) never pass either
    float gate = MODE == 0 ? d2_gate(1e10f) : d2max;
    long long hits = 0;

    if (tid == 0) {
        mbar_init(bar0, 1);
        mbar_init(bar0 + 8, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const long long n_tiles = (n_slots + kTilePts - 1) / kTilePts;
    auto issue = [&](long long t) {
        const long long first = t * kTilePts;
        const unsigned bytes = (unsigned)((n_slots - first < kTilePts ? n_slots - first : kTilePts) * 16);
        const unsigned bar = bar0 + (unsigned)(t & 1) * 8u;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy reads of the buffer before the async write
        mbar_expect_tx(bar, bytes);
        tma_load_1d(s_base + (unsigned)(t & 1) * kTileBytes, pts2 + first, bytes, bar);
    };
    if (tid == 0 && n_tiles > 0) issue(0);
    for (long long t = 0; t < n_tiles; ++t) {
        if (tid == 0 && t + 1 < n_tiles) issue(t + 1);           // (its buffer was released by the barrier below)
        mbar_wait(bar0 + (unsigned)(t & 1) * 8u, (unsigned)((t >> 1) & 1));
        const long long first = t * kTilePts;
        const int n_ent = (int)(n_slots - first < kTilePts ? n_slots - first : kTilePts);   // float4 entries (2 per pair)
        const unsigned buf = s_base + (unsigned)(t & 1) * kTileBytes;
#pragma unroll 1
        for (int e0 = 0; e0 < n_ent; e0 += 8) {                  // n_slots is a multiple of 32: whole chunks
            float4 p[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) p[j] = lds_f4(buf + (unsigned)(e0 + j) * 16u);
            float d[8];
#pragma unroll
            for (int m = 0; m < 4; ++m)
                ref_d2_pair(QX, QY, QZ, pk2(p[2 * m].x, p[2 * m].y), pk2(p[2 * m].z, p[2 * m].w),
                            pk2(p[2 * m + 1].x, p[2 * m + 1].y), d[2 * m], d[2 * m + 1]);
            if (MODE == 1) {
#pragma unroll
                for (int j = 0; j < 8; ++j) hits += d[j] <= gate ? 1 : 0;
            } else {
                const float dm = min3f(min3f(d[0], d[1], d[2]), min3f(d[3], d[4], d[5]), fminf(d[6], d[7]));
                if (dm <= gate) {
#pragma unroll
                    for (int m = 0; m < 4; ++m) {
                        if (d[2 * m] <= gate) top.insert(pack_key(__fsqrt_rn(d[2 * m]), __float_as_uint(p[2 * m + 1].z)));
                        if (d[2 * m + 1] <= gate) top.insert(pack_key(__fsqrt_rn(d[2 * m + 1]), __float_as_uint(p[2 * m + 1].w)));
                    }
                    // candidates whose float32 sqrt can still tie the k-th distance stay admissible (canonical order)
                    const unsigned long long kth = top.key[K - 1];
                    if (key_index(kth) != 0xffffffffu) gate = fminf(gate, d2_gate(key_dist(kth)));
                }
            }
        }
        __syncthreads();                                         // everyone is done with this buffer
    }
    if (!active) return;
    if (MODE == 1) {
        out_count64[qid] = hits;
        return;
    }
    int cnt = 0;
#pragma unroll
    for (int i = 0; i < K; ++i) {
        if (i >= K - k) {
            const unsigned id = key_index(top.key[i]);
            const bool filled = id != 0xffffffffu;
            cnt += filled ? 1 : 0;
            out_idx[qid * k + (i - (K - k))] = filled ? (int)id : 0;
            out_dist[qid * k + (i - (K - k))] = filled ? key_dist(top.key[i]) : PCP_KNN_EMPTY_DIST;
        }
    }
    if (out_cnt) out_cnt[qid] = cnt;
}

}  // namespace

int knn_brute_device(pcp_index *ix, const float *d_q, int64_t nq, int64_t stride, int k, int32_t *d_idx, float *d_dist,
                     int32_t *d_count) {
    if (nq == 0) return PCP_OK;
    pcp_ctx *ctx = ix->ctx;
    const IndexView v = ix->n > 0 ? make_view(ix) : IndexView();
    const long long n_slots = ix->n > 0 ? v.n_leaf * kLeaf : 0;
    const unsigned grid = (unsigned)((nq + kBruteThreads - 1) / kBruteThreads);
    const size_t smem = 2 * kTileBytes + 16;
#define PCP_BRUTE(KK)                                                                                                  \
    do {                                                                                                               \
        static bool attr_dev[kMaxDevices] = {false};   /* function attributes are per device */                          \
        bool &attr = attr_dev[ctx->device & (kMaxDevices - 1)];                                                          \
        if (!attr) {                                                                                                   \
            PCP_CUDA(cudaFuncSetAttribute(brute_kernel<KK, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
            attr = true;                                                                                               \
        }                                                                                                              \
        brute_kernel<KK, 0><<<grid, kBruteThreads, smem, ctx->stream>>>(v.pts2, n_slots, d_q, nq, stride, k, 0.f,      \
                                                                        d_idx, d_dist, d_count, nullptr);              \
    } while (0)
    if (k <= 8) PCP_BRUTE(8);
    else if (k <= 16) PCP_BRUTE(16);
    else PCP_BRUTE(32);
#undef PCP_BRUTE
    ctx->launches += 1;
    PCP_CUDA(cudaGetLastError());
    return PCP_OK;
}

int radius_count_brute_device(pcp_index *ix, const float *d_q, int64_t nq, int64_t stride, float d2max, long long *d_counts) {
    if (nq == 0) return PCP_OK;
    pcp_ctx *ctx = ix->ctx;
    const IndexView v = ix->n > 0 ? make_view(ix) : IndexView();
    const long long n_slots = ix->n > 0 ? v.n_leaf * kLeaf : 0;
    const unsigned grid = (unsigned)((nq + kBruteThreads - 1) / kBruteThreads);
    const size_t smem = 2 * kTileBytes + 16;
    static bool attr_dev[kMaxDevices] = {false};
    bool &attr = attr_dev[ctx->device & (kMaxDevices - 1)];
    if (!attr) {
        PCP_CUDA(cudaFuncSetAttribute(brute_kernel<8, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr = true;
    }
    brute_kernel<8, 1><<<grid, kBruteThreads, smem, ctx->stream>>>(v.pts2, n_slots, d_q, nq, stride, 1, d2max, nullptr,
                                                                   nullptr, nullptr, d_counts);
    ctx->launches += 1;
    PCP_CUDA(cudaGetLastError());
    return PCP_OK;
}


// ---- brute-force kNN without any allocation (pcp_knn_bruteforce_noalloc, include/pcp.h) -------------------------
// For callers that cannot call cudaMalloc (stream capture, serving loops, harnesses that forbid it).  The one piece of
// scratch memory the brute-force kernel needs -- the pair-layout copy of the cloud that the TMA tiles are cut from --
// lives in a file-scope device array of fixed size instead of a context buffer: every GPU that loads the library holds
// PCP_NOALLOC_MAX_POINTS x 16 B = 4 MiB until the process exits.  No context, no index, no stream of its own.
// The array is SHARED by every call on a device and keeps the last call's contents: the pack kernel rewrites every
// slot the search kernel reads (pads included), nothing accumulates, so nothing has to start at zero; calls that
// overlap on the GPU (two streams, two host threads) race on it.
namespace {

constexpr long long kNoAllocMaxPoints = PCP_NOALLOC_MAX_POINTS;
static_assert(kNoAllocMaxPoints % kLeaf == 0, "the scratch holds whole groups of 32 slots");
static __device__ __align__(256) float4 g_noalloc_pairs[kNoAllocMaxPoints];

// raw rows -> pair layout in input order: entries 2m / 2m+1 = (x0, x1, y0, y1) / (z0, z1, w0, w1) of points 2m, 2m+1,
// w = the row number; slots beyond the cloud are pads (x = +inf: d2 = +inf passes no gate; index -1)
__global__ void __launch_bounds__(256) noalloc_pack_kernel(const float *__restrict__ xyz, long long n, long long stride,
                                                            long long n_slots, float4 *__restrict__ pairs) {
    const long long m = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (2 * m >= n_slots) return;
    const float inf = __uint_as_float(0x7f800000u);
    float x[2] = {inf, inf}, y[2] = {0.f, 0.f}, z[2] = {0.f, 0.f};
    unsigned w[2] = {0xffffffffu, 0xffffffffu};
#pragma unroll
    for (int j = 0; j < 2; ++j) {
        const long long i = 2 * m + j;
        if (i < n) {
            const float *p = xyz + i * stride;
            x[j] = __ldg(p); y[j] = __ldg(p + 1); z[j] = __ldg(p + 2);
            w[j] = (unsigned)i;
        }
    }
    pairs[2 * m] = make_float4(x[0], x[1], y[0], y[1]);
    pairs[2 * m + 1] = make_float4(z[0], z[1], __uint_as_float(w[0]), __uint_as_float(w[1]));
}

constexpr int kNoAllocMaxDevices = 64;
std::mutex g_noalloc_mutex;
float4 *g_noalloc_addr[kNoAllocMaxDevices] = {nullptr};

// Once per device, with that device current: the DEVICE address of the array (in host code its name is only a
// host-side stand-in) and the shared-memory opt-in of the three kernel widths.  None of these calls allocates.
int noalloc_setup(int device, float4 **out) {
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_addr[device]) {
        void *ptr = nullptr;
        PCP_CUDA(cudaGetSymbolAddress(&ptr, g_noalloc_pairs));
        if (!ptr) return fail(PCP_ERR_CUDA, "pcp_knn_bruteforce_noalloc: cudaGetSymbolAddress returned NULL");
        const int smem = (int)(2 * kTileBytes + 16);
        PCP_CUDA(cudaFuncSetAttribute(brute_kernel<8, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        PCP_CUDA(cudaFuncSetAttribute(brute_kernel<16, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        PCP_CUDA(cudaFuncSetAttribute(brute_kernel<32, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        g_noalloc_addr[device] = static_cast<float4 *>(ptr);
    }
    *out = g_noalloc_addr[device];
    return PCP_OK;
}

}  // namespace

}  // namespace pcp

extern "C" PCP_API int pcp_knn_bruteforce_noalloc(int device, const float *d_points, int64_t n, int64_t point_stride,
                                                  const float *d_queries, int64_t nq, int64_t query_stride, int k,
                                                  int32_t *d_idx, float *d_dist, int32_t *d_count, void *stream) {
    using namespace pcp;
    if (n < 0 || nq < 0) return fail(PCP_ERR_INVALID, "pcp_knn_bruteforce_noalloc: negative size");
    if (k < 1 || k > PCP_MAX_K) return fail(PCP_ERR_INVALID, "pcp_knn_bruteforce_noalloc: k must be in 1.." + std::to_string(PCP_MAX_K));
    if (n > kNoAllocMaxPoints)
        return fail(PCP_ERR_UNSUPPORTED, "pcp_knn_bruteforce_noalloc: " + std::to_string(n) + " points exceed the static scratch of " +
                                             std::to_string(kNoAllocMaxPoints) + " points (PCP_NOALLOC_MAX_POINTS); use pcp_knn_bruteforce");
    if ((n > 0 && (!d_points || point_stride < 3)) || (nq > 0 && (!d_queries || query_stride < 3)))
        return fail(PCP_ERR_INVALID, "pcp_knn_bruteforce_noalloc: NULL cloud or stride < 3");
    if (nq > 0 && (!d_idx || !d_dist)) return fail(PCP_ERR_INVALID, "pcp_knn_bruteforce_noalloc: idx/dist is NULL");
    if (device < 0 || device >= kNoAllocMaxDevices) return fail(PCP_ERR_INVALID, "pcp_knn_bruteforce_noalloc: bad device");
    if (nq == 0) return PCP_OK;
    int prev = -1;
    PCP_CUDA(cudaGetDevice(&prev));
    if (prev != device) PCP_CUDA(cudaSetDevice(device));
    struct Restore {
        int prev, device;
        ~Restore() { if (prev != device) cudaSetDevice(prev); }
    } restore{prev, device};
    float4 *pairs = nullptr;
    PCP_TRY(noalloc_setup(device, &pairs));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const long long n_slots = (n + kLeaf - 1) / kLeaf * kLeaf;
    if (n_slots > 0) {
        noalloc_pack_kernel<<<(unsigned)((n_slots / 2 + 255) / 256), 256, 0, st>>>(d_points, n, point_stride, n_slots, pairs);
        PCP_CUDA(cudaGetLastError());
    }
    const unsigned grid = (unsigned)((nq + kBruteThreads - 1) / kBruteThreads);
    const size_t smem = 2 * kTileBytes + 16;
    if (k <= 8)
        brute_kernel<8, 0><<<grid, kBruteThreads, smem, st>>>(pairs, n_slots, d_queries, nq, query_stride, k, 0.f, d_idx, d_dist, d_count, nullptr);
    else if (k <= 16)
        brute_kernel<16, 0><<<grid, kBruteThreads, smem, st>>>(pairs, n_slots, d_queries, nq, query_stride, k, 0.f, d_idx, d_dist, d_count, nullptr);
    else
        brute_kernel<32, 0><<<grid, kBruteThreads, smem, st>>>(pairs, n_slots, d_queries, nq, query_stride, k, 0.f, d_idx, d_dist, d_count, nullptr);
    PCP_CUDA(cudaGetLastError());
    return PCP_OK;
}
