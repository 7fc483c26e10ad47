// Multi-GPU entry points of the C ABI (include/pcp.h, "multi-GPU"): one process per GPU, work sharded where the
// path shards.
//
//   query-sharded search over a REPLICATED reference cloud (BASELINE config 5): every rank holds a shard of the
//     reference; ONE ncclAllGather over NVLink rebuilds the whole cloud on every GPU, each rank builds its own
//     (deterministic, hence identical) index and searches its own queries.  That all-gather is the only collective
//     of the library.
//   registration pairs (configs 2 and 4): independent -> a contiguous range of pairs per rank, balanced by
//     points, NO collective; every rank writes the rows of its own range.
//
// NCCL is bound at run time (dlopen of libnccl.so.2 -- the copy already loaded into the process if there is one,
// e.g. the one torch ships, else the system library): libpcp_b200.so has no link-time dependency on it and the
// single-GPU entry points work on a machine without NCCL.
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#include "pcp_internal.cuh"

namespace pcp {

namespace {

struct NcclApi {
    void *lib = nullptr;
    decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
    decltype(&ncclCommInitRank) CommInitRank = nullptr;
    decltype(&ncclCommDestroy) CommDestroy = nullptr;
    decltype(&ncclAllGather) AllGather = nullptr;
    decltype(&ncclGetErrorString) GetErrorString = nullptr;
    decltype(&ncclGetVersion) GetVersion = nullptr;
    std::string error;
};

NcclApi &nccl() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        // RTLD_NOLOAD first: reuse the NCCL the process already carries (two copies must not coexist)
        for (const char *name : {"libnccl.so.2", "libnccl.so"}) {
            api.lib = dlopen(name, RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
            if (api.lib) break;
        }
        // PCP_NCCL_LIB: an explicit path (multi.py points it at the NCCL wheel next to torch, so that a later
        // `import torch` in the same process finds its own, newer libnccl.so.2 already loaded instead of an older
        // system copy under the same soname)
        if (!api.lib)
            if (const char *path = getenv("PCP_NCCL_LIB")) api.lib = dlopen(path, RTLD_NOW | RTLD_GLOBAL);
        if (!api.lib)
            for (const char *name : {"libnccl.so.2", "libnccl.so"}) {
                api.lib = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
                if (api.lib) break;
            }
        if (!api.lib) {
            api.error = std::string("cannot load libnccl.so.2: ") + (dlerror() ? dlerror() : "?");
            return;
        }
#define PCP_SYM(field, name)                                                            \
    api.field = reinterpret_cast<decltype(api.field)>(dlsym(api.lib, name));            \
    if (!api.field && api.error.empty()) api.error = std::string("libnccl: missing symbol ") + name
        PCP_SYM(GetUniqueId, "ncclGetUniqueId");
        PCP_SYM(CommInitRank, "ncclCommInitRank");
        PCP_SYM(CommDestroy, "ncclCommDestroy");
        PCP_SYM(AllGather, "ncclAllGather");
        PCP_SYM(GetErrorString, "ncclGetErrorString");
        PCP_SYM(GetVersion, "ncclGetVersion");
#undef PCP_SYM
    });
    return api;
}

int nccl_fail(ncclResult_t r, const char *what) {
    NcclApi &api = nccl();
    return fail(PCP_ERR_CUDA, std::string(what) + ": NCCL error " + std::to_string((int)r) + " (" +
                                  (api.GetErrorString ? api.GetErrorString(r) : "?") + ")");
}

#define PCP_NCCL(expr)                                          \
    do {                                                        \
        ncclResult_t _r = (expr);                               \
        if (_r != ncclSuccess) return nccl_fail(_r, #expr);     \
    } while (0)

}  // namespace

}  // namespace pcp

struct pcp_multi {
    pcp_ctx *ctx = nullptr;
    int rank = 0, world = 1;
    ncclComm_t comm = nullptr;
    pcp_index *index = nullptr;              // over the gathered reference cloud
    pcp::DevBuf gathered, pad, sizes;
    std::vector<int64_t> shard_rows;         // rows contributed by every rank
    int64_t n_total = 0, stride = 3;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};   // gather begin / end, build end, spare
    float gather_ms = 0.f, build_ms = 0.f;
};

using namespace pcp;

extern "C" {

int pcp_multi_unique_id(void *id128) {
    if (!id128) return fail(PCP_ERR_INVALID, "pcp_multi_unique_id: id is NULL");
    NcclApi &api = nccl();
    if (!api.error.empty()) return fail(PCP_ERR_UNSUPPORTED, "pcp_multi_unique_id: " + api.error);
    static_assert(sizeof(ncclUniqueId) == PCP_MULTI_ID_BYTES, "unique id size");
    ncclUniqueId id;
    PCP_NCCL(api.GetUniqueId(&id));
    memcpy(id128, &id, sizeof id);
    return PCP_OK;
}

int pcp_multi_create(pcp_ctx *ctx, int rank, int world, const void *id128, pcp_multi **out) {
    if (!ctx || !out) return fail(PCP_ERR_INVALID, "pcp_multi_create: ctx/out is NULL");
    if (world < 1 || rank < 0 || rank >= world) return fail(PCP_ERR_INVALID, "pcp_multi_create: rank/world out of range");
    *out = nullptr;
    pcp_multi *m = new (std::nothrow) pcp_multi();
    if (!m) return fail(PCP_ERR_OOM, "pcp_multi_create: out of host memory");
    m->ctx = ctx;
    m->rank = rank;
    m->world = world;
    cudaSetDevice(ctx->device);
    for (auto &e : m->ev) cudaEventCreate(&e);
    if (world > 1) {
        NcclApi &api = nccl();
        if (!api.error.empty()) { delete m; return fail(PCP_ERR_UNSUPPORTED, "pcp_multi_create: " + api.error); }
        if (!id128) { delete m; return fail(PCP_ERR_INVALID, "pcp_multi_create: world > 1 needs the unique id of rank 0"); }
        ncclUniqueId id;
        memcpy(&id, id128, sizeof id);
        ncclResult_t r = api.CommInitRank(&m->comm, world, id, rank);
        if (r != ncclSuccess) { delete m; return nccl_fail(r, "ncclCommInitRank"); }
    }
    int s = pcp_index_build(ctx, nullptr, 0, 3, PCP_DEVICE, &m->index);
    if (s != PCP_OK) { pcp_multi_destroy(m); return s; }
    *out = m;
    return PCP_OK;
}

int pcp_multi_destroy(pcp_multi *m) {
    if (!m) return PCP_OK;
    cudaSetDevice(m->ctx->device);
    cudaStreamSynchronize(m->ctx->stream);
    if (m->index) pcp_index_destroy(m->index);
    if (m->comm) nccl().CommDestroy(m->comm);
    for (auto &e : m->ev) if (e) cudaEventDestroy(e);
    m->gathered.release();
    m->pad.release();
    m->sizes.release();
    delete m;
    return PCP_OK;
}

int pcp_multi_info(const pcp_multi *m, int *rank, int *world, int *nccl_version, int *comm_nranks) {
    if (!m) return fail(PCP_ERR_INVALID, "pcp_multi_info: handle is NULL");
    if (rank) *rank = m->rank;
    if (world) *world = m->world;
    if (nccl_version) {
        *nccl_version = 0;
        if (m->world > 1) {                      // (a single rank never loads NCCL)
            NcclApi &api = nccl();
            if (api.GetVersion) api.GetVersion(nccl_version);
        }
    }
    if (comm_nranks) *comm_nranks = m->comm ? m->world : 1;
    return PCP_OK;
}

// All-gather of the reference shards + index build over the gathered cloud.
int pcp_multi_set_reference(pcp_multi *m, const float *shard_xyz, int64_t n_shard, int64_t stride, int mem,
                            int64_t *n_total) {
    if (!m) return fail(PCP_ERR_INVALID, "pcp_multi_set_reference: handle is NULL");
    if (n_shard < 0 || stride < 3) return fail(PCP_ERR_INVALID, "pcp_multi_set_reference: bad shard shape");
    if (n_shard > 0 && !shard_xyz) return fail(PCP_ERR_INVALID, "pcp_multi_set_reference: shard is NULL");
    if (mem != PCP_HOST && mem != PCP_DEVICE) return fail(PCP_ERR_INVALID, "pcp_multi_set_reference: bad mem");
    pcp_ctx *ctx = m->ctx;
    PCP_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const int W = m->world;
    const size_t row = sizeof(float) * (size_t)stride;
    m->stride = stride;
    m->shard_rows.assign((size_t)W, 0);
    const float *d_shard = shard_xyz;
    if (mem == PCP_HOST && n_shard > 0) {
        PCP_TRY(m->pad.ensure(row * (size_t)n_shard));
        PCP_CUDA(cudaMemcpyAsync(m->pad.p, shard_xyz, row * (size_t)n_shard, cudaMemcpyHostToDevice, st));
        d_shard = m->pad.as<float>();
    }
    PCP_CUDA(cudaEventRecord(m->ev[0], st));
    if (W == 1) {
        m->shard_rows[0] = n_shard;
        m->n_total = n_shard;
        PCP_TRY(m->gathered.ensure(std::max<size_t>(row * (size_t)n_shard, 16)));
        if (n_shard > 0) PCP_CUDA(cudaMemcpyAsync(m->gathered.p, d_shard, row * (size_t)n_shard, cudaMemcpyDeviceToDevice, st));
    } else {
        NcclApi &api = nccl();
        // shard sizes (8 bytes per rank) first: shards may differ in length
        PCP_TRY(m->sizes.ensure(sizeof(int64_t) * (size_t)(W + 1)));
        int64_t *d_sizes = m->sizes.as<int64_t>();
        PCP_CUDA(cudaMemcpyAsync(d_sizes + W, &n_shard, sizeof(int64_t), cudaMemcpyHostToDevice, st));
        PCP_NCCL(api.AllGather(d_sizes + W, d_sizes, sizeof(int64_t), ncclChar, m->comm, st));
        PCP_CUDA(cudaMemcpyAsync(m->shard_rows.data(), d_sizes, sizeof(int64_t) * (size_t)W, cudaMemcpyDeviceToHost, st));
        PCP_CUDA(cudaStreamSynchronize(st));
        int64_t n_max = 0, total = 0;
        for (int64_t v : m->shard_rows) { n_max = std::max(n_max, v); total += v; }
        m->n_total = total;
        const bool even = std::all_of(m->shard_rows.begin(), m->shard_rows.end(), [&](int64_t v) { return v == n_max; });
        PCP_CUDA(cudaEventRecord(m->ev[0], st));
        PCP_TRY(m->gathered.ensure(std::max<size_t>(row * (size_t)n_max * (size_t)W, 16)));
        if (n_max > 0) {
            if (even) {
                // THE collective: every rank contributes its rows, every rank receives the whole cloud
                PCP_NCCL(api.AllGather(d_shard, m->gathered.p, row * (size_t)n_max, ncclChar, m->comm, st));
            } else {
                // ragged shards: gather slots of n_max rows, then close the gaps in place (front to back)
                DevBuf &tmp = ctx->scratch[2];
                PCP_TRY(tmp.ensure(row * (size_t)n_max));
                PCP_CUDA(cudaMemsetAsync(tmp.p, 0, row * (size_t)n_max, st));
                if (n_shard > 0) PCP_CUDA(cudaMemcpyAsync(tmp.p, d_shard, row * (size_t)n_shard, cudaMemcpyDeviceToDevice, st));
                PCP_NCCL(api.AllGather(tmp.p, m->gathered.p, row * (size_t)n_max, ncclChar, m->comm, st));
                size_t at = (size_t)m->shard_rows[0];
                for (int r = 1; r < W; ++r) {
                    const size_t rows = (size_t)m->shard_rows[(size_t)r];
                    // (destination never overlaps a later source: at <= r * n_max)
                    if (rows > 0 && at != (size_t)r * (size_t)n_max)
                        PCP_CUDA(cudaMemcpyAsync(static_cast<char *>(m->gathered.p) + at * row,
                                                 static_cast<char *>(m->gathered.p) + (size_t)r * (size_t)n_max * row,
                                                 rows * row, cudaMemcpyDeviceToDevice, st));
                    at += rows;
                }
            }
        }
    }
    PCP_CUDA(cudaEventRecord(m->ev[1], st));
    PCP_TRY(pcp_index_rebuild(m->index, m->gathered.as<float>(), m->n_total, stride, PCP_DEVICE));
    PCP_CUDA(cudaEventRecord(m->ev[2], st));
    PCP_CUDA(cudaStreamSynchronize(st));
    PCP_CUDA(cudaEventElapsedTime(&m->gather_ms, m->ev[0], m->ev[1]));
    PCP_CUDA(cudaEventElapsedTime(&m->build_ms, m->ev[1], m->ev[2]));
    if (n_total) *n_total = m->n_total;
    return PCP_OK;
}

int pcp_multi_last_timing(const pcp_multi *m, double *allgather_ms, double *build_ms) {
    if (!m) return fail(PCP_ERR_INVALID, "pcp_multi_last_timing: handle is NULL");
    if (allgather_ms) *allgather_ms = m->gather_ms;
    if (build_ms) *build_ms = m->build_ms;
    return PCP_OK;
}

int pcp_multi_reference(const pcp_multi *m, const float **device_xyz, int64_t *n_total, int64_t *stride,
                        int64_t *shard_rows) {
    if (!m) return fail(PCP_ERR_INVALID, "pcp_multi_reference: handle is NULL");
    if (device_xyz) *device_xyz = m->gathered.as<float>();
    if (n_total) *n_total = m->n_total;
    if (stride) *stride = m->stride;
    if (shard_rows) for (int r = 0; r < m->world; ++r) shard_rows[r] = r < (int)m->shard_rows.size() ? m->shard_rows[(size_t)r] : 0;
    return PCP_OK;
}

pcp_index *pcp_multi_index(pcp_multi *m) { return m ? m->index : nullptr; }

int pcp_multi_knn(pcp_multi *m, const float *queries, int64_t nq, int64_t stride, int k, int mem, int32_t *idx,
                  float *dist, int32_t *count) {
    if (!m) return fail(PCP_ERR_INVALID, "pcp_multi_knn: handle is NULL");
    return pcp_knn(m->index, queries, nq, stride, k, mem, idx, dist, count);
}

int pcp_multi_pair_range(const int64_t *src_offsets, const int64_t *tgt_offsets, int64_t n_pairs, int rank, int world,
                         int64_t *first, int64_t *last) {
    if (!src_offsets || !tgt_offsets || !first || !last) return fail(PCP_ERR_INVALID, "pcp_multi_pair_range: NULL argument");
    if (n_pairs < 0 || world < 1 || rank < 0 || rank >= world) return fail(PCP_ERR_INVALID, "pcp_multi_pair_range: bad range");
    // boundary r = first pair whose cumulative point count reaches r / world of the total (non-decreasing)
    const double total = (double)((src_offsets[n_pairs] - src_offsets[0]) + (tgt_offsets[n_pairs] - tgt_offsets[0]));
    auto boundary = [&](int r) -> int64_t {
        if (r <= 0) return 0;
        if (r >= world) return n_pairs;
        const double target = total * r / world;
        int64_t lo = 0, hi = n_pairs;                 // first i with cum(i) >= target, cum(i) = points of pairs [0, i)
        while (lo < hi) {
            const int64_t mid = (lo + hi) / 2;
            const double cum = (double)((src_offsets[mid] - src_offsets[0]) + (tgt_offsets[mid] - tgt_offsets[0]));
            if (cum >= target) hi = mid; else lo = mid + 1;
        }
        return lo;
    };
    int64_t b0 = 0;
    for (int r = 1; r <= rank; ++r) b0 = std::max(b0, boundary(r));
    int64_t b1 = rank + 1 >= world ? n_pairs : std::max(b0, boundary(rank + 1));
    *first = b0;
    *last = b1;
    return PCP_OK;
}

int pcp_multi_icp_batch(pcp_multi *m, const pcp_icp_params *params, int64_t n_pairs, const float *src_xyz,
                        const int64_t *src_offsets, const float *tgt_xyz, const int64_t *tgt_offsets, int64_t stride,
                        int mem, const double *T_init, double *T_last, double *T_total, int32_t *iterations,
                        int32_t *status, double *diff_log, int64_t *first_pair, int64_t *last_pair) {
    if (!m) return fail(PCP_ERR_INVALID, "pcp_multi_icp_batch: handle is NULL");
    if (!params) return fail(PCP_ERR_INVALID, "pcp_multi_icp_batch: params is NULL");
    int64_t lo = 0, hi = 0;
    PCP_TRY(pcp_multi_pair_range(src_offsets, tgt_offsets, n_pairs, m->rank, m->world, &lo, &hi));
    if (first_pair) *first_pair = lo;
    if (last_pair) *last_pair = hi;
    const int64_t P = hi - lo;
    if (P == 0) return PCP_OK;
    std::vector<int64_t> so((size_t)P + 1), to((size_t)P + 1);
    for (int64_t i = 0; i <= P; ++i) {
        so[(size_t)i] = src_offsets[lo + i] - src_offsets[lo];
        to[(size_t)i] = tgt_offsets[lo + i] - tgt_offsets[lo];
    }
    const int32_t mi = params->max_iteration;
    return pcp_icp_batch(m->ctx, params, P, src_xyz + src_offsets[lo] * stride, so.data(), tgt_xyz + tgt_offsets[lo] * stride,
                         to.data(), stride, mem, T_init ? T_init + lo * 16 : nullptr, T_last ? T_last + lo * 16 : nullptr,
                         T_total ? T_total + lo * 16 : nullptr, iterations ? iterations + lo : nullptr,
                         status ? status + lo : nullptr, diff_log ? diff_log + lo * (int64_t)mi * 2 : nullptr);
}

}  // extern "C"
