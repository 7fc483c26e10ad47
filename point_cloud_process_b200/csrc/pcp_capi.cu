// C ABI of libpcp_b200.so (include/pcp.h): argument checking, host<->device staging, error state.
#include <cmath>
#include <cstdio>
#include <cstring>
#include <map>
#include <mutex>
#include <new>
#include <unordered_map>

#include "pcp_internal.cuh"

namespace pcp {

static thread_local std::string g_last_error;

void set_error(const std::string &msg) { g_last_error = msg; }

int fail(int code, const std::string &msg) {
    g_last_error = msg;
    return code;
}

int cuda_fail(cudaError_t e, const char *what, const char *file, int line) {
    char buf[512];
    snprintf(buf, sizeof buf, "CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), file, line, what);
    g_last_error = buf;
    cudaGetLastError();   // clear sticky-free errors
    return e == cudaErrorMemoryAllocation ? PCP_ERR_OOM : PCP_ERR_CUDA;
}

static thread_local bool g_capture_on = false, g_capture_grew = false;
void capture_note_begin() { g_capture_on = true; g_capture_grew = false; }
bool capture_note_grew() { g_capture_on = false; return g_capture_grew; }

int DevBuf::ensure(size_t bytes) {
    if (bytes <= cap) return PCP_OK;
    if (g_capture_on) g_capture_grew = true;
    size_t want = bytes + bytes / 8 + 256;
    void *np = nullptr;
    cudaError_t e = cudaMalloc(&np, want);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(PCP_ERR_OOM, "device allocation of " + std::to_string(want) + " bytes failed: " +
                                     cudaGetErrorString(e));
    }
    if (p) cudaFree(p);     // contents are scratch: never preserved across a grow
    p = np;
    cap = want;
    return PCP_OK;
}
void DevBuf::release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
}
int PinBuf::ensure(size_t bytes) {
    if (bytes <= cap) return PCP_OK;
    size_t want = bytes + bytes / 8 + 256;
    void *np = nullptr;
    cudaError_t e = cudaMallocHost(&np, want);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(PCP_ERR_OOM, "pinned host allocation of " + std::to_string(want) + " bytes failed");
    }
    if (p) cudaFreeHost(p);
    p = np;
    cap = want;
    return PCP_OK;
}
void PinBuf::release() {
    if (p) cudaFreeHost(p);
    p = nullptr;
    cap = 0;
}

namespace {

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) {
        cudaGetDevice(&prev);
        if (prev != dev) cudaSetDevice(dev);
        else prev = -1;
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

// host -> device through a pinned bounce buffer (pageable memcpyAsync would serialise anyway;
// pinned staging keeps the copy on the context stream at full PCIe rate)
int upload(pcp_ctx *ctx, int pin_slot, DevBuf &dst, const void *src, size_t bytes) {
    PCP_TRY(dst.ensure(bytes));
    if (bytes == 0) return PCP_OK;
    cudaPointerAttributes attr;
    bool pinned = cudaPointerGetAttributes(&attr, src) == cudaSuccess && attr.type == cudaMemoryTypeHost;
    cudaGetLastError();
    if (pinned) {
        PCP_CUDA(cudaMemcpyAsync(dst.p, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
        return PCP_OK;
    }
    PCP_TRY(ctx->pin[pin_slot].ensure(bytes));
    memcpy(ctx->pin[pin_slot].p, src, bytes);
    PCP_CUDA(cudaMemcpyAsync(dst.p, ctx->pin[pin_slot].p, bytes, cudaMemcpyHostToDevice, ctx->stream));
    return PCP_OK;
}

// Host -> device copy of a QUERY set (scratch destination that no queued kernel reads: every PCP_HOST search
// call synchronises before it returns).  A page-locked source is copied on the side stream, so that the copy
// overlaps whatever is still queued on the main stream (typically the index build of a pcp_index_rebuild that
// returned as soon as its own upload was done); the main stream then waits for the copy.
int upload_side(pcp_ctx *ctx, int pin_slot, DevBuf &dst, const void *src, size_t bytes, bool *on_side = nullptr,
                bool main_waits = true) {
    if (on_side) *on_side = false;
    PCP_TRY(dst.ensure(bytes));
    if (bytes == 0) return PCP_OK;
    cudaPointerAttributes attr;
    const bool pinned = cudaPointerGetAttributes(&attr, src) == cudaSuccess && attr.type == cudaMemoryTypeHost;
    cudaGetLastError();
    if (!pinned) return upload(ctx, pin_slot, dst, src, bytes);
    PCP_CUDA(cudaMemcpyAsync(dst.p, src, bytes, cudaMemcpyHostToDevice, ctx->copy_stream));
    PCP_CUDA(cudaEventRecord(ctx->ev_copy, ctx->copy_stream));
    if (main_waits) PCP_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->ev_copy, 0));   // else: the caller orders on ev_copy
    if (on_side) *on_side = true;
    return PCP_OK;
}

bool host_is_pinned(const void *p) {
    cudaPointerAttributes attr;
    const bool pinned = cudaPointerGetAttributes(&attr, p) == cudaSuccess && attr.type == cudaMemoryTypeHost;
    cudaGetLastError();
    return pinned;
}

// points (source + target) per run of pairs of the pipelined pcp_icp_batch(PCP_HOST)
constexpr int64_t kIcpChunkPoints = 12 << 20;

// device -> host, synchronous on return
int download(pcp_ctx *ctx, void *dst, const void *src, size_t bytes) {
    if (bytes == 0) return PCP_OK;
    PCP_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    return PCP_OK;
}

// kNN results device slab (idx | dist | count, contiguous) -> host.  One copy when the caller's three arrays are laid
// out the same way (the pybind11 layer allocates them as one page-locked slab): every separate copy costs ~10 us of
// fixed latency on top of its bytes.
int download_knn(pcp_ctx *ctx, int32_t *idx, float *dist, int32_t *count, const int32_t *d_idx, size_t nres, size_t nq) {
    const char *d = reinterpret_cast<const char *>(d_idx);
    const bool slab = reinterpret_cast<char *>(dist) == reinterpret_cast<char *>(idx) + nres * 4 &&
                      (!count || reinterpret_cast<char *>(count) == reinterpret_cast<char *>(dist) + nres * 4);
    if (slab) return download(ctx, idx, d, nres * 8 + (count ? nq * 4 : 0));
    PCP_TRY(download(ctx, idx, d, nres * 4));
    PCP_TRY(download(ctx, dist, d + nres * 4, nres * 4));
    if (count) PCP_TRY(download(ctx, count, d + nres * 8, nq * 4));
    return PCP_OK;
}

bool bad_mem(int mem) { return mem != PCP_HOST && mem != PCP_DEVICE; }

}  // namespace
}  // namespace pcp

using namespace pcp;

extern "C" {

int pcp_version(void) { return PCP_VERSION; }

const char *pcp_last_error(void) { return g_last_error.c_str(); }

// ---------------------------------------------------------------------------------------------- pinned pool
namespace {
struct HostPool {
    std::mutex mu;
    std::map<size_t, std::vector<void *>> free_by_class;     // size class -> cached buffers
    std::unordered_map<void *, size_t> live;                 // buffer -> size class
    size_t cached_bytes = 0;
    static constexpr size_t kMaxCached = size_t(4) << 30;
    static size_t size_class(size_t bytes) {
        const size_t big = size_t(64) << 20;                 // powers of two up to 64 MB, then multiples of 64 MB
        if (bytes > big) return (bytes + big - 1) / big * big;
        size_t c = 4096;
        while (c < bytes) c <<= 1;
        return c;
    }
};
HostPool &host_pool() {
    static HostPool *p = new HostPool();   // never destroyed: buffers may outlive static destruction order
    return *p;
}
}  // namespace

int pcp_host_alloc(size_t bytes, void **out) {
    if (!out) return fail(PCP_ERR_INVALID, "pcp_host_alloc: out is NULL");
    *out = nullptr;
    HostPool &hp = host_pool();
    const size_t cls = HostPool::size_class(bytes);
    {
        std::lock_guard<std::mutex> lk(hp.mu);
        auto it = hp.free_by_class.find(cls);
        if (it != hp.free_by_class.end() && !it->second.empty()) {
            *out = it->second.back();
            it->second.pop_back();
            hp.cached_bytes -= cls;
            hp.live[*out] = cls;
            return PCP_OK;
        }
    }
    void *p = nullptr;
    cudaError_t e = cudaHostAlloc(&p, cls, cudaHostAllocPortable);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(e == cudaErrorMemoryAllocation ? PCP_ERR_OOM : PCP_ERR_CUDA,
                    std::string("pcp_host_alloc: ") + cudaGetErrorString(e));
    }
    std::lock_guard<std::mutex> lk(hp.mu);
    hp.live[p] = cls;
    *out = p;
    return PCP_OK;
}

int pcp_host_free(void *ptr) {
    if (!ptr) return PCP_OK;
    HostPool &hp = host_pool();
    size_t cls = 0;
    {
        std::lock_guard<std::mutex> lk(hp.mu);
        auto it = hp.live.find(ptr);
        if (it == hp.live.end()) return fail(PCP_ERR_INVALID, "pcp_host_free: pointer was not returned by pcp_host_alloc");
        cls = it->second;
        hp.live.erase(it);
        if (hp.cached_bytes + cls <= HostPool::kMaxCached) {
            hp.free_by_class[cls].push_back(ptr);
            hp.cached_bytes += cls;
            return PCP_OK;
        }
    }
    cudaFreeHost(ptr);
    cudaGetLastError();
    return PCP_OK;
}

int pcp_device_count(int *count) {
    if (!count) return fail(PCP_ERR_INVALID, "pcp_device_count: count is NULL");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        cudaGetLastError();
        *count = 0;
        return fail(PCP_ERR_CUDA, std::string("no CUDA device: ") + cudaGetErrorString(e));
    }
    *count = n;
    return PCP_OK;
}

int pcp_ctx_create(int device, pcp_ctx **out) {
    if (!out) return fail(PCP_ERR_INVALID, "pcp_ctx_create: out is NULL");
    *out = nullptr;
    int n = 0;
    PCP_TRY(pcp_device_count(&n));
    if (device < 0 || device >= n)
        return fail(PCP_ERR_CUDA, "pcp_ctx_create: device " + std::to_string(device) + " not available (" +
                                      std::to_string(n) + " CUDA devices); this library has no CPU fallback");
    cudaDeviceProp prop;
    PCP_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10)
        return fail(PCP_ERR_CUDA, std::string("pcp_ctx_create: ") + prop.name + " is sm_" + std::to_string(prop.major) +
                                      std::to_string(prop.minor) + "; libpcp_b200 is built for sm_100a only");
    DeviceGuard g(device);
    pcp_ctx *ctx = new (std::nothrow) pcp_ctx();
    if (!ctx) return fail(PCP_ERR_OOM, "pcp_ctx_create: out of host memory");
    ctx->device = device;
    ctx->sm_count = prop.multiProcessorCount;
    cudaError_t e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) {
        delete ctx;
        return cuda_fail(e, "cudaStreamCreateWithFlags", __FILE__, __LINE__);
    }
    if (cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&ctx->ev_copy, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&ctx->ev_q_ready, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&ctx->ev_cap_bbox, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&ctx->ev_cap_q, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&ctx->ev_search_end, cudaEventDisableTiming) != cudaSuccess) {
        cudaGetLastError();
        pcp_ctx_destroy(ctx);
        return fail(PCP_ERR_CUDA, "pcp_ctx_create: cannot create the copy stream");
    }
    *out = ctx;
    return PCP_OK;
}

int pcp_ctx_destroy(pcp_ctx *ctx) {
    if (!ctx) return PCP_OK;
    DeviceGuard g(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    ctx->cub_tmp.release();
    for (auto &b : ctx->ss_state) b.release();
    ctx->counters.release();
    ctx->radius_perm.release();
    ctx->voxel_out.release();
    ctx->voxel_meta.release();
    ctx->normals_nn.release();
    for (auto &b : ctx->iss_buf) b.release();
    ctx->icp_targets.release();
    ctx->icp_sources.release();
    for (auto &b : ctx->icp_up) b.release();
    ctx->small_stage.release();
    for (auto &e : ctx->ev_icp_up) if (e) cudaEventDestroy(e);
    for (auto &b : ctx->scratch) b.release();
    for (auto &b : ctx->pin) b.release();
    for (auto &e : ctx->timing_events) { cudaEventDestroy(e.first); cudaEventDestroy(e.second); }
    if (ctx->copy_stream) cudaStreamSynchronize(ctx->copy_stream);
    for (auto &b : ctx->q_keys) b.release();
    ctx->q_perm.release();
    ctx->cub_tmp_side.release();
    if (ctx->ev_copy) cudaEventDestroy(ctx->ev_copy);
    if (ctx->ev_q_ready) cudaEventDestroy(ctx->ev_q_ready);
    if (ctx->ev_search_end) cudaEventDestroy(ctx->ev_search_end);
    if (ctx->ev_cap_bbox) cudaEventDestroy(ctx->ev_cap_bbox);
    if (ctx->ev_cap_q) cudaEventDestroy(ctx->ev_cap_q);
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    cudaStreamDestroy(ctx->stream);
    delete ctx;
    return PCP_OK;
}

int pcp_ctx_synchronize(pcp_ctx *ctx) {
    if (!ctx) return fail(PCP_ERR_INVALID, "pcp_ctx_synchronize: ctx is NULL");
    DeviceGuard g(ctx->device);
    PCP_CUDA(cudaStreamSynchronize(ctx->stream));
    return PCP_OK;
}

void *pcp_ctx_stream(pcp_ctx *ctx) { return ctx ? (void *)ctx->stream : nullptr; }

int64_t pcp_ctx_launch_count(pcp_ctx *ctx) { return ctx ? ctx->launches : 0; }

int pcp_ctx_enable_kernel_timing(pcp_ctx *ctx, int enable) {
    if (!ctx) return fail(PCP_ERR_INVALID, "pcp_ctx_enable_kernel_timing: ctx is NULL");
    ctx->time_kernels = enable != 0;
    ctx->timing_used = 0;
    return PCP_OK;
}

int pcp_ctx_read_kernel_timing(pcp_ctx *ctx, double *total_ms, int64_t *launches) {
    if (!ctx || !total_ms || !launches) return fail(PCP_ERR_INVALID, "pcp_ctx_read_kernel_timing: NULL argument");
    DeviceGuard g(ctx->device);
    PCP_CUDA(cudaStreamSynchronize(ctx->stream));
    double sum = 0.0;
    for (size_t i = 0; i < ctx->timing_used; ++i) {
        float ms = 0.f;
        PCP_CUDA(cudaEventElapsedTime(&ms, ctx->timing_events[i].first, ctx->timing_events[i].second));
        sum += ms;
    }
    *total_ms = sum;
    *launches = (int64_t)ctx->timing_used;
    ctx->timing_used = 0;
    return PCP_OK;
}

int pcp_ctx_enable_counters(pcp_ctx *ctx, int enable) {
    if (!ctx) return fail(PCP_ERR_INVALID, "pcp_ctx_enable_counters: ctx is NULL");
    DeviceGuard g(ctx->device);
    if (enable) {
        PCP_TRY(ctx->counters.ensure(8 * sizeof(unsigned long long)));
        PCP_CUDA(cudaMemsetAsync(ctx->counters.p, 0, 8 * sizeof(unsigned long long), ctx->stream));
    }
    ctx->count_work = enable != 0;
    return PCP_OK;
}

int pcp_ctx_read_counters(pcp_ctx *ctx, uint64_t out[8]) {
    if (!ctx || !out) return fail(PCP_ERR_INVALID, "pcp_ctx_read_counters: NULL argument");
    memset(out, 0, 8 * sizeof(uint64_t));
    if (!ctx->counters.p) return PCP_OK;
    DeviceGuard g(ctx->device);
    PCP_CUDA(cudaMemcpyAsync(out, ctx->counters.p, 8 * sizeof(uint64_t), cudaMemcpyDeviceToHost, ctx->stream));
    PCP_CUDA(cudaMemsetAsync(ctx->counters.p, 0, 8 * sizeof(unsigned long long), ctx->stream));
    PCP_CUDA(cudaStreamSynchronize(ctx->stream));
    return PCP_OK;
}

// ---------------------------------------------------------------------------------------------- index

static int check_cloud(const char *fn, const float *xyz, int64_t n, int64_t stride, int mem) {
    if (n < 0) return fail(PCP_ERR_INVALID, std::string(fn) + ": negative point count");
    if (n > 0 && !xyz) return fail(PCP_ERR_INVALID, std::string(fn) + ": xyz is NULL");
    if (stride < 3)
        return fail(PCP_ERR_INVALID, std::string(fn) + ": stride_floats must be >= 3 (rows are x,y,z[,...]); the "
                                                       "search structure is 3-D only (dim != 3 is rejected)");
    if (bad_mem(mem)) return fail(PCP_ERR_INVALID, std::string(fn) + ": mem must be PCP_HOST or PCP_DEVICE");
    return PCP_OK;
}

int pcp_index_rebuild(pcp_index *ix, const float *xyz, int64_t n, int64_t stride, int mem) {
    if (!ix) return fail(PCP_ERR_INVALID, "pcp_index_rebuild: index is NULL");
    PCP_TRY(check_cloud("pcp_index_build", xyz, n, stride, mem));
    pcp_ctx *ctx = ix->ctx;
    DeviceGuard g(ctx->device);
    ix->xyz64_n = -1;                                   // float64 rows of an earlier cloud do not describe this one
    const float *d_xyz = xyz;
    if (mem == PCP_HOST && n > 0) {
        PCP_TRY(upload(ctx, 0, ix->src_copy, xyz, sizeof(float) * (size_t)n * (size_t)stride));
        d_xyz = ix->src_copy.as<float>();
    }
    if (mem == PCP_HOST && n > 0) PCP_CUDA(cudaEventRecord(ctx->ev_copy, ctx->stream));
    PCP_TRY(index_build_device(ix, d_xyz, n, stride));
    // PCP_HOST: the caller's buffer is free once the upload is done; the build itself stays queued (every later
    // call on this context is ordered behind it on the stream)
    if (mem == PCP_HOST && n > 0) PCP_CUDA(cudaEventSynchronize(ctx->ev_copy));
    return PCP_OK;
}

int pcp_index_attach_f64(pcp_index *ix, const double *xyz, int64_t n, int64_t stride, int mem) {
    if (!ix) return fail(PCP_ERR_INVALID, "pcp_index_attach_f64: index is NULL");
    if (bad_mem(mem)) return fail(PCP_ERR_INVALID, "pcp_index_attach_f64: mem must be PCP_HOST or PCP_DEVICE");
    if (n != ix->n) return fail(PCP_ERR_INVALID, "pcp_index_attach_f64: row count differs from the indexed cloud");
    if (stride < 3) return fail(PCP_ERR_INVALID, "pcp_index_attach_f64: stride_doubles must be >= 3");
    if (n > 0 && !xyz) return fail(PCP_ERR_INVALID, "pcp_index_attach_f64: xyz is NULL");
    pcp_ctx *ctx = ix->ctx;
    DeviceGuard g(ctx->device);
    std::vector<double> rows((size_t)n * (size_t)stride), packed((size_t)n * 3);
    if (n > 0) {
        if (mem == PCP_DEVICE) {
            PCP_CUDA(cudaStreamSynchronize(ctx->stream));
            PCP_CUDA(cudaMemcpy(rows.data(), xyz, sizeof(double) * rows.size(), cudaMemcpyDeviceToHost));
            xyz = rows.data();
        }
        double mx = 0.0;
        for (int64_t i = 0; i < n; ++i)
            for (int a = 0; a < 3; ++a) {
                const double v = xyz[i * stride + a];
                packed[(size_t)i * 3 + a] = v;
                mx = std::max(mx, std::fabs(v));
            }
        ix->xyz64_max_abs = mx;
        PCP_TRY(ix->xyz64.ensure(sizeof(double) * packed.size()));
        PCP_CUDA(cudaMemcpyAsync(ix->xyz64.p, packed.data(), sizeof(double) * packed.size(), cudaMemcpyHostToDevice, ctx->stream));
        PCP_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    ix->xyz64_n = n;
    return PCP_OK;
}

int pcp_index_build(pcp_ctx *ctx, const float *xyz, int64_t n, int64_t stride, int mem, pcp_index **out) {
    if (!ctx || !out) return fail(PCP_ERR_INVALID, "pcp_index_build: ctx/out is NULL");
    *out = nullptr;
    PCP_TRY(check_cloud("pcp_index_build", xyz, n, stride, mem));
    pcp_index *ix = new (std::nothrow) pcp_index();
    if (!ix) return fail(PCP_ERR_OOM, "pcp_index_build: out of host memory");
    ix->ctx = ctx;
    int s = pcp_index_rebuild(ix, xyz, n, stride, mem);
    if (s != PCP_OK) {
        pcp_index_destroy(ix);
        return s;
    }
    *out = ix;
    return PCP_OK;
}

int pcp_index_destroy(pcp_index *ix) {
    if (!ix) return PCP_OK;
    DeviceGuard g(ix->ctx->device);
    cudaStreamSynchronize(ix->ctx->stream);
    ix->b.release();
    ix->src_copy.release();
    ix->xyz64.release();
    delete ix;
    return PCP_OK;
}

int64_t pcp_index_size(const pcp_index *ix) { return ix ? ix->n : 0; }

int pcp_index_stats(const pcp_index *ix, int64_t *n_buckets, int *n_levels) {
    if (!ix) return fail(PCP_ERR_INVALID, "pcp_index_stats: index is NULL");
    if (n_buckets) *n_buckets = ix->b.h_views.empty() ? 0 : ix->b.h_views[0].n_leaf;
    if (n_levels) *n_levels = ix->b.h_views.empty() ? 0 : ix->b.h_views[0].n_levels;
    return PCP_OK;
}

// ---------------------------------------------------------------------------------------------- kNN

static int check_k(int k) {
    if (k < 1) return fail(PCP_ERR_INVALID, "pcp_knn: k must be >= 1");
    if (k > PCP_MAX_K) return fail(PCP_ERR_UNSUPPORTED, "pcp_knn: k > PCP_MAX_K (32) is not implemented");
    return PCP_OK;
}

int pcp_knn(pcp_index *ix, const float *queries, int64_t nq, int64_t stride, int k, int mem, int32_t *idx,
            float *dist, int32_t *count) {
    if (!ix) return fail(PCP_ERR_INVALID, "pcp_knn: index is NULL");
    PCP_TRY(check_cloud("pcp_knn", queries, nq, stride, mem));
    PCP_TRY(check_k(k));
    if (nq > 0 && (!idx || !dist)) return fail(PCP_ERR_INVALID, "pcp_knn: idx/dist is NULL");
    if (nq == 0) return PCP_OK;
    pcp_ctx *ctx = ix->ctx;
    DeviceGuard g(ctx->device);
    const float *d_q = queries;
    int32_t *d_idx = idx, *d_cnt = count;
    float *d_dist = dist;
    const size_t nres = (size_t)nq * (size_t)k;
    bool side_ok = false;      // queries were uploaded on the side stream: their regrouping may run there too
    if (mem == PCP_HOST) {
        PCP_TRY(upload_side(ctx, 1, ctx->scratch[6], queries, sizeof(float) * (size_t)nq * (size_t)stride, &side_ok));
        d_q = ctx->scratch[6].as<float>();
        // one result slab: idx | dist | count
        PCP_TRY(ctx->scratch[7].ensure(nres * 8 + sizeof(int32_t) * (size_t)nq));
        d_idx = ctx->scratch[7].as<int32_t>();
        d_dist = reinterpret_cast<float *>(d_idx + nres);
        d_cnt = reinterpret_cast<int32_t *>(d_dist + nres);
    }
    const uint32_t *d_perm = nullptr;
    PCP_TRY(knn_device(ix, d_q, nq, stride, d_perm, k, d_idx, d_dist, d_cnt, side_ok));
    if (mem == PCP_HOST) {
        PCP_TRY(download_knn(ctx, idx, dist, count, d_idx, nres, (size_t)nq));
        PCP_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    return PCP_OK;
}

int pcp_knn_bruteforce(pcp_index *ix, const float *queries, int64_t nq, int64_t stride, int k, int mem, int32_t *idx,
                       float *dist, int32_t *count) {
    if (!ix) return fail(PCP_ERR_INVALID, "pcp_knn_bruteforce: index is NULL");
    PCP_TRY(check_cloud("pcp_knn_bruteforce", queries, nq, stride, mem));
    PCP_TRY(check_k(k));
    if (nq > 0 && (!idx || !dist)) return fail(PCP_ERR_INVALID, "pcp_knn_bruteforce: idx/dist is NULL");
    if (nq == 0) return PCP_OK;
    pcp_ctx *ctx = ix->ctx;
    DeviceGuard g(ctx->device);
    const float *d_q = queries;
    int32_t *d_idx = idx, *d_cnt = count;
    float *d_dist = dist;
    const size_t nres = (size_t)nq * (size_t)k;
    if (mem == PCP_HOST) {
        PCP_TRY(upload(ctx, 1, ctx->scratch[6], queries, sizeof(float) * (size_t)nq * (size_t)stride));
        d_q = ctx->scratch[6].as<float>();
        PCP_TRY(ctx->scratch[7].ensure(nres * 8 + sizeof(int32_t) * (size_t)nq));
        d_idx = ctx->scratch[7].as<int32_t>();
        d_dist = reinterpret_cast<float *>(d_idx + nres);
        d_cnt = reinterpret_cast<int32_t *>(d_dist + nres);
    }
    PCP_TRY(knn_brute_device(ix, d_q, nq, stride, k, d_idx, d_dist, d_cnt));
    if (mem == PCP_HOST) {
        PCP_TRY(download_knn(ctx, idx, dist, count, d_idx, nres, (size_t)nq));
        PCP_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    return PCP_OK;
}

int pcp_radius_count_bruteforce(pcp_index *ix, const float *queries, int64_t nq, int64_t stride, float radius, int mem,
                                int64_t *counts) {
    if (!ix) return fail(PCP_ERR_INVALID, "pcp_radius_count_bruteforce: index is NULL");
    PCP_TRY(check_cloud("pcp_radius_count_bruteforce", queries, nq, stride, mem));
    if (nq > 0 && !counts) return fail(PCP_ERR_INVALID, "pcp_radius_count_bruteforce: counts is NULL");
    if (std::isnan(radius)) return fail(PCP_ERR_INVALID, "pcp_radius_count_bruteforce: radius is NaN");
    if (nq == 0) return PCP_OK;
    pcp_ctx *ctx = ix->ctx;
    DeviceGuard g(ctx->device);
    const float *d_q = queries;
    int64_t *d_cnt = counts;
    if (mem == PCP_HOST) {
        PCP_TRY(upload(ctx, 1, ctx->scratch[6], queries, sizeof(float) * (size_t)nq * (size_t)stride));
        d_q = ctx->scratch[6].as<float>();
        PCP_TRY(ctx->scratch[7].ensure(sizeof(int64_t) * (size_t)nq));
        d_cnt = ctx->scratch[7].as<int64_t>();
    }
    PCP_TRY(radius_count_brute_device(ix, d_q, nq, stride, exact_d2_threshold(radius), reinterpret_cast<long long *>(d_cnt)));
    if (mem == PCP_HOST) {
        PCP_TRY(download(ctx, counts, d_cnt, sizeof(int64_t) * (size_t)nq));
        PCP_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    return PCP_OK;
}

int pcp_knn_self(pcp_index *ix, int k, int mem, int32_t *idx, float *dist, int32_t *count) {
    if (!ix) return fail(PCP_ERR_INVALID, "pcp_knn_self: index is NULL");
    if (bad_mem(mem)) return fail(PCP_ERR_INVALID, "pcp_knn_self: mem must be PCP_HOST or PCP_DEVICE");
    PCP_TRY(check_k(k));
    const int64_t nq = ix->n;
    if (nq == 0) return PCP_OK;
    if (!idx || !dist) return fail(PCP_ERR_INVALID, "pcp_knn_self: idx/dist is NULL");
    pcp_ctx *ctx = ix->ctx;
    DeviceGuard g(ctx->device);
    int32_t *d_idx = idx, *d_cnt = count;
    float *d_dist = dist;
    const size_t nres = (size_t)nq * (size_t)k;
    if (mem == PCP_HOST) {
        PCP_TRY(ctx->scratch[7].ensure(nres * 8 + sizeof(int32_t) * (size_t)nq));
        d_idx = ctx->scratch[7].as<int32_t>();
        d_dist = reinterpret_cast<float *>(d_idx + nres);
        d_cnt = reinterpret_cast<int32_t *>(d_dist + nres);
    }
    PCP_TRY(knn_self_device(ix, k, d_idx, d_dist, d_cnt));
    if (mem == PCP_HOST) {
        PCP_TRY(download_knn(ctx, idx, dist, count, d_idx, nres, (size_t)nq));
        PCP_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    return PCP_OK;
}

// ---------------------------------------------------------------------------------------------- radius

int pcp_radius_count(pcp_index *ix, const float *queries, int64_t nq, int64_t stride, float radius, int mem,
                     int64_t *offsets) {
    if (!ix) return fail(PCP_ERR_INVALID, "pcp_radius_count: index is NULL");
    PCP_TRY(check_cloud("pcp_radius_count", queries, nq, stride, mem));
    if (!offsets) return fail(PCP_ERR_INVALID, "pcp_radius_count: offsets is NULL");
    if (std::isnan(radius)) return fail(PCP_ERR_INVALID, "pcp_radius_count: radius is NaN");
    pcp_ctx *ctx = ix->ctx;
    DeviceGuard g(ctx->device);
    const float *d_q = queries;
    int64_t *d_off = offsets;
    if (mem == PCP_HOST) {
        PCP_TRY(upload(ctx, 1, ctx->scratch[6], queries, sizeof(float) * (size_t)nq * (size_t)stride));
        d_q = ctx->scratch[6].as<float>();
        PCP_TRY(ctx->scratch[7].ensure(sizeof(int64_t) * (size_t)(nq + 1)));
        d_off = ctx->scratch[7].as<int64_t>();
    }
    PCP_TRY(radius_count_device(ix, d_q, nq, stride, radius, d_off));
    if (mem == PCP_HOST) {
        PCP_TRY(download(ctx, offsets, d_off, sizeof(int64_t) * (size_t)(nq + 1)));
        PCP_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    return PCP_OK;
}

int pcp_radius_fill(pcp_index *ix, const float *queries, int64_t nq, int64_t stride, float radius, int mem,
                    const int64_t *offsets, int32_t *idx, float *dist, int sorted) {
    if (!ix) return fail(PCP_ERR_INVALID, "pcp_radius_fill: index is NULL");
    PCP_TRY(check_cloud("pcp_radius_fill", queries, nq, stride, mem));
    if (!offsets) return fail(PCP_ERR_INVALID, "pcp_radius_fill: offsets is NULL");
    if (std::isnan(radius)) return fail(PCP_ERR_INVALID, "pcp_radius_fill: radius is NaN");
    if (nq == 0) return PCP_OK;
    pcp_ctx *ctx = ix->ctx;
    DeviceGuard g(ctx->device);
    if (mem == PCP_DEVICE) return radius_fill_device(ix, queries, nq, stride, radius, offsets, idx, dist, sorted);
    const int64_t m = offsets[nq];
    if (m < 0) return fail(PCP_ERR_INVALID, "pcp_radius_fill: offsets[nq] is negative");
    if (m > 0 && (!idx || !dist)) return fail(PCP_ERR_INVALID, "pcp_radius_fill: idx/dist is NULL");
    PCP_TRY(upload(ctx, 1, ctx->scratch[6], queries, sizeof(float) * (size_t)nq * (size_t)stride));
    PCP_TRY(upload(ctx, 2, ctx->scratch[7], offsets, sizeof(int64_t) * (size_t)(nq + 1)));
    PCP_TRY(ctx->scratch[2].ensure((size_t)m * 8 + 16));
    int32_t *d_idx = ctx->scratch[2].as<int32_t>();
    float *d_dist = reinterpret_cast<float *>(d_idx + m);
    PCP_TRY(radius_fill_device(ix, ctx->scratch[6].as<float>(), nq, stride, radius, ctx->scratch[7].as<int64_t>(),
                               d_idx, d_dist, sorted));
    PCP_TRY(download(ctx, idx, d_idx, (size_t)m * 4));
    PCP_TRY(download(ctx, dist, d_dist, (size_t)m * 4));
    PCP_CUDA(cudaStreamSynchronize(ctx->stream));
    return PCP_OK;
}

// ---------------------------------------------------------------------------------------------- ICP

int pcp_icp_default_params(pcp_icp_params *p) {
    if (!p) return fail(PCP_ERR_INVALID, "pcp_icp_default_params: p is NULL");
    p->max_iteration = 100;        // main.py:98
    p->R_diff_thres = 0.5;         // main.py:101
    p->t_diff_thres = 0.5;         // main.py:102
    p->dist_thres = 5.0;           // main.py:103
    p->first_iter_broadcast = 1;   // main.py:150 as built
    return PCP_OK;
}

int pcp_icp_batch(pcp_ctx *ctx, const pcp_icp_params *params, int64_t n_pairs, const float *src_xyz,
                  const int64_t *src_offsets, const float *tgt_xyz, const int64_t *tgt_offsets, int64_t stride, int mem,
                  const double *T_init, double *T_last, double *T_total, int32_t *iterations, int32_t *status,
                  double *diff_log) {
    if (!ctx) return fail(PCP_ERR_INVALID, "pcp_icp_batch: ctx is NULL");
    if (n_pairs < 0) return fail(PCP_ERR_INVALID, "pcp_icp_batch: negative pair count");
    if (n_pairs == 0) return PCP_OK;
    if (!src_offsets || !tgt_offsets) return fail(PCP_ERR_INVALID, "pcp_icp_batch: offsets are NULL");
    if (src_offsets[0] != 0 || tgt_offsets[0] != 0) return fail(PCP_ERR_INVALID, "pcp_icp_batch: offsets[0] must be 0");
    const int64_t ns = src_offsets[n_pairs], nt = tgt_offsets[n_pairs];
    PCP_TRY(check_cloud("pcp_icp_batch", src_xyz, ns, stride, mem));
    PCP_TRY(check_cloud("pcp_icp_batch", tgt_xyz, nt, stride, mem));
    pcp_icp_params prm;
    pcp_icp_default_params(&prm);
    if (params) prm = *params;
    if (prm.max_iteration < 0) return fail(PCP_ERR_INVALID, "pcp_icp_batch: max_iteration < 0");
    if (std::isnan(prm.dist_thres) || std::isnan(prm.R_diff_thres) || std::isnan(prm.t_diff_thres))
        return fail(PCP_ERR_INVALID, "pcp_icp_batch: NaN threshold");
    DeviceGuard g(ctx->device);
    const float *d_src = src_xyz, *d_tgt = tgt_xyz;
    cudaEvent_t src_ready = nullptr;
    if (mem == PCP_HOST && host_is_pinned(src_xyz) && host_is_pinned(tgt_xyz) &&
        !(getenv("PCP_ICP_PIPELINE") && atoi(getenv("PCP_ICP_PIPELINE")) == 0)) {
        // Page-locked clouds: the batch is cut into runs of pairs of ~kIcpChunkPoints points; the side stream copies
        // run c + 1 into the second staging set while the main stream registers run c, so that all but the first
        // copy and the last registration overlap (C4: 1.97 GB over PCIe against ~31 ms of device work).
        if (!ctx->ev_icp_up[0]) {
            for (int i = 0; i < 4; ++i) PCP_CUDA(cudaEventCreateWithFlags(&ctx->ev_icp_up[i], cudaEventDisableTiming));
        }
        std::vector<int64_t> cuts(1, 0);
        for (int64_t p = 0; p < n_pairs;) {
            int64_t q = p + 1;
            while (q < n_pairs && (src_offsets[q + 1] - src_offsets[p]) + (tgt_offsets[q + 1] - tgt_offsets[p]) <= kIcpChunkPoints) ++q;
            cuts.push_back(q);
            p = q;
        }
        const size_t row = sizeof(float) * (size_t)stride;
        auto enqueue_copy = [&](size_t c) -> int {
            const int64_t p0 = cuts[c], p1 = cuts[c + 1];
            const int set = (int)(c & 1);
            const size_t tb = row * (size_t)(tgt_offsets[p1] - tgt_offsets[p0]), sb = row * (size_t)(src_offsets[p1] - src_offsets[p0]);
            PCP_TRY(ctx->icp_up[2 * set].ensure(tb));
            PCP_TRY(ctx->icp_up[2 * set + 1].ensure(sb));
            if (tb) PCP_CUDA(cudaMemcpyAsync(ctx->icp_up[2 * set].p, tgt_xyz + tgt_offsets[p0] * stride, tb, cudaMemcpyHostToDevice, ctx->copy_stream));
            PCP_CUDA(cudaEventRecord(ctx->ev_icp_up[2 * set], ctx->copy_stream));
            if (sb) PCP_CUDA(cudaMemcpyAsync(ctx->icp_up[2 * set + 1].p, src_xyz + src_offsets[p0] * stride, sb, cudaMemcpyHostToDevice, ctx->copy_stream));
            PCP_CUDA(cudaEventRecord(ctx->ev_icp_up[2 * set + 1], ctx->copy_stream));
            return PCP_OK;
        };
        int64_t launches = 0, syncs = 0, enqueued = 0;
        std::vector<int64_t> so, to;
        int64_t max_run = 0;
        for (size_t c = 0; c + 1 < cuts.size(); ++c) max_run = std::max(max_run, cuts[c + 1] - cuts[c]);
        PCP_TRY(ctx->small_stage.ensure((size_t)max_run * (2 * sizeof(IndexView) + 512) + 4096));   // (previous calls are synchronised)
        ctx->h2d_by_kernel = true;
        const size_t n_chunks = cuts.size() - 1;
        int rc = enqueue_copy(0);
        for (size_t c = 0; c < n_chunks && rc == PCP_OK; ++c) {
            // (set (c + 1) & 1 was last read by run c - 1, whose call has returned: synchronised)
            if (c + 1 < n_chunks) rc = enqueue_copy(c + 1);
            if (rc != PCP_OK) break;
            const int64_t p0 = cuts[c], p1 = cuts[c + 1], np = p1 - p0;
            const int set = (int)(c & 1);
            ctx->small_used = 0;                               // (run c - 1 has been waited for)
            so.resize((size_t)np + 1);
            to.resize((size_t)np + 1);
            for (int64_t i = 0; i <= np; ++i) { so[(size_t)i] = src_offsets[p0 + i] - src_offsets[p0]; to[(size_t)i] = tgt_offsets[p0 + i] - tgt_offsets[p0]; }
            if (cudaStreamWaitEvent(ctx->stream, ctx->ev_icp_up[2 * set], 0) != cudaSuccess) { rc = fail(PCP_ERR_CUDA, "pcp_icp_batch: cudaStreamWaitEvent failed"); break; }
            rc = icp_batch_device(ctx, prm, np, ctx->icp_up[2 * set + 1].as<float>(), so.data(), ctx->icp_up[2 * set].as<float>(), to.data(),
                                  stride, T_init ? T_init + p0 * 16 : nullptr, T_last ? T_last + p0 * 16 : nullptr,
                                  T_total ? T_total + p0 * 16 : nullptr, iterations ? iterations + p0 : nullptr,
                                  status ? status + p0 : nullptr, diff_log ? diff_log + p0 * 2 * (int64_t)prm.max_iteration : nullptr,
                                  ctx->ev_icp_up[2 * set + 1]);
            launches += ctx->icp_last_launches;
            syncs += ctx->icp_last_syncs;
            enqueued = std::max(enqueued, ctx->icp_last_enqueued);
        }
        ctx->h2d_by_kernel = false;
        cudaStreamSynchronize(ctx->copy_stream);               // the caller's buffers are free on return
        ctx->icp_last_launches = launches;
        ctx->icp_last_syncs = syncs;
        ctx->icp_last_enqueued = enqueued;
        return rc;
    }
    if (mem == PCP_HOST) {
        // Targets first, on the main stream: their bucket build starts as soon as they are on the device.  The
        // sources follow on the side stream (page-locked input), behind the target copy and behind whatever the
        // main stream still runs on the scratch buffer, and the main stream only waits for them before the
        // source build -- the target build runs beside the second half of the transfer.
        PCP_TRY(upload(ctx, 1, ctx->scratch[6], tgt_xyz, sizeof(float) * (size_t)nt * (size_t)stride));
        PCP_CUDA(cudaEventRecord(ctx->ev_search_end, ctx->stream));
        PCP_CUDA(cudaStreamWaitEvent(ctx->copy_stream, ctx->ev_search_end, 0));
        bool src_side = false;
        PCP_TRY(upload_side(ctx, 0, ctx->scratch[5], src_xyz, sizeof(float) * (size_t)ns * (size_t)stride, &src_side, false));
        if (src_side) src_ready = ctx->ev_copy;
        d_src = ctx->scratch[5].as<float>();
        d_tgt = ctx->scratch[6].as<float>();
    }
    const int rc = icp_batch_device(ctx, prm, n_pairs, d_src, src_offsets, d_tgt, tgt_offsets, stride, T_init, T_last,
                                    T_total, iterations, status, diff_log, src_ready);
    if (rc != PCP_OK && src_ready) cudaStreamSynchronize(ctx->copy_stream);   // the caller's buffer is free on return
    return rc;
}

// ---------------------------------------------------------------------------------------------- CUDA graphs

struct pcp_graph {
    pcp_ctx *ctx = nullptr;
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    size_t nodes = 0;
};

int pcp_graph_begin(pcp_ctx *ctx) {
    if (!ctx) return fail(PCP_ERR_INVALID, "pcp_graph_begin: ctx is NULL");
    if (ctx->capturing) return fail(PCP_ERR_INVALID, "pcp_graph_begin: a capture is already open on this context");
    DeviceGuard g(ctx->device);
    PCP_CUDA(cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeRelaxed));
    ctx->capturing = true;
    ctx->cap_has_bbox = false;
    capture_note_begin();
    return PCP_OK;
}

int pcp_graph_end(pcp_ctx *ctx, pcp_graph **out) {
    if (!ctx || !out) return fail(PCP_ERR_INVALID, "pcp_graph_end: ctx/out is NULL");
    if (!ctx->capturing) return fail(PCP_ERR_INVALID, "pcp_graph_end: no capture is open on this context");
    *out = nullptr;
    DeviceGuard g(ctx->device);
    ctx->capturing = false;
    const bool grew = capture_note_grew();
    cudaGraph_t graph = nullptr;
    cudaError_t e = cudaStreamEndCapture(ctx->stream, &graph);
    if (e != cudaSuccess || !graph) {
        cudaGetLastError();
        return fail(PCP_ERR_CUDA, std::string("pcp_graph_end: capture failed (") + cudaGetErrorString(e) +
                                      "): only PCP_DEVICE calls of this context may be captured");
    }
    if (grew) {
        cudaGraphDestroy(graph);
        return fail(PCP_ERR_INVALID, "pcp_graph_end: a library buffer had to grow during the capture; run the same "
                                     "calls once before capturing them");
    }
    pcp_graph *pg = new (std::nothrow) pcp_graph();
    if (!pg) { cudaGraphDestroy(graph); return fail(PCP_ERR_OOM, "pcp_graph_end: out of host memory"); }
    pg->ctx = ctx;
    pg->graph = graph;
    cudaGraphGetNodes(graph, nullptr, &pg->nodes);
    e = cudaGraphInstantiate(&pg->exec, graph, 0);
    if (e != cudaSuccess) {
        cudaGraphDestroy(graph);
        delete pg;
        return cuda_fail(e, "cudaGraphInstantiate", __FILE__, __LINE__);
    }
    *out = pg;
    return PCP_OK;
}

int pcp_graph_launch(pcp_graph *g) {
    if (!g) return fail(PCP_ERR_INVALID, "pcp_graph_launch: graph is NULL");
    DeviceGuard dg(g->ctx->device);
    PCP_CUDA(cudaGraphLaunch(g->exec, g->ctx->stream));
    g->ctx->launches += (int64_t)g->nodes;
    return PCP_OK;
}

int64_t pcp_graph_nodes(const pcp_graph *g) { return g ? (int64_t)g->nodes : 0; }

int pcp_graph_destroy(pcp_graph *g) {
    if (!g) return PCP_OK;
    DeviceGuard dg(g->ctx->device);
    cudaStreamSynchronize(g->ctx->stream);
    if (g->exec) cudaGraphExecDestroy(g->exec);
    if (g->graph) cudaGraphDestroy(g->graph);
    delete g;
    return PCP_OK;
}

int pcp_icp_last_stats(pcp_ctx *ctx, int64_t *launches, int64_t *host_syncs, int64_t *iterations_enqueued) {
    if (!ctx) return fail(PCP_ERR_INVALID, "pcp_icp_last_stats: ctx is NULL");
    if (launches) *launches = ctx->icp_last_launches;
    if (host_syncs) *host_syncs = ctx->icp_last_syncs;
    if (iterations_enqueued) *iterations_enqueued = ctx->icp_last_enqueued;
    return PCP_OK;
}

int pcp_procrustes(pcp_ctx *ctx, const double *A, const double *B, int64_t n, int mem, double *R, double *t, double *cost) {
    if (!ctx) return fail(PCP_ERR_INVALID, "pcp_procrustes: ctx is NULL");
    if (n < 1 || !A || !B) return fail(PCP_ERR_INVALID, "pcp_procrustes: need n >= 1 correspondences and non-NULL A, B");
    if (bad_mem(mem)) return fail(PCP_ERR_INVALID, "pcp_procrustes: mem must be PCP_HOST or PCP_DEVICE");
    DeviceGuard g(ctx->device);
    const double *d_A = A, *d_B = B;
    if (mem == PCP_HOST) {
        PCP_TRY(upload(ctx, 0, ctx->scratch[5], A, sizeof(double) * 3 * (size_t)n));
        PCP_TRY(upload(ctx, 1, ctx->scratch[6], B, sizeof(double) * 3 * (size_t)n));
        d_A = ctx->scratch[5].as<double>();
        d_B = ctx->scratch[6].as<double>();
    }
    return procrustes_device(ctx, d_A, d_B, n, R, t, cost);
}

// ---------------------------------------------------------------------------------------------- ICP front end

int pcp_voxel_filter(pcp_ctx *ctx, const float *xyz, int64_t n, int64_t stride, double leaf_size, int mode,
                     uint64_t seed, int mem, int64_t *n_out) {
    if (!ctx || !n_out) return fail(PCP_ERR_INVALID, "pcp_voxel_filter: ctx/n_out is NULL");
    *n_out = 0;
    PCP_TRY(check_cloud("pcp_voxel_filter", xyz, n, stride, mem));
    if (n == 0) return fail(PCP_ERR_INVALID, "pcp_voxel_filter: empty cloud (the reference raises on np.max of an empty array, voxel_filter.py:20)");
    if (n > 0x7fffffffLL) return fail(PCP_ERR_UNSUPPORTED, "pcp_voxel_filter: more than 2^31-1 points");
    if (!(leaf_size > 0.0) || !std::isfinite(leaf_size)) return fail(PCP_ERR_INVALID, "pcp_voxel_filter: leaf_size must be finite and > 0");
    if (mode != PCP_VOXEL_CENTROID && mode != PCP_VOXEL_RANDOM) return fail(PCP_ERR_INVALID, "pcp_voxel_filter: mode must be PCP_VOXEL_CENTROID or PCP_VOXEL_RANDOM");
    DeviceGuard g(ctx->device);
    const float *d_xyz = xyz;
    if (mem == PCP_HOST) {
        PCP_TRY(upload(ctx, 0, ctx->scratch[6], xyz, sizeof(float) * (size_t)n * (size_t)stride));
        d_xyz = ctx->scratch[6].as<float>();
    }
    return voxel_filter_device(ctx, d_xyz, n, stride, leaf_size, mode, seed, n_out);
}

int pcp_voxel_filter_read(pcp_ctx *ctx, int mem, double *out_xyz) {
    if (!ctx) return fail(PCP_ERR_INVALID, "pcp_voxel_filter_read: ctx is NULL");
    if (bad_mem(mem)) return fail(PCP_ERR_INVALID, "pcp_voxel_filter_read: mem must be PCP_HOST or PCP_DEVICE");
    if (ctx->voxel_n_out < 0) return fail(PCP_ERR_INVALID, "pcp_voxel_filter_read: no pcp_voxel_filter result on this context");
    if (ctx->voxel_n_out == 0) return PCP_OK;
    if (!out_xyz) return fail(PCP_ERR_INVALID, "pcp_voxel_filter_read: out_xyz is NULL");
    DeviceGuard g(ctx->device);
    const size_t bytes = sizeof(double) * 3 * (size_t)ctx->voxel_n_out;
    PCP_CUDA(cudaMemcpyAsync(out_xyz, ctx->voxel_out.p, bytes, mem == PCP_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice, ctx->stream));
    if (mem == PCP_HOST) PCP_CUDA(cudaStreamSynchronize(ctx->stream));
    return PCP_OK;
}

int pcp_normals(pcp_index *ix, int k, int mem, double *normals, double *eigenvalues) {
    if (!ix) return fail(PCP_ERR_INVALID, "pcp_normals: index is NULL");
    if (bad_mem(mem)) return fail(PCP_ERR_INVALID, "pcp_normals: mem must be PCP_HOST or PCP_DEVICE");
    PCP_TRY(check_k(k));
    if (k < 2) return fail(PCP_ERR_INVALID, "pcp_normals: k must be >= 2 (np.cov of one sample is undefined)");
    const int64_t n = ix->n;
    if (n == 0) return PCP_OK;
    if (!normals) return fail(PCP_ERR_INVALID, "pcp_normals: normals is NULL");
    pcp_ctx *ctx = ix->ctx;
    DeviceGuard g(ctx->device);
    double *d_nrm = normals, *d_ev = eigenvalues;
    const size_t row = sizeof(double) * 3 * (size_t)n;
    if (mem == PCP_HOST) {
        PCP_TRY(ctx->scratch[7].ensure(row * (eigenvalues ? 2 : 1)));
        d_nrm = ctx->scratch[7].as<double>();
        d_ev = eigenvalues ? d_nrm + 3 * (size_t)n : nullptr;
    }
    PCP_TRY(normals_device(ix, k, d_nrm, d_ev));
    if (mem == PCP_HOST) {
        PCP_TRY(download(ctx, normals, d_nrm, row));
        if (eigenvalues) PCP_TRY(download(ctx, eigenvalues, d_ev, row));
        PCP_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    return PCP_OK;
}

int pcp_pca(pcp_ctx *ctx, const float *xyz, int64_t n, int64_t stride, int mem, double *eigenvalues, double *eigenvectors) {
    if (!ctx || !eigenvalues || !eigenvectors) return fail(PCP_ERR_INVALID, "pcp_pca: ctx/eigenvalues/eigenvectors is NULL");
    PCP_TRY(check_cloud("pcp_pca", xyz, n, stride, mem));
    if (n < 2) return fail(PCP_ERR_INVALID, "pcp_pca: at least 2 points are needed (np.cov divides by n - 1)");
    DeviceGuard g(ctx->device);
    const float *d_xyz = xyz;
    if (mem == PCP_HOST) {
        PCP_TRY(upload(ctx, 0, ctx->scratch[6], xyz, sizeof(float) * (size_t)n * (size_t)stride));
        d_xyz = ctx->scratch[6].as<float>();
    }
    return pca_device(ctx, d_xyz, n, stride, eigenvalues, eigenvectors);
}

int pcp_iss_keypoints(pcp_index *ix, double radius, double lambda21, double lambda32, double non_max_radius, int iss_count,
                      int32_t *keypoints, int32_t *n_keypoints, double *lambda3) {
    if (!ix) return fail(PCP_ERR_INVALID, "pcp_iss_keypoints: index is NULL");
    if (!keypoints || !n_keypoints) return fail(PCP_ERR_INVALID, "pcp_iss_keypoints: keypoints/n_keypoints is NULL");
    if (!(radius > 0.0) || !(non_max_radius >= 0.0) || std::isnan(lambda21) || std::isnan(lambda32) || iss_count < 0)
        return fail(PCP_ERR_INVALID, "pcp_iss_keypoints: radius must be > 0, non_max_radius >= 0, iss_count >= 0");
    *n_keypoints = 0;
    const int64_t n = ix->n;
    if (n == 0) return PCP_OK;
    pcp_ctx *ctx = ix->ctx;
    DeviceGuard g(ctx->device);
    double *d_l3 = nullptr;
    if (lambda3) {
        PCP_TRY(ctx->scratch[7].ensure(sizeof(double) * (size_t)n));
        d_l3 = ctx->scratch[7].as<double>();
    }
    PCP_TRY(iss_device(ix, radius, lambda21, lambda32, non_max_radius, iss_count, keypoints, n_keypoints, d_l3));
    if (lambda3) {
        PCP_TRY(download(ctx, lambda3, d_l3, sizeof(double) * (size_t)n));
        PCP_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    return PCP_OK;
}

int pcp_normals_hybrid(pcp_index *ix, double radius, int max_nn, int mem, double *normals) {
    if (!ix) return fail(PCP_ERR_INVALID, "pcp_normals_hybrid: index is NULL");
    if (bad_mem(mem)) return fail(PCP_ERR_INVALID, "pcp_normals_hybrid: mem must be PCP_HOST or PCP_DEVICE");
    if (!(radius > 0.0) || max_nn < 1) return fail(PCP_ERR_INVALID, "pcp_normals_hybrid: radius > 0 and max_nn >= 1 required");
    const int64_t n = ix->n;
    if (n == 0) return PCP_OK;
    if (!normals) return fail(PCP_ERR_INVALID, "pcp_normals_hybrid: normals is NULL");
    pcp_ctx *ctx = ix->ctx;
    DeviceGuard g(ctx->device);
    double *d_n = normals;
    if (mem == PCP_HOST) {
        PCP_TRY(ctx->scratch[7].ensure(sizeof(double) * 3 * (size_t)n));
        d_n = ctx->scratch[7].as<double>();
    }
    PCP_TRY(normals_hybrid_device(ix, radius, max_nn, d_n));
    if (mem == PCP_HOST) {
        PCP_TRY(download(ctx, normals, d_n, sizeof(double) * 3 * (size_t)n));
        PCP_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    return PCP_OK;
}

int pcp_fpfh(pcp_index *ix, const double *normals, double radius, int max_nn, int mem, double *features) {
    if (!ix) return fail(PCP_ERR_INVALID, "pcp_fpfh: index is NULL");
    if (bad_mem(mem)) return fail(PCP_ERR_INVALID, "pcp_fpfh: mem must be PCP_HOST or PCP_DEVICE");
    if (!(radius > 0.0) || max_nn < 1) return fail(PCP_ERR_INVALID, "pcp_fpfh: radius > 0 and max_nn >= 1 required");
    const int64_t n = ix->n;
    if (n == 0) return PCP_OK;
    if (!normals || !features) return fail(PCP_ERR_INVALID, "pcp_fpfh: normals/features is NULL");
    pcp_ctx *ctx = ix->ctx;
    DeviceGuard g(ctx->device);
    const double *d_n = normals;
    double *d_f = features;
    if (mem == PCP_HOST) {
        PCP_TRY(upload(ctx, 1, ctx->scratch[6], normals, sizeof(double) * 3 * (size_t)n));
        PCP_TRY(ctx->scratch[7].ensure(sizeof(double) * 33 * (size_t)n));
        d_n = ctx->scratch[6].as<double>();
        d_f = ctx->scratch[7].as<double>();
    }
    PCP_TRY(fpfh_device(ix, d_n, radius, max_nn, d_f));
    if (mem == PCP_HOST) {
        PCP_TRY(download(ctx, features, d_f, sizeof(double) * 33 * (size_t)n));
        PCP_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    return PCP_OK;
}

int pcp_feature_match(pcp_ctx *ctx, const double *src, int64_t ns, const double *tgt, int64_t nt, int dim, int mem,
                      int32_t *nearest, double *dist2) {
    if (!ctx) return fail(PCP_ERR_INVALID, "pcp_feature_match: ctx is NULL");
    if (bad_mem(mem)) return fail(PCP_ERR_INVALID, "pcp_feature_match: mem must be PCP_HOST or PCP_DEVICE");
    if (ns < 0 || nt < 0 || dim < 1) return fail(PCP_ERR_INVALID, "pcp_feature_match: bad sizes");
    if (ns == 0) return PCP_OK;
    if (!src || !nearest || (nt > 0 && !tgt)) return fail(PCP_ERR_INVALID, "pcp_feature_match: NULL array");
    DeviceGuard g(ctx->device);
    const double *d_s = src, *d_t = tgt;
    int32_t *d_i = nearest;
    double *d_d = dist2;
    if (mem == PCP_HOST) {
        PCP_TRY(upload(ctx, 1, ctx->scratch[6], src, sizeof(double) * (size_t)ns * (size_t)dim));
        if (nt) PCP_TRY(upload(ctx, 2, ctx->scratch[5], tgt, sizeof(double) * (size_t)nt * (size_t)dim));
        else PCP_TRY(ctx->scratch[5].ensure(16));
        PCP_TRY(ctx->scratch[7].ensure((size_t)ns * 16));
        d_s = ctx->scratch[6].as<double>();
        d_t = ctx->scratch[5].as<double>();
        d_d = ctx->scratch[7].as<double>();
        d_i = reinterpret_cast<int32_t *>(d_d + ns);
    }
    PCP_TRY(feature_match_device(ctx, d_s, ns, d_t, nt, dim, d_i, d_d));
    if (mem == PCP_HOST) {
        PCP_TRY(download(ctx, nearest, d_i, sizeof(int32_t) * (size_t)ns));
        if (dist2) PCP_TRY(download(ctx, dist2, d_d, sizeof(double) * (size_t)ns));
        PCP_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    return PCP_OK;
}

int pcp_ransac_init(pcp_ctx *ctx, const pcp_ransac_params *prm, const float *src, int64_t ns, int64_t stride,
                    pcp_index *tgt, const int32_t *corr, int64_t n_corr, int mem, double *T, double *fitness,
                    double *rmse, int64_t *iterations, int64_t *evaluated) {
    if (!ctx || !prm || !tgt || !T) return fail(PCP_ERR_INVALID, "pcp_ransac_init: NULL argument");
    if (tgt->ctx != ctx) return fail(PCP_ERR_INVALID, "pcp_ransac_init: the target index belongs to another context");
    PCP_TRY(check_cloud("pcp_ransac_init", src, ns, stride, mem));
    if (n_corr < 0 || (n_corr > 0 && !corr)) return fail(PCP_ERR_INVALID, "pcp_ransac_init: bad correspondences");
    if (prm->ransac_n != 3) return fail(PCP_ERR_UNSUPPORTED, "pcp_ransac_init: ransac_n must be 3");
    if (!(prm->max_correspondence_distance > 0.0)) return fail(PCP_ERR_INVALID, "pcp_ransac_init: max_correspondence_distance must be > 0");
    DeviceGuard g(ctx->device);
    const float *d_s = src;
    const int32_t *d_c = corr;
    if (mem == PCP_HOST) {
        if (ns) PCP_TRY(upload(ctx, 1, ctx->scratch[6], src, sizeof(float) * (size_t)ns * (size_t)stride));
        if (n_corr) PCP_TRY(upload(ctx, 2, ctx->scratch[7], corr, sizeof(int32_t) * 2 * (size_t)n_corr));
        d_s = ctx->scratch[6].as<float>();
        d_c = ctx->scratch[7].as<int32_t>();
    }
    double fit = 0, rm = 0;
    int64_t it = 0, ev = 0;
    PCP_TRY(ransac_device(ctx, *prm, d_s, ns, stride, tgt, d_c, n_corr, T, &fit, &rm, &it, &ev));
    if (fitness) *fitness = fit;
    if (rmse) *rmse = rm;
    if (iterations) *iterations = it;
    if (evaluated) *evaluated = ev;
    return PCP_OK;
}

int pcp_ball_query(pcp_ctx *ctx, int b, int n, int m, float radius, int nsample, const float *new_xyz, const float *xyz,
                   int mem, int32_t *idx) {
    if (!ctx) return fail(PCP_ERR_INVALID, "pcp_ball_query: ctx is NULL");
    if (b < 0 || n < 0 || m < 0 || nsample < 0) return fail(PCP_ERR_INVALID, "pcp_ball_query: negative size");
    if (bad_mem(mem)) return fail(PCP_ERR_INVALID, "pcp_ball_query: mem must be PCP_HOST or PCP_DEVICE");
    const size_t nc = (size_t)b * (size_t)m * 3, np = (size_t)b * (size_t)n * 3, ni = (size_t)b * (size_t)m * (size_t)nsample;
    if ((nc && !new_xyz) || (np && !xyz) || (ni && !idx)) return fail(PCP_ERR_INVALID, "pcp_ball_query: NULL array");
    if (ni == 0) return PCP_OK;
    DeviceGuard g(ctx->device);
    const float *d_c = new_xyz, *d_p = xyz;
    int32_t *d_i = idx;
    if (mem == PCP_HOST) {
        PCP_TRY(upload(ctx, 1, ctx->scratch[6], new_xyz, sizeof(float) * nc));
        if (np) PCP_TRY(upload(ctx, 2, ctx->scratch[5], xyz, sizeof(float) * np));
        else PCP_TRY(ctx->scratch[5].ensure(16));
        PCP_TRY(ctx->scratch[7].ensure(sizeof(int32_t) * ni));
        d_c = ctx->scratch[6].as<float>();
        d_p = ctx->scratch[5].as<float>();
        d_i = ctx->scratch[7].as<int32_t>();
    }
    PCP_TRY(ball_query_device(ctx, b, n, m, radius, nsample, d_c, d_p, d_i));
    if (mem == PCP_HOST) {
        PCP_TRY(download(ctx, idx, d_i, sizeof(int32_t) * ni));
        PCP_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    return PCP_OK;
}

int pcp_three_nn(pcp_ctx *ctx, int b, int n, int m, const float *unknown, const float *known, int mem, float *dist2,
                 int32_t *idx) {
    if (!ctx) return fail(PCP_ERR_INVALID, "pcp_three_nn: ctx is NULL");
    if (b < 0 || n < 0 || m < 0) return fail(PCP_ERR_INVALID, "pcp_three_nn: negative size");
    if (bad_mem(mem)) return fail(PCP_ERR_INVALID, "pcp_three_nn: mem must be PCP_HOST or PCP_DEVICE");
    const size_t nu = (size_t)b * (size_t)n * 3, nk = (size_t)b * (size_t)m * 3;
    if ((nu && (!unknown || !dist2 || !idx)) || (nk && !known)) return fail(PCP_ERR_INVALID, "pcp_three_nn: NULL array");
    if (nu == 0) return PCP_OK;
    DeviceGuard g(ctx->device);
    const float *d_u = unknown, *d_k = known;
    float *d_d = dist2;
    int32_t *d_i = idx;
    if (mem == PCP_HOST) {
        PCP_TRY(upload(ctx, 1, ctx->scratch[6], unknown, sizeof(float) * nu));
        if (nk) PCP_TRY(upload(ctx, 2, ctx->scratch[5], known, sizeof(float) * nk));
        else PCP_TRY(ctx->scratch[5].ensure(16));
        PCP_TRY(ctx->scratch[7].ensure(nu * 8));
        d_u = ctx->scratch[6].as<float>();
        d_k = ctx->scratch[5].as<float>();
        d_d = ctx->scratch[7].as<float>();
        d_i = reinterpret_cast<int32_t *>(d_d + nu);
    }
    PCP_TRY(three_nn_device(ctx, b, n, m, d_u, d_k, d_d, d_i));
    if (mem == PCP_HOST) {
        PCP_TRY(download(ctx, dist2, d_d, sizeof(float) * nu));
        PCP_TRY(download(ctx, idx, d_i, sizeof(int32_t) * nu));
        PCP_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    return PCP_OK;
}

int pcp_neighbor_counts_f64(pcp_index *ix, double radius, int64_t *counts) {
    if (!ix) return fail(PCP_ERR_INVALID, "pcp_neighbor_counts_f64: index is NULL");
    if (!(radius >= 0.0)) return fail(PCP_ERR_INVALID, "pcp_neighbor_counts_f64: radius must be >= 0");
    const int64_t n = ix->n;
    if (n == 0) return PCP_OK;
    if (!counts) return fail(PCP_ERR_INVALID, "pcp_neighbor_counts_f64: counts is NULL");
    pcp_ctx *ctx = ix->ctx;
    DeviceGuard g(ctx->device);
    PCP_TRY(ctx->scratch[7].ensure(sizeof(int64_t) * (size_t)(n + 1)));
    int64_t *d_off = ctx->scratch[7].as<int64_t>();
    PCP_TRY(radius_count_device(ix, ix->built_from, n, ix->built_stride, (float)radius, d_off, true, radius));
    std::vector<int64_t> off((size_t)n + 1);
    PCP_CUDA(cudaMemcpyAsync(off.data(), d_off, sizeof(int64_t) * off.size(), cudaMemcpyDeviceToHost, ctx->stream));
    PCP_CUDA(cudaStreamSynchronize(ctx->stream));
    for (int64_t i = 0; i < n; ++i) counts[i] = off[(size_t)i + 1] - off[(size_t)i];
    return PCP_OK;
}

int pcp_dbscan(pcp_index *ix, double radius, int min_pts, int mem, int32_t *labels, int32_t *n_clusters) {
    if (!ix) return fail(PCP_ERR_INVALID, "pcp_dbscan: index is NULL");
    if (bad_mem(mem)) return fail(PCP_ERR_INVALID, "pcp_dbscan: mem must be PCP_HOST or PCP_DEVICE");
    if (!(radius >= 0.0)) return fail(PCP_ERR_INVALID, "pcp_dbscan: radius must be >= 0");
    int32_t dummy = 0;
    if (!n_clusters) n_clusters = &dummy;
    *n_clusters = 0;
    const int64_t n = ix->n;
    if (n == 0) return PCP_OK;
    if (!labels) return fail(PCP_ERR_INVALID, "pcp_dbscan: labels is NULL");
    pcp_ctx *ctx = ix->ctx;
    DeviceGuard g(ctx->device);
    int32_t *d_lab = labels;
    if (mem == PCP_HOST) {
        PCP_TRY(ctx->scratch[7].ensure(sizeof(int32_t) * (size_t)n));
        d_lab = ctx->scratch[7].as<int32_t>();
    }
    PCP_TRY(dbscan_device(ix, radius, min_pts, d_lab, n_clusters));
    if (mem == PCP_HOST) {
        PCP_TRY(download(ctx, labels, d_lab, sizeof(int32_t) * (size_t)n));
        PCP_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    return PCP_OK;
}

// ---------------------------------------------------------------------------------------------- pose helpers

// rotmat2quaternion / copysign / homo2tq of main.py:158-180 (host arithmetic, float64)
int pcp_homo2tq(const double *m, double *tq) {
    if (!m || !tq) return fail(PCP_ERR_INVALID, "pcp_homo2tq: NULL argument");
    auto at = [&](int r, int c) { return m[r * 4 + c]; };
    auto half_sqrt = [](double v) { return std::sqrt(v > 0.0 ? v : 0.0) / 2.0; };     // np.sqrt(max(0, v)) / 2
    auto copysign_ref = [](double v, double s) { return v * s < 0 ? -v : v; };       // main.py:176-180
    const double trace = at(0, 0) + at(1, 1) + at(2, 2);
    double qw = half_sqrt(trace + 1);
    double qx = half_sqrt(1 + at(0, 0) - at(1, 1) - at(2, 2));
    double qy = half_sqrt(1 - at(0, 0) + at(1, 1) - at(2, 2));
    double qz = half_sqrt(1 - at(0, 0) - at(1, 1) + at(2, 2));
    qx = copysign_ref(qx, at(2, 1) - at(1, 2));
    qy = copysign_ref(qy, at(0, 2) - at(2, 0));
    qz = copysign_ref(qz, at(1, 0) - at(0, 1));
    tq[0] = at(0, 3); tq[1] = at(1, 3); tq[2] = at(2, 3);
    tq[3] = qw; tq[4] = qx; tq[5] = qy; tq[6] = qz;
    return PCP_OK;
}

}  // extern "C"
