// pybind11 module `_pcp`: thin Python binding over the C ABI of include/pcp.h and NOTHING else
// (the PCLKeypoints/src/keypoints.cpp:240-298 pattern: one PYBIND11_MODULE, m.def with py::arg
// defaults, __version__ attribute, C++ failures surface as Python exceptions).
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <algorithm>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "pcp.h"

namespace py = pybind11;

namespace {

using f32_array = py::array_t<float, py::array::c_style | py::array::forcecast>;
using f64_array = py::array_t<double, py::array::c_style | py::array::forcecast>;
using i64_array = py::array_t<int64_t, py::array::c_style | py::array::forcecast>;

void check(int status, const char *what) {
    if (status == PCP_OK) return;
    std::string msg = std::string(what) + ": " + pcp_last_error();
    if (status == PCP_ERR_INVALID) throw py::value_error(msg);
    if (status == PCP_ERR_OOM) throw std::bad_alloc();
    throw std::runtime_error(msg);
}

// Result arrays: page-locked buffers from the library's pool (DMA target of the device->host copy), owned by
// the returned numpy array and handed back to the pool when it is collected.  Small results use numpy's own
// allocator.
template <class T>
py::array_t<T> result_array(std::vector<py::ssize_t> shape) {
    size_t count = 1;
    for (py::ssize_t d : shape) count *= (size_t)d;
    const size_t bytes = count * sizeof(T);
    if (bytes < (64u << 10)) return py::array_t<T>(shape);
    void *p = nullptr;
    check(pcp_host_alloc(bytes, &p), "pcp_host_alloc");
    py::capsule owner(p, [](void *q) { pcp_host_free(q); });
    return py::array_t<T>(shape, static_cast<T *>(p), owner);
}

// The three result arrays of a kNN call as views of ONE page-locked slab (idx | dist | count): the library then
// brings them back with a single device -> host copy.
struct KnnResult {
    py::array_t<int32_t> idx, cnt;
    py::array_t<float> dist;
    KnnResult(int64_t n, int k) {
        const size_t nres = (size_t)n * (size_t)k, bytes = nres * 8 + (size_t)n * 4;
        if (bytes < (64u << 10)) {
            idx = py::array_t<int32_t>({(py::ssize_t)n, (py::ssize_t)k});
            dist = py::array_t<float>({(py::ssize_t)n, (py::ssize_t)k});
            cnt = py::array_t<int32_t>((py::ssize_t)n);
            return;
        }
        void *p = nullptr;
        check(pcp_host_alloc(bytes, &p), "pcp_host_alloc");
        py::capsule owner(p, [](void *q) { pcp_host_free(q); });
        char *c = static_cast<char *>(p);
        idx = py::array_t<int32_t>({(py::ssize_t)n, (py::ssize_t)k}, reinterpret_cast<int32_t *>(c), owner);
        dist = py::array_t<float>({(py::ssize_t)n, (py::ssize_t)k}, reinterpret_cast<float *>(c + nres * 4), owner);
        cnt = py::array_t<int32_t>({(py::ssize_t)n}, reinterpret_cast<int32_t *>(c + nres * 8), owner);
    }
    py::tuple tuple() const { return py::make_tuple(idx, dist, cnt); }
};

// rows x cols float32 view; the reference's clouds are (N, 3) (benchmark.py:19); (N, >=3) rows with
// extra columns (KITTI intensity, Oxford normals) are accepted and only x,y,z are read.
struct Cloud {
    f32_array arr;
    int64_t n, stride;
    explicit Cloud(const py::object &o, const char *name) : arr(f32_array::ensure(o)) {
        if (!arr) throw py::value_error(std::string(name) + ": cannot convert to a float32 array");
        if (arr.ndim() == 1 && arr.shape(0) == 3) arr = arr.reshape({(py::ssize_t)1, (py::ssize_t)3});
        if (arr.ndim() != 2 || arr.shape(1) < 3)
            throw py::value_error(std::string(name) + ": expected shape (N, 3) [row per point], got ndim=" +
                                  std::to_string(arr.ndim()) +
                                  (arr.ndim() == 2 ? " cols=" + std::to_string(arr.shape(1)) : "") +
                                  "; the search structure is 3-D only");
        n = arr.shape(0);
        stride = arr.shape(1);
    }
    const float *data() const { return arr.data(); }
};

// CUDA graph of captured PCP_DEVICE calls (pcp_graph_*)
class Graph {
   public:
    explicit Graph(pcp_graph *g) : g_(g) {}
    ~Graph() { pcp_graph_destroy(g_); }
    Graph(const Graph &) = delete;
    void launch() { check(pcp_graph_launch(g_), "pcp_graph_launch"); }
    int64_t nodes() const { return pcp_graph_nodes(g_); }

   private:
    pcp_graph *g_;
};

class Context {
   public:
    explicit Context(int device) { check(pcp_ctx_create(device, &ctx_), "pcp_ctx_create"); }
    ~Context() { pcp_ctx_destroy(ctx_); }
    Context(const Context &) = delete;
    pcp_ctx *get() const { return ctx_; }
    void synchronize() { check(pcp_ctx_synchronize(ctx_), "pcp_ctx_synchronize"); }
    int64_t launch_count() const { return pcp_ctx_launch_count(ctx_); }
    uintptr_t stream() const { return reinterpret_cast<uintptr_t>(pcp_ctx_stream(ctx_)); }
    void enable_counters(bool on) { check(pcp_ctx_enable_counters(ctx_, on ? 1 : 0), "pcp_ctx_enable_counters"); }
    void enable_kernel_timing(bool on) { check(pcp_ctx_enable_kernel_timing(ctx_, on ? 1 : 0), "pcp_ctx_enable_kernel_timing"); }
    py::tuple read_kernel_timing() {
        double ms = 0.0;
        int64_t n = 0;
        check(pcp_ctx_read_kernel_timing(ctx_, &ms, &n), "pcp_ctx_read_kernel_timing");
        return py::make_tuple(ms, n);
    }
    void graph_begin() { check(pcp_graph_begin(ctx_), "pcp_graph_begin"); }
    std::unique_ptr<Graph> graph_end() {
        pcp_graph *g = nullptr;
        check(pcp_graph_end(ctx_, &g), "pcp_graph_end");
        return std::unique_ptr<Graph>(new Graph(g));
    }
    py::tuple icp_last_stats() {
        int64_t l = 0, s = 0, e = 0;
        check(pcp_icp_last_stats(ctx_, &l, &s, &e), "pcp_icp_last_stats");
        return py::make_tuple(l, s, e);
    }
    std::vector<uint64_t> read_counters() {
        std::vector<uint64_t> v(8);
        check(pcp_ctx_read_counters(ctx_, v.data()), "pcp_ctx_read_counters");
        return v;
    }

   private:
    pcp_ctx *ctx_ = nullptr;
};

class Index {
   public:
    Index(std::shared_ptr<Context> ctx, const py::object &points) : ctx_(std::move(ctx)) {
        Cloud c(points, "points");
        check(pcp_index_build(ctx_->get(), c.data(), c.n, c.stride, PCP_HOST, &ix_), "pcp_index_build");
    }
    // device-resident cloud: raw device pointer (e.g. torch tensor .data_ptr()) + shape
    Index(std::shared_ptr<Context> ctx, uintptr_t dptr, int64_t n, int64_t stride) : ctx_(std::move(ctx)) {
        check(pcp_index_build(ctx_->get(), reinterpret_cast<const float *>(dptr), n, stride, PCP_DEVICE, &ix_),
              "pcp_index_build");
    }
    ~Index() { pcp_index_destroy(ix_); }
    Index(const Index &) = delete;

    void rebuild(const py::object &points) {
        Cloud c(points, "points");
        check(pcp_index_rebuild(ix_, c.data(), c.n, c.stride, PCP_HOST), "pcp_index_rebuild");
    }
    void rebuild_device(uintptr_t dptr, int64_t n, int64_t stride) {
        check(pcp_index_rebuild(ix_, reinterpret_cast<const float *>(dptr), n, stride, PCP_DEVICE), "pcp_index_rebuild");
    }
    int64_t size() const { return pcp_index_size(ix_); }
    // global initialisation: normals / FPFH over hybrid neighbourhoods (main.py:33-47)
    py::array_t<double> normals_hybrid(double radius, int max_nn) {
        py::array_t<double> nrm({(py::ssize_t)size(), (py::ssize_t)3});
        check(pcp_normals_hybrid(ix_, radius, max_nn, PCP_HOST, nrm.mutable_data()), "pcp_normals_hybrid");
        return nrm;
    }
    py::array_t<double> fpfh(const f64_array &normals, double radius, int max_nn) {
        if (normals.ndim() != 2 || normals.shape(0) != size() || normals.shape(1) != 3)
            throw py::value_error("fpfh: normals must be (N, 3) float64 rows of the indexed cloud");
        py::array_t<double> f({(py::ssize_t)size(), (py::ssize_t)33});
        check(pcp_fpfh(ix_, normals.data(), radius, max_nn, PCP_HOST, f.mutable_data()), "pcp_fpfh");
        return f;
    }
    pcp_index *raw() const { return ix_; }
    py::array_t<int64_t> neighbor_counts_f64(double radius) {
        py::array_t<int64_t> c((py::ssize_t)size());
        check(pcp_neighbor_counts_f64(ix_, radius, c.mutable_data()), "pcp_neighbor_counts_f64");
        return c;
    }
    void attach_f64(const f64_array &xyz) {
        if (xyz.ndim() != 2 || xyz.shape(1) < 3) throw py::value_error("attach_f64: expected (N, >= 3) float64 rows");
        check(pcp_index_attach_f64(ix_, xyz.data(), xyz.shape(0), xyz.shape(1), PCP_HOST), "pcp_index_attach_f64");
    }
    py::tuple stats() const {
        int64_t nb = 0;
        int nl = 0;
        check(pcp_index_stats(ix_, &nb, &nl), "pcp_index_stats");
        return py::make_tuple(nb, nl);
    }

    py::tuple knn(const py::object &queries, int k) {
        Cloud q(queries, "queries");
        if (k < 1) throw py::value_error("k must be >= 1");
        KnnResult r(q.n, k);
        {
            py::gil_scoped_release rel;
            check_nogil(pcp_knn(ix_, q.data(), q.n, q.stride, k, PCP_HOST, r.idx.mutable_data(), r.dist.mutable_data(),
                                r.cnt.mutable_data()), "pcp_knn");
        }
        rethrow();
        return r.tuple();
    }
    // exhaustive kernel (kdtree.py:234-238 as a device kernel): the exhaustive cross-check
    py::tuple knn_bruteforce(const py::object &queries, int k) {
        Cloud q(queries, "queries");
        if (k < 1) throw py::value_error("k must be >= 1");
        auto idx = result_array<int32_t>({(py::ssize_t)q.n, (py::ssize_t)k});
        auto dist = result_array<float>({(py::ssize_t)q.n, (py::ssize_t)k});
        auto cnt = result_array<int32_t>({(py::ssize_t)q.n});
        {
            py::gil_scoped_release rel;
            check_nogil(pcp_knn_bruteforce(ix_, q.data(), q.n, q.stride, k, PCP_HOST, idx.mutable_data(),
                                           dist.mutable_data(), cnt.mutable_data()), "pcp_knn_bruteforce");
        }
        rethrow();
        return py::make_tuple(idx, dist, cnt);
    }
    py::array_t<int64_t> radius_count_bruteforce(const py::object &queries, double radius) {
        Cloud q(queries, "queries");
        py::array_t<int64_t> counts((py::ssize_t)q.n);
        check(pcp_radius_count_bruteforce(ix_, q.data(), q.n, q.stride, (float)radius, PCP_HOST, counts.mutable_data()),
              "pcp_radius_count_bruteforce");
        return counts;
    }
    void knn_bruteforce_device(uintptr_t q, int64_t nq, int64_t stride, int k, uintptr_t idx, uintptr_t dist, uintptr_t cnt) {
        check(pcp_knn_bruteforce(ix_, reinterpret_cast<const float *>(q), nq, stride, k, PCP_DEVICE,
                                 reinterpret_cast<int32_t *>(idx), reinterpret_cast<float *>(dist),
                                 reinterpret_cast<int32_t *>(cnt)), "pcp_knn_bruteforce");
    }
    void radius_count_bruteforce_device(uintptr_t q, int64_t nq, int64_t stride, double radius, uintptr_t counts) {
        check(pcp_radius_count_bruteforce(ix_, reinterpret_cast<const float *>(q), nq, stride, (float)radius, PCP_DEVICE,
                                          reinterpret_cast<int64_t *>(counts)), "pcp_radius_count_bruteforce");
    }
    py::tuple knn_self(int k) {
        if (k < 1) throw py::value_error("k must be >= 1");
        const int64_t n = size();
        KnnResult r(n, k);
        {
            py::gil_scoped_release rel;
            check_nogil(pcp_knn_self(ix_, k, PCP_HOST, r.idx.mutable_data(), r.dist.mutable_data(), r.cnt.mutable_data()),
                        "pcp_knn_self");
        }
        rethrow();
        return r.tuple();
    }
    // normals by k-NN PCA (pca_normal.py:85-90): (normals (N,3) f64, eigenvalues (N,3) f64 descending)
    py::tuple normals(int k) {
        const int64_t n = size();
        auto nrm = result_array<double>({(py::ssize_t)n, (py::ssize_t)3});
        auto ev = result_array<double>({(py::ssize_t)n, (py::ssize_t)3});
        check(pcp_normals(ix_, k, PCP_HOST, nrm.mutable_data(), ev.mutable_data()), "pcp_normals");
        return py::make_tuple(nrm, ev);
    }
    // ISS.py:40-73 -> (keypoint indices in selection order, lambda3 (N,) with NaN for non-candidates)
    py::tuple iss_keypoints(double radius, double lambda21, double lambda32, double non_max_radius, int iss_count) {
        if (iss_count < 0) throw py::value_error("iss_count must be >= 0");
        std::vector<int32_t> kp((size_t)iss_count + 1);
        int32_t nk = 0;
        auto l3 = result_array<double>({(py::ssize_t)size()});
        check(pcp_iss_keypoints(ix_, radius, lambda21, lambda32, non_max_radius, iss_count, kp.data(), &nk, l3.mutable_data()),
              "pcp_iss_keypoints");
        py::array_t<int32_t> out(nk);
        std::copy(kp.begin(), kp.begin() + nk, out.mutable_data());
        return py::make_tuple(out, l3);
    }
    // DBSCAN.fit + predict of Cluster_dbscan/dbscan.py:10-38 -> (labels (N,) int32, number of clusters)
    py::tuple dbscan(double radius, int min_pts) {
        auto lab = result_array<int32_t>({(py::ssize_t)size()});
        int32_t nc = 0;
        check(pcp_dbscan(ix_, radius, min_pts, PCP_HOST, lab.mutable_data(), &nc), "pcp_dbscan");
        return py::make_tuple(lab, nc);
    }
    int dbscan_device(double radius, int min_pts, uintptr_t labels) {
        int32_t nc = 0;
        check(pcp_dbscan(ix_, radius, min_pts, PCP_DEVICE, reinterpret_cast<int32_t *>(labels), &nc), "pcp_dbscan");
        return nc;
    }
    void normals_device(int k, uintptr_t nrm, uintptr_t ev) {
        check(pcp_normals(ix_, k, PCP_DEVICE, reinterpret_cast<double *>(nrm), reinterpret_cast<double *>(ev)), "pcp_normals");
    }
    // device in / device out (raw pointers), asynchronous on the context stream
    void knn_device(uintptr_t q, int64_t nq, int64_t stride, int k, uintptr_t idx, uintptr_t dist, uintptr_t cnt) {
        check(pcp_knn(ix_, reinterpret_cast<const float *>(q), nq, stride, k, PCP_DEVICE,
                      reinterpret_cast<int32_t *>(idx), reinterpret_cast<float *>(dist),
                      reinterpret_cast<int32_t *>(cnt)), "pcp_knn");
    }
    void knn_self_device(int k, uintptr_t idx, uintptr_t dist, uintptr_t cnt) {
        check(pcp_knn_self(ix_, k, PCP_DEVICE, reinterpret_cast<int32_t *>(idx), reinterpret_cast<float *>(dist),
                           reinterpret_cast<int32_t *>(cnt)), "pcp_knn_self");
    }

    py::tuple radius(const py::object &queries, double radius, bool sorted) {
        Cloud q(queries, "queries");
        py::array_t<int64_t> offsets((py::ssize_t)(q.n + 1));
        check(pcp_radius_count(ix_, q.data(), q.n, q.stride, (float)radius, PCP_HOST, offsets.mutable_data()),
              "pcp_radius_count");
        const int64_t m = offsets.data()[q.n];
        auto idx = result_array<int32_t>({(py::ssize_t)m});
        auto dist = result_array<float>({(py::ssize_t)m});
        check(pcp_radius_fill(ix_, q.data(), q.n, q.stride, (float)radius, PCP_HOST, offsets.data(),
                              idx.mutable_data(), dist.mutable_data(), sorted ? 1 : 0), "pcp_radius_fill");
        return py::make_tuple(offsets, idx, dist);
    }
    void radius_count_device(uintptr_t q, int64_t nq, int64_t stride, double radius, uintptr_t offsets) {
        check(pcp_radius_count(ix_, reinterpret_cast<const float *>(q), nq, stride, (float)radius, PCP_DEVICE,
                               reinterpret_cast<int64_t *>(offsets)), "pcp_radius_count");
    }
    void radius_fill_device(uintptr_t q, int64_t nq, int64_t stride, double radius, uintptr_t offsets, uintptr_t idx,
                            uintptr_t dist, bool sorted) {
        check(pcp_radius_fill(ix_, reinterpret_cast<const float *>(q), nq, stride, (float)radius, PCP_DEVICE,
                              reinterpret_cast<const int64_t *>(offsets), reinterpret_cast<int32_t *>(idx),
                              reinterpret_cast<float *>(dist), sorted ? 1 : 0), "pcp_radius_fill");
    }

   private:
    // pcp_last_error() is thread-local and we stay on the same thread, so it is safe to defer
    void check_nogil(int status, const char *what) {
        pending_ = status;
        what_ = what;
        if (status != PCP_OK) msg_ = pcp_last_error();
    }
    void rethrow() {
        if (pending_ == PCP_OK) return;
        int s = pending_;
        pending_ = PCP_OK;
        std::string msg = std::string(what_) + ": " + msg_;
        if (s == PCP_ERR_INVALID) throw py::value_error(msg);
        throw std::runtime_error(msg);
    }
    std::shared_ptr<Context> ctx_;
    pcp_index *ix_ = nullptr;
    int pending_ = PCP_OK;
    const char *what_ = "";
    std::string msg_;
};

// Multi-GPU handle (one process per GPU): pcp_multi_* of include/pcp.h
class Multi {
   public:
    Multi(std::shared_ptr<Context> ctx, int rank, int world, const py::object &id) : ctx_(std::move(ctx)) {
        std::string raw;
        if (!id.is_none()) raw = py::cast<std::string>(py::bytes(id));
        if (world > 1 && raw.size() != PCP_MULTI_ID_BYTES) throw py::value_error("unique id must be 128 bytes");
        check(pcp_multi_create(ctx_->get(), rank, world, world > 1 ? raw.data() : nullptr, &m_), "pcp_multi_create");
    }
    ~Multi() { pcp_multi_destroy(m_); }
    Multi(const Multi &) = delete;
    static py::bytes unique_id() {
        char id[PCP_MULTI_ID_BYTES];
        check(pcp_multi_unique_id(id), "pcp_multi_unique_id");
        return py::bytes(id, sizeof id);
    }
    py::dict info() const {
        int r = 0, w = 0, v = 0, n = 0;
        check(pcp_multi_info(m_, &r, &w, &v, &n), "pcp_multi_info");
        py::dict d;
        d["rank"] = r; d["world"] = w; d["nccl_version"] = v; d["comm_nranks"] = n;
        return d;
    }
    int64_t set_reference(const py::object &shard) {
        Cloud c(shard, "reference shard");
        int64_t n = 0;
        int rc;
        std::string msg;
        {
            py::gil_scoped_release rel;
            rc = pcp_multi_set_reference(m_, c.data(), c.n, c.stride, PCP_HOST, &n);
            if (rc != PCP_OK) msg = pcp_last_error();
        }
        if (rc != PCP_OK) throw std::runtime_error("pcp_multi_set_reference: " + msg);
        return n;
    }
    int64_t set_reference_device(uintptr_t p, int64_t n_shard, int64_t stride) {
        int64_t n = 0;
        check(pcp_multi_set_reference(m_, reinterpret_cast<const float *>(p), n_shard, stride, PCP_DEVICE, &n),
              "pcp_multi_set_reference");
        return n;
    }
    py::tuple last_timing() const {
        double a = 0, b = 0;
        check(pcp_multi_last_timing(m_, &a, &b), "pcp_multi_last_timing");
        return py::make_tuple(a, b);
    }
    py::tuple reference() const {
        const float *p = nullptr;
        int64_t n = 0, st = 0;
        int w = 0;
        check(pcp_multi_info(m_, nullptr, &w, nullptr, nullptr), "pcp_multi_info");
        std::vector<int64_t> rows((size_t)w);
        check(pcp_multi_reference(m_, &p, &n, &st, rows.data()), "pcp_multi_reference");
        return py::make_tuple(reinterpret_cast<uintptr_t>(p), n, st, rows);
    }
    py::tuple knn(const py::object &queries, int k) {
        Cloud q(queries, "queries");
        if (k < 1) throw py::value_error("k must be >= 1");
        auto idx = result_array<int32_t>({(py::ssize_t)q.n, (py::ssize_t)k});
        auto dist = result_array<float>({(py::ssize_t)q.n, (py::ssize_t)k});
        auto cnt = result_array<int32_t>({(py::ssize_t)q.n});
        check(pcp_multi_knn(m_, q.data(), q.n, q.stride, k, PCP_HOST, idx.mutable_data(), dist.mutable_data(),
                            cnt.mutable_data()), "pcp_multi_knn");
        return py::make_tuple(idx, dist, cnt);
    }
    void knn_device(uintptr_t q, int64_t nq, int64_t stride, int k, uintptr_t idx, uintptr_t dist, uintptr_t cnt) {
        check(pcp_multi_knn(m_, reinterpret_cast<const float *>(q), nq, stride, k, PCP_DEVICE,
                            reinterpret_cast<int32_t *>(idx), reinterpret_cast<float *>(dist),
                            reinterpret_cast<int32_t *>(cnt)), "pcp_multi_knn");
    }
    void knn_bruteforce_device(uintptr_t q, int64_t nq, int64_t stride, int k, uintptr_t idx, uintptr_t dist, uintptr_t cnt) {
        check(pcp_knn_bruteforce(pcp_multi_index(m_), reinterpret_cast<const float *>(q), nq, stride, k, PCP_DEVICE,
                                 reinterpret_cast<int32_t *>(idx), reinterpret_cast<float *>(dist),
                                 reinterpret_cast<int32_t *>(cnt)), "pcp_knn_bruteforce");
    }
    // global batch (packed clouds + offsets) -> dict with this rank's range and its rows of the results
    py::dict icp_batch(const py::object &src, const i64_array &src_off, const py::object &tgt, const i64_array &tgt_off,
                       const f64_array &T_init, int max_iteration, double R_diff_thres, double t_diff_thres,
                       double dist_thres, bool first_iter_broadcast) {
        Cloud s(src, "source"), t(tgt, "target");
        if (s.stride != t.stride) throw py::value_error("source and target must have the same number of columns");
        const int64_t P = src_off.size() - 1;
        if (P < 0 || tgt_off.size() != src_off.size()) throw py::value_error("offset arrays must both have n_pairs+1 entries");
        if (T_init.size() != P * 16) throw py::value_error("T_init must be (n_pairs, 4, 4)");
        pcp_icp_params prm;
        pcp_icp_default_params(&prm);
        prm.max_iteration = max_iteration;
        prm.R_diff_thres = R_diff_thres;
        prm.t_diff_thres = t_diff_thres;
        prm.dist_thres = dist_thres;
        prm.first_iter_broadcast = first_iter_broadcast ? 1 : 0;
        py::array_t<double> T_last({(py::ssize_t)P, (py::ssize_t)4, (py::ssize_t)4}), T_total({(py::ssize_t)P, (py::ssize_t)4, (py::ssize_t)4});
        py::array_t<int32_t> iters((py::ssize_t)P), status((py::ssize_t)P);
        std::fill_n(T_last.mutable_data(), P * 16, 0.0);
        std::fill_n(T_total.mutable_data(), P * 16, 0.0);
        std::fill_n(iters.mutable_data(), P, 0);
        std::fill_n(status.mutable_data(), P, -1);
        int64_t lo = 0, hi = 0;
        int rc;
        std::string msg;
        {
            py::gil_scoped_release rel;
            rc = pcp_multi_icp_batch(m_, &prm, P, s.data(), src_off.data(), t.data(), tgt_off.data(), s.stride, PCP_HOST,
                                     T_init.data(), T_last.mutable_data(), T_total.mutable_data(), iters.mutable_data(),
                                     status.mutable_data(), nullptr, &lo, &hi);
            if (rc != PCP_OK) msg = pcp_last_error();
        }
        if (rc != PCP_OK) throw std::runtime_error("pcp_multi_icp_batch: " + msg);
        py::dict d;
        d["first_pair"] = lo; d["last_pair"] = hi;
        d["T_last"] = T_last; d["T_total"] = T_total; d["iterations"] = iters; d["status"] = status;
        return d;
    }

   private:
    std::shared_ptr<Context> ctx_;
    pcp_multi *m_ = nullptr;
};

py::tuple multi_pair_range(const i64_array &src_off, const i64_array &tgt_off, int rank, int world) {
    const int64_t P = src_off.size() - 1;
    if (P < 0 || tgt_off.size() != src_off.size()) throw py::value_error("offset arrays must both have n_pairs+1 entries");
    int64_t lo = 0, hi = 0;
    check(pcp_multi_pair_range(src_off.data(), tgt_off.data(), P, rank, world, &lo, &hi), "pcp_multi_pair_range");
    return py::make_tuple(lo, hi);
}

// nearest target feature of every source feature: rows = points, columns = feature dimensions
py::tuple feature_match(std::shared_ptr<Context> ctx, const f64_array &src, const f64_array &tgt) {
    if (src.ndim() != 2 || tgt.ndim() != 2 || src.shape(1) != tgt.shape(1))
        throw py::value_error("feature_match: expected (Ns, D) and (Nt, D) float64 arrays");
    py::array_t<int32_t> idx((py::ssize_t)src.shape(0));
    py::array_t<double> d2((py::ssize_t)src.shape(0));
    check(pcp_feature_match(ctx->get(), src.data(), src.shape(0), tgt.data(), tgt.shape(0), (int)src.shape(1), PCP_HOST,
                            idx.mutable_data(), d2.mutable_data()), "pcp_feature_match");
    return py::make_tuple(idx, d2);
}
// feature-matching RANSAC (main.py:68-84) -> dict(transformation 4x4, fitness, inlier_rmse, iterations, evaluated)
py::dict ransac_init(std::shared_ptr<Context> ctx, const py::object &src_points, const Index &target,
                     const py::array_t<int32_t, py::array::c_style | py::array::forcecast> &corr, double max_corr_dist,
                     double edge_ratio, int max_iteration, double confidence, uint64_t seed) {
    Cloud s(src_points, "source");
    if (corr.ndim() != 2 || corr.shape(1) != 2) throw py::value_error("ransac_init: correspondences must be (M, 2) int32");
    pcp_ransac_params prm;
    prm.max_correspondence_distance = max_corr_dist;
    prm.edge_length_threshold = edge_ratio;
    prm.confidence = confidence;
    prm.max_iteration = max_iteration;
    prm.ransac_n = 3;
    prm.seed = seed;
    py::array_t<double> T({(py::ssize_t)4, (py::ssize_t)4});
    double fit = 0, rmse = 0;
    int64_t it = 0, ev = 0;
    check(pcp_ransac_init(ctx->get(), &prm, s.data(), s.n, s.stride, target.raw(), corr.data(), corr.shape(0), PCP_HOST,
                          T.mutable_data(), &fit, &rmse, &it, &ev), "pcp_ransac_init");
    py::dict d;
    d["transformation"] = T; d["fitness"] = fit; d["inlier_rmse"] = rmse; d["iterations"] = it; d["evaluated"] = ev;
    return d;
}

// PointNet++ layouts: (b, m, 3) / (b, n, 3) float32 arrays
py::array_t<int32_t> ball_query(std::shared_ptr<Context> ctx, double radius, int nsample, const f32_array &xyz,
                                const f32_array &new_xyz) {
    if (xyz.ndim() != 3 || new_xyz.ndim() != 3 || xyz.shape(2) != 3 || new_xyz.shape(2) != 3 || xyz.shape(0) != new_xyz.shape(0))
        throw py::value_error("ball_query: expected xyz (B, N, 3) and new_xyz (B, npoint, 3)");
    if (nsample < 0) throw py::value_error("ball_query: nsample must be >= 0");
    const int b = (int)xyz.shape(0), n = (int)xyz.shape(1), m = (int)new_xyz.shape(1);
    py::array_t<int32_t> idx({(py::ssize_t)b, (py::ssize_t)m, (py::ssize_t)nsample});
    check(pcp_ball_query(ctx->get(), b, n, m, (float)radius, nsample, new_xyz.data(), xyz.data(), PCP_HOST, idx.mutable_data()),
          "pcp_ball_query");
    return idx;
}
py::tuple three_nn(std::shared_ptr<Context> ctx, const f32_array &unknown, const f32_array &known) {
    if (unknown.ndim() != 3 || known.ndim() != 3 || unknown.shape(2) != 3 || known.shape(2) != 3 || unknown.shape(0) != known.shape(0))
        throw py::value_error("three_nn: expected unknown (B, n, 3) and known (B, m, 3)");
    const int b = (int)unknown.shape(0), n = (int)unknown.shape(1), m = (int)known.shape(1);
    py::array_t<float> d2({(py::ssize_t)b, (py::ssize_t)n, (py::ssize_t)3});
    py::array_t<int32_t> idx({(py::ssize_t)b, (py::ssize_t)n, (py::ssize_t)3});
    check(pcp_three_nn(ctx->get(), b, n, m, unknown.data(), known.data(), PCP_HOST, d2.mutable_data(), idx.mutable_data()),
          "pcp_three_nn");
    return py::make_tuple(d2, idx);
}

// icp_batch: packed clouds + offsets -> dict(T_last, T_total, iterations, status[, diff_log])
py::dict icp_batch(std::shared_ptr<Context> ctx, const py::object &src, const i64_array &src_off, const py::object &tgt,
                   const i64_array &tgt_off, const f64_array &T_init, int max_iteration, double R_diff_thres,
                   double t_diff_thres, double dist_thres, bool first_iter_broadcast, bool want_log) {
    Cloud s(src, "source"), t(tgt, "target");
    if (s.stride != t.stride) throw py::value_error("source and target must have the same number of columns");
    const int64_t P = src_off.size() - 1;
    if (P < 0 || tgt_off.size() != src_off.size()) throw py::value_error("offset arrays must both have n_pairs+1 entries");
    if (T_init.size() != P * 16) throw py::value_error("T_init must be (n_pairs, 4, 4)");
    if (P > 0 && (src_off.data()[P] != s.n || tgt_off.data()[P] != t.n))
        throw py::value_error("offsets do not match the packed clouds");
    pcp_icp_params prm;
    pcp_icp_default_params(&prm);
    prm.max_iteration = max_iteration;
    prm.R_diff_thres = R_diff_thres;
    prm.t_diff_thres = t_diff_thres;
    prm.dist_thres = dist_thres;
    prm.first_iter_broadcast = first_iter_broadcast ? 1 : 0;
    py::array_t<double> T_last({(py::ssize_t)P, (py::ssize_t)4, (py::ssize_t)4}), T_total({(py::ssize_t)P, (py::ssize_t)4, (py::ssize_t)4});
    py::array_t<int32_t> iters((py::ssize_t)P), status((py::ssize_t)P);
    py::array_t<double> log;
    if (want_log) log = py::array_t<double>({(py::ssize_t)P, (py::ssize_t)std::max(max_iteration, 0), (py::ssize_t)2});
    int rc;
    std::string msg;
    {
        py::gil_scoped_release rel;
        rc = pcp_icp_batch(ctx->get(), &prm, P, s.data(), src_off.data(), t.data(), tgt_off.data(), s.stride, PCP_HOST,
                           T_init.data(), T_last.mutable_data(), T_total.mutable_data(), iters.mutable_data(),
                           status.mutable_data(), want_log ? log.mutable_data() : nullptr);
        if (rc != PCP_OK) msg = pcp_last_error();
    }
    if (rc == PCP_ERR_INVALID) throw py::value_error("pcp_icp_batch: " + msg);
    if (rc != PCP_OK) throw std::runtime_error("pcp_icp_batch: " + msg);
    py::dict out;
    out["T_last"] = T_last;
    out["T_total"] = T_total;
    out["iterations"] = iters;
    out["status"] = status;
    if (want_log) out["diff_log"] = log;
    return out;
}

// device-resident variant for benches: raw device pointers for the packed clouds
py::dict icp_batch_device(std::shared_ptr<Context> ctx, uintptr_t src, const i64_array &src_off, uintptr_t tgt,
                          const i64_array &tgt_off, int64_t stride, int max_iteration, double R_diff_thres,
                          double t_diff_thres, double dist_thres, bool first_iter_broadcast) {
    const int64_t P = src_off.size() - 1;
    if (P < 0 || tgt_off.size() != src_off.size()) throw py::value_error("offset arrays must both have n_pairs+1 entries");
    pcp_icp_params prm;
    pcp_icp_default_params(&prm);
    prm.max_iteration = max_iteration;
    prm.R_diff_thres = R_diff_thres;
    prm.t_diff_thres = t_diff_thres;
    prm.dist_thres = dist_thres;
    prm.first_iter_broadcast = first_iter_broadcast ? 1 : 0;
    py::array_t<double> T_last({(py::ssize_t)P, (py::ssize_t)4, (py::ssize_t)4}), T_total({(py::ssize_t)P, (py::ssize_t)4, (py::ssize_t)4});
    py::array_t<int32_t> iters((py::ssize_t)P), status((py::ssize_t)P);
    check(pcp_icp_batch(ctx->get(), &prm, P, reinterpret_cast<const float *>(src), src_off.data(),
                        reinterpret_cast<const float *>(tgt), tgt_off.data(), stride, PCP_DEVICE, nullptr,
                        T_last.mutable_data(), T_total.mutable_data(), iters.mutable_data(), status.mutable_data(), nullptr),
          "pcp_icp_batch");
    py::dict out;
    out["T_last"] = T_last;
    out["T_total"] = T_total;
    out["iterations"] = iters;
    out["status"] = status;
    return out;
}

py::tuple procrustes(std::shared_ptr<Context> ctx, const f64_array &A, const f64_array &B) {
    if (A.ndim() != 2 || A.shape(0) != 3 || B.ndim() != 2 || B.shape(0) != 3 || A.shape(1) != B.shape(1) || A.shape(1) < 1)
        throw py::value_error("procrustes: A and B must both be (3, N) with N >= 1");
    py::array_t<double> R({(py::ssize_t)3, (py::ssize_t)3}), t({(py::ssize_t)3, (py::ssize_t)1});
    double cost = 0.0;
    check(pcp_procrustes(ctx->get(), A.data(), B.data(), A.shape(1), PCP_HOST, R.mutable_data(), t.mutable_data(), &cost),
          "pcp_procrustes");
    return py::make_tuple(R, t, cost);
}

// voxel_filter(point_cloud, leaf_size, type) of Pca_and_Voxel_filter/voxel_filter.py:10-66 -> (M, 3) float64
py::array_t<double> voxel_filter(std::shared_ptr<Context> ctx, const py::object &points, double leaf_size,
                                 const std::string &type, uint64_t seed) {
    Cloud c(points, "point_cloud");
    int mode;
    if (type == "centroid") mode = PCP_VOXEL_CENTROID;
    else if (type == "random") mode = PCP_VOXEL_RANDOM;
    else throw py::value_error("voxel_filter: type must be 'centroid' or 'random'");
    int64_t m = 0;
    check(pcp_voxel_filter(ctx->get(), c.data(), c.n, c.stride, leaf_size, mode, seed, PCP_HOST, &m), "pcp_voxel_filter");
    auto out = result_array<double>({(py::ssize_t)m, (py::ssize_t)3});
    check(pcp_voxel_filter_read(ctx->get(), PCP_HOST, out.mutable_data()), "pcp_voxel_filter_read");
    return out;
}

// PCA(data) of pca_normal.py:10-36 -> (eigenvalues (3,) descending, eigenvectors (3,3) columns)
py::tuple pca(std::shared_ptr<Context> ctx, const py::object &points) {
    Cloud c(points, "data");
    py::array_t<double> w(3), v({(py::ssize_t)3, (py::ssize_t)3});
    check(pcp_pca(ctx->get(), c.data(), c.n, c.stride, PCP_HOST, w.mutable_data(), v.mutable_data()), "pcp_pca");
    return py::make_tuple(w, v);
}
py::tuple pca_device(std::shared_ptr<Context> ctx, uintptr_t dptr, int64_t n, int64_t stride) {
    py::array_t<double> w(3), v({(py::ssize_t)3, (py::ssize_t)3});
    check(pcp_pca(ctx->get(), reinterpret_cast<const float *>(dptr), n, stride, PCP_DEVICE, w.mutable_data(), v.mutable_data()), "pcp_pca");
    return py::make_tuple(w, v);
}

// device-resident cloud: returns the number of output rows (the rows stay in the context until read)
int64_t voxel_filter_device(std::shared_ptr<Context> ctx, uintptr_t dptr, int64_t n, int64_t stride, double leaf_size,
                            int mode, uint64_t seed) {
    int64_t m = 0;
    check(pcp_voxel_filter(ctx->get(), reinterpret_cast<const float *>(dptr), n, stride, leaf_size, mode, seed,
                           PCP_DEVICE, &m), "pcp_voxel_filter");
    return m;
}
void voxel_filter_read_device(std::shared_ptr<Context> ctx, uintptr_t out) {
    check(pcp_voxel_filter_read(ctx->get(), PCP_DEVICE, reinterpret_cast<double *>(out)), "pcp_voxel_filter_read");
}

py::array_t<double> homo2tq(const f64_array &homo) {
    if (homo.size() != 16) throw py::value_error("homo2tq: expected a 4x4 matrix");
    py::array_t<double> tq(7);
    check(pcp_homo2tq(homo.data(), tq.mutable_data()), "pcp_homo2tq");
    return tq;
}

}  // namespace

PYBIND11_MODULE(_pcp, m) {
    m.doc() = "B200-native exact kNN / radius search and point-to-point ICP (C ABI: include/pcp.h)";
    m.attr("__version__") = std::to_string(pcp_version() / 100) + "." + std::to_string(pcp_version() % 100);
    m.attr("MAX_K") = PCP_MAX_K;
    m.attr("KNN_EMPTY_DIST") = (double)PCP_KNN_EMPTY_DIST;
    m.def("device_count", []() {
        int n = 0;
        int s = pcp_device_count(&n);
        return s == PCP_OK ? n : 0;
    });
    m.def("last_error", []() { return std::string(pcp_last_error()); });

    py::class_<Context, std::shared_ptr<Context>>(m, "Context")
        .def(py::init<int>(), py::arg("device") = 0)
        .def("synchronize", &Context::synchronize)
        .def("launch_count", &Context::launch_count)
        .def("stream", &Context::stream)
        .def("enable_counters", &Context::enable_counters, py::arg("on") = true)
        .def("read_counters", &Context::read_counters)
        .def("enable_kernel_timing", &Context::enable_kernel_timing)
        .def("read_kernel_timing", &Context::read_kernel_timing)
        .def("icp_last_stats", &Context::icp_last_stats)
        .def("graph_begin", &Context::graph_begin)
        .def("graph_end", &Context::graph_end);
    py::class_<Graph>(m, "Graph").def("launch", &Graph::launch).def("nodes", &Graph::nodes);

    py::class_<Index>(m, "Index")
        .def(py::init<std::shared_ptr<Context>, const py::object &>(), py::arg("ctx"), py::arg("points"))
        .def(py::init<std::shared_ptr<Context>, uintptr_t, int64_t, int64_t>(), py::arg("ctx"), py::arg("device_ptr"),
             py::arg("n"), py::arg("stride_floats") = 3)
        .def("rebuild", &Index::rebuild, py::arg("points"))
        .def("rebuild_device", &Index::rebuild_device, py::arg("device_ptr"), py::arg("n"), py::arg("stride_floats") = 3)
        .def("size", &Index::size)
        .def("attach_f64", &Index::attach_f64, py::arg("xyz"))
        .def("normals_hybrid", &Index::normals_hybrid, py::arg("radius"), py::arg("max_nn"))
        .def("fpfh", &Index::fpfh, py::arg("normals"), py::arg("radius"), py::arg("max_nn"))
        .def("neighbor_counts_f64", &Index::neighbor_counts_f64, py::arg("radius"))
        .def("stats", &Index::stats)
        .def("knn", &Index::knn, py::arg("queries"), py::arg("k"))
        .def("knn_self", &Index::knn_self, py::arg("k"))
        .def("knn_bruteforce", &Index::knn_bruteforce, py::arg("queries"), py::arg("k"))
        .def("radius_count_bruteforce", &Index::radius_count_bruteforce, py::arg("queries"), py::arg("radius"))
        .def("knn_bruteforce_device", &Index::knn_bruteforce_device)
        .def("radius_count_bruteforce_device", &Index::radius_count_bruteforce_device)
        .def("normals", &Index::normals, py::arg("k") = 5)
        .def("normals_device", &Index::normals_device)
        .def("dbscan", &Index::dbscan, py::arg("radius") = 0.5f, py::arg("min_pts") = 10)
        .def("dbscan_device", &Index::dbscan_device)
        .def("iss_keypoints", &Index::iss_keypoints, py::arg("radius") = 0.5f, py::arg("lambda21") = 0.5,
             py::arg("lambda32") = 0.5, py::arg("non_max_radius") = 0.5f, py::arg("iss_count") = 20)
        .def("knn_device", &Index::knn_device)
        .def("knn_self_device", &Index::knn_self_device)
        .def("radius", &Index::radius, py::arg("queries"), py::arg("radius"), py::arg("sorted") = false)
        .def("radius_count_device", &Index::radius_count_device)
        .def("radius_fill_device", &Index::radius_fill_device);

    py::class_<Multi>(m, "Multi")
        .def(py::init<std::shared_ptr<Context>, int, int, const py::object &>(), py::arg("ctx"), py::arg("rank"),
             py::arg("world"), py::arg("unique_id") = py::none())
        .def_static("unique_id", &Multi::unique_id)
        .def("info", &Multi::info)
        .def("set_reference", &Multi::set_reference, py::arg("shard"))
        .def("set_reference_device", &Multi::set_reference_device)
        .def("last_timing", &Multi::last_timing)
        .def("reference", &Multi::reference)
        .def("knn", &Multi::knn, py::arg("queries"), py::arg("k"))
        .def("knn_device", &Multi::knn_device)
        .def("knn_bruteforce_device", &Multi::knn_bruteforce_device)
        .def("icp_batch", &Multi::icp_batch, py::arg("source"), py::arg("source_offsets"), py::arg("target"),
             py::arg("target_offsets"), py::arg("T_init"), py::arg("max_iteration") = 100, py::arg("R_diff_thres") = 0.5,
             py::arg("t_diff_thres") = 0.5, py::arg("dist_thres") = 5.0, py::arg("first_iter_broadcast") = true);
    m.def("multi_pair_range", &multi_pair_range, py::arg("source_offsets"), py::arg("target_offsets"), py::arg("rank"),
          py::arg("world"));

    m.def("feature_match", &feature_match, py::arg("ctx"), py::arg("src_features"), py::arg("tgt_features"));
    m.def("ransac_init", &ransac_init, py::arg("ctx"), py::arg("source"), py::arg("target_index"), py::arg("correspondences"),
          py::arg("max_correspondence_distance"), py::arg("edge_length_threshold") = 0.9, py::arg("max_iteration") = 100000,
          py::arg("confidence") = 0.999, py::arg("seed") = 0);
    m.def("ball_query", &ball_query, py::arg("ctx"), py::arg("radius"), py::arg("nsample"), py::arg("xyz"), py::arg("new_xyz"));
    m.def("three_nn", &three_nn, py::arg("ctx"), py::arg("unknown"), py::arg("known"));
    m.def("icp_batch", &icp_batch, py::arg("ctx"), py::arg("source"), py::arg("source_offsets"), py::arg("target"),
          py::arg("target_offsets"), py::arg("T_init"), py::arg("max_iteration") = 100, py::arg("R_diff_thres") = 0.5,
          py::arg("t_diff_thres") = 0.5, py::arg("dist_thres") = 5.0, py::arg("first_iter_broadcast") = true,
          py::arg("want_log") = false);
    m.def("icp_batch_device", &icp_batch_device);
    m.def("homo2tq", &homo2tq, py::arg("homo"));
    m.def("pca", &pca, py::arg("ctx"), py::arg("data"));
    m.def("pca_device", &pca_device);
    m.def("voxel_filter_device", &voxel_filter_device);
    m.def("voxel_filter_read_device", &voxel_filter_read_device);
    m.def("voxel_filter", &voxel_filter, py::arg("ctx"), py::arg("point_cloud"), py::arg("leaf_size"),
          py::arg("type") = "centroid", py::arg("seed") = 0);
    m.def("procrustes", &procrustes, py::arg("ctx"), py::arg("A"), py::arg("B"));

}
