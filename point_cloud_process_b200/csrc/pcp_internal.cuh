// Internal declarations shared by the CUDA translation units of libpcp_b200.so (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <utility>
#include <vector>

#include "pcp.h"

namespace pcp { struct RadiusTest; }

namespace pcp {

// ------------------------------------------------------------------------------------------------
// errors
void set_error(const std::string &msg);
// graph capture on this thread: DevBuf::ensure notes a buffer that had to grow (the captured graph would then
// hold pointers of a call pattern that was never run before: capture is refused)
void capture_note_begin();
bool capture_note_grew();
int fail(int code, const std::string &msg);
int cuda_fail(cudaError_t e, const char *what, const char *file, int line);

#define PCP_CUDA(expr)                                                        \
    do {                                                                      \
        cudaError_t _e = (expr);                                              \
        if (_e != cudaSuccess) return ::pcp::cuda_fail(_e, #expr, __FILE__, __LINE__); \
    } while (0)

#define PCP_TRY(expr)            \
    do {                         \
        int _s = (expr);         \
        if (_s != PCP_OK) return _s; \
    } while (0)

// ------------------------------------------------------------------------------------------------
// grow-only device / pinned-host buffers: steady-state calls never hit cudaMalloc
struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    int ensure(size_t bytes);
    void release();
    template <class T> T *as() const { return reinterpret_cast<T *>(p); }
};

struct PinBuf {
    void *p = nullptr;
    size_t cap = 0;
    int ensure(size_t bytes);
    void release();
    template <class T> T *as() const { return reinterpret_cast<T *>(p); }
};

constexpr int kMaxDevices = 64;    // per-device caches of function attributes / occupancy (power of two)
constexpr int kLeaf = 32;          // points per bucket == warp width
constexpr int kFan = 32;           // children per hierarchy node == warp width
constexpr int kMaxLevels = 8;      // 32^7 buckets > 2^31 points

}  // namespace pcp


namespace pcp {

// One indexed cloud ("segment") as the kernels see it.  All pointers are already offset to the
// segment's own sub-arrays, so the same device code serves a single index and one pair of a batch.
//   pts       float4[n_leaf*32]  Morton-sorted points, .w = index bits (row within the segment); pad lanes NaN / -1
//   box_lo/hi float4[nodes]      AABB per node, level 0 (buckets) first, then level 1 ... root last
//   half      float4[n_leaf*4]   AABBs of the two 16-point halves of every bucket (the kd refinement's last split)
//   level_off[l] first node of level l in box_*; level_cnt[l] nodes on level l
struct IndexView {
    const float4 *pts;
    // the same buckets in PAIR layout for the packed-FP32 (f32x2) kernels: bucket b = float4[32] at pts2 + 32 b,
    // float4 2m = (x[2m], x[2m+1], y[2m], y[2m+1]), float4 2m+1 = (z[2m], z[2m+1], w[2m], w[2m+1]);
    // pad lanes x = +inf, y = z = 0, w = -1 (every distance against them is +inf, above every finite gate)
    const float4 *pts2;
    const float4 *box_lo;
    const float4 *box_hi;
    const float4 *half;     // float4[n_leaf*4]: per bucket {lo, hi} of points 0..15 and {lo, hi} of points 16..31 (empty: +inf/-inf)
    const float *bbox;      // lo.xyz, cells per unit, hi.xyz, extent
    long long src_off;      // first row of this segment in the source array handed to the build
    long long n;
    long long n_leaf;
    int n_levels;           // number of levels incl. the bucket level; root = level n_levels-1
    long long level_off[kMaxLevels];
    long long level_cnt[kMaxLevels];
};

// Search structure over P independent clouds built in ONE pass (segmented Morton sort): P = 1 for
// pcp_index, P = number of registration pairs for the ICP targets.  DESIGN.md section 3.
struct BatchIndex {
    pcp_ctx *ctx = nullptr;
    int64_t P = 0;
    int64_t total_n = 0, total_leaf = 0, total_nodes = 0;
    int64_t max_n = 0, max_leaf = 0;
    int max_levels = 0;
    int64_t level_max_cnt[kMaxLevels] = {0};
    std::vector<IndexView> h_views;
    cudaEvent_t ev_bbox = nullptr;   // optional: recorded on the build stream once the segment bounding boxes are final
    cudaEvent_t ev_bbox_record = nullptr;   // the event the running build records (ev_bbox, or the capture's own)
    DevBuf pts, pts2, box_lo, box_hi, half, bbox, views, keys_a, keys_b, vals_a, vals_b;
    const IndexView *d_views() const { return views.as<IndexView>(); }
    void release();
};

// Host -> device copy of a small host structure, ordered on st.  Normally a cudaMemcpyAsync from pageable memory.
// While pcp_icp_batch(PCP_HOST) streams the next run of clouds over the copy engine (ctx->h2d_by_kernel) such a copy
// would queue behind hundreds of MB on that engine and stall the kernels after it: the bytes are then placed in a
// page-locked ring (ctx->small_stage; the caller sizes it and resets small_used between synchronised calls) and a
// kernel fetches them over PCIe itself.  bytes must be a multiple of 4.
int small_h2d(pcp_ctx *ctx, void *dst, const void *src, size_t bytes, cudaStream_t st);

// Build over rows [offsets[p], offsets[p+1]) of d_xyz for p < P (offsets: P+1 host values).
int batch_build_device(BatchIndex *b, const float *d_xyz, int64_t stride, int64_t P, const int64_t *h_offsets);

}  // namespace pcp

struct pcp_ctx {
    int device = 0;
    int sm_count = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t copy_stream = nullptr;   // host->device copies of query sets: overlap the work already queued on `stream`
    cudaEvent_t ev_copy = nullptr;        // end of the last copy on copy_stream / of the last cloud upload on stream
    // Query regrouping on the side stream (Morton keys, radix sort, kd regrouping of a kNN query set), concurrent
    // with an index build still queued on `stream`: own key buffers and CUB storage, ordered by two events
    pcp::DevBuf q_keys[3], q_perm, cub_tmp_side;
    cudaEvent_t ev_q_ready = nullptr;     // side stream: permutation of the current query set complete
    cudaEvent_t ev_search_end = nullptr;  // main stream: last search that read q_perm has finished
    // CUDA-graph capture (pcp_graph_begin / pcp_graph_end): the stream is capturing; the fork / join of the side
    // stream uses events of its own (an event recorded inside a capture must not be waited on outside of it)
    bool capturing = false, cap_has_bbox = false;
    cudaEvent_t ev_cap_bbox = nullptr, ev_cap_q = nullptr;
    int64_t launches = 0;
    // last pcp_icp_batch: kernels launched in its iteration loop, host synchronisations, iterations enqueued
    int64_t icp_last_launches = 0, icp_last_syncs = 0, icp_last_enqueued = 0;
    pcp::DevBuf icp_up[4];                // pcp_icp_batch(PCP_HOST): two staging sets {targets, sources}
    cudaEvent_t ev_icp_up[4] = {nullptr, nullptr, nullptr, nullptr};
    pcp::PinBuf small_stage;              // small_h2d(): page-locked ring the copy kernel reads from
    size_t small_used = 0;
    bool h2d_by_kernel = false;           // a large upload occupies the copy engine: small structures go by kernel
    pcp::DevBuf cub_tmp;           // CUB temporary storage
    pcp::DevBuf ss_state[2];       // small_sort_kernel: barrier counters + histograms (main stream / side stream)
    pcp::DevBuf scratch[8];        // per-call device scratch (keys, permutations, staged queries ...)
    pcp::PinBuf pin[4];            // pinned staging for host<->device copies
    pcp::DevBuf counters;          // 8 x u64 work counters (pcp_ctx_enable_counters)
    pcp::BatchIndex icp_sources;   // Morton buckets of the ICP source clouds (query groups of the tiled association)
    pcp::BatchIndex icp_targets;   // hierarchies of the ICP target clouds (pcp_icp_batch), reused across calls
    // Morton/kd regrouping of the query set of the last pcp_radius_count, reused by the pcp_radius_fill that
    // follows it (the two-phase API passes the same queries twice): valid while the tag matches
    pcp::DevBuf radius_perm;
    struct { const void *q = nullptr; int64_t nq = 0, stride = 0; const void *ix = nullptr; uint64_t serial = 0; } radius_perm_tag;
    // ICP front end (pcp_front.cu): result of the last pcp_voxel_filter (read by pcp_voxel_filter_read), its run
    // count / bounding-box accumulators, and the neighbour lists of pcp_normals
    pcp::DevBuf voxel_out, voxel_meta, normals_nn;
    pcp::DevBuf ransac_buf[4];     // pcp_ransac_init: hypotheses, counter, transformed queries, per-query results
    pcp::DevBuf iss_buf[4];        // pcp_iss_keypoints: CSR of the radius search (offsets, indices, distances) and its work arrays
    int64_t voxel_n_out = -1;
    bool count_work = false;
    // optional timing of the dominant kernel (pcp_ctx_enable_kernel_timing): event pairs around every
    // knn_tile_kernel launch, summed by pcp_ctx_read_kernel_timing
    bool time_kernels = false;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> timing_events;
    size_t timing_used = 0;
    unsigned long long *ctr() const { return count_work ? counters.as<unsigned long long>() : nullptr; }
};

struct pcp_index {
    pcp_ctx *ctx = nullptr;
    int64_t n = 0;
    uint64_t serial = 0;     // bumped by every (re)build
    const float *built_from = nullptr;   // device pointer / shape of the cloud of the last build
    int64_t built_stride = 0;
    pcp::BatchIndex b;       // P == 1
    pcp::DevBuf src_copy;
    // float64 originals of the indexed rows (pcp_index_attach_f64): n x 3 doubles, dropped by the next rebuild
    pcp::DevBuf xyz64;
    int64_t xyz64_n = -1;
    double xyz64_max_abs = 0.0;
};

namespace pcp {

inline const IndexView &make_view(const pcp_index *ix) { return ix->b.h_views[0]; }

// build (pcp_build.cu)
int index_build_device(pcp_index *ix, const float *d_xyz, int64_t n, int64_t stride);
// Morton keys of arbitrary query points in the index's frame, sorted: writes perm (nq) to d_perm
int sort_queries(pcp_index *ix, const float *d_q, int64_t nq, int64_t stride, uint32_t *d_perm);
// the same on an arbitrary stream with caller-provided key buffers (3) and CUB storage
int sort_queries_on(pcp_index *ix, const float *d_q, int64_t nq, int64_t stride, uint32_t *d_perm, cudaStream_t st,
                    DevBuf *keys, DevBuf &cub_tmp);

// search (pcp_knn.cu / pcp_radius.cu); all pointers are device pointers
// side_ok: the query data is known to be complete in the side stream's order (uploaded there by the library) --
// the regrouping may then run beside an index build that is still queued
int knn_device(pcp_index *ix, const float *d_q, int64_t nq, int64_t stride, const uint32_t *d_perm, int k,
               int32_t *d_idx, float *d_dist, int32_t *d_count, bool side_ok = false);
int knn_self_device(pcp_index *ix, int k, int32_t *d_idx, float *d_dist, int32_t *d_count);
// tiled kernel (pcp_knn_tile.cu): one warp per group of 32 Morton-adjacent queries
int knn_tile_device(pcp_index *ix, const float *d_q, int64_t nq, int64_t stride, const uint32_t *d_perm, int k,
                    int32_t *d_idx, float *d_dist, int32_t *d_count);
int knn_tile_self_device(pcp_index *ix, int k, int32_t *d_idx, float *d_dist, int32_t *d_count);
int knn_warp_list_device(pcp_index *ix, bool self, const float *d_q, int64_t nq, int64_t stride, const uint32_t *d_perm,
                         const uint32_t *d_list, const uint32_t *d_sched, uint32_t n_groups, const float *d_hints,
                         uint32_t hint_groups, int k, int32_t *d_idx, float *d_dist, int32_t *d_count, bool overlap);
// 0 = choose by query count, 1 = warp-per-query kernel, 2 = tiled kernel (PCP_KNN_KERNEL=warp|tile, for tests)
int knn_kernel_choice();
// exact64: borderline pairs are decided in float64 (RadiusTest, pcp_device.cuh) -- scipy's semantics for ISS / DBSCAN
int radius_count_device(pcp_index *ix, const float *d_q, int64_t nq, int64_t stride, float radius,
                        int64_t *d_offsets, bool exact64 = false, double radius64 = -1.0);
int radius_fill_device(pcp_index *ix, const float *d_q, int64_t nq, int64_t stride, float radius,
                       const int64_t *d_offsets, int32_t *d_idx, float *d_dist, int sorted, bool exact64 = false,
                       double radius64 = -1.0);
// tiled kernel (pcp_radius_tile.cu): COUNT when d_counts != nullptr, else FILL
int radius_warp_list_device(pcp_index *ix, const float *d_q, int64_t nq, int64_t stride, const uint32_t *d_perm,
                            const uint32_t *d_list, const RadiusTest &d2max, long long *d_counts, const long long *d_offsets,
                            int32_t *d_idx, float *d_dist);
int radius_tile_device(pcp_index *ix, const float *d_q, int64_t nq, int64_t stride, const uint32_t *d_perm, const RadiusTest &d2max,
                       long long *d_counts, const long long *d_offsets, int32_t *d_idx, float *d_dist);

// brute force over the whole cloud (pcp_brute.cu): the on-device exhaustive cross-check
float exact_d2_threshold(float r);
int knn_brute_device(pcp_index *ix, const float *d_q, int64_t nq, int64_t stride, int k, int32_t *d_idx, float *d_dist,
                     int32_t *d_count);
int radius_count_brute_device(pcp_index *ix, const float *d_q, int64_t nq, int64_t stride, float d2max, long long *d_counts);

// global initialisation (pcp_global.cu, pcp_global_ransac.cu)
int normals_hybrid_device(pcp_index *ix, double radius, int max_nn, double *d_normals);
int fpfh_device(pcp_index *ix, const double *d_normals, double radius, int max_nn, double *d_fpfh);
int feature_match_device(pcp_ctx *ctx, const double *d_src, int64_t ns, const double *d_tgt, int64_t nt, int dim,
                         int32_t *d_idx, double *d_d2);
int ransac_device(pcp_ctx *ctx, const pcp_ransac_params &prm, const float *d_src, int64_t ns, int64_t stride,
                  pcp_index *tgt, const int32_t *d_corr, int64_t n_corr, double *h_T, double *h_fitness, double *h_rmse,
                  int64_t *h_iterations, int64_t *h_evaluated);

// PointNet++ neighbourhood layouts (pcp_pointnet.cu)
int ball_query_device(pcp_ctx *ctx, int b, int n, int m, float radius, int nsample, const float *d_new_xyz,
                      const float *d_xyz, int32_t *d_idx);
int three_nn_device(pcp_ctx *ctx, int b, int n, int m, const float *d_unknown, const float *d_known, float *d_dist2,
                    int32_t *d_idx);

// icp (pcp_icp.cu)
int icp_batch_device(pcp_ctx *ctx, const pcp_icp_params &prm, int64_t n_pairs, const float *d_src,
                     const int64_t *h_src_off, const float *d_tgt, const int64_t *h_tgt_off, int64_t stride,
                     const double *h_T_init, double *h_T_last, double *h_T_total, int32_t *h_iters,
                     int32_t *h_status, double *h_log, cudaEvent_t src_ready = nullptr);

// ICP front end (pcp_front.cu)
int voxel_filter_device(pcp_ctx *ctx, const float *d_xyz, int64_t n, int64_t stride, double leaf, int mode,
                        unsigned long long seed, int64_t *n_out);
int normals_device(pcp_index *ix, int k, double *d_normals, double *d_evals);
int iss_device(pcp_index *ix, double radius, double lambda21, double lambda32, double nms_radius, int iss_count,
               int32_t *h_keypoints, int32_t *h_n_keypoints, double *d_lambda3_out);
// DBSCAN on the radius CSR (pcp_dbscan.cu)
int dbscan_device(pcp_index *ix, double radius, int min_pts, int32_t *d_labels, int32_t *h_n_clusters);
int pca_device(pcp_ctx *ctx, const float *d_xyz, int64_t n, int64_t stride, double *h_w, double *h_V);

int procrustes_device(pcp_ctx *ctx, const double *d_A, const double *d_B, int64_t n, double *h_R, double *h_t,
                      double *h_cost);

}  // namespace pcp
