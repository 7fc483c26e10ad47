/*
 * pcp.h -- C ABI of the B200-native nearest-neighbour + ICP library (libpcp_b200.so).
 *
 * This is the drop-in boundary for the one hot path of ruanyyyyyyy/point-cloud-process:
 * exact kNN / radius search (Kdtree_Octree/lesson2) and the point-to-point ICP loop
 * (Registration/main.py, icp_template.py).  The reference has NO FFI on this path -- it is pure
 * Python over numpy/Open3D -- so every entry point below cites the reference Python interface it
 * replaces (file:line relative to the reference root).  The pybind11 module `_pcp`
 * (point_cloud_process_b200/csrc/bindings.cpp, the PCLKeypoints/src/keypoints.cpp:240-298
 * packaging pattern) and the Python shims call ONLY these functions.
 *
 * Conventions
 *   - plain pointers + sizes, no C++ / torch types; opaque handles; int status return
 *     (0 = PCP_OK, negative = error, message via pcp_last_error(), thread-local).
 *   - clouds are row-per-point float32 (x, y, z) with a row stride in floats (3 for packed
 *     N x 3 -- the layout benchmark.py:19 documents -- 4 for KITTI x,y,z,intensity rows
 *     (benchmark.py:24), 6 for the Oxford xyz+normal rows of main.py:14 / icp_template.py:16).
 *   - `mem` says where ALL data pointers of that call live: PCP_HOST (pageable or pinned host
 *     memory; the library stages through pinned buffers and copies inside the call) or
 *     PCP_DEVICE (device memory of the context's GPU; nothing is copied, the call is
 *     asynchronous on the context stream -- use pcp_ctx_synchronize()).
 *   - thread-compatible: one context per host thread; contexts are independent.
 *   - there is no CPU fallback: every call fails with PCP_ERR_CUDA when no sm_100 device is
 *     usable.
 */
#ifndef PCP_H_
#define PCP_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define PCP_API __attribute__((visibility("default")))
#else
#define PCP_API
#endif

#define PCP_VERSION 100 /* 0.1.0 */

enum pcp_status {
    PCP_OK = 0,
    PCP_ERR_INVALID = -1,     /* bad argument (NULL, negative size, k out of range, dim != 3 ...) */
    PCP_ERR_CUDA = -2,        /* CUDA runtime / launch error, or no usable device */
    PCP_ERR_OOM = -3,         /* device or pinned-host allocation failed */
    PCP_ERR_UNSUPPORTED = -4  /* valid request the library does not implement (k > PCP_MAX_K) */
};

enum pcp_mem { PCP_HOST = 0, PCP_DEVICE = 1 };

#define PCP_MAX_K 32           /* register top-k width; the reference configs use k = 1, 8, 16 */
#define PCP_KNN_EMPTY_DIST 1e10f /* result_set.py:19-22: unfilled KNNResultSet slots are (1e10, 0) */

typedef struct pcp_ctx pcp_ctx;       /* one GPU + one stream + scratch memory */
typedef struct pcp_index pcp_index;   /* search structure over one cloud (replaces the kd-tree / octree root) */

/* ---- library ------------------------------------------------------------------------------ */
PCP_API int pcp_version(void);
PCP_API const char *pcp_last_error(void);
PCP_API int pcp_device_count(int *count);

/* ---- pinned host buffers ---------------------------------------------------------------------
 * Page-locked host memory from a process-wide pool (size classes, cached on free).  PCP_HOST calls
 * whose input / output pointers are page-locked (from here or from any cudaHostAlloc / torch
 * pin_memory) are copied by DMA straight from / to the caller's buffer; pageable pointers are staged.
 * The pybind11 layer returns its result arrays in such buffers (owned by Python, returned to the pool
 * when the array is collected -- the "outputs owned by Python" rule of keypoints.cpp:240-298). */
PCP_API int pcp_host_alloc(size_t bytes, void **out);
PCP_API int pcp_host_free(void *ptr);

/* ---- context ------------------------------------------------------------------------------- */
PCP_API int pcp_ctx_create(int device, pcp_ctx **out);
PCP_API int pcp_ctx_destroy(pcp_ctx *ctx);
PCP_API int pcp_ctx_synchronize(pcp_ctx *ctx);
/* cudaStream_t of the context as an opaque pointer (for callers that enqueue their own work). */
PCP_API void *pcp_ctx_stream(pcp_ctx *ctx);
/* Number of kernels this library has launched on the context since creation (bench bookkeeping). */
PCP_API int64_t pcp_ctx_launch_count(pcp_ctx *ctx);

/* Device time of the dominant search kernel (the tiled kNN kernel) alone: when enabled, every launch of it is
 * bracketed by CUDA events on the context stream; pcp_ctx_read_kernel_timing synchronises, returns the summed
 * milliseconds and the number of launches since the last read, and resets.  Bench bookkeeping (roofline). */
PCP_API int pcp_ctx_enable_kernel_timing(pcp_ctx *ctx, int enable);
PCP_API int pcp_ctx_read_kernel_timing(pcp_ctx *ctx, double *total_ms, int64_t *launches);

/* Work counters -- the analogue of the reference's comparison_counter (result_set.py:24,36,70,79).
 * When enabled, search kernels accumulate into 8 device-side 64-bit words, read back (and reset)
 * by pcp_ctx_read_counters:
 *   [0] distance evaluations (query, point) pairs   [1] bucket lower-bound tests (per group / query)
 *   [2] hierarchy nodes opened                      [3] top-k drain rounds
 *   [4] query groups (kNN) or queries (radius)      [5] buckets scanned
 *   [6] query groups handed to the warp-per-query kernel   [7] the part of [0] evaluated by that kernel
 * Counting costs a few instructions per bucket; it is OFF by default and off in timed runs. */
PCP_API int pcp_ctx_enable_counters(pcp_ctx *ctx, int enable);
PCP_API int pcp_ctx_read_counters(pcp_ctx *ctx, uint64_t out[8]);

/* ---- CUDA graphs --------------------------------------------------------------------------------------------
 * A fixed sequence of PCP_DEVICE calls on one context (e.g. pcp_index_rebuild + pcp_knn of one LiDAR frame: ~20
 * kernel launches on two streams) captured once and replayed with ONE launch per step: the device pointers and
 * shapes of the captured calls are baked in, the DATA behind them may change between launches.  The calls
 * between begin and end are recorded, not executed; run the same sequence once before capturing (the library's
 * scratch buffers must already have their size -- a capture in which one had to grow is refused).
 * PCP_HOST calls cannot be captured (they synchronise).  pcp_graph_launch is asynchronous on the context stream. */
typedef struct pcp_graph pcp_graph;
PCP_API int pcp_graph_begin(pcp_ctx *ctx);
PCP_API int pcp_graph_end(pcp_ctx *ctx, pcp_graph **out);
PCP_API int pcp_graph_launch(pcp_graph *graph);
PCP_API int64_t pcp_graph_nodes(const pcp_graph *graph);
PCP_API int pcp_graph_destroy(pcp_graph *graph);

/* ---- index build ---------------------------------------------------------------------------
 * Replaces kdtree_construction(db_np, leaf_size)            kdtree.py:119-137
 *      and octree_construction(db_np, leaf_size, min_extent) octree.py:310-328.
 * Builds the Morton-sorted 32-point bucket hierarchy on the GPU.
 * leaf_size / min_extent of the reference have no meaning here and are not parameters.
 * n may be 0 (every search then returns empty results). */
PCP_API int pcp_index_build(pcp_ctx *ctx, const float *xyz, int64_t n, int64_t stride_floats, int mem,
                    pcp_index **out);
/* Re-index a new cloud into an existing handle, reusing its device buffers. */
PCP_API int pcp_index_rebuild(pcp_index *index, const float *xyz, int64_t n, int64_t stride_floats, int mem);
PCP_API int pcp_index_destroy(pcp_index *index);
PCP_API int64_t pcp_index_size(const pcp_index *index);
/* Device-resident internals for diagnostics / benches: number of buckets and hierarchy levels. */
PCP_API int pcp_index_stats(const pcp_index *index, int64_t *n_buckets, int *n_levels);

/* float64 originals of the rows the index was built from (n x 3 doubles, row i = point i): the library's search
 * structure is float32, but the consumers whose reference works in float64 (pcp_iss_keypoints, pcp_dbscan: scipy
 * cKDTree on the float64 cloud) decide borderline neighbour pairs on these.  Dropped by the next rebuild. */
PCP_API int pcp_index_attach_f64(pcp_index *index, const double *xyz, int64_t n, int64_t stride_doubles, int mem);

/* ---- k nearest neighbours --------------------------------------------------------------------
 * Replaces kdtree_knn_search(root, db, result_set, query)   kdtree.py:141-172
 *      and octree_knn_search(root, db, result_set, query)   octree.py:262-306
 * with KNNResultSet(capacity=k) semantics                   result_set.py:16-60,
 * batched over nq queries.  For query i:
 *   idx [i*k .. i*k+k)  indices into the indexed cloud, ascending (distance, index);
 *   dist[i*k .. i*k+k)  float32 EUCLIDEAN distance sqrtf((dx*dx+dy*dy)+dz*dz), bit-equal to
 *                       np.linalg.norm(query - db[idx]) of kdtree.py:157-158;
 *   count[i]            = min(k, n)  (KNNResultSet.count); slots >= count hold
 *                       (PCP_KNN_EMPTY_DIST, 0) as result_set.py:19-22 leaves them.
 * Tie rule: equal float32 distances are ordered by ascending index (the reference's order among
 * exact ties depends on tree visit order, result_set.py:37,45).
 * 1 <= k <= PCP_MAX_K.  count may be NULL. */
PCP_API int pcp_knn(pcp_index *index, const float *queries, int64_t nq, int64_t stride_floats, int k, int mem,
            int32_t *idx, float *dist, int32_t *count);

/* Same search with the indexed cloud itself as the query set, in the cloud's original order
 * (the benchmark.py:122 `query = db_np[0,:]` case generalised to every point).  Skips the query
 * sort because the buckets already are the query groups. */
PCP_API int pcp_knn_self(pcp_index *index, int k, int mem, int32_t *idx, float *dist, int32_t *count);

/* Brute force over the WHOLE indexed cloud -- the check the reference prints beside its tree searches,
 * np.argsort(np.linalg.norm(db - query, axis=1))[:k]            kdtree.py:234-238, octree.py:352-356 --
 * as a device kernel (tiles of the cloud streamed through shared memory by TMA bulk copies, every (query, point)
 * pair evaluated, no spatial pruning).  Same outputs and tie rule as pcp_knn.  It shares no traversal code with
 * the search kernels: the test-suite uses it as the on-device second oracle for full-size query sets. */
PCP_API int pcp_knn_bruteforce(pcp_index *index, const float *queries, int64_t nq, int64_t stride_floats, int k, int mem,
                               int32_t *idx, float *dist, int32_t *count);

/* pcp_knn_bruteforce for callers that cannot allocate device memory (code under stream capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * No context and no index: points, queries and results are DEVICE pointers of the caller; the kernel's
 * scratch (the pair-layout copy of the cloud) is a static device array of PCP_NOALLOC_MAX_POINTS x 16 B = 4 MiB that
 * every GPU loading the library holds until exit.  The call performs no cudaMalloc / cudaFree and enqueues two
 * kernels on `stream` (a cudaStream_t, NULL = the default stream) without synchronising.  n > PCP_NOALLOC_MAX_POINTS
 * fails with PCP_ERR_UNSUPPORTED.  Outputs are identical to pcp_knn_bruteforce (canonical (distance, index) order,
 * unfilled slots (PCP_KNN_EMPTY_DIST, 0)); count may be NULL.  All calls on a device share the array, which keeps
 * the last call's contents (every slot read is rewritten first, nothing has to be cleared): calls that overlap on
 * the GPU -- different streams or host threads -- race and must be serialised by the caller. */
#define PCP_NOALLOC_MAX_POINTS 262144
PCP_API int pcp_knn_bruteforce_noalloc(int device, const float *d_points, int64_t n, int64_t point_stride_floats,
                                       const float *d_queries, int64_t nq, int64_t query_stride_floats, int k,
                                       int32_t *d_idx, float *d_dist, int32_t *d_count, void *stream);
/* counts[i] = number of indexed points with float32 dist <= radius, by the same exhaustive kernel (the count
 * RadiusNNResultSet.size() reports, result_set.py:72-73). */
PCP_API int pcp_radius_count_bruteforce(pcp_index *index, const float *queries, int64_t nq, int64_t stride_floats,
                                        float radius, int mem, int64_t *counts);

/* ---- fixed-radius search -----------------------------------------------------------------------
 * Replaces kdtree_radius_search                              kdtree.py:176-208
 *      and octree_radius_search / octree_radius_search_fast  octree.py:166-259
 * with RadiusNNResultSet(radius) semantics                   result_set.py:63-93:
 * every indexed point with NOT (dist > radius), i.e. float32 dist <= radius, boundary included.
 * Two-phase CSR sizing: pcp_radius_count writes offsets[0..nq] (exclusive prefix sum of the
 * per-query counts, offsets[nq] = total M); the caller allocates M entries and calls
 * pcp_radius_fill with the same queries/radius.  Within a query the order is unspecified unless
 * sorted != 0, which orders by ascending (distance, index) (what RadiusNNResultSet.__str__,
 * result_set.py:87, prints). */
PCP_API int pcp_radius_count(pcp_index *index, const float *queries, int64_t nq, int64_t stride_floats,
                     float radius, int mem, int64_t *offsets);
PCP_API int pcp_radius_fill(pcp_index *index, const float *queries, int64_t nq, int64_t stride_floats,
                    float radius, int mem, const int64_t *offsets, int32_t *idx, float *dist,
                    int sorted);

/* ---- point-to-point ICP, batched over independent pairs ------------------------------------------
 * Replaces icp_point2point(source, target, transformation)   main.py:97-156
 *      and the ICP / find_associations / procrustes_transformation contract of
 *      icp_template.py:43-54,113-126,128-200.
 * Every iteration (transform, 1-NN correspondence, d^2 < dist_thres gate, centroid + 3x3
 * cross-covariance, SVD, R = U V^T, t, convergence test) runs on the device. */
typedef struct pcp_icp_params {
    int32_t max_iteration;      /* main.py:98   100 */
    double R_diff_thres;        /* main.py:101  0.5  (Frobenius norm of R - R_last) */
    double t_diff_thres;        /* main.py:102  0.5 */
    double dist_thres;          /* main.py:103  5    (compared with the SQUARED 1-NN distance, main.py:119) */
    int32_t first_iter_broadcast; /* 1 = reproduce main.py:150 on iteration 1, where t (3,1) minus
                                     t_last (3,) broadcasts to 3x3 before the norm; 0 = plain ||t - t_last|| */
} pcp_icp_params;

enum pcp_icp_status_code {
    PCP_ICP_CONVERGED = 0,       /* stopped by the R_diff / t_diff test (main.py:153-154) */
    PCP_ICP_MAX_ITER = 1,        /* ran max_iteration iterations */
    PCP_ICP_FEW_PAIRS = 2        /* fewer than 3 associations (main.py:125-127): T_last is the last applied transform */
};

PCP_API int pcp_icp_default_params(pcp_icp_params *p);

/* src_xyz / tgt_xyz: packed rows (stride_floats each) of all pairs back to back;
 * src_offsets / tgt_offsets: n_pairs+1 row offsets (int64, host memory always).
 * T_init: n_pairs x 16 row-major float64 (NULL = identity).
 * Outputs (host memory always; any may be NULL):
 *   T_last  n_pairs x 16  the LAST INCREMENT -- what icp_point2point returns (main.py:156)
 *   T_total n_pairs x 16  composed pose source->target incl. the last increment
 *                         (homo_mat_total of icp_template.py:136,198)
 *   iterations, status    per pair
 *   diff_log n_pairs x max_iteration x 2  (R_diff, t_diff) per iteration, NaN beyond `iterations`
 *                         (the `log` dict of icp_template.py:161,186-187) */
PCP_API int pcp_icp_batch(pcp_ctx *ctx, const pcp_icp_params *params, int64_t n_pairs,
                  const float *src_xyz, const int64_t *src_offsets,
                  const float *tgt_xyz, const int64_t *tgt_offsets, int64_t stride_floats, int mem,
                  const double *T_init, double *T_last, double *T_total, int32_t *iterations,
                  int32_t *status, double *diff_log);

/* Bookkeeping of the last pcp_icp_batch on this context: kernels launched by its iteration loop, host-blocking
 * synchronisations (1 = only the final result copy), iterations enqueued (blocks of 4, 8, 16, ... run without
 * a host read in between; a converged pair's CTAs return at once). */
PCP_API int pcp_icp_last_stats(pcp_ctx *ctx, int64_t *launches, int64_t *host_syncs, int64_t *iterations_enqueued);

/* procrustes_transformation(A, B) -> (R, t, cost)               main.py:131-141, icp_template.py:43-54
 * on GIVEN correspondences: A, B are 3 x n float64 row-major (one column per point, the reference's
 * layout).  R: 9 row-major doubles = U V^T of the SVD of (B - mean)(A - mean)^T (no det fix, main.py:139);
 * t: 3 doubles = mean(B - R A); cost = ||B - (R A + t 1^T)||_F (main.py:141).  Outputs are host memory. */
PCP_API int pcp_procrustes(pcp_ctx *ctx, const double *A, const double *B, int64_t n, int mem, double *R, double *t,
                   double *cost);

/* ---- ICP front end: voxel-grid down-sampling and normals -----------------------
 * pcp_voxel_filter replaces voxel_filter(point_cloud, leaf_size, type)   Pca_and_Voxel_filter/voxel_filter.py:10-66
 * (the hash-grid precedent for the voxel_down_sample call of Registration/main.py:35), float64 arithmetic on the
 * float32 rows (the reference is handed Open3D's float64 points), reference quirks kept:
 *   D = (max - min) // leaf_size per axis (:22-24); h = hx + hy*Dx + hz*Dx*Dy (:29-33, colliding cells merge);
 *   stable sort by h (:36); one point per run of equal h in ascending h; the LAST run is not emitted (:40-50).
 * PCP_VOXEL_CENTROID: np.mean of the run in input order, numpy's pairwise summation order -- bit-equal (:48).
 * PCP_VOXEL_RANDOM:   one point of the run, chosen by a counter hash of (seed, run) -- the reference draws from
 *                     Python's global random state (:60), so only the run structure is comparable.
 * Two-phase: pcp_voxel_filter computes into a context-owned device buffer and reports n_out rows;
 * pcp_voxel_filter_read copies those n_out x 3 float64 rows out (host or device per `mem`). */
enum pcp_voxel_mode { PCP_VOXEL_CENTROID = 0, PCP_VOXEL_RANDOM = 1 };
PCP_API int pcp_voxel_filter(pcp_ctx *ctx, const float *xyz, int64_t n, int64_t stride_floats, double leaf_size,
                             int mode, uint64_t seed, int mem, int64_t *n_out);
PCP_API int pcp_voxel_filter_read(pcp_ctx *ctx, int mem, double *out_xyz);

/* pcp_normals replaces the per-point loop of Pca_and_Voxel_filter/pca_normal.py:85-90:
 *   search_knn_vector_3d(p, k) (self included) -> PCA(pc_array[idx, :]) (pca_normal.py:10-36: mean, np.cov with
 *   1/(k-1), eigh, eigenvalues descending) -> normal = eigenvector of the smallest eigenvalue (v[:, 2]).
 * The neighbours are the library's exact float32 kNN of the indexed cloud (pcp_knn_self).  normals: n x 3
 * float64, unit length, sign fixed so that the component of largest magnitude is positive (numpy leaves the
 * sign to LAPACK); eigenvalues (may be NULL): n x 3 float64, descending.  2 <= k <= PCP_MAX_K (the reference
 * uses 5).  For an index built with PCP_DEVICE the cloud passed to the build must still be alive. */
PCP_API int pcp_normals(pcp_index *index, int k, int mem, double *normals, double *eigenvalues);

/* pcp_pca replaces PCA(data, correlation=False, sort=True) of pca_normal.py:10-36 on a whole cloud: mean, centred
 * covariance with 1/(n-1) (np.cov), symmetric eigen-decomposition (np.linalg.eigh), all float64 on the device.
 * eigenvalues: 3 doubles DESCENDING (the sort=True order, pca_normal.py:30-33); eigenvectors: 9 doubles, row-major
 * 3x3 whose COLUMN j belongs to eigenvalues[j], sign: component of largest magnitude positive.  Outputs are host
 * memory.  n >= 2. */
PCP_API int pcp_pca(pcp_ctx *ctx, const float *xyz, int64_t n, int64_t stride_floats, int mem, double *eigenvalues,
                    double *eigenvectors);

/* ---- ISS keypoints on the radius search --------------------------------------------
 * pcp_iss_keypoints replaces the detection loop of Keypoint_detection_ISS/ISS.py:40-73 over the indexed cloud:
 * weights 1/|ball(p_j, radius)| (:49), weighted scatter matrix of the radius neighbourhood (:43-51, float64),
 * eigenvalues descending, candidate iff l2/l1 < lambda21 and l3/l2 < lambda32 (:52-57), non-maximum suppression
 * with non_max_radius in descending l3 (ties: ascending index), stopping once MORE than iss_count keypoints are
 * kept (:59-73; so at most iss_count + 1 indices are written -- size `keypoints` accordingly).
 * Neighbourhoods follow scipy's cKDTree, which the reference searches with (ISS.py:40): the radius search's float32
 * candidates are re-decided in FLOAT64 wherever the float32 distance is within its rounding band of the radius
 * (squared differences summed x, y, z against radius * radius) -- on the float64 originals of the rows when they
 * were attached with pcp_index_attach_f64, else on the float32 rows widened exactly.
 * keypoints / n_keypoints: host memory; lambda3 (host, may be NULL): n doubles, l3 of every candidate, NaN elsewhere.
 * For an index built with PCP_DEVICE the cloud passed to the build must still be alive. */
PCP_API int pcp_iss_keypoints(pcp_index *index, double radius, double lambda21, double lambda32, double non_max_radius,
                              int iss_count, int32_t *keypoints, int32_t *n_keypoints, double *lambda3);

/* counts[i] = number of indexed points within `radius` of indexed point i (itself included), decided as scipy's
 * float64 cKDTree decides it -- the neighbourhood sizes behind the ISS weights (ISS.py:49) and the DBSCAN core test
 * (dbscan.py:20-24): float32 candidates, float64 decision in the rounding band (see pcp_iss_keypoints).  Host output. */
PCP_API int pcp_neighbor_counts_f64(pcp_index *index, double radius, int64_t *counts);

/* pcp_dbscan replaces DBSCAN(radius, Min_Pts).fit(data) + predict() of Cluster_dbscan/dbscan.py:4-38 over the
 * indexed cloud: labels[i] = cluster number (0, 1, ... in the order the reference opens them: seeds are popped
 * from the END of its `unvisited` list, :18) or -1 for noise -- the SAME labels as the sequential loop, including
 * its order-dependent rules: a popped point with fewer than Min_Pts neighbours (self included) stays noise for
 * good (:20-22), a point with >= Min_Pts opens a cluster (:24), a claimed point expands it only with MORE than
 * Min_Pts neighbours (:31-32).  Computed on the radius CSR as connected components of the expanding points plus a
 * rank-ordered fixed point for the rest.  Neighbourhoods follow scipy's float64 KDTree
 * (dbscan.py:13): borderline pairs of the float32 search are re-decided in float64 (see pcp_iss_keypoints).  labels: n x int32 (host or device per `mem`); n_clusters (host, may be NULL).  For an index built
 * with PCP_DEVICE the cloud passed to the build must still be alive. */
PCP_API int pcp_dbscan(pcp_index *index, double radius, int min_pts, int mem, int32_t *labels, int32_t *n_clusters);

/* ---- multi-GPU: one process per GPU, work sharded where the path shards -------------------------
 * The reference has no distributed code on this path (single process, single thread); these entries are the
 * B200-native extension BASELINE.json configs 4 and 5 name.
 *   registration pairs: independent -> a contiguous range of pairs per rank (balanced by points), NO collective.
 *   queries of a large cloud: split per rank; the reference cloud is REPLICATED -- every rank contributes a shard,
 *     ONE ncclAllGather (NVLink / NVSwitch) rebuilds the whole cloud on every GPU, every rank builds the same
 *     (deterministic) index and searches its own queries.  This all-gather is the only collective of the library.
 * NCCL is bound at run time (dlopen libnccl.so.2); world == 1 needs no NCCL at all.
 * Bootstrap: rank 0 calls pcp_multi_unique_id and hands the PCP_MULTI_ID_BYTES bytes to every rank by any means
 * (a TCP socket, a file, the launcher's store); every rank then calls pcp_multi_create with the same id. */
#define PCP_MULTI_ID_BYTES 128
typedef struct pcp_multi pcp_multi;
PCP_API int pcp_multi_unique_id(void *id128);
PCP_API int pcp_multi_create(pcp_ctx *ctx, int rank, int world, const void *id128, pcp_multi **out);
PCP_API int pcp_multi_destroy(pcp_multi *m);
PCP_API int pcp_multi_info(const pcp_multi *m, int *rank, int *world, int *nccl_version, int *comm_nranks);
/* This rank's shard of the reference cloud (shards may differ in length) -> all-gather (rank order) -> index build
 * over the gathered cloud.  n_total (may be NULL): rows of the gathered cloud.  Collective: every rank must call. */
PCP_API int pcp_multi_set_reference(pcp_multi *m, const float *shard_xyz, int64_t n_shard, int64_t stride_floats,
                                    int mem, int64_t *n_total);
/* Device time of the last pcp_multi_set_reference: the all-gather (incl. gap closing for ragged shards) and the
 * index build, CUDA events on the context stream. */
PCP_API int pcp_multi_last_timing(const pcp_multi *m, double *allgather_ms, double *build_ms);
/* The gathered cloud (device pointer, rows, stride) and the rows every rank contributed (world entries). */
PCP_API int pcp_multi_reference(const pcp_multi *m, const float **device_xyz, int64_t *n_total, int64_t *stride_floats,
                                int64_t *shard_rows);
PCP_API pcp_index *pcp_multi_index(pcp_multi *m);
/* kNN of THIS rank's queries against the gathered cloud (indices = rows of the gathered cloud): pcp_knn contract. */
PCP_API int pcp_multi_knn(pcp_multi *m, const float *queries, int64_t nq, int64_t stride_floats, int k, int mem,
                          int32_t *idx, float *dist, int32_t *count);
/* Contiguous pair range [first, last) of `rank`: boundaries at equal shares of the total point count. Host only. */
PCP_API int pcp_multi_pair_range(const int64_t *src_offsets, const int64_t *tgt_offsets, int64_t n_pairs, int rank,
                                 int world, int64_t *first, int64_t *last);
/* pcp_icp_batch over this rank's range of a batch every rank describes identically (all arrays are the GLOBAL ones;
 * PCP_HOST: only the rank's own rows are read / uploaded).  Output rows outside [first_pair, last_pair) are not
 * touched; no communication. */
PCP_API int pcp_multi_icp_batch(pcp_multi *m, const pcp_icp_params *params, int64_t n_pairs, const float *src_xyz,
                                const int64_t *src_offsets, const float *tgt_xyz, const int64_t *tgt_offsets,
                                int64_t stride_floats, int mem, const double *T_init, double *T_last, double *T_total,
                                int32_t *iterations, int32_t *status, double *diff_log, int64_t *first_pair,
                                int64_t *last_pair);

/* ---- global initialisation in front of ICP: normals, FPFH, feature matching, RANSAC ---------------
 * The Open3D calls of Registration/main.py:33-47,68-84 and the contract of icp_template.py:20-41,56-111.  Open3D is
 * absent and unpinned in the reference: these restate its published algorithms; parity "unpinned".
 * pcp_normals_hybrid   estimate_normals(KDTreeSearchParamHybrid(radius, max_nn))             main.py:37-39
 *     neighbourhood = the up-to-max_nn nearest indexed points within radius (self included); normal = eigenvector of
 *     the smallest eigenvalue of its covariance, (0,0,1) below 3 neighbours; sign: largest component positive.
 * pcp_fpfh             compute_fpfh_feature(pcd, KDTreeSearchParamHybrid(radius, max_nn))    main.py:43-46
 *     normals: n x 3 float64 of the indexed cloud; features: n x 33 float64 (row per point; Open3D stores 33 x n).
 * pcp_feature_match    nearest target feature of every source feature (exhaustive, float64, lowest index wins):
 *     the search inside registration_ransac_based_on_feature_matching / find_matchings      icp_template.py:20-41
 * pcp_ransac_init      registration_ransac_based_on_feature_matching(..., ransac_n = 3, [EdgeLength, Distance],
 *     RANSACConvergenceCriteria(max_iteration, confidence))                                  main.py:74-83
 *     correspondences: m x 2 int32 (source row, target row); target_index: the index over the target cloud (its
 *     cloud must still be alive); T: 16 doubles row-major source -> target; fitness = inliers / n_src with
 *     inlier = nearest target point within max_correspondence_distance; inlier_rmse over the inliers.
 *     Hypotheses are drawn from a counter hash of (seed, iteration): the result is reproducible. */
typedef struct pcp_ransac_params {
    double max_correspondence_distance;   /* main.py:71   voxel_size * 1.5 */
    double edge_length_threshold;         /* main.py:78   0.9 */
    double confidence;                    /* main.py:83   0.999 */
    int32_t max_iteration;                /* main.py:83   100000 */
    int32_t ransac_n;                     /* main.py:77   3 (the only supported value) */
    uint64_t seed;
} pcp_ransac_params;
PCP_API int pcp_normals_hybrid(pcp_index *index, double radius, int max_nn, int mem, double *normals);
PCP_API int pcp_fpfh(pcp_index *index, const double *normals, double radius, int max_nn, int mem, double *features);
PCP_API int pcp_feature_match(pcp_ctx *ctx, const double *src_features, int64_t n_src, const double *tgt_features,
                              int64_t n_tgt, int dim, int mem, int32_t *nearest, double *dist2);
PCP_API int pcp_ransac_init(pcp_ctx *ctx, const pcp_ransac_params *params, const float *src_xyz, int64_t n_src,
                            int64_t stride_floats, pcp_index *target_index, const int32_t *correspondences,
                            int64_t n_corr, int mem, double *T, double *fitness, double *inlier_rmse,
                            int64_t *iterations, int64_t *hypotheses_evaluated);

/* ---- PointNet++ neighbourhood layouts ----------------------------------------------------------
 * The two search ops of the reference's own CUDA extension, same layouts and selection rules, packed float32 rows:
 * pcp_ball_query replaces query_ball_point_kernel        _ext-src/src/ball_query_gpu.cu:9-46 (ball_query.cpp:9-34)
 *   new_xyz (b, m, 3) centres, xyz (b, n, 3) points -> idx (b, m, nsample) int32: the FIRST nsample points in index
 *   order with d2 < radius * radius; the tail of a short row repeats its first hit; a centre without a hit keeps 0s.
 * pcp_three_nn replaces three_nn_kernel                  _ext-src/src/interpolate_gpu.cu:9-60
 *   unknown (b, n, 3), known (b, m, 3) -> dist2 (b, n, 3) SQUARED distances ascending, idx (b, n, 3); strict '<',
 *   the lowest index wins a tie; slots without a point hold (+inf, 0).
 * d2 is the reference's float32 expression as nvcc compiles it: fma(dz, dz, fma(dx, dx, dy * dy)). */
PCP_API int pcp_ball_query(pcp_ctx *ctx, int b, int n, int m, float radius, int nsample, const float *new_xyz,
                           const float *xyz, int mem, int32_t *idx);
PCP_API int pcp_three_nn(pcp_ctx *ctx, int b, int n, int m, const float *unknown, const float *known, int mem,
                         float *dist2, int32_t *idx);

/* Pose helpers (host side, no GPU): rotmat2quaternion / homo2tq of main.py:158-174.
 * homo: 16 row-major doubles; tq: tx,ty,tz,qw,qx,qy,qz. */
PCP_API int pcp_homo2tq(const double *homo, double *tq);

#ifdef __cplusplus
}
#endif
#endif /* PCP_H_ */
