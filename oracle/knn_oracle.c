/*
 * oracle/knn_oracle.c -- TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement (plain C, float32, no FMA contraction) of the reference's exact
 * nearest-neighbour path.  It is the checker for the CUDA product and the "port" CPU
 * baseline of bench.py; nothing under point_cloud_process_b200/ may link or call it.
 *
 * Reference files restated here (paths relative to /root/reference):
 *   Kdtree_Octree/lesson2/result_set.py:16-60   KNNResultSet      -> oracle_knn_add()
 *   Kdtree_Octree/lesson2/result_set.py:63-93   RadiusNNResultSet -> radius append below
 *   Kdtree_Octree/lesson2/kdtree.py:59-97       kdtree_recursive_build
 *   Kdtree_Octree/lesson2/kdtree.py:119-137     kdtree_construction
 *   Kdtree_Octree/lesson2/kdtree.py:141-172     kdtree_knn_search
 *   Kdtree_Octree/lesson2/kdtree.py:176-208     kdtree_radius_search
 *   Kdtree_Octree/lesson2/octree.py:51-103,310-328  octree build
 *   Kdtree_Octree/lesson2/octree.py:106-164     inside / overlaps / contains
 *   Kdtree_Octree/lesson2/octree.py:216-306     octree_radius_search / octree_knn_search
 *   leaf distance (kdtree.py:157-158): np.linalg.norm(q[None]-P, axis=1) in float32 ==
 *       sqrtf((dx*dx + dy*dy) + dz*dz), dx = q - p, every operation rounded separately.
 *
 * Parity pin: tests/test_oracle_golden.py checks every entry point against vectors produced
 * by the reference's own Python modules (tests/golden/make_golden_knn.py, run in the build
 * container where /root/reference is importable).
 *
 * Build: gcc -O2 -ffp-contract=off -fopenmp -shared -fPIC  (see oracle/build_oracle.py)
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORACLE_API __attribute__((visibility("default")))

/* ---------------------------------------------------------------- distance */

/* kdtree.py:157-158 / octree.py:276-277: float32 Euclidean distance, sequential add order. */
static inline float ref_dist(const float *q, const float *p) {
    volatile float dx = q[0] - p[0];
    volatile float dy = q[1] - p[1];
    volatile float dz = q[2] - p[2];
    volatile float xx = dx * dx;
    volatile float yy = dy * dy;
    volatile float zz = dz * dz;
    volatile float s = xx + yy;
    s = s + zz;
    return sqrtf(s);
}

/* ---------------------------------------------------------------- result sets */

typedef struct {
    int capacity, count;
    float worst;      /* result_set.py:19  worst_dist = 1e10 */
    float *dist;      /* result_set.py:21-22 slots pre-filled (1e10, 0) */
    int64_t *index;
    int64_t comparisons;
} knn_set;

static void knn_set_init(knn_set *s, int k, float *dist, int64_t *index) {
    s->capacity = k; s->count = 0; s->worst = 1e10f; s->dist = dist; s->index = index;
    s->comparisons = 0;
    for (int i = 0; i < k; ++i) { dist[i] = 1e10f; index[i] = 0; }
}

/* result_set.py:35-53 */
static void oracle_knn_add(knn_set *s, float d, int64_t idx) {
    s->comparisons++;
    if (d > s->worst) return;                       /* :37 */
    if (s->count < s->capacity) s->count++;         /* :40-41 */
    int i = s->count - 1;                           /* :43 */
    while (i > 0) {                                 /* :44-49 */
        if (s->dist[i - 1] > d) { s->dist[i] = s->dist[i - 1]; s->index[i] = s->index[i - 1]; --i; }
        else break;
    }
    s->dist[i] = d; s->index[i] = idx;              /* :51-52 */
    s->worst = s->dist[s->capacity - 1];            /* :53 */
}

typedef struct {
    float radius;
    int64_t count, cap;
    float *dist;
    int64_t *index;
    int64_t comparisons;
} rad_set;

/* result_set.py:78-84: keep iff NOT dist > radius; appended in visit order. */
static void rad_add(rad_set *s, float d, int64_t idx) {
    s->comparisons++;
    if (d > s->radius) return;
    if (s->count == s->cap) {
        s->cap = s->cap ? s->cap * 2 : 64;
        s->dist = (float *)realloc(s->dist, sizeof(float) * (size_t)s->cap);
        s->index = (int64_t *)realloc(s->index, sizeof(int64_t) * (size_t)s->cap);
    }
    s->dist[s->count] = d; s->index[s->count] = idx; s->count++;
}

/* ---------------------------------------------------------------- kd-tree */

typedef struct {
    int axis;
    int is_leaf;
    float value;            /* split value, kdtree.py:85 */
    int64_t left, right;    /* node ids, -1 = None */
    int64_t begin, end;     /* range in perm[] (point_indices of this node) */
} kd_node;

typedef struct {
    const float *db;        /* N x 3 row-major (not owned) */
    int64_t n;
    int leaf_size;
    int64_t *perm;          /* point indices; every node owns a contiguous slice */
    kd_node *nodes;
    int64_t n_nodes, cap_nodes;
} kd_tree;

typedef struct { float v; int64_t idx; } kv;

static int kv_cmp(const void *a, const void *b) {
    const kv *x = (const kv *)a, *y = (const kv *)b;
    if (x->v < y->v) return -1;
    if (x->v > y->v) return 1;
    /* np.argsort (kdtree.py:46) is not stable; ties are broken here by position, which only
     * matters for which side receives equal keys -- search results are unaffected. */
    return (x->idx > y->idx) - (x->idx < y->idx);
}

static int64_t kd_new_node(kd_tree *t) {
    if (t->n_nodes == t->cap_nodes) {
        t->cap_nodes = t->cap_nodes ? t->cap_nodes * 2 : 1024;
        t->nodes = (kd_node *)realloc(t->nodes, sizeof(kd_node) * (size_t)t->cap_nodes);
    }
    return t->n_nodes++;
}

/* kdtree.py:59-97 */
static int64_t kd_build(kd_tree *t, int64_t begin, int64_t end, int axis, kv *scratch) {
    int64_t id = kd_new_node(t);
    kd_node nd; nd.axis = axis; nd.is_leaf = 1; nd.value = 0.f; nd.left = nd.right = -1;
    nd.begin = begin; nd.end = end;
    int64_t n = end - begin;
    if (n > t->leaf_size) {                                     /* :74 */
        /* :76 sort this node's indices by their axis value (sort_key_by_vale, kdtree.py:35-49) */
        int64_t *saved = (int64_t *)malloc(sizeof(int64_t) * (size_t)n);
        memcpy(saved, t->perm + begin, sizeof(int64_t) * (size_t)n);
        for (int64_t i = 0; i < n; ++i) { scratch[i].idx = i; scratch[i].v = t->db[3 * saved[i] + axis]; }
        qsort(scratch, (size_t)n, sizeof(kv), kv_cmp);
        for (int64_t i = 0; i < n; ++i) t->perm[begin + i] = saved[scratch[i].idx];
        free(saved);
        int64_t mid_left = (n + 1) / 2 - 1;                     /* :77 ceil(n/2)-1 */
        int64_t mid_right = mid_left + 1;                       /* :81 */
        float vl = t->db[3 * t->perm[begin + mid_left] + axis];
        float vr = t->db[3 * t->perm[begin + mid_right] + axis];
        volatile float sum = vr + vl;                           /* :85 float32 arithmetic */
        nd.value = sum * 0.5f;
        nd.is_leaf = 0;
        int next = (axis == 2) ? 0 : axis + 1;                  /* :52-56 axis_round_robin, dim=3 */
        t->nodes[id] = nd;
        int64_t l = kd_build(t, begin, begin + mid_right, next, scratch);
        int64_t r = kd_build(t, begin + mid_right, end, next, scratch);
        t->nodes[id].left = l; t->nodes[id].right = r;
        return id;
    }
    t->nodes[id] = nd;
    return id;
}

/* kdtree.py:119-137 */
ORACLE_API void *oracle_kdtree_construction(const float *db, int64_t n, int leaf_size) {
    kd_tree *t = (kd_tree *)calloc(1, sizeof(kd_tree));
    t->db = db; t->n = n; t->leaf_size = leaf_size;
    t->perm = (int64_t *)malloc(sizeof(int64_t) * (size_t)(n > 0 ? n : 1));
    for (int64_t i = 0; i < n; ++i) t->perm[i] = i;
    kv *scratch = (kv *)malloc(sizeof(kv) * (size_t)(n > 0 ? n : 1));
    kd_build(t, 0, n, 0, scratch);
    free(scratch);
    return t;
}

ORACLE_API void oracle_kdtree_destroy(void *h) {
    kd_tree *t = (kd_tree *)h;
    if (!t) return;
    free(t->perm); free(t->nodes); free(t);
}

/* kdtree.py:141-172 */
static void kd_knn(const kd_tree *t, int64_t id, knn_set *rs, const float *q) {
    if (id < 0) return;
    const kd_node *nd = &t->nodes[id];
    if (nd->is_leaf) {
        for (int64_t i = nd->begin; i < nd->end; ++i)
            oracle_knn_add(rs, ref_dist(q, t->db + 3 * t->perm[i]), t->perm[i]);
        return;
    }
    volatile float gap = q[nd->axis] - nd->value;
    if (q[nd->axis] < nd->value) {                              /* :163 */
        kd_knn(t, nd->left, rs, q);
        if (fabsf(gap) < rs->worst) kd_knn(t, nd->right, rs, q); /* :165 */
    } else {
        kd_knn(t, nd->right, rs, q);
        if (fabsf(gap) < rs->worst) kd_knn(t, nd->left, rs, q);  /* :169 */
    }
}

/* kdtree.py:176-208 */
static void kd_radius(const kd_tree *t, int64_t id, rad_set *rs, const float *q) {
    if (id < 0) return;
    const kd_node *nd = &t->nodes[id];
    if (nd->is_leaf) {
        for (int64_t i = nd->begin; i < nd->end; ++i)
            rad_add(rs, ref_dist(q, t->db + 3 * t->perm[i]), t->perm[i]);
        return;
    }
    volatile float gap = q[nd->axis] - nd->value;
    if (q[nd->axis] < nd->value) {
        kd_radius(t, nd->left, rs, q);
        if (fabsf(gap) < rs->radius) kd_radius(t, nd->right, rs, q);
    } else {
        kd_radius(t, nd->right, rs, q);
        if (fabsf(gap) < rs->radius) kd_radius(t, nd->left, rs, q);
    }
}

/* Batched drivers.  out_dist/out_index are (nq, k); unfilled slots keep (1e10, 0) exactly as
 * KNNResultSet leaves them; out_count (nq) = rs.count; comparisons (nq, may be NULL) =
 * rs.comparison_counter.  threads<=1 -> serial. */
ORACLE_API void oracle_kdtree_knn(void *h, const float *queries, int64_t nq, int k,
                                  float *out_dist, int64_t *out_index, int32_t *out_count,
                                  int64_t *comparisons, int threads) {
    const kd_tree *t = (const kd_tree *)h;
#ifdef _OPENMP
    if (threads < 1) threads = 1;
#pragma omp parallel for schedule(dynamic, 64) num_threads(threads)
#endif
    for (int64_t i = 0; i < nq; ++i) {
        knn_set rs;
        knn_set_init(&rs, k, out_dist + i * k, out_index + i * k);
        kd_knn(t, t->n_nodes ? 0 : -1, &rs, queries + 3 * i);
        out_count[i] = rs.count;
        if (comparisons) comparisons[i] = rs.comparisons;
    }
}

/* Two-call radius search: first with offsets only (counts), then with buffers.
 * offsets has nq+1 entries; results of query i are written in tree-visit order. */
ORACLE_API void oracle_kdtree_radius(void *h, const float *queries, int64_t nq, float radius,
                                     int64_t *offsets, float *out_dist, int64_t *out_index,
                                     int threads) {
    const kd_tree *t = (const kd_tree *)h;
    int fill = (out_dist != NULL);
#ifdef _OPENMP
    if (threads < 1) threads = 1;
#pragma omp parallel for schedule(dynamic, 64) num_threads(threads)
#endif
    for (int64_t i = 0; i < nq; ++i) {
        rad_set rs; memset(&rs, 0, sizeof rs); rs.radius = radius;
        kd_radius(t, t->n_nodes ? 0 : -1, &rs, queries + 3 * i);
        if (fill) {
            memcpy(out_dist + offsets[i], rs.dist, sizeof(float) * (size_t)rs.count);
            memcpy(out_index + offsets[i], rs.index, sizeof(int64_t) * (size_t)rs.count);
        } else {
            offsets[i + 1] = rs.count;
        }
        free(rs.dist); free(rs.index);
    }
    if (!fill) {
        offsets[0] = 0;
        for (int64_t i = 0; i < nq; ++i) offsets[i + 1] += offsets[i];
    }
}

/* ---------------------------------------------------------------- octree */

/* Arithmetic note (octree.py:87-94,106-164,318-322): with a float32 db every quantity the
 * reference's octree touches is a float32 -- np.mean/np.amin/np.amax of a float32 array are
 * float32, "factor * extent" is python-float * np.float32 = float32 (NEP 50 weak scalars),
 * np.asarray of float32 scalars is a float32 array, and query - center is float32 - float32.
 * The restatement therefore carries centres, extents and all predicate arithmetic in float
 * with one rounding per operation (volatile + -ffp-contract=off). */
typedef struct {
    int64_t child[8];       /* -1 = None (octree.py:63-64) */
    float center[3];
    float extent;
    int is_leaf;
    int64_t begin, end;     /* range in perm */
} oc_node;

typedef struct {
    const float *db; int64_t n; int leaf_size; float min_extent;
    int64_t *perm; oc_node *nodes; int64_t n_nodes, cap_nodes;
} oc_tree;

static int64_t oc_new_node(oc_tree *t) {
    if (t->n_nodes == t->cap_nodes) {
        t->cap_nodes = t->cap_nodes ? t->cap_nodes * 2 : 1024;
        t->nodes = (oc_node *)realloc(t->nodes, sizeof(oc_node) * (size_t)t->cap_nodes);
    }
    return t->n_nodes++;
}

static inline int oc_code(const float *p, const float *c) {     /* octree.py:77-84, 285-291 */
    int code = 0;
    if (p[0] > c[0]) code |= 1;
    if (p[1] > c[1]) code |= 2;
    if (p[2] > c[2]) code |= 4;
    return code;
}

/* octree.py:51-103 */
static int64_t oc_build(oc_tree *t, int64_t begin, int64_t end, const float c[3], float extent,
                        int64_t *scratch) {
    if (end == begin) return -1;                                    /* :60-61 */
    int64_t id = oc_new_node(t);
    oc_node nd; memset(&nd, 0, sizeof nd);
    for (int i = 0; i < 8; ++i) nd.child[i] = -1;
    nd.center[0] = c[0]; nd.center[1] = c[1]; nd.center[2] = c[2]; nd.extent = extent;
    nd.begin = begin; nd.end = end; nd.is_leaf = 1;
    int64_t n = end - begin;
    if (n <= t->leaf_size || extent <= t->min_extent) {             /* :70 */
        t->nodes[id] = nd; return id;
    }
    nd.is_leaf = 0;
    int64_t cnt[9] = {0};                                           /* :73-85 stable 8-way partition */
    for (int64_t i = begin; i < end; ++i) cnt[oc_code(t->db + 3 * t->perm[i], c) + 1]++;
    for (int i = 0; i < 8; ++i) cnt[i + 1] += cnt[i];
    int64_t pos[8]; for (int i = 0; i < 8; ++i) pos[i] = cnt[i];
    for (int64_t i = begin; i < end; ++i) {
        int code = oc_code(t->db + 3 * t->perm[i], c);
        scratch[begin + pos[code]++] = t->perm[i];
    }
    memcpy(t->perm + begin, scratch + begin, sizeof(int64_t) * (size_t)n);
    t->nodes[id] = nd;
    volatile float half = 0.5f * extent;                            /* :93 child_extent; also factor*extent :90-92 */
    for (int i = 0; i < 8; ++i) {                                   /* :87-101 */
        volatile float cx = c[0] + ((i & 1) ? half : -half);
        volatile float cy = c[1] + ((i & 2) ? half : -half);
        volatile float cz = c[2] + ((i & 4) ? half : -half);
        float cc[3] = {cx, cy, cz};
        int64_t ch = oc_build(t, begin + cnt[i], begin + cnt[i + 1], cc, half, scratch);
        t->nodes[id].child[i] = ch;
    }
    return id;
}

/* octree.py:310-328 */
ORACLE_API void *oracle_octree_construction(const float *db, int64_t n, int leaf_size, double min_extent) {
    oc_tree *t = (oc_tree *)calloc(1, sizeof(oc_tree));
    t->db = db; t->n = n; t->leaf_size = leaf_size; t->min_extent = (float)min_extent;
    t->perm = (int64_t *)malloc(sizeof(int64_t) * (size_t)(n > 0 ? n : 1));
    for (int64_t i = 0; i < n; ++i) t->perm[i] = i;
    if (n == 0) return t;
    float mn[3], mx[3];
    for (int a = 0; a < 3; ++a) { mn[a] = db[a]; mx[a] = db[a]; }
    for (int64_t i = 1; i < n; ++i)
        for (int a = 0; a < 3; ++a) {
            float v = db[3 * i + a];
            if (v < mn[a]) mn[a] = v;
            if (v > mx[a]) mx[a] = v;
        }
    float ext = 0.f;
    for (int a = 0; a < 3; ++a) { volatile float d = mx[a] - mn[a]; if (a == 0 || d > ext) ext = d; }
    volatile float ext_half = ext * 0.5f;                           /* :321 */
    /* :322 np.mean(float32, axis=0) adds float32 values in numpy's own blocked order; that
     * order is not restated.  The root centre only shapes the tree (any centre yields the same
     * exact search answers), so the mean is accumulated in double and rounded. */
    double cm[3] = {0, 0, 0};
    for (int64_t i = 0; i < n; ++i) for (int a = 0; a < 3; ++a) cm[a] += db[3 * i + a];
    float c[3];
    for (int a = 0; a < 3; ++a) c[a] = (float)(cm[a] / (double)n);
    int64_t *scratch = (int64_t *)malloc(sizeof(int64_t) * (size_t)n);
    oc_build(t, 0, n, c, ext_half, scratch);
    free(scratch);
    return t;
}

ORACLE_API void oracle_octree_destroy(void *h) {
    oc_tree *t = (oc_tree *)h;
    if (!t) return;
    free(t->perm); free(t->nodes); free(t);
}

/* octree.py:106-117 */
static int oc_inside(const float *q, float radius, const oc_node *o) {
    for (int a = 0; a < 3; ++a) {
        volatile float off = q[a] - o->center[a];
        volatile float space = fabsf(off) + radius;
        if (!(space < o->extent)) return 0;
    }
    return 1;
}

/* octree.py:120-148 */
static int oc_overlaps(const float *q, float radius, const oc_node *o) {
    float off[3];
    for (int a = 0; a < 3; ++a) { volatile float d = q[a] - o->center[a]; off[a] = fabsf(d); }
    volatile float max_dist = radius + o->extent;                   /* :132 */
    for (int a = 0; a < 3; ++a) if (off[a] > max_dist) return 0;    /* :133-134 */
    int c = 0;
    for (int a = 0; a < 3; ++a) if (off[a] < o->extent) ++c;        /* :137-138 */
    if (c >= 2) return 1;
    volatile float x = off[0] - o->extent; if (!(x > 0.f)) x = 0.f; /* :143-145 */
    volatile float y = off[1] - o->extent; if (!(y > 0.f)) y = 0.f;
    volatile float z = off[2] - o->extent; if (!(z > 0.f)) z = 0.f;
    volatile float xx = x * x, yy = y * y, zz = z * z;
    volatile float s = xx + yy; s = s + zz;
    volatile float rr = radius * radius;
    return s < rr;                                                  /* :147 */
}

/* octree.py:262-306 */
static int oc_knn(const oc_tree *t, int64_t id, knn_set *rs, const float *q) {
    if (id < 0) return 0;
    const oc_node *o = &t->nodes[id];
    if (o->is_leaf && o->end > o->begin) {
        for (int64_t i = o->begin; i < o->end; ++i)
            oracle_knn_add(rs, ref_dist(q, t->db + 3 * t->perm[i]), t->perm[i]);
        return oc_inside(q, rs->worst, o);
    }
    int code = oc_code(q, o->center);
    if (oc_knn(t, o->child[code], rs, q)) return 1;
    for (int c = 0; c < 8; ++c) {
        if (c == code || o->child[c] < 0) continue;
        if (!oc_overlaps(q, rs->worst, &t->nodes[o->child[c]])) continue;
        if (oc_knn(t, o->child[c], rs, q)) return 1;
    }
    return oc_inside(q, rs->worst, o);
}

/* octree.py:216-259 */
static int oc_radius(const oc_tree *t, int64_t id, rad_set *rs, const float *q) {
    if (id < 0) return 0;
    const oc_node *o = &t->nodes[id];
    if (o->is_leaf && o->end > o->begin) {
        for (int64_t i = o->begin; i < o->end; ++i)
            rad_add(rs, ref_dist(q, t->db + 3 * t->perm[i]), t->perm[i]);
        return oc_inside(q, rs->radius, o);
    }
    int code = oc_code(q, o->center);
    if (oc_radius(t, o->child[code], rs, q)) return 1;
    for (int c = 0; c < 8; ++c) {
        if (c == code || o->child[c] < 0) continue;
        if (!oc_overlaps(q, rs->radius, &t->nodes[o->child[c]])) continue;
        if (oc_radius(t, o->child[c], rs, q)) return 1;
    }
    return oc_inside(q, rs->radius, o);
}

ORACLE_API void oracle_octree_knn(void *h, const float *queries, int64_t nq, int k,
                                  float *out_dist, int64_t *out_index, int32_t *out_count,
                                  int64_t *comparisons, int threads) {
    const oc_tree *t = (const oc_tree *)h;
#ifdef _OPENMP
    if (threads < 1) threads = 1;
#pragma omp parallel for schedule(dynamic, 64) num_threads(threads)
#endif
    for (int64_t i = 0; i < nq; ++i) {
        knn_set rs;
        knn_set_init(&rs, k, out_dist + i * k, out_index + i * k);
        oc_knn(t, t->n_nodes ? 0 : -1, &rs, queries + 3 * i);
        out_count[i] = rs.count;
        if (comparisons) comparisons[i] = rs.comparisons;
    }
}

ORACLE_API void oracle_octree_radius(void *h, const float *queries, int64_t nq, float radius,
                                     int64_t *offsets, float *out_dist, int64_t *out_index,
                                     int threads) {
    const oc_tree *t = (const oc_tree *)h;
    int fill = (out_dist != NULL);
#ifdef _OPENMP
    if (threads < 1) threads = 1;
#pragma omp parallel for schedule(dynamic, 64) num_threads(threads)
#endif
    for (int64_t i = 0; i < nq; ++i) {
        rad_set rs; memset(&rs, 0, sizeof rs); rs.radius = radius;
        oc_radius(t, t->n_nodes ? 0 : -1, &rs, queries + 3 * i);
        if (fill) {
            memcpy(out_dist + offsets[i], rs.dist, sizeof(float) * (size_t)rs.count);
            memcpy(out_index + offsets[i], rs.index, sizeof(int64_t) * (size_t)rs.count);
        } else {
            offsets[i + 1] = rs.count;
        }
        free(rs.dist); free(rs.index);
    }
    if (!fill) {
        offsets[0] = 0;
        for (int64_t i = 0; i < nq; ++i) offsets[i + 1] += offsets[i];
    }
}

/* ---------------------------------------------------------------- brute force (canonical order) */

/* kdtree.py:234-236 restated with the canonical tie rule of DESIGN.md: ascending
 * (float32 distance, index).  Used to check tie cases where tree visit order decides. */
ORACLE_API void oracle_bruteforce_knn(const float *db, int64_t n, const float *queries, int64_t nq,
                                      int k, float *out_dist, int64_t *out_index, int32_t *out_count,
                                      int threads) {
#ifdef _OPENMP
    if (threads < 1) threads = 1;
#pragma omp parallel for schedule(dynamic, 16) num_threads(threads)
#endif
    for (int64_t qi = 0; qi < nq; ++qi) {
        float *od = out_dist + qi * k; int64_t *oi = out_index + qi * k;
        int cnt = 0;
        for (int j = 0; j < k; ++j) { od[j] = 1e10f; oi[j] = 0; }
        for (int64_t i = 0; i < n; ++i) {
            float d = ref_dist(queries + 3 * qi, db + 3 * i);
            if (cnt == k && !(d < od[k - 1])) continue;   /* scan is index-ascending: ties keep the lower index */
            int j = cnt < k ? cnt++ : k - 1;
            while (j > 0 && od[j - 1] > d) { od[j] = od[j - 1]; oi[j] = oi[j - 1]; --j; }
            od[j] = d; oi[j] = i;
        }
        out_count[qi] = cnt;
    }
}

/* Distances of explicit (query, point) pairs -- lets tests check returned distances bit-for-bit. */
ORACLE_API void oracle_pair_dist(const float *db, const float *queries, const int64_t *qidx,
                                 const int64_t *pidx, int64_t m, float *out) {
    for (int64_t i = 0; i < m; ++i) out[i] = ref_dist(queries + 3 * qidx[i], db + 3 * pidx[i]);
}
