// Feature-matching RANSAC (Registration/main.py:68-84: registration_ransac_based_on_feature_matching with ransac_n = 3,
// CorrespondenceCheckerBasedOnEdgeLength(0.9), CorrespondenceCheckerBasedOnDistance, RANSACConvergenceCriteria(100000,
// 0.999); icp_template.py:56-111 ransac_init) as a batch problem on the device:
//
//   ransac_hypothesis_kernel  one thread per iteration: 3 correspondences from a counter hash of (seed, iteration),
//                             edge-length check, 3-point least-squares rigid transform (Kabsch / Umeyama without
//                             scale, proper rotation), distance check; survivors are appended to a list
//   ransac_transform_kernel   survivors x source points -> transformed float32 query points
//   pcp_knn (k = 1)           the library's own nearest-neighbour kernel over the target index
//   ransac_score_kernel       per survivor: inliers (d <= max_correspondence_distance) and sum of d^2
//   host                      best = (fitness desc, RMSE asc, iteration asc) -- independent of the survivors' order;
//                             iterations needed for the confidence = log(1 - c) / log(1 - fitness^3), checked per round
#include <algorithm>
#include <cmath>
#include <vector>

#include "pcp_device.cuh"
#include "pcp_internal.cuh"

namespace pcp {

namespace {

constexpr int kRound = 8192;                  // iterations per round (one hypothesis kernel launch)
constexpr long long kMaxQueries = 1LL << 24;  // transformed points evaluated per chunk (192 MB of queries)

__device__ __forceinline__ unsigned long long mix64(unsigned long long x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

struct Hyp {
    double T[12];            // 3x4 row-major: x -> R x + t
    long long iteration;
    long long pad;
};

// Least-squares rigid transform of 3 point pairs: R = U diag(1, 1, det(U) det(V)) V^T of the SVD of
// H = sum (t - tbar)(s - sbar)^T (Umeyama without scale), t = tbar - R sbar.  One-sided Jacobi; the direction of the
// smallest singular value is completed by cross products, which makes R a proper rotation also for the rank-2 H of
// three points.  false: degenerate sample (collinear or coincident points).
__device__ bool rigid3(const double s[3][3], const double t[3][3], double T[12]) {
    double sb[3], tb[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) { sb[a] = (s[0][a] + s[1][a] + s[2][a]) / 3.0; tb[a] = (t[0][a] + t[1][a] + t[2][a]) / 3.0; }
    double A[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, V[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
#pragma unroll
    for (int k = 0; k < 3; ++k)
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) A[i * 3 + j] += (t[k][i] - tb[i]) * (s[k][j] - sb[j]);
    for (int sweep = 0; sweep < 30; ++sweep) {
        double off = 0.0;
#pragma unroll
        for (int pq = 0; pq < 3; ++pq) {
            const int p = pq == 2 ? 1 : 0, q = pq == 0 ? 1 : 2;
            double alpha = 0, beta = 0, gamma = 0;
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                alpha += A[r * 3 + p] * A[r * 3 + p];
                beta += A[r * 3 + q] * A[r * 3 + q];
                gamma += A[r * 3 + p] * A[r * 3 + q];
            }
            if (gamma == 0.0 || fabs(gamma) <= 1e-15 * sqrt(alpha * beta)) continue;
            off = fmax(off, fabs(gamma) / fmax(sqrt(alpha * beta), 1e-300));
            const double zeta = (beta - alpha) / (2.0 * gamma);
            const double tt = copysign(1.0, zeta) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
            const double c = 1.0 / sqrt(1.0 + tt * tt), sn = c * tt;
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                const double ap = A[r * 3 + p], aq = A[r * 3 + q];
                A[r * 3 + p] = c * ap - sn * aq;
                A[r * 3 + q] = sn * ap + c * aq;
                const double vp = V[r * 3 + p], vq = V[r * 3 + q];
                V[r * 3 + p] = c * vp - sn * vq;
                V[r * 3 + q] = sn * vp + c * vq;
            }
        }
        if (off < 1e-15) break;
    }
    double nrm[3];
#pragma unroll
    for (int j = 0; j < 3; ++j) nrm[j] = sqrt(A[j] * A[j] + A[3 + j] * A[3 + j] + A[6 + j] * A[6 + j]);
    int c = 0;
    if (nrm[1] < nrm[c]) c = 1;
    if (nrm[2] < nrm[c]) c = 2;
    const int a = (c + 1) % 3, b = (c + 2) % 3;
    const double big = fmax(nrm[a], nrm[b]);
    if (!(fmin(nrm[a], nrm[b]) > 1e-12 * big) || !(big > 0.0)) return false;
    double U[9];
#pragma unroll
    for (int r = 0; r < 3; ++r) { U[r * 3 + a] = A[r * 3 + a] / nrm[a]; U[r * 3 + b] = A[r * 3 + b] / nrm[b]; }
    U[0 * 3 + c] = U[1 * 3 + a] * U[2 * 3 + b] - U[2 * 3 + a] * U[1 * 3 + b];
    U[1 * 3 + c] = U[2 * 3 + a] * U[0 * 3 + b] - U[0 * 3 + a] * U[2 * 3 + b];
    U[2 * 3 + c] = U[0 * 3 + a] * U[1 * 3 + b] - U[1 * 3 + a] * U[0 * 3 + b];
    V[0 * 3 + c] = V[1 * 3 + a] * V[2 * 3 + b] - V[2 * 3 + a] * V[1 * 3 + b];
    V[1 * 3 + c] = V[2 * 3 + a] * V[0 * 3 + b] - V[0 * 3 + a] * V[2 * 3 + b];
    V[2 * 3 + c] = V[0 * 3 + a] * V[1 * 3 + b] - V[1 * 3 + a] * V[0 * 3 + b];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
#pragma unroll
        for (int j = 0; j < 3; ++j)
            T[i * 4 + j] = U[i * 3 + 0] * V[j * 3 + 0] + U[i * 3 + 1] * V[j * 3 + 1] + U[i * 3 + 2] * V[j * 3 + 2];
        T[i * 4 + 3] = 0.0;
    }
#pragma unroll
    for (int i = 0; i < 3; ++i) T[i * 4 + 3] = tb[i] - (T[i * 4 + 0] * sb[0] + T[i * 4 + 1] * sb[1] + T[i * 4 + 2] * sb[2]);
    return true;
}

__global__ void __launch_bounds__(128)
ransac_hypothesis_kernel(const float *__restrict__ src, const float *__restrict__ tgt, long long stride,
                         const int *__restrict__ corr, long long n_corr, unsigned long long seed, long long first_it,
                         int n_it, double edge_ratio, double max_dist, Hyp *__restrict__ out, unsigned *__restrict__ n_out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_it) return;
    const long long it = first_it + i;
    double s[3][3], t[3][3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const long long c = (long long)(mix64(seed ^ mix64((unsigned long long)it * 3ull + (unsigned long long)k)) % (unsigned long long)n_corr);
        const float *ps = src + (long long)corr[c * 2] * stride, *pt = tgt + (long long)corr[c * 2 + 1] * stride;
#pragma unroll
        for (int a = 0; a < 3; ++a) { s[k][a] = ps[a]; t[k][a] = pt[a]; }
    }
    // CorrespondenceCheckerBasedOnEdgeLength: every edge of the sampled triangle keeps its length up to the ratio
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const int l = (k + 1) % 3;
        double ds = 0.0, dt = 0.0;
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            ds += (s[k][a] - s[l][a]) * (s[k][a] - s[l][a]);
            dt += (t[k][a] - t[l][a]) * (t[k][a] - t[l][a]);
        }
        ds = sqrt(ds); dt = sqrt(dt);
        if (ds < dt * edge_ratio || dt < ds * edge_ratio) return;
    }
    Hyp h;
    if (!rigid3(s, t, h.T)) return;
    // CorrespondenceCheckerBasedOnDistance: the sampled pairs agree under the estimated transform
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        double d2 = 0.0;
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const double v = h.T[a * 4 + 0] * s[k][0] + h.T[a * 4 + 1] * s[k][1] + h.T[a * 4 + 2] * s[k][2] + h.T[a * 4 + 3] - t[k][a];
            d2 += v * v;
        }
        if (d2 > max_dist * max_dist) return;
    }
    h.iteration = it;
    h.pad = 0;
    out[atomicAdd(n_out, 1u)] = h;
}

__global__ void __launch_bounds__(256)
ransac_transform_kernel(const float *__restrict__ src, long long ns, long long stride, const Hyp *__restrict__ hyp,
                        float *__restrict__ q) {
    const long long h = blockIdx.y;
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= ns) return;
    const double *T = hyp[h].T;
    const float *p = src + i * stride;
    const double x = p[0], y = p[1], z = p[2];
    float *o = q + (h * ns + i) * 3;
    o[0] = (float)(T[0] * x + T[1] * y + T[2] * z + T[3]);
    o[1] = (float)(T[4] * x + T[5] * y + T[6] * z + T[7]);
    o[2] = (float)(T[8] * x + T[9] * y + T[10] * z + T[11]);
}

// One CTA per hypothesis: inlier count and sum of squared distances over its ns nearest-neighbour distances,
// fixed-order tree reduction (deterministic).
__global__ void __launch_bounds__(256)
ransac_score_kernel(const float *__restrict__ dist, const int *__restrict__ cnt, long long ns, float max_dist,
                    int *__restrict__ inliers, double *__restrict__ sum_d2) {
    __shared__ int s_n[256];
    __shared__ double s_d[256];
    const long long h = blockIdx.x;
    int n = 0;
    double d2 = 0.0;
    for (long long i = threadIdx.x; i < ns; i += 256) {
        const float d = dist[h * ns + i];
        if (cnt[h * ns + i] > 0 && d <= max_dist) { ++n; d2 += (double)d * (double)d; }
    }
    s_n[threadIdx.x] = n;
    s_d[threadIdx.x] = d2;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o) { s_n[threadIdx.x] += s_n[threadIdx.x + o]; s_d[threadIdx.x] += s_d[threadIdx.x + o]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) { inliers[h] = s_n[0]; sum_d2[h] = s_d[0]; }
}

}  // namespace

int ransac_device(pcp_ctx *ctx, const pcp_ransac_params &prm, const float *d_src, int64_t ns, int64_t stride,
                  pcp_index *tgt, const int32_t *d_corr, int64_t n_corr, double *h_T, double *h_fitness, double *h_rmse,
                  int64_t *h_iterations, int64_t *h_evaluated) {
    cudaStream_t st = ctx->stream;
    for (int i = 0; i < 16; ++i) h_T[i] = (i % 5 == 0) ? 1.0 : 0.0;
    *h_fitness = 0.0;
    *h_rmse = 0.0;
    *h_iterations = 0;
    *h_evaluated = 0;
    if (ns == 0 || tgt->n == 0 || n_corr < 3 || prm.max_iteration <= 0) return PCP_OK;
    const float *d_tgt = tgt->built_from;
    const int64_t tstride = tgt->built_stride;
    if (tstride != stride) return fail(PCP_ERR_INVALID, "pcp_ransac_init: source and target strides differ");
    // (own buffers: the kNN call below uses the context's scratch slots for its query regrouping)
    DevBuf &b_hyp = ctx->ransac_buf[0], &b_cnt = ctx->ransac_buf[1], &b_q = ctx->ransac_buf[2], &b_res = ctx->ransac_buf[3];
    PCP_TRY(b_hyp.ensure(sizeof(Hyp) * kRound));
    PCP_TRY(b_cnt.ensure(sizeof(unsigned) * 4));
    const long long per_chunk = std::max<long long>(1, std::min<long long>(kRound, kMaxQueries / ns));
    PCP_TRY(b_q.ensure(sizeof(float) * 3 * (size_t)(per_chunk * ns)));
    // per query: idx (int), dist (float), count (int); per hypothesis: inliers (int), sum d2 (double)
    PCP_TRY(b_res.ensure((size_t)(per_chunk * ns) * 12 + (size_t)per_chunk * 16 + 64));
    int32_t *d_idx = b_res.as<int32_t>();
    float *d_dist = reinterpret_cast<float *>(d_idx + per_chunk * ns);
    int32_t *d_count = reinterpret_cast<int32_t *>(d_dist + per_chunk * ns);
    double *d_sum = reinterpret_cast<double *>((reinterpret_cast<uintptr_t>(d_count + per_chunk * ns) + 15) & ~(uintptr_t)15);
    int *d_inl = reinterpret_cast<int *>(d_sum + per_chunk);

    double best_fit = 0.0, best_rmse = 0.0;
    long long best_it = -1, done = 0, needed = prm.max_iteration, evaluated = 0;
    std::vector<Hyp> h_hyp(kRound);
    std::vector<int> h_inl((size_t)per_chunk);
    std::vector<double> h_sum((size_t)per_chunk);
    while (done < std::min<long long>(needed, prm.max_iteration)) {
        const int n_it = (int)std::min<long long>(kRound, prm.max_iteration - done);
        PCP_CUDA(cudaMemsetAsync(b_cnt.p, 0, sizeof(unsigned), st));
        ransac_hypothesis_kernel<<<(n_it + 127) / 128, 128, 0, st>>>(d_src, d_tgt, stride, d_corr, n_corr, prm.seed, done, n_it,
                                                                      prm.edge_length_threshold, prm.max_correspondence_distance,
                                                                      b_hyp.as<Hyp>(), b_cnt.as<unsigned>());
        ctx->launches += 1;
        unsigned H = 0;
        PCP_CUDA(cudaMemcpyAsync(&H, b_cnt.p, sizeof(unsigned), cudaMemcpyDeviceToHost, st));
        PCP_CUDA(cudaStreamSynchronize(st));
        if (H > 0) PCP_CUDA(cudaMemcpyAsync(h_hyp.data(), b_hyp.p, sizeof(Hyp) * H, cudaMemcpyDeviceToHost, st));
        for (long long h0 = 0; h0 < (long long)H; h0 += per_chunk) {
            const long long hc = std::min<long long>(per_chunk, (long long)H - h0);
            ransac_transform_kernel<<<dim3((unsigned)((ns + 255) / 256), (unsigned)hc), 256, 0, st>>>(d_src, ns, stride,
                                                                                                        b_hyp.as<Hyp>() + h0, b_q.as<float>());
            ctx->launches += 1;
            PCP_TRY(knn_device(tgt, b_q.as<float>(), hc * ns, 3, nullptr, 1, d_idx, d_dist, d_count));
            ransac_score_kernel<<<(unsigned)hc, 256, 0, st>>>(d_dist, d_count, ns, (float)prm.max_correspondence_distance, d_inl, d_sum);
            ctx->launches += 1;
            PCP_CUDA(cudaMemcpyAsync(h_inl.data(), d_inl, sizeof(int) * (size_t)hc, cudaMemcpyDeviceToHost, st));
            PCP_CUDA(cudaMemcpyAsync(h_sum.data(), d_sum, sizeof(double) * (size_t)hc, cudaMemcpyDeviceToHost, st));
            PCP_CUDA(cudaStreamSynchronize(st));
            for (long long j = 0; j < hc; ++j) {
                if (h_inl[(size_t)j] == 0) continue;
                const double fit = (double)h_inl[(size_t)j] / (double)ns, rmse = std::sqrt(h_sum[(size_t)j] / h_inl[(size_t)j]);
                const long long it = h_hyp[(size_t)(h0 + j)].iteration;
                const bool better = fit > best_fit || (fit == best_fit && (rmse < best_rmse || (rmse == best_rmse && it < best_it)));
                if (better || best_it < 0) {
                    best_fit = fit; best_rmse = rmse; best_it = it;
                    const double *T = h_hyp[(size_t)(h0 + j)].T;
                    for (int i = 0; i < 12; ++i) h_T[i] = T[i];
                    h_T[12] = h_T[13] = h_T[14] = 0.0; h_T[15] = 1.0;
                }
            }
            evaluated += hc;
        }
        done += n_it;
        // RANSACConvergenceCriteria: iterations needed to have drawn an all-inlier sample with the given confidence
        if (best_fit > 0.0 && prm.confidence > 0.0 && prm.confidence < 1.0) {
            const double p3 = std::pow(std::min(best_fit, 1.0), 3.0);
            if (p3 >= 1.0) needed = done;
            else needed = (long long)std::ceil(std::log(1.0 - prm.confidence) / std::log(1.0 - p3));
        }
    }
    *h_fitness = best_fit;
    *h_rmse = best_rmse;
    *h_iterations = done;
    *h_evaluated = evaluated;
    PCP_CUDA(cudaGetLastError());
    return PCP_OK;
}

}  // namespace pcp
