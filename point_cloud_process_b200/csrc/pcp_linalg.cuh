// Small dense float64 linear algebra shared by the device code: symmetric 3x3 eigen-decomposition (normals, PCA,
// ISS) and the orthogonal polar factor of a 3x3 matrix (ICP / Procrustes, main.py:138-139).
#pragma once

#include <cuda_runtime.h>

namespace pcp {

// cyclic Jacobi on a symmetric 3x3 (float64): eigenvalues w[3], eigenvectors as the COLUMNS of V (row-major 3x3)
__device__ inline void sym_eig3(double a00, double a01, double a02, double a11, double a12, double a22, double w[3],
                         double V[9]) {
    double A[3][3] = {{a00, a01, a02}, {a01, a11, a12}, {a02, a12, a22}};
    double E[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
    for (int sweep = 0; sweep < 24; ++sweep) {
        const double off = fabs(A[0][1]) + fabs(A[0][2]) + fabs(A[1][2]);
        const double diag = fabs(A[0][0]) + fabs(A[1][1]) + fabs(A[2][2]);
        if (off <= 1e-300 || off <= 1e-18 * diag) break;
#pragma unroll
        for (int pq = 0; pq < 3; ++pq) {
            const int p = pq == 2 ? 1 : 0, q = pq == 0 ? 1 : 2;
            const double apq = A[p][q];
            if (apq == 0.0) continue;
            const double theta = (A[q][q] - A[p][p]) / (2.0 * apq);
            const double t = copysign(1.0, theta) / (fabs(theta) + sqrt(theta * theta + 1.0));
            const double c = 1.0 / sqrt(t * t + 1.0), s = t * c;
#pragma unroll
            for (int r = 0; r < 3; ++r) {   // A <- A J
                const double arp = A[r][p], arq = A[r][q];
                A[r][p] = c * arp - s * arq;
                A[r][q] = s * arp + c * arq;
            }
#pragma unroll
            for (int r = 0; r < 3; ++r) {   // A <- J^T A
                const double apr = A[p][r], aqr = A[q][r];
                A[p][r] = c * apr - s * aqr;
                A[q][r] = s * apr + c * aqr;
            }
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                const double erp = E[r][p], erq = E[r][q];
                E[r][p] = c * erp - s * erq;
                E[r][q] = s * erp + c * erq;
            }
        }
    }
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        w[i] = A[i][i];
#pragma unroll
        for (int j = 0; j < 3; ++j) V[i * 3 + j] = E[i][j];
    }
}

// One-sided (Hestenes) Jacobi SVD of a 3x3 float64 matrix: A V = U S.  Returns R = U V^T, the
// orthogonal polar factor u @ vt of main.py:138-139 (independent of the SVD's sign conventions
// whenever the smallest singular value is > 0).
__device__ inline void polar_uvt(const double H[9], double R[9]) {
    double A[9], V[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
#pragma unroll
    for (int i = 0; i < 9; ++i) A[i] = H[i];
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
            const double lim = 1e-15 * sqrt(alpha * beta);
            if (fabs(gamma) <= lim || gamma == 0.0) continue;
            off = fmax(off, fabs(gamma) / fmax(sqrt(alpha * beta), 1e-300));
            const double zeta = (beta - alpha) / (2.0 * gamma);
            const double t = copysign(1.0, zeta) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
            const double c = 1.0 / sqrt(1.0 + t * t), s = c * t;
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                const double ap = A[r * 3 + p], aq = A[r * 3 + q];
                A[r * 3 + p] = c * ap - s * aq;
                A[r * 3 + q] = s * ap + c * aq;
                const double vp = V[r * 3 + p], vq = V[r * 3 + q];
                V[r * 3 + p] = c * vp - s * vq;
                V[r * 3 + q] = s * vp + c * vq;
            }
        }
        if (off < 1e-15) break;
    }
    // U columns = normalised A columns; a vanishing column (rank-deficient H: collinear / planar
    // correspondences) is completed by the cross product of the other two, as any SVD would pick
    // some unit vector orthogonal to them.
    double U[9], nrm[3];
#pragma unroll
    for (int j = 0; j < 3; ++j) nrm[j] = sqrt(A[j] * A[j] + A[3 + j] * A[3 + j] + A[6 + j] * A[6 + j]);
    const double big = fmax(nrm[0], fmax(nrm[1], nrm[2]));
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        const double inv = nrm[j] > 1e-14 * big && nrm[j] > 0.0 ? 1.0 / nrm[j] : 0.0;
        U[j] = A[j] * inv; U[3 + j] = A[3 + j] * inv; U[6 + j] = A[6 + j] * inv;
    }
    for (int j = 0; j < 3; ++j) {
        if (nrm[j] > 1e-14 * big && nrm[j] > 0.0) continue;
        const int a = (j + 1) % 3, b = (j + 2) % 3;
        U[j] = U[3 + a] * U[6 + b] - U[6 + a] * U[3 + b];
        U[3 + j] = U[6 + a] * U[b] - U[a] * U[6 + b];
        U[6 + j] = U[a] * U[3 + b] - U[3 + a] * U[b];
    }
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            R[i * 3 + j] = U[i * 3 + 0] * V[j * 3 + 0] + U[i * 3 + 1] * V[j * 3 + 1] + U[i * 3 + 2] * V[j * 3 + 2];
}

}  // namespace pcp
