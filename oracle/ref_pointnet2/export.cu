// TEST INFRASTRUCTURE (oracle/): C entry points over the reference's own kernel wrappers (compiled from
// /root/reference/Classification_ModelNet40/pointnet2_ops_lib/pointnet2_ops/_ext-src/src/{ball_query,interpolate}_gpu.cu),
// host arrays in / out, the output buffers zero-initialised as ball_query.cpp:20-22 / interpolate.cpp do (torch::zeros).
#include <cuda_runtime.h>

void query_ball_point_kernel_wrapper(int b, int n, int m, float radius, int nsample, const float *new_xyz,
                                     const float *xyz, int *idx);
void three_nn_kernel_wrapper(int b, int n, int m, const float *unknown, const float *known, float *dist2, int *idx);

extern "C" int ref_ball_query(int b, int n, int m, float radius, int nsample, const float *new_xyz, const float *xyz,
                              int *idx) {
    float *d_c = nullptr, *d_p = nullptr;
    int *d_i = nullptr;
    const size_t nc = sizeof(float) * 3 * b * m, np = sizeof(float) * 3 * b * n, ni = sizeof(int) * (size_t)b * m * nsample;
    cudaMalloc(&d_c, nc + 16); cudaMalloc(&d_p, np + 16); cudaMalloc(&d_i, ni + 16);
    cudaMemcpy(d_c, new_xyz, nc, cudaMemcpyHostToDevice);
    cudaMemcpy(d_p, xyz, np, cudaMemcpyHostToDevice);
    cudaMemset(d_i, 0, ni);
    if (b > 0 && m > 0) query_ball_point_kernel_wrapper(b, n, m, radius, nsample, d_c, d_p, d_i);
    cudaError_t e = cudaDeviceSynchronize();
    cudaMemcpy(idx, d_i, ni, cudaMemcpyDeviceToHost);
    cudaFree(d_c); cudaFree(d_p); cudaFree(d_i);
    return (int)e;
}

extern "C" int ref_three_nn(int b, int n, int m, const float *unknown, const float *known, float *dist2, int *idx) {
    float *d_u = nullptr, *d_k = nullptr, *d_d = nullptr;
    int *d_i = nullptr;
    const size_t nu = sizeof(float) * 3 * b * n, nk = sizeof(float) * 3 * b * m;
    cudaMalloc(&d_u, nu + 16); cudaMalloc(&d_k, nk + 16); cudaMalloc(&d_d, nu + 16); cudaMalloc(&d_i, nu + 16);
    cudaMemcpy(d_u, unknown, nu, cudaMemcpyHostToDevice);
    cudaMemcpy(d_k, known, nk, cudaMemcpyHostToDevice);
    cudaMemset(d_d, 0, nu); cudaMemset(d_i, 0, nu);
    if (b > 0 && n > 0) three_nn_kernel_wrapper(b, n, m, d_u, d_k, d_d, d_i);
    cudaError_t e = cudaDeviceSynchronize();
    cudaMemcpy(dist2, d_d, nu, cudaMemcpyDeviceToHost);
    cudaMemcpy(idx, d_i, nu, cudaMemcpyDeviceToHost);
    cudaFree(d_u); cudaFree(d_k); cudaFree(d_d); cudaFree(d_i);
    return (int)e;
}
