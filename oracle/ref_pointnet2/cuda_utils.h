// TEST INFRASTRUCTURE (oracle/): stand-in for the reference's _ext-src/include/cuda_utils.h, found first on the
// include path when oracle/build_oracle.py compiles the reference's OWN ball_query_gpu.cu / interpolate_gpu.cu
// where they lie under /root/reference.  The original header pulls in ATen only to fetch the current torch stream;
// the kernels themselves are plain CUDA.  This header provides the three names those two sources use, without torch:
// the launch-width helper, the error-check macro, and at::cuda::getCurrentCUDAStream() (the default stream).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>

using std::max;
using std::min;

#define TOTAL_THREADS 512

// largest power of two <= work_size, clamped to [1, TOTAL_THREADS]
inline int opt_n_threads(int work_size) {
    int t = 1;
    while (t * 2 <= work_size && t * 2 <= TOTAL_THREADS) t *= 2;
    return t;
}

inline dim3 opt_block_config(int x, int y) {
    const int xt = opt_n_threads(x);
    const int yt = std::max(std::min(opt_n_threads(y), TOTAL_THREADS / xt), 1);
    return dim3(xt, yt, 1);
}

namespace at { namespace cuda { inline cudaStream_t getCurrentCUDAStream() { return nullptr; } } }

#define CUDA_CHECK_ERRORS()                                                                      \
    do {                                                                                         \
        cudaError_t e_ = cudaGetLastError();                                                     \
        if (e_ != cudaSuccess) { fprintf(stderr, "reference kernel failed: %s\n", cudaGetErrorString(e_)); exit(-1); } \
    } while (0)
