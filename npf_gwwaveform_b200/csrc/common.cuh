// Shared helpers for the npfk kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/npfk.h"

namespace npfk {

// Thread-local last error text, read through npfk_last_error().
void set_error(const char* fmt, ...);

inline int check_launch(const char* what) {
    cudaError_t e = cudaPeekAtLastError();
    if (e != cudaSuccess) {
        (void)cudaGetLastError();
        set_error("%s: %s", what, cudaGetErrorString(e));
        return NPFK_ERR_CUDA;
    }
    return NPFK_OK;
}

#define NPFK_REQUIRE(cond, ...)            \
    do {                                   \
        if (!(cond)) {                     \
            npfk::set_error(__VA_ARGS__);  \
            return NPFK_ERR_INVALID;       \
        }                                  \
    } while (0)

inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

constexpr int kNumSMs = 148;  // B200

__host__ __device__ inline int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// torch's F.softplus (beta=1, threshold=20)
__device__ __forceinline__ float softplus_f(float x) { return x > 20.f ? x : log1pf(expf(x)); }
__device__ __forceinline__ float sigmoid_f(float x) { return 1.f / (1.f + expf(-x)); }

}  // namespace npfk

namespace npfk {
void count_launch(int n = 1);
}
// launch + count (gpu_launches evidence for bench.py)
#define NPFK_LAUNCH(kernel, grid, block, smem, stream, ...)          \
    do {                                                             \
        kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__);  \
        npfk::count_launch();                                        \
    } while (0)
