// Context / target split on the device: the step right before the model in the reference's training loop
// (npf/utils/datasplit.py:60-255, utils/data/dataloader.py:6-37).  The reference shuffles an index matrix with numpy
// on the host for every batch (npf/utils/helpers.py:82-96) and copies the gathered points H2D; here a batch that is
// already resident in HBM is split in place:
//   npfk_random_subset   idx[b, 0:n] = n distinct indices of [lo, lo + N) per task, uniformly at random
//                        (partial Fisher-Yates in shared memory, counter-based RNG keyed on (seed, task, step))
//   npfk_gather_rows     dst[b, i, :] = src[b, idx[b, i], :]            (CntxtTrgtGetter.select, datasplit.py:246-255)
#include "common.cuh"

namespace npfk {

// splitmix64 finaliser: a counter-based generator, one 64-bit draw per (seed, task, step)
__device__ __forceinline__ uint64_t mix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

// one CTA per task (or one CTA in total when `shared`: every task gets the same subset)
__global__ void __launch_bounds__(256)
random_subset_kernel(int64_t* __restrict__ idx, int B, int N, int n, int64_t lo, uint64_t seed, int shared) {
    extern __shared__ int perm[];
    const int task = blockIdx.x;
    for (int k = threadIdx.x; k < N; k += blockDim.x) perm[k] = k;
    __syncthreads();
    if (threadIdx.x == 0) {
        const uint64_t key = mix64(seed ^ mix64((uint64_t)task + 0x51ED270B0ull));
        for (int i = 0; i < n; ++i) {
            const uint64_t r = mix64(key + (uint64_t)i);
            const int j = i + (int)__umul64hi(r, (uint64_t)(N - i));      // uniform in [i, N), bias < N / 2^64
            const int a = perm[i], c = perm[j];
            perm[i] = c;
            perm[j] = a;
        }
    }
    __syncthreads();
    if (shared) {
        for (int64_t e = threadIdx.x; e < (int64_t)B * n; e += blockDim.x) idx[e] = lo + perm[e % n];
    } else {
        for (int i = threadIdx.x; i < n; i += blockDim.x) idx[(int64_t)task * n + i] = lo + perm[i];
    }
}

__global__ void gather_rows_kernel(const float* __restrict__ src, const int64_t* __restrict__ idx,
                                   float* __restrict__ dst, int B, int64_t N, int64_t n, int D, int* __restrict__ bad) {
    const int64_t total = (int64_t)B * n * D;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const int d = (int)(e % D);
        const int64_t bi = e / D;
        const int64_t b = bi / n;
        const int64_t j = idx[bi];
        if (j < 0 || j >= N) {          // torch.gather raises on the host; here the flag is read by the caller
            if (bad) atomicExch(bad, 1);
            dst[e] = 0.f;
        } else {
            dst[e] = __ldg(src + (b * N + j) * D + d);
        }
    }
}

}  // namespace npfk

using namespace npfk;

extern "C" {

int npfk_random_subset(int64_t* idx, int B, int N, int n, int64_t lo, int64_t seed, int shared, void* stream) {
    NPFK_REQUIRE(B > 0 && N >= 0 && n >= 0 && n <= N, "random_subset: need 0 <= n <= N (B=%d N=%d n=%d)", B, N, n);
    if (n == 0) return NPFK_OK;
    NPFK_REQUIRE(idx, "random_subset: null pointer");
    NPFK_REQUIRE(N <= 57344, "random_subset: N=%d exceeds the shared-memory permutation (57344 points)", N);
    const size_t smem = (size_t)N * sizeof(int);
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(random_subset_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 57344 * 4);
        if (e != cudaSuccess) {
            set_error("random_subset: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
            return NPFK_ERR_CUDA;
        }
        attr_set = true;
    }
    NPFK_LAUNCH(random_subset_kernel, shared ? 1 : B, 256, smem, as_stream(stream), idx, B, N, n, lo, (uint64_t)seed, shared);
    return check_launch("random_subset");
}

int npfk_gather_rows(const float* src, const int64_t* idx, float* dst, int B, int64_t N, int64_t n, int D, int* bad,
                     void* stream) {
    NPFK_REQUIRE(B >= 0 && N >= 0 && n >= 0 && D > 0, "gather_rows: bad shape");
    const int64_t total = (int64_t)B * n * D;
    if (total == 0) return NPFK_OK;
    NPFK_REQUIRE(src && idx && dst, "gather_rows: null pointer");
    int64_t g = ceil_div64(total, 256);
    if (g > (int64_t)kNumSMs * 16) g = (int64_t)kNumSMs * 16;
    NPFK_LAUNCH(gather_rows_kernel, (unsigned)g, 256, 0, as_stream(stream), src, idx, dst, B, N, n, D, bad);
    return check_launch("gather_rows");
}

}  // extern "C"
