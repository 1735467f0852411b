// fp32 (FFMA) GEMM family behind npfk_linear_* : the NPFK_MODE_FP32 parity path.
//
// One register-tiled kernel serves the three products of a Linear layer by describing each operand
// with two element strides:
//   forward  Y[M,N]  = X[M,K] . W[N,K]^T          (reduction over K, both operands reduction-contiguous)
//   dgrad    dX[M,K] = G[M,N] . W[N,K]            (reduction over N)
//   wgrad    dW[N,K] = G[M,N]^T . X[M,K]          (reduction over M, split across CTAs, atomicAdd)
// Replaces cuBLAS addmm + separate bias/ReLU kernels of MLP.forward (reference npf/architectures/mlp.py:95-109).
#include "common.cuh"

namespace npfk {

struct GemmArgs {
    const float* A; int64_t a_row, a_red;   // A(i, r) = A[i*a_row + r*a_red]
    const float* B; int64_t b_red, b_col;   // B(r, j) = B[r*b_red + j*b_col]
    float* C; int64_t ldc;                  // C(i, j) = C[i*ldc + j]
    int64_t I; int J; int64_t R;            // output rows / cols, reduction length
    const float* bias;                      // [J] or null
    const float* res; int64_t ldres, res_rows;
    const float* mask; int64_t ldmask;      // output zeroed where mask(i, j) <= 0 (ReLU backward of the layer below)
    int relu; int atomic;
    int64_t r_chunk;                        // reduction elements handled per blockIdx.z
};

template <int BI, int BJ, int BR, int TI, int TJ, bool A_ROW_CONTIG, bool B_COL_CONTIG>
__global__ void __launch_bounds__((BI / TI) * (BJ / TJ), 2)
sgemm_kernel(const GemmArgs p) {
    constexpr int NT = (BI / TI) * (BJ / TJ);
    constexpr int LDA = BI + 4, LDB = BJ + 4;
    constexpr int A_PER = (BI * BR) / NT, B_PER = (BJ * BR + NT - 1) / NT;
    static_assert((BI * BR) % NT == 0, "A tile must divide evenly");
    __shared__ __align__(16) float As[BR * LDA];
    __shared__ __align__(16) float Bs[BR * LDB];

    const int tid = threadIdx.x;
    const int tx = tid % (BJ / TJ), ty = tid / (BJ / TJ);
    const int64_t i0 = (int64_t)blockIdx.x * BI;
    const int j0 = blockIdx.y * BJ;
    const int64_t r_begin = (int64_t)blockIdx.z * p.r_chunk;
    const int64_t r_end = min(p.R, r_begin + p.r_chunk);
    if (r_begin >= r_end && p.atomic) return;

    float acc[TI][TJ];
#pragma unroll
    for (int a = 0; a < TI; ++a)
#pragma unroll
        for (int b = 0; b < TJ; ++b) acc[a][b] = 0.f;

    float ra[A_PER], rb[B_PER];

    auto load_tiles = [&](int64_t r0) {
#pragma unroll
        for (int it = 0; it < A_PER; ++it) {
            const int e = tid + it * NT;
            const int r = A_ROW_CONTIG ? e / BI : e % BR;
            const int i = A_ROW_CONTIG ? e % BI : e / BR;
            const int64_t gi = i0 + i, gr = r0 + r;
            ra[it] = (gi < p.I && gr < r_end) ? __ldg(p.A + gi * p.a_row + gr * p.a_red) : 0.f;
        }
#pragma unroll
        for (int it = 0; it < B_PER; ++it) {
            const int e = tid + it * NT;
            if (e < BJ * BR) {
                const int r = B_COL_CONTIG ? e / BJ : e % BR;
                const int j = B_COL_CONTIG ? e % BJ : e / BR;
                const int64_t gr = r0 + r;
                const int gj = j0 + j;
                rb[it] = (gj < p.J && gr < r_end) ? __ldg(p.B + gr * p.b_red + (int64_t)gj * p.b_col) : 0.f;
            }
        }
    };
    auto store_tiles = [&]() {
#pragma unroll
        for (int it = 0; it < A_PER; ++it) {
            const int e = tid + it * NT;
            const int r = A_ROW_CONTIG ? e / BI : e % BR;
            const int i = A_ROW_CONTIG ? e % BI : e / BR;
            As[r * LDA + i] = ra[it];
        }
#pragma unroll
        for (int it = 0; it < B_PER; ++it) {
            const int e = tid + it * NT;
            if (e < BJ * BR) {
                const int r = B_COL_CONTIG ? e / BJ : e % BR;
                const int j = B_COL_CONTIG ? e % BJ : e / BR;
                Bs[r * LDB + j] = rb[it];
            }
        }
    };
    // local row / column owned by accumulator (a, b)
    auto row_of = [&](int a) { return (TI == 8) ? (a / 4) * (BI / 2) + ty * 4 + (a % 4) : ty * TI + a; };
    auto col_of = [&](int b) { return (TJ == 8) ? (b / 4) * (BJ / 2) + tx * 4 + (b % 4) : tx * TJ + b; };

    load_tiles(r_begin);
    store_tiles();
    __syncthreads();
    for (int64_t r0 = r_begin; r0 < r_end; r0 += BR) {
        const bool more = r0 + BR < r_end;
        if (more) load_tiles(r0 + BR);
#pragma unroll
        for (int r = 0; r < BR; ++r) {
            float av[TI], bv[TJ];
            if constexpr (TI == 8) {
                const float4 a0 = *reinterpret_cast<const float4*>(&As[r * LDA + ty * 4]);
                const float4 a1 = *reinterpret_cast<const float4*>(&As[r * LDA + BI / 2 + ty * 4]);
                av[0] = a0.x; av[1] = a0.y; av[2] = a0.z; av[3] = a0.w;
                av[4] = a1.x; av[5] = a1.y; av[6] = a1.z; av[7] = a1.w;
            } else {
#pragma unroll
                for (int a = 0; a < TI; ++a) av[a] = As[r * LDA + ty * TI + a];
            }
            if constexpr (TJ == 8) {
                const float4 b0 = *reinterpret_cast<const float4*>(&Bs[r * LDB + tx * 4]);
                const float4 b1 = *reinterpret_cast<const float4*>(&Bs[r * LDB + BJ / 2 + tx * 4]);
                bv[0] = b0.x; bv[1] = b0.y; bv[2] = b0.z; bv[3] = b0.w;
                bv[4] = b1.x; bv[5] = b1.y; bv[6] = b1.z; bv[7] = b1.w;
            } else {
#pragma unroll
                for (int b = 0; b < TJ; ++b) bv[b] = Bs[r * LDB + tx * TJ + b];
            }
#pragma unroll
            for (int a = 0; a < TI; ++a)
#pragma unroll
                for (int b = 0; b < TJ; ++b) acc[a][b] = fmaf(av[a], bv[b], acc[a][b]);
        }
        __syncthreads();
        if (more) {
            store_tiles();
            __syncthreads();
        }
    }

    const int64_t res_rows = p.res_rows > 0 ? p.res_rows : p.I;
#pragma unroll
    for (int a = 0; a < TI; ++a) {
        const int64_t gi = i0 + row_of(a);
        if (gi >= p.I) continue;
#pragma unroll
        for (int b = 0; b < TJ; ++b) {
            const int gj = j0 + col_of(b);
            if (gj >= p.J) continue;
            float v = acc[a][b];
            if (p.atomic) {
                atomicAdd(p.C + gi * p.ldc + gj, v);
            } else {
                if (p.bias) v += __ldg(p.bias + gj);
                if (p.res) v += __ldg(p.res + (gi % res_rows) * p.ldres + gj);
                if (p.relu) v = fmaxf(v, 0.f);
                if (p.mask && !(__ldg(p.mask + gi * p.ldmask + gj) > 0.f)) v = 0.f;
                p.C[gi * p.ldc + gj] = v;
            }
        }
    }
}

template <bool A_ROW_CONTIG, bool B_COL_CONTIG>
static int launch_gemm(GemmArgs& a, bool split_reduction, cudaStream_t st) {
    if (a.I <= 0 || a.J <= 0) return NPFK_OK;
    const bool narrow = a.J <= 32;
    const int BI = 128, BJ = narrow ? 32 : 128, BR = 16;
    const int64_t gx = ceil_div64(a.I, BI);
    const int gy = (a.J + BJ - 1) / BJ;
    int64_t splits = 1;
    if (split_reduction) {
        const int64_t want = (4 * kNumSMs + gx * gy - 1) / (gx * gy);
        splits = max((int64_t)1, min(want, ceil_div64(a.R, 8 * BR)));
    }
    a.r_chunk = ceil_div64(ceil_div64(a.R, splits), BR) * BR;
    splits = ceil_div64(a.R, a.r_chunk);
    if (splits < 1) splits = 1;
    a.atomic = split_reduction ? 1 : 0;
    NPFK_REQUIRE(gx <= 2147483647LL && gy <= 65535 && splits <= 65535, "gemm grid too large");
    dim3 grid((unsigned)gx, (unsigned)gy, (unsigned)splits);
    if (narrow) {
        NPFK_LAUNCH((sgemm_kernel<128, 32, 16, 8, 2, A_ROW_CONTIG, B_COL_CONTIG>), grid, 256, 0, st, a);
    } else {
        NPFK_LAUNCH((sgemm_kernel<128, 128, 16, 8, 8, A_ROW_CONTIG, B_COL_CONTIG>), grid, 256, 0, st, a);
    }
    return check_launch("sgemm");
}

// ---- elementwise / reduction helpers ------------------------------------------------------------
__global__ void relu_bwd_kernel(const float* __restrict__ dy, int64_t lddy, const float* __restrict__ y, int64_t ldy,
                                float* __restrict__ g, int64_t ldg, int64_t M, int N) {
    const int64_t total = M * N;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const int64_t m = e / N;
        const int n = (int)(e - m * N);
        g[m * ldg + n] = y[m * ldy + n] > 0.f ? dy[m * lddy + n] : 0.f;
    }
}

// db[n] += sum_m g[m,n] : 32 columns per warp row, 8 row lanes per CTA, ROWS_PER_CTA rows per CTA
constexpr int kColsumRows = 2048;
__global__ void colsum_kernel(const float* __restrict__ g, int64_t ldg, float* __restrict__ out, int64_t M, int N) {
    __shared__ float sm[8][33];
    const int lane = threadIdx.x & 31, wl = threadIdx.x >> 5;
    const int64_t m0 = (int64_t)blockIdx.x * kColsumRows;
    const int64_t m1 = min(M, m0 + kColsumRows);
    for (int n0 = blockIdx.y * 32; n0 < N; n0 += gridDim.y * 32) {
        const int n = n0 + lane;
        float s = 0.f;
        if (n < N)
            for (int64_t m = m0 + wl; m < m1; m += 8) s += g[m * ldg + n];
        sm[wl][lane] = s;
        __syncthreads();
        if (wl == 0) {
            float t = 0.f;
#pragma unroll
            for (int w = 0; w < 8; ++w) t += sm[w][lane];
            if (n < N) atomicAdd(out + n, t);
        }
        __syncthreads();
    }
}

__global__ void rows_fold_sum_kernel(const float* __restrict__ g, int64_t ldg, float* __restrict__ out, int64_t ldo,
                                     int64_t rows, int64_t reps, int N) {
    const int64_t total = rows * N;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = e / N;
        const int n = (int)(e - r * N);
        float s = 0.f;
        for (int64_t j = 0; j < reps; ++j) s += g[(j * rows + r) * ldg + n];
        out[r * ldo + n] = s;
    }
}

inline unsigned grid_for(int64_t total, int block, int max_blocks = kNumSMs * 16) {
    int64_t b = ceil_div64(total, block);
    if (b < 1) b = 1;
    if (b > max_blocks) b = max_blocks;
    return (unsigned)b;
}

// tc_gemm.cu: tcgen05 kernels for NPFK_MODE_FAST
namespace tc {
bool nt_eligible(const float* a, int64_t lda, const float* b, int64_t ldb, const float* c, int64_t ldc, int64_t M, int N,
                 int K, const float* res, int64_t ldres, const float* mask, int64_t ldmask);
bool tn_eligible(const float* g, int64_t ldg, const float* x, int64_t ldx, int64_t M, int N, int K);
int tc_linear_fwd(const float* x, int64_t ldx, const float* w, int64_t ldw, const float* bias, const float* res,
                  int64_t ldres, int64_t res_rows, float* y, int64_t ldy, int64_t M, int N, int K, int relu,
                  int round_out, cudaStream_t st);
int tc_linear_bwd_dx(const float* g, int64_t ldg, const float* w, int64_t ldw, const float* mask, int64_t ldmask,
                     float* dx, int64_t lddx, int64_t M, int N, int K, cudaStream_t st);
int tc_linear_bwd_dw(const float* g, int64_t ldg, const float* x, int64_t ldx, float* dw, int64_t lddw, float* db,
                     int64_t M, int N, int K, cudaStream_t st);
bool fused_bwd_eligible(const float* g, int64_t ldg, const float* h, int64_t ldh, const float* w, int64_t ldw,
                        const float* dx, int64_t lddx, const float* dw, int64_t lddw, int64_t M, int N, int K);
int tc_linear_bwd_fused(const float* g, int64_t ldg, const float* h, int64_t ldh, const float* w, int64_t ldw, int masked,
                        float* dx, int64_t lddx, float* dw, int64_t lddw, float* db, int64_t M, cudaStream_t st);
}  // namespace tc

// skinny.cu: streaming kernels for layers with K <= 4 or N <= 4 (return 1 = handled, 0 = not applicable)
int skinny_linear_fwd(const float* x, int64_t ldx, const float* w, int64_t ldw, const float* bias, const float* res,
                      int64_t ldres, int64_t res_rows, float* y, int64_t ldy, int64_t M, int N, int K, int relu,
                      cudaStream_t st);
int skinny_linear_bwd_dx(const float* g, int64_t ldg, const float* w, int64_t ldw, const float* mask, int64_t ldmask,
                         float* dx, int64_t lddx, int64_t M, int N, int K, cudaStream_t st);
int skinny_linear_bwd_dw(const float* g, int64_t ldg, const float* x, int64_t ldx, float* dw, int64_t lddw, float* db,
                         int64_t M, int N, int K, cudaStream_t st);

__global__ void round_tf32_kernel(const float* __restrict__ src, float* __restrict__ dst, int64_t n) {
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (int64_t)gridDim.x * blockDim.x) {
        uint32_t r;
        asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(src[e]));
        dst[e] = __uint_as_float(r);
    }
}

}  // namespace npfk

using namespace npfk;

static bool skinny_enabled() {       // NPFK_NO_SKINNY=1 routes the thin layers through the tiled GEMM (debugging)
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("NPFK_NO_SKINNY");
        v = (e && e[0] == '1') ? 0 : 1;
    }
    return v != 0;
}

extern "C" {

int npfk_linear_fwd(const float* x, int64_t ldx, const float* w, int64_t ldw, const float* bias, const float* res,
                    int64_t ldres, int64_t res_rows, float* y, int64_t ldy, int64_t M, int N, int K, int relu,
                    int mode, void* stream) {
    NPFK_REQUIRE(x && w && y, "linear_fwd: null pointer");
    NPFK_REQUIRE(M >= 0 && N > 0 && K > 0, "linear_fwd: bad shape M=%lld N=%d K=%d", (long long)M, N, K);
    NPFK_REQUIRE(ldx >= K && ldw >= K && ldy >= N, "linear_fwd: leading dimension too small");
    NPFK_REQUIRE(mode == NPFK_MODE_FP32 || mode == NPFK_MODE_FAST, "linear_fwd: unknown mode %d", mode);
    if ((K <= 4 || N <= 4) && skinny_enabled()) {
        const int rc = skinny_linear_fwd(x, ldx, w, ldw, bias, res, ldres, res_rows, y, ldy, M, N, K, relu, as_stream(stream));
        if (rc != 0) return rc < 0 ? rc : NPFK_OK;
    }
    if (mode == NPFK_MODE_FAST && tc::nt_eligible(x, ldx, w, ldw, y, ldy, M, N, K, res, ldres, nullptr, 0))
        return tc::tc_linear_fwd(x, ldx, w, ldw, bias, res, ldres, res_rows, y, ldy, M, N, K, relu, /*round_out=*/relu,
                                 as_stream(stream));
    GemmArgs a{};
    a.A = x; a.a_row = ldx; a.a_red = 1;
    a.B = w; a.b_red = 1; a.b_col = ldw;
    a.C = y; a.ldc = ldy;
    a.I = M; a.J = N; a.R = K;
    a.bias = bias; a.res = res; a.ldres = ldres; a.res_rows = res_rows;
    a.relu = relu;
    return launch_gemm<false, false>(a, false, as_stream(stream));
}

int npfk_relu_bwd(const float* dy, int64_t lddy, const float* y, int64_t ldy, float* g, int64_t ldg, int64_t M, int N,
                  void* stream) {
    NPFK_REQUIRE(dy && y && g, "relu_bwd: null pointer");
    if (M * N == 0) return NPFK_OK;
    NPFK_LAUNCH(relu_bwd_kernel, grid_for(M * N, 256), 256, 0, as_stream(stream), dy, lddy, y, ldy, g, ldg, M, N);
    return check_launch("relu_bwd");
}

int npfk_linear_bwd_dx(const float* g, int64_t ldg, const float* w, int64_t ldw, const float* mask, int64_t ldmask,
                       float* dx, int64_t lddx, int64_t M, int N, int K, int mode, void* stream) {
    NPFK_REQUIRE(g && w && dx, "linear_bwd_dx: null pointer");
    NPFK_REQUIRE(mode == NPFK_MODE_FP32 || mode == NPFK_MODE_FAST, "linear_bwd_dx: unknown mode %d", mode);
    if ((K <= 4 || N <= 4) && skinny_enabled()) {
        const int rc = skinny_linear_bwd_dx(g, ldg, w, ldw, mask, ldmask, dx, lddx, M, N, K, as_stream(stream));
        if (rc != 0) return rc < 0 ? rc : NPFK_OK;
    }
    if (mode == NPFK_MODE_FAST && tc::nt_eligible(g, ldg, w, ldw, dx, lddx, M, K, N, nullptr, 0, mask, ldmask))
        return tc::tc_linear_bwd_dx(g, ldg, w, ldw, mask, ldmask, dx, lddx, M, N, K, as_stream(stream));
    GemmArgs a{};
    a.mask = mask; a.ldmask = ldmask;
    a.A = g; a.a_row = ldg; a.a_red = 1;
    a.B = w; a.b_red = ldw; a.b_col = 1;
    a.C = dx; a.ldc = lddx;
    a.I = M; a.J = K; a.R = N;
    return launch_gemm<false, true>(a, false, as_stream(stream));
}

int npfk_linear_bwd_dw(const float* g, int64_t ldg, const float* x, int64_t ldx, float* dw, int64_t lddw, float* db,
                       int64_t M, int N, int K, int mode, void* stream) {
    NPFK_REQUIRE(g && x && dw, "linear_bwd_dw: null pointer");
    NPFK_REQUIRE(mode == NPFK_MODE_FP32 || mode == NPFK_MODE_FAST, "linear_bwd_dw: unknown mode %d", mode);
    if (M <= 0) return NPFK_OK;
    if ((K <= 4 || N <= 4) && skinny_enabled()) {
        const int rc = skinny_linear_bwd_dw(g, ldg, x, ldx, dw, lddw, db, M, N, K, as_stream(stream));
        if (rc != 0) return rc < 0 ? rc : NPFK_OK;
    }
    if (mode == NPFK_MODE_FAST && tc::tn_eligible(g, ldg, x, ldx, M, N, K))
        return tc::tc_linear_bwd_dw(g, ldg, x, ldx, dw, lddw, db, M, N, K, as_stream(stream));
    GemmArgs a{};
    a.A = g; a.a_row = 1; a.a_red = ldg;
    a.B = x; a.b_red = ldx; a.b_col = 1;
    a.C = dw; a.ldc = lddw;
    a.I = N; a.J = K; a.R = M;
    int rc = launch_gemm<true, true>(a, true, as_stream(stream));
    if (rc != NPFK_OK) return rc;
    if (db) {
        dim3 grid((unsigned)ceil_div64(M, kColsumRows), (unsigned)min(64, (N + 31) / 32));
        NPFK_LAUNCH(colsum_kernel, grid, 256, 0, as_stream(stream), g, ldg, db, M, N);
        rc = check_launch("colsum");
    }
    return rc;
}

int npfk_round_tf32(const float* src, float* dst, int64_t n, void* stream) {
    NPFK_REQUIRE(src && dst, "round_tf32: null pointer");
    if (n == 0) return NPFK_OK;
    NPFK_LAUNCH(round_tf32_kernel, grid_for(n, 256), 256, 0, as_stream(stream), src, dst, n);
    return check_launch("round_tf32");
}

int npfk_rows_fold_sum(const float* g, int64_t ldg, float* out, int64_t ldo, int64_t rows, int64_t reps, int N,
                       void* stream) {
    NPFK_REQUIRE(g && out, "rows_fold_sum: null pointer");
    if (rows * N == 0) return NPFK_OK;
    NPFK_LAUNCH(rows_fold_sum_kernel, grid_for(rows * N, 256), 256, 0, as_stream(stream), g, ldg, out, ldo, rows, reps,
                N);
    return check_launch("rows_fold_sum");
}


int npfk_linear_bwd_fused(const float* g, int64_t ldg, const float* x, int64_t ldx, const float* w, int64_t ldw,
                          int masked, float* dx, int64_t lddx, float* dw, int64_t lddw, float* db, int64_t M, int N,
                          int K, int mode, void* stream) {
    NPFK_REQUIRE(g && x && w && dx && dw, "linear_bwd_fused: null pointer");
    NPFK_REQUIRE(mode == NPFK_MODE_FP32 || mode == NPFK_MODE_FAST, "linear_bwd_fused: unknown mode %d", mode);
    if (M <= 0) return NPFK_OK;
    if (mode == NPFK_MODE_FAST && tc::fused_bwd_eligible(g, ldg, x, ldx, w, ldw, dx, lddx, dw, lddw, M, N, K))
        return tc::tc_linear_bwd_fused(g, ldg, x, ldx, w, ldw, masked, dx, lddx, dw, lddw, db, M, as_stream(stream));
    if (masked & NPFK_BWD_DX_ACCUMULATE) {
        set_error("linear_bwd_fused: dx accumulation needs the fused tensor-core kernel (fast mode, N = K = 128, M >= 128, "
                  "16-byte aligned operands); got M=%lld N=%d K=%d mode=%d", (long long)M, N, K, mode);
        return NPFK_ERR_UNSUPPORTED;
    }
    masked &= 1;
    int rc = npfk_linear_bwd_dw(g, ldg, x, ldx, dw, lddw, db, M, N, K, mode, stream);
    if (rc != NPFK_OK) return rc;
    return npfk_linear_bwd_dx(g, ldg, w, ldw, masked ? x : nullptr, masked ? ldx : 0, dx, lddx, M, N, K, mode, stream);
}

}  // extern "C"
