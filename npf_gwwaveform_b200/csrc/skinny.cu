// Skinny Linear layers: one side of the layer is <= 4 wide (x_dim = y_dim = 1 inputs, the 2-wide (loc, scale)
// output).  Replaces nn.Linear(1 -> 128), nn.Linear(1 -> 32), nn.Linear(128 -> 2) of the reference MLPs
// (npf/architectures/mlp.py:95-109 as instantiated by XEncoder / XYEncoder resizer / Decoder) and their autograd.
// These are HBM-bound streaming passes over the wide [M, W] matrix, not GEMMs: a tiled GEMM wastes its tile on the
// thin dimension.  fp32 FFMA arithmetic in both numeric modes.
//
//   expand  : out[m, :W] = act/mask( bias + res + sum_s T[m,s] * Wt[s, :W] )      forward K<=4, dgrad N<=4
//   rowdot  : out[m, :S] = act( bias + sum_w X[m,w] * Wt[s,w] )                    forward N<=4, dgrad K<=4
//   reduce  : out[s, w] += sum_m T[m,s] * X[m,w]  (+ column sums)                   wgrad with N<=4 or K<=4
#include <stdlib.h>

#include "common.cuh"

namespace npfk {

constexpr int kSkMaxThin = 4;
constexpr int kSkMaxWide = 1024;

struct ExpandArgs {
    const float* T; int64_t ldt;          // thin input  [M, S]
    const float* w; int64_t w_ss, w_sw;   // weight element (s, wcol) at w[s*w_ss + wcol*w_sw]
    const float* bias;                    // [W] or null
    const float* res; int64_t ldres, res_rows;
    const float* mask; int64_t ldmask;    // output zeroed where mask <= 0
    float* out; int64_t ldo;
    int64_t M; int W, S, relu;
};

__global__ void __launch_bounds__(256) skinny_expand_kernel(const ExpandArgs a) {
    __shared__ __align__(16) float sW[kSkMaxThin][kSkMaxWide];
    __shared__ __align__(16) float sB[kSkMaxWide];
    for (int i = threadIdx.x; i < a.S * a.W; i += blockDim.x) {
        const int s = i / a.W, wc = i % a.W;
        sW[s][wc] = __ldg(a.w + s * a.w_ss + wc * a.w_sw);
    }
    for (int i = threadIdx.x; i < a.W; i += blockDim.x) sB[i] = a.bias ? __ldg(a.bias + i) : 0.f;
    __syncthreads();
    const int G = a.W >> 2;
    const int gsh = (G & (G - 1)) == 0 ? 31 - __clz(G) : -1;
    const int64_t total = a.M * G;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        // G is a power of two for every layer of the path (W = 32 / 128 / 256): shift and mask instead of a 64-bit divide
        const int64_t m = gsh >= 0 ? (e >> gsh) : e / G;
        const int c = (int)(gsh >= 0 ? (e & (G - 1)) : (e % G)) << 2;
        // products first, bias afterwards: the rounding order of addmm in the reference (matters for parity on
        // ill-conditioned gradients, e.g. the pretrained AttnCNP key transform)
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int s = 0; s < a.S; ++s) {
            const float t = __ldg(a.T + m * a.ldt + s);
            const float4 wv = *reinterpret_cast<const float4*>(&sW[s][c]);
            acc.x = fmaf(t, wv.x, acc.x);
            acc.y = fmaf(t, wv.y, acc.y);
            acc.z = fmaf(t, wv.z, acc.z);
            acc.w = fmaf(t, wv.w, acc.w);
        }
        {
            const float4 bv = *reinterpret_cast<const float4*>(&sB[c]);
            acc.x += bv.x; acc.y += bv.y; acc.z += bv.z; acc.w += bv.w;
        }
        if (a.res) {
            const float4 r = __ldg(reinterpret_cast<const float4*>(a.res + (m % a.res_rows) * a.ldres + c));
            acc.x += r.x; acc.y += r.y; acc.z += r.z; acc.w += r.w;
        }
        if (a.relu) {
            acc.x = fmaxf(acc.x, 0.f); acc.y = fmaxf(acc.y, 0.f); acc.z = fmaxf(acc.z, 0.f); acc.w = fmaxf(acc.w, 0.f);
        }
        if (a.mask) {
            const float4 k = __ldg(reinterpret_cast<const float4*>(a.mask + m * a.ldmask + c));
            acc.x = k.x > 0.f ? acc.x : 0.f;
            acc.y = k.y > 0.f ? acc.y : 0.f;
            acc.z = k.z > 0.f ? acc.z : 0.f;
            acc.w = k.w > 0.f ? acc.w : 0.f;
        }
        *reinterpret_cast<float4*>(a.out + m * a.ldo + c) = acc;
    }
}

struct RowdotArgs {
    const float* X; int64_t ldx;          // wide input [M, W]
    const float* w; int64_t w_ss, w_sw;
    const float* bias;                    // [S] or null
    const float* res; int64_t ldres, res_rows;
    float* out; int64_t ldo;
    int64_t M; int W, S, relu;
};

// one warp per row, 4 rows in flight per warp
template <int S>
__global__ void __launch_bounds__(256) skinny_rowdot_kernel(const RowdotArgs a) {
    __shared__ __align__(16) float sW[S][kSkMaxWide];
    for (int i = threadIdx.x; i < S * a.W; i += blockDim.x) {
        const int s = i / a.W, wc = i % a.W;
        sW[s][wc] = __ldg(a.w + s * a.w_ss + wc * a.w_sw);
    }
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int G = a.W >> 2;
    constexpr int R = 4;
    for (int64_t m0 = warp * R; m0 < a.M; m0 += nwarps * R) {
        float acc[R][S];
#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
            for (int s = 0; s < S; ++s) acc[r][s] = 0.f;
        for (int g = lane; g < G; g += 32) {
            float4 xv[R];
#pragma unroll
            for (int r = 0; r < R; ++r)
                xv[r] = (m0 + r < a.M) ? __ldg(reinterpret_cast<const float4*>(a.X + (m0 + r) * a.ldx) + g)
                                       : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int s = 0; s < S; ++s) {
                const float4 wv = *reinterpret_cast<const float4*>(&sW[s][g << 2]);
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    acc[r][s] = fmaf(xv[r].x, wv.x, acc[r][s]);
                    acc[r][s] = fmaf(xv[r].y, wv.y, acc[r][s]);
                    acc[r][s] = fmaf(xv[r].z, wv.z, acc[r][s]);
                    acc[r][s] = fmaf(xv[r].w, wv.w, acc[r][s]);
                }
            }
        }
#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
            for (int s = 0; s < S; ++s) acc[r][s] = warp_sum(acc[r][s]);
        if (lane < R * S) {
            const int r = lane / S, s = lane % S;
            float v = 0.f;
#pragma unroll
            for (int rr = 0; rr < R; ++rr)
#pragma unroll
                for (int ss = 0; ss < S; ++ss)
                    if (rr == r && ss == s) v = acc[rr][ss];
            const int64_t m = m0 + r;
            if (m < a.M) {
                if (a.bias) v += __ldg(a.bias + s);
                if (a.res) v += __ldg(a.res + (m % a.res_rows) * a.ldres + s);
                if (a.relu) v = fmaxf(v, 0.f);
                a.out[m * a.ldo + s] = v;
            }
        }
    }
}

struct ReduceArgs {
    const float* T; int64_t ldt;          // thin  [M, S]
    const float* X; int64_t ldx;          // wide  [M, W]
    float* out; int64_t o_ss, o_sw;       // out element (s, wcol) at out[s*o_ss + wcol*o_sw]   "+="
    float* thin_sum;                      // [S] "+=" or null  (column sums of T)
    float* wide_sum;                      // [W] "+=" or null  (column sums of X)
    int64_t M, rows_per_cta; int W, S;
};

// CTA = 256 threads = (W/4 column groups) x (row lanes); G = W/4 must divide 256.
template <int S>
__global__ void __launch_bounds__(256) skinny_reduce_kernel(const ReduceArgs a) {
    extern __shared__ float sAcc[];       // [(S + 1) * W + S]
    const int G = a.W >> 2, RL = 256 / G;
    const int cg = threadIdx.x % G, rl = threadIdx.x / G;
    for (int i = threadIdx.x; i < (S + 1) * a.W + S; i += 256) sAcc[i] = 0.f;
    __syncthreads();
    const int64_t r0 = (int64_t)blockIdx.x * a.rows_per_cta;
    const int64_t r1 = min(a.M, r0 + a.rows_per_cta);
    float4 acc[S];
    float4 ws = make_float4(0.f, 0.f, 0.f, 0.f);
    float ts[S];
#pragma unroll
    for (int s = 0; s < S; ++s) {
        acc[s] = make_float4(0.f, 0.f, 0.f, 0.f);
        ts[s] = 0.f;
    }
#pragma unroll 4
    for (int64_t m = r0 + rl; m < r1; m += RL) {
        const float4 x = __ldg(reinterpret_cast<const float4*>(a.X + m * a.ldx) + cg);
        ws.x += x.x; ws.y += x.y; ws.z += x.z; ws.w += x.w;
#pragma unroll
        for (int s = 0; s < S; ++s) {
            const float t = __ldg(a.T + m * a.ldt + s);
            ts[s] += t;
            acc[s].x = fmaf(t, x.x, acc[s].x);
            acc[s].y = fmaf(t, x.y, acc[s].y);
            acc[s].z = fmaf(t, x.z, acc[s].z);
            acc[s].w = fmaf(t, x.w, acc[s].w);
        }
    }
    const int c = cg << 2;
#pragma unroll
    for (int s = 0; s < S; ++s) {
        atomicAdd(&sAcc[s * a.W + c], acc[s].x);
        atomicAdd(&sAcc[s * a.W + c + 1], acc[s].y);
        atomicAdd(&sAcc[s * a.W + c + 2], acc[s].z);
        atomicAdd(&sAcc[s * a.W + c + 3], acc[s].w);
        if (cg == 0) atomicAdd(&sAcc[(S + 1) * a.W + s], ts[s]);
    }
    atomicAdd(&sAcc[S * a.W + c], ws.x);
    atomicAdd(&sAcc[S * a.W + c + 1], ws.y);
    atomicAdd(&sAcc[S * a.W + c + 2], ws.z);
    atomicAdd(&sAcc[S * a.W + c + 3], ws.w);
    __syncthreads();
    for (int i = threadIdx.x; i < S * a.W; i += 256) {
        const int s = i / a.W, wc = i % a.W;
        atomicAdd(a.out + s * a.o_ss + wc * a.o_sw, sAcc[i]);
    }
    if (a.wide_sum)
        for (int i = threadIdx.x; i < a.W; i += 256) atomicAdd(a.wide_sum + i, sAcc[S * a.W + i]);
    if (a.thin_sum && threadIdx.x < S) atomicAdd(a.thin_sum + threadIdx.x, sAcc[(S + 1) * a.W + threadIdx.x]);
}

// W == 128 without residual or mask (the 1 -> 128 / 2 -> 128 first layers of the x / xy encoders): a lane owns four columns
// with their weights and bias in registers, a warp writes four rows per iteration (their inputs are loaded together) with
// streaming stores.  Same arithmetic order as skinny_expand_kernel.  The generic kernel's 21 KB of static shared memory and
// one dependent load per 16 bytes written left it at half the write bandwidth (151 us for 537 MB; a fill takes 74 us).
template <int S>
__global__ void __launch_bounds__(256) skinny_expand128_kernel(const ExpandArgs a) {
    const int lane = threadIdx.x & 31, c = lane * 4;
    float4 w[S];
#pragma unroll
    for (int s = 0; s < S; ++s)
        w[s] = make_float4(__ldg(a.w + s * a.w_ss + (c + 0) * a.w_sw), __ldg(a.w + s * a.w_ss + (c + 1) * a.w_sw),
                           __ldg(a.w + s * a.w_ss + (c + 2) * a.w_sw), __ldg(a.w + s * a.w_ss + (c + 3) * a.w_sw));
    const float4 b = a.bias ? __ldg(reinterpret_cast<const float4*>(a.bias) + lane) : make_float4(0.f, 0.f, 0.f, 0.f);
    const int64_t nwarps = (int64_t)gridDim.x * 8;
    for (int64_t m0 = ((int64_t)blockIdx.x * 8 + (threadIdx.x >> 5)) * 4; m0 < a.M; m0 += nwarps * 4) {
        float t[4][S];
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int s = 0; s < S; ++s) t[r][s] = m0 + r < a.M ? __ldg(a.T + (m0 + r) * a.ldt + s) : 0.f;
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            if (m0 + r >= a.M) break;
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int s = 0; s < S; ++s) {
                acc.x = fmaf(t[r][s], w[s].x, acc.x);
                acc.y = fmaf(t[r][s], w[s].y, acc.y);
                acc.z = fmaf(t[r][s], w[s].z, acc.z);
                acc.w = fmaf(t[r][s], w[s].w, acc.w);
            }
            acc.x += b.x; acc.y += b.y; acc.z += b.z; acc.w += b.w;
            if (a.relu) {
                acc.x = fmaxf(acc.x, 0.f); acc.y = fmaxf(acc.y, 0.f); acc.z = fmaxf(acc.z, 0.f); acc.w = fmaxf(acc.w, 0.f);
            }
            __stcs(reinterpret_cast<float4*>(a.out + (m0 + r) * a.ldo + c), acc);
        }
    }
}

static inline bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
static inline unsigned sk_grid(int64_t items, int block) {
    int64_t b = ceil_div64(items, block);
    if (b < 1) b = 1;
    if (b > (int64_t)kNumSMs * 8) b = (int64_t)kNumSMs * 8;
    return (unsigned)b;
}

// ---- dispatchers: return 1 when the shape was handled, 0 when the caller should use the general GEMM, <0 on error
int skinny_linear_fwd(const float* x, int64_t ldx, const float* w, int64_t ldw, const float* bias, const float* res,
                      int64_t ldres, int64_t res_rows, float* y, int64_t ldy, int64_t M, int N, int K, int relu,
                      cudaStream_t st) {
    if (M <= 0) return 0;
    if (K <= kSkMaxThin && N % 4 == 0 && N <= kSkMaxWide && ldy % 4 == 0 && al16(y) &&
        (!res || (ldres % 4 == 0 && al16(res)))) {
        ExpandArgs a{};
        a.T = x; a.ldt = ldx;
        a.w = w; a.w_ss = 1; a.w_sw = ldw;               // Wt[s=k][wcol=n] = w[n*ldw + k]
        a.bias = bias; a.res = res; a.ldres = ldres; a.res_rows = res_rows > 0 ? res_rows : M;
        a.out = y; a.ldo = ldy; a.M = M; a.W = N; a.S = K; a.relu = relu;
        if (N == 128 && !res && K <= 2 && (!bias || al16(bias))) {
            const unsigned grid = sk_grid(ceil_div64(M, 4) * 32, 256);
            if (K == 1) NPFK_LAUNCH(skinny_expand128_kernel<1>, grid, 256, 0, st, a);
            else NPFK_LAUNCH(skinny_expand128_kernel<2>, grid, 256, 0, st, a);
        } else {
            NPFK_LAUNCH(skinny_expand_kernel, sk_grid(M * (N / 4), 256), 256, 0, st, a);
        }
        int rc = check_launch("skinny_expand(fwd)");
        return rc ? rc : 1;
    }
    if (N <= kSkMaxThin && K % 4 == 0 && K <= kSkMaxWide && ldx % 4 == 0 && al16(x)) {
        RowdotArgs a{};
        a.X = x; a.ldx = ldx;
        a.w = w; a.w_ss = ldw; a.w_sw = 1;               // Wt[s=n][wcol=k] = w[n*ldw + k]
        a.bias = bias; a.res = res; a.ldres = ldres; a.res_rows = res_rows > 0 ? res_rows : M;
        a.out = y; a.ldo = ldy; a.M = M; a.W = K; a.S = N; a.relu = relu;
        const unsigned grid = sk_grid(ceil_div64(M, 4) * 32, 256);
        switch (N) {
            case 1: NPFK_LAUNCH(skinny_rowdot_kernel<1>, grid, 256, 0, st, a); break;
            case 2: NPFK_LAUNCH(skinny_rowdot_kernel<2>, grid, 256, 0, st, a); break;
            case 3: NPFK_LAUNCH(skinny_rowdot_kernel<3>, grid, 256, 0, st, a); break;
            default: NPFK_LAUNCH(skinny_rowdot_kernel<4>, grid, 256, 0, st, a); break;
        }
        int rc = check_launch("skinny_rowdot(fwd)");
        return rc ? rc : 1;
    }
    return 0;
}

int skinny_linear_bwd_dx(const float* g, int64_t ldg, const float* w, int64_t ldw, const float* mask, int64_t ldmask,
                         float* dx, int64_t lddx, int64_t M, int N, int K, cudaStream_t st) {
    if (M <= 0) return 0;
    if (N <= kSkMaxThin && K % 4 == 0 && K <= kSkMaxWide && lddx % 4 == 0 && al16(dx) &&
        (!mask || (ldmask % 4 == 0 && al16(mask)))) {
        ExpandArgs a{};
        a.T = g; a.ldt = ldg;
        a.w = w; a.w_ss = ldw; a.w_sw = 1;               // Wt[s=n][wcol=k] = w[n*ldw + k]
        a.res_rows = M;
        a.mask = mask; a.ldmask = ldmask;
        a.out = dx; a.ldo = lddx; a.M = M; a.W = K; a.S = N;
        NPFK_LAUNCH(skinny_expand_kernel, sk_grid(M * (K / 4), 256), 256, 0, st, a);
        int rc = check_launch("skinny_expand(dgrad)");
        return rc ? rc : 1;
    }
    if (K <= kSkMaxThin && !mask && N % 4 == 0 && N <= kSkMaxWide && ldg % 4 == 0 && al16(g)) {
        RowdotArgs a{};
        a.X = g; a.ldx = ldg;
        a.w = w; a.w_ss = 1; a.w_sw = ldw;               // Wt[s=k][wcol=n] = w[n*ldw + k]
        a.res_rows = M;
        a.out = dx; a.ldo = lddx; a.M = M; a.W = N; a.S = K;
        const unsigned grid = sk_grid(ceil_div64(M, 4) * 32, 256);
        switch (K) {
            case 1: NPFK_LAUNCH(skinny_rowdot_kernel<1>, grid, 256, 0, st, a); break;
            case 2: NPFK_LAUNCH(skinny_rowdot_kernel<2>, grid, 256, 0, st, a); break;
            case 3: NPFK_LAUNCH(skinny_rowdot_kernel<3>, grid, 256, 0, st, a); break;
            default: NPFK_LAUNCH(skinny_rowdot_kernel<4>, grid, 256, 0, st, a); break;
        }
        int rc = check_launch("skinny_rowdot(dgrad)");
        return rc ? rc : 1;
    }
    return 0;
}

static int sk_reduce_ctas_per_sm() {
    static int v = -1;      // NPFK_SKINNY_REDUCE_CTAS (developer knob)
    if (v < 0) {
        const char* e = getenv("NPFK_SKINNY_REDUCE_CTAS");
        v = e ? atoi(e) : 6;      // 40 registers x 256 threads: six CTAs are resident per SM, so 6 x 148 CTAs are exactly one wave
        if (v < 1) v = 1;      // (4: 124 us, 6: 96 us, 8: 129 us -- a third of a second wave --, 12: 99 us for 537 MB)
    }
    return v;
}

int skinny_linear_bwd_dw(const float* g, int64_t ldg, const float* x, int64_t ldx, float* dw, int64_t lddw, float* db,
                         int64_t M, int N, int K, cudaStream_t st) {
    if (M <= 0) return 0;
    ReduceArgs a{};
    a.M = M;
    int S;
    auto wide_ok = [](int W, int64_t ld, const float* p) {
        const int G = W / 4;
        return W % 4 == 0 && W <= kSkMaxWide && G >= 1 && 256 % G == 0 && ld % 4 == 0 && al16(p);
    };
    if (N <= kSkMaxThin && wide_ok(K, ldx, x)) {          // dw[n, k] = sum_m g[m,n] x[m,k] ; db = thin sums
        a.T = g; a.ldt = ldg; a.X = x; a.ldx = ldx;
        a.out = dw; a.o_ss = lddw; a.o_sw = 1;
        a.thin_sum = db; a.W = K; S = N;
    } else if (K <= kSkMaxThin && wide_ok(N, ldg, g)) {   // dw[n, k] = sum_m x[m,k] g[m,n] ; db = wide sums
        a.T = x; a.ldt = ldx; a.X = g; a.ldx = ldg;
        a.out = dw; a.o_ss = 1; a.o_sw = lddw;
        a.wide_sum = db; a.W = N; S = K;
    } else {
        return 0;
    }
    a.S = S;
    int64_t ctas = ceil_div64(M, 64);         // at least 64 rows per CTA, at most kSkReduceCtasPerSm CTAs per SM
    if (ctas > (int64_t)kNumSMs * sk_reduce_ctas_per_sm()) ctas = (int64_t)kNumSMs * sk_reduce_ctas_per_sm();
    a.rows_per_cta = ceil_div64(M, ctas);
    ctas = ceil_div64(M, a.rows_per_cta);
    const size_t smem = sizeof(float) * ((size_t)(S + 1) * a.W + S);
    switch (S) {
        case 1: NPFK_LAUNCH(skinny_reduce_kernel<1>, (unsigned)ctas, 256, smem, st, a); break;
        case 2: NPFK_LAUNCH(skinny_reduce_kernel<2>, (unsigned)ctas, 256, smem, st, a); break;
        case 3: NPFK_LAUNCH(skinny_reduce_kernel<3>, (unsigned)ctas, 256, smem, st, a); break;
        default: NPFK_LAUNCH(skinny_reduce_kernel<4>, (unsigned)ctas, 256, smem, st, a); break;
    }
    int rc = check_launch("skinny_reduce(wgrad)");
    return rc ? rc : 1;
}

}  // namespace npfk
