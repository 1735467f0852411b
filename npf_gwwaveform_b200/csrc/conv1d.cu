// 1-D depth-separable residual conv blocks on channel-last tensors [B, L, C] (NPFK_MODE_FP32 parity path).
//
// Replaces BatchNorm1d -> ReLU -> Conv1d(groups=C) (+ residual) of ResConvBlock.forward
// (reference npf/architectures/cnn.py:204-215; depth-separable wrapper npf/utils/helpers.py:354-403) and the
// AvgPool1d / F.interpolate / cat glue of UnetCNN.apply_convs (cnn.py:451-490).  The reference permutes
// NLC -> NCL around the CNN (cnn.py:363-375); here everything stays NLC so the channel index is the
// coalesced one and the 1x1 convolutions are plain row-major Linear layers (npfk_linear_*).
#include <mutex>

#include "common.cuh"

namespace npfk {

// ------------------------------------------------------------------------- batch statistics
// One kernel for the batch statistics: grid (row chunks of kBnRows, channel tiles of 128), 8 warps per block; a lane
// owns 4 consecutive channels (16-byte loads), a warp walks every 8th row of the chunk, the warps' partial sums are
// reduced in shared memory and added to the double-precision scratch with ONE atomic per (block, channel).  The last
// block to finish (ticket in scratch[2C]) turns the sums into mean / biased variance and updates the running
// statistics, so no second launch is needed.  (Was: one thread per (32 rows, channel) with two double atomics each --
// 192-way contention per address at the UNet shapes, 13 us -- plus a finalize launch of 4 us.)
constexpr int kBnRows = 128;
__global__ void __launch_bounds__(256)
bn_stats_kernel(const float* __restrict__ x, int64_t rows, int C, double* __restrict__ scratch, float* __restrict__ stats,
                float* __restrict__ running_mean, float* __restrict__ running_var, float momentum,
                long long* __restrict__ num_batches_tracked) {
    __shared__ double red[8][128][2];
    __shared__ bool is_last;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int c0 = blockIdx.y * 128 + 4 * lane;
    const int64_t r0 = (int64_t)blockIdx.x * kBnRows, r1 = min(rows, r0 + kBnRows);
    double s[4] = {0.0, 0.0, 0.0, 0.0}, ss[4] = {0.0, 0.0, 0.0, 0.0};      // double throughout, as the statistics feed 1e-4 parity gates
    if (c0 + 3 < C && (C & 3) == 0) {
        for (int64_t r = r0 + warp; r < r1; r += 8) {
            const float4 v = __ldg(reinterpret_cast<const float4*>(x + r * C + c0));
            const double d0 = v.x, d1 = v.y, d2 = v.z, d3 = v.w;
            s[0] += d0; s[1] += d1; s[2] += d2; s[3] += d3;
            ss[0] = fma(d0, d0, ss[0]); ss[1] = fma(d1, d1, ss[1]);
            ss[2] = fma(d2, d2, ss[2]); ss[3] = fma(d3, d3, ss[3]);
        }
    } else {
        for (int64_t r = r0 + warp; r < r1; r += 8)
            for (int j = 0; j < 4; ++j)
                if (c0 + j < C) {
                    const double v = (double)__ldg(x + r * C + c0 + j);
                    s[j] += v;
                    ss[j] = fma(v, v, ss[j]);
                }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        red[warp][4 * lane + j][0] = s[j];
        red[warp][4 * lane + j][1] = ss[j];
    }
    __syncthreads();
    {
        const int ch = threadIdx.x & 127, which = threadIdx.x >> 7;      // 256 threads = 128 channels x {sum, sum of squares}
        double t = 0.0;
#pragma unroll
        for (int w = 0; w < 8; ++w) t += red[w][ch][which];
        const int c = blockIdx.y * 128 + ch;
        if (c < C) atomicAdd(scratch + which * C + c, t);
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned total = gridDim.x * gridDim.y;
        const unsigned ticket = atomicAdd(reinterpret_cast<unsigned*>(scratch + 2 * C), 1u);
        is_last = ticket == total - 1;
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    if (threadIdx.x == 0 && num_batches_tracked) *num_batches_tracked += 1;
    const double n = (double)rows;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        const double mean = __ldcg(scratch + c) / n;
        double var = __ldcg(scratch + C + c) / n - mean * mean;
        if (var < 0.0) var = 0.0;
        stats[c] = (float)mean;
        stats[C + c] = (float)var;
        if (running_mean) {
            const double unbiased = rows > 1 ? var * n / (n - 1.0) : var;
            running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * (float)mean;
            running_var[c] = (1.f - momentum) * running_var[c] + momentum * (float)unbiased;
        }
    }
}

// ------------------------------------------------------------------------- depthwise conv forward
constexpr int kDwMaxKs = 31;
struct NormArgs {
    const float* mean;    // NULL = no normalisation
    const float* var;
    const float* gamma;
    const float* beta;
    float eps;
};

__device__ __forceinline__ float pre_act(float x, float mean, float a, float beta, int relu) {
    const float h = (x - mean) * a + beta;
    return relu ? fmaxf(h, 0.f) : h;
}

// one thread per output element, channel fastest
template <int KS>
__global__ void __launch_bounds__(256)
dwconv_fwd_kernel(const float* __restrict__ x, const NormArgs nrm, const float* __restrict__ w,
                  const float* __restrict__ bias, const float* __restrict__ res, float* __restrict__ y, int B, int L,
                  int C, int ks_rt, int relu) {
    const int ks = KS > 0 ? KS : ks_rt;
    const int pad = ks / 2;
    const int64_t total = (int64_t)B * L * C;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const int c = (int)(e % C);
        const int64_t bl = e / C;
        const int l = (int)(bl % L);
        const int64_t b = bl / L;
        float mean = 0.f, a = 1.f, beta = 0.f;
        if (nrm.mean) {
            mean = __ldg(nrm.mean + c);
            a = __ldg(nrm.gamma + c) * rsqrtf(__ldg(nrm.var + c) + nrm.eps);
            beta = __ldg(nrm.beta + c);
        }
        const float* xb = x + b * L * C + c;
        float acc = bias ? __ldg(bias + c) : 0.f;
#pragma unroll
        for (int j = 0; j < (KS > 0 ? KS : kDwMaxKs); ++j) {
            if (j < ks) {
                const int li = l + j - pad;
                if (li >= 0 && li < L)
                    acc = fmaf(__ldg(w + c * ks + j), pre_act(__ldg(xb + (int64_t)li * C), mean, a, beta, relu), acc);
            }
        }
        if (res) acc += __ldg(res + e);
        y[e] = acc;
    }
}

// Register-window variant for the common kernel sizes: a thread owns (channel, LC consecutive positions), loads and
// activates its LC + KS - 1 inputs once and does LC * KS FMAs from registers (same tap order as the kernel above,
// so the sums are bit-identical).  3.4 instead of 19 loads per output at KS = 19, LC = 8, no index divisions.
template <int KS, int LC>
__global__ void __launch_bounds__(128)
dwconv_fwd_win_kernel(const float* __restrict__ x, const NormArgs nrm, const float* __restrict__ w,
                      const float* __restrict__ bias, const float* __restrict__ res, float* __restrict__ y, int L, int C,
                      int relu) {
    constexpr int PAD = KS / 2, WN = LC + KS - 1;
    const int c = blockIdx.x * 128 + threadIdx.x;
    if (c >= C) return;
    const int b = blockIdx.z;
    const int l0 = blockIdx.y * LC;
    float mean = 0.f, a = 1.f, beta = 0.f;
    if (nrm.mean) {
        mean = __ldg(nrm.mean + c);
        a = __ldg(nrm.gamma + c) * rsqrtf(__ldg(nrm.var + c) + nrm.eps);
        beta = __ldg(nrm.beta + c);
    }
    const float* xb = x + (int64_t)b * L * C + c;
    float hw[WN];
#pragma unroll
    for (int i = 0; i < WN; ++i) {
        const int li = l0 - PAD + i;
        const bool ok = li >= 0 && li < L;
        const float xv = ok ? __ldg(xb + (int64_t)li * C) : 0.f;
        hw[i] = ok ? pre_act(xv, mean, a, beta, relu) : 0.f;
    }
    float rs[LC];
#pragma unroll
    for (int l = 0; l < LC; ++l)
        rs[l] = (res && l0 + l < L) ? __ldg(res + ((int64_t)b * L + l0 + l) * C + c) : 0.f;
    float wr[KS];
#pragma unroll
    for (int j = 0; j < KS; ++j) wr[j] = __ldg(w + c * KS + j);
    const float bv = bias ? __ldg(bias + c) : 0.f;
    float* yb = y + (int64_t)b * L * C + c;
#pragma unroll
    for (int l = 0; l < LC; ++l) {
        if (l0 + l < L) {
            float acc = bv;
#pragma unroll
            for (int j = 0; j < KS; ++j) acc = fmaf(wr[j], hw[l + j], acc);   // hw is 0 outside [0, L): exact no-op
            if (res) acc += rs[l];
            yb[(int64_t)(l0 + l) * C] = acc;
        }
    }
}

// ------------------------------------------------------------------------- depthwise conv backward
// grid (C tiles of 128, L chunks, B); thread <-> channel, loops over the positions of its chunk.
constexpr int kDwLChunk = 64;
__global__ void __launch_bounds__(128)
dwconv_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ x, const NormArgs nrm,
                  const float* __restrict__ w, float* __restrict__ dh, float* __restrict__ dw,
                  float* __restrict__ dbias, float* __restrict__ sums, int L, int C, int ks, int relu) {
    const int c = blockIdx.x * 128 + threadIdx.x;
    if (c >= C) return;
    const int b = blockIdx.z;
    const int l0 = blockIdx.y * kDwLChunk, l1 = min(L, l0 + kDwLChunk);
    const int pad = ks / 2;
    float mean = 0.f, a = 1.f, beta = 0.f, invstd = 1.f;
    if (nrm.mean) {
        mean = __ldg(nrm.mean + c);
        invstd = rsqrtf(__ldg(nrm.var + c) + nrm.eps);
        a = __ldg(nrm.gamma + c) * invstd;
        beta = __ldg(nrm.beta + c);
    }
    float wr[kDwMaxKs], dwacc[kDwMaxKs];
#pragma unroll
    for (int j = 0; j < kDwMaxKs; ++j) {
        wr[j] = j < ks ? __ldg(w + c * ks + j) : 0.f;
        dwacc[j] = 0.f;
    }
    const float* xb = x + (int64_t)b * L * C + c;
    const float* dyb = dy + (int64_t)b * L * C + c;
    float* dhb = dh + (int64_t)b * L * C + c;
    float sb = 0.f, s1 = 0.f, s2 = 0.f;
    for (int l = l0; l < l1; ++l) {
        // gradient w.r.t. the activated input h[l]: correlate dy with the flipped kernel
        float g = 0.f;
#pragma unroll
        for (int j = 0; j < kDwMaxKs; ++j) {
            if (j < ks) {
                const int lo = l - j + pad;   // output position that read h[l] through tap j
                if (lo >= 0 && lo < L) g = fmaf(wr[j], __ldg(dyb + (int64_t)lo * C), g);
            }
        }
        const float xv = __ldg(xb + (int64_t)l * C);
        const float hpre = (xv - mean) * a + beta;
        if (relu && !(hpre > 0.f)) g = 0.f;
        dhb[(int64_t)l * C] = g;
        s1 += g;
        s2 = fmaf(g, (xv - mean) * invstd, s2);
        // weight gradient: dw[j] += dy[l] * h[l + j - pad]
        const float dyv = __ldg(dyb + (int64_t)l * C);
        sb += dyv;
#pragma unroll
        for (int j = 0; j < kDwMaxKs; ++j) {
            if (j < ks) {
                const int li = l + j - pad;
                if (li >= 0 && li < L)
                    dwacc[j] = fmaf(dyv, pre_act(__ldg(xb + (int64_t)li * C), mean, a, beta, relu), dwacc[j]);
            }
        }
    }
#pragma unroll
    for (int j = 0; j < kDwMaxKs; ++j)
        if (j < ks) atomicAdd(dw + c * ks + j, dwacc[j]);
    if (dbias) atomicAdd(dbias + c, sb);
    if (sums) {
        atomicAdd(sums + c, s1);
        atomicAdd(sums + C + c, s2);
    }
}

// Register-window variant for the common kernel sizes: a thread owns (channel, LC consecutive positions), loads the
// LC + KS - 1 window of dy and of the activated input once (all loads independent -> deep memory parallelism) and
// does the 2 * LC * KS FMAs from registers.  CTA = 32 channels x kDwSlices (task, position-chunk) slices; the slices'
// dw / dbias / BN sums are added up in shared memory first, so an address sees B*L/(LC*kDwSlices) atomics instead of
// B*L/LC (same-address atomics serialise in L2: that was most of the kernel's time).
constexpr int kDwSlices = 8;
template <int KS, int LC>
__global__ void __launch_bounds__(32 * kDwSlices, 2)
dwconv_bwd_win_kernel(const float* __restrict__ dy, const float* __restrict__ x, const NormArgs nrm,
                      const float* __restrict__ w, float* __restrict__ dh, float* __restrict__ dw,
                      float* __restrict__ dbias, float* __restrict__ sums, int B, int L, int C, int relu) {
    constexpr int PAD = KS / 2, WN = LC + KS - 1, NR = KS + 3;
    __shared__ float sRed[kDwSlices][NR][32];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + tx;
    const int n_lc = (L + LC - 1) / LC;
    const int64_t slice = (int64_t)blockIdx.y * kDwSlices + ty;
    const bool live = c < C && slice < (int64_t)n_lc * B;
    float dwacc[KS];
#pragma unroll
    for (int j = 0; j < KS; ++j) dwacc[j] = 0.f;
    float sb = 0.f, s1 = 0.f, s2 = 0.f;
    if (live) {
        const int b = (int)(slice / n_lc);
        const int l0 = (int)(slice % n_lc) * LC;
        float mean = 0.f, a = 1.f, beta = 0.f, invstd = 1.f;
        if (nrm.mean) {
            mean = __ldg(nrm.mean + c);
            invstd = rsqrtf(__ldg(nrm.var + c) + nrm.eps);
            a = __ldg(nrm.gamma + c) * invstd;
            beta = __ldg(nrm.beta + c);
        }
        const float* xb = x + (int64_t)b * L * C + c;
        const float* dyb = dy + (int64_t)b * L * C + c;
        float* dhb = dh + (int64_t)b * L * C + c;
        float dyw[WN], hw[WN], xc[LC];
#pragma unroll
        for (int i = 0; i < WN; ++i) {
            const int li = l0 - PAD + i;
            const bool ok = li >= 0 && li < L;
            dyw[i] = ok ? __ldg(dyb + (int64_t)li * C) : 0.f;
            const float xv = ok ? __ldg(xb + (int64_t)li * C) : 0.f;
            hw[i] = ok ? pre_act(xv, mean, a, beta, relu) : 0.f;
            if (i >= PAD && i < PAD + LC) xc[i - PAD] = xv;
        }
        float wr[KS];
#pragma unroll
        for (int j = 0; j < KS; ++j) wr[j] = __ldg(w + c * KS + j);
#pragma unroll
        for (int l = 0; l < LC; ++l) {
            if (l0 + l < L) {
                float g = 0.f;
#pragma unroll
                for (int j = 0; j < KS; ++j) g = fmaf(wr[j], dyw[l - j + 2 * PAD], g);
                const float hpre = (xc[l] - mean) * a + beta;
                if (relu && !(hpre > 0.f)) g = 0.f;
                dhb[(int64_t)(l0 + l) * C] = g;
                s1 += g;
                s2 = fmaf(g, (xc[l] - mean) * invstd, s2);
                const float dyv = dyw[l + PAD];
                sb += dyv;
#pragma unroll
                for (int j = 0; j < KS; ++j) dwacc[j] = fmaf(dyv, hw[l + j], dwacc[j]);
            }
        }
    }
#pragma unroll
    for (int j = 0; j < KS; ++j) sRed[ty][j][tx] = dwacc[j];
    sRed[ty][KS][tx] = sb;
    sRed[ty][KS + 1][tx] = s1;
    sRed[ty][KS + 2][tx] = s2;
    __syncthreads();
    for (int e = threadIdx.x; e < NR * 32; e += 32 * kDwSlices) {
        const int j = e >> 5, cx = e & 31;
        const int cc = blockIdx.x * 32 + cx;
        if (cc >= C) continue;
        float t = 0.f;
#pragma unroll
        for (int sl = 0; sl < kDwSlices; ++sl) t += sRed[sl][j][cx];
        if (j < KS) atomicAdd(dw + cc * KS + j, t);
        else if (j == KS) { if (dbias) atomicAdd(dbias + cc, t); }
        else if (sums) atomicAdd(sums + (j - KS - 1) * C + cc, t);
    }
}

__global__ void bn_bwd_kernel(const float* dh, const float* __restrict__ x, const float* __restrict__ mean,
                              const float* __restrict__ var, const float* __restrict__ gamma, float eps,
                              const float* __restrict__ sums, float* dx, float* __restrict__ dgamma,
                              float* __restrict__ dbeta, int64_t rows, int C, int training) {
    const int64_t total = rows * C;
    const float inv_n = 1.f / (float)rows;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const int c = (int)(e % C);
        const float invstd = rsqrtf(__ldg(var + c) + eps);
        const float a = __ldg(gamma + c) * invstd;
        float g = dh[e];
        if (training) {
            const float xhat = (x[e] - __ldg(mean + c)) * invstd;
            g = g - __ldg(sums + c) * inv_n - xhat * __ldg(sums + C + c) * inv_n;
        }
        dx[e] = a * g;
    }
    if (blockIdx.x == 0) {
        for (int c = threadIdx.x; c < C; c += blockDim.x) {
            dbeta[c] += sums[c];
            dgamma[c] += sums[C + c];
        }
    }
}

// ------------------------------------------------------------------------- pooling / upsampling / copies
__global__ void avgpool2_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, int B, int L, int C) {
    const int Lo = L / 2;
    const int64_t total = (int64_t)B * Lo * C;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const int c = (int)(e % C);
        const int64_t bl = e / C;
        const int lo = (int)(bl % Lo);
        const int64_t b = bl / Lo;
        const float* p = x + (b * L + 2 * lo) * C + c;
        y[e] = (p[0] + p[C]) * 0.5f;
    }
}

__global__ void avgpool2_bwd_kernel(const float* __restrict__ dy, float* __restrict__ dx, int B, int L, int C) {
    const int Lo = L / 2;
    const int64_t total = (int64_t)B * L * C;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const int c = (int)(e % C);
        const int64_t bl = e / C;
        const int l = (int)(bl % L);
        const int64_t b = bl / L;
        const int lo = l / 2;
        dx[e] = lo < Lo ? 0.5f * dy[(b * Lo + lo) * C + c] : 0.f;
    }
}

__device__ __forceinline__ void upsample_src(int lo, int L, int& i0, int& i1, float& lam) {
    // align_corners=True: src = dst * (L-1)/(2L-1), evaluated in fp32 like ATen's area_pixel_compute_scale
    const float scale = (2 * L > 1) ? (float)(L - 1) / (float)(2 * L - 1) : 0.f;
    const float src = scale * (float)lo;
    i0 = min((int)src, L - 1);
    i1 = min(i0 + 1, L - 1);
    lam = src - (float)i0;
}

__global__ void upsample2_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, int64_t ldy, int B, int L,
                                     int C) {
    const int Lo = 2 * L;
    const int64_t total = (int64_t)B * Lo * C;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const int c = (int)(e % C);
        const int64_t bl = e / C;
        const int lo = (int)(bl % Lo);
        const int64_t b = bl / Lo;
        int i0, i1;
        float lam;
        upsample_src(lo, L, i0, i1, lam);
        const float* xb = x + b * L * C + c;
        y[bl * ldy + c] = (1.f - lam) * xb[(int64_t)i0 * C] + lam * xb[(int64_t)i1 * C];
    }
}

__global__ void upsample2_bwd_kernel(const float* __restrict__ dy, int64_t lddy, float* __restrict__ dx, int B, int L,
                                     int C) {
    const int Lo = 2 * L;
    const int64_t total = (int64_t)B * Lo * C;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const int c = (int)(e % C);
        const int64_t bl = e / C;
        const int lo = (int)(bl % Lo);
        const int64_t b = bl / Lo;
        int i0, i1;
        float lam;
        upsample_src(lo, L, i0, i1, lam);
        const float g = dy[bl * lddy + c];
        float* dxb = dx + b * L * C + c;
        atomicAdd(dxb + (int64_t)i0 * C, (1.f - lam) * g);
        atomicAdd(dxb + (int64_t)i1 * C, lam * g);
    }
}

__global__ void copy2d_kernel(const float* __restrict__ src, int64_t lds, float* __restrict__ dst, int64_t ldd,
                              int64_t rows, int cols) {
    const int64_t total = rows * cols;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = e / cols;
        const int c = (int)(e - r * cols);
        dst[r * ldd + c] = src[r * lds + c];
    }
}

inline unsigned ew_grid2(int64_t total, int block = 256) {
    int64_t g = ceil_div64(total, block);
    return (unsigned)max((int64_t)1, min(g, (int64_t)kNumSMs * 16));
}

// ------------------------------------------------------------------------- npfk_bn_stats_static: scratch inside the library
// For callers that cannot allocate (no cudaMalloc inside their loop, no torch): the [2C + 1] doubles bn_stats_kernel
// accumulates into live in this file-scope array, 64 KB + 256 B per GPU that loads the library, held until exit.
// On the host `g_bn_static_scratch` is only a handle for cudaGetSymbolAddress, never a device pointer: the address is
// looked up once per device (with that device current) and that is what the kernel receives.
constexpr int kBnStaticMaxC = 4096;
constexpr int kBnStaticMaxDevices = 64;
static __device__ __align__(256) double g_bn_static_scratch[2 * kBnStaticMaxC + 32];
static double* g_bn_static_ptr[kBnStaticMaxDevices];      // device address per ordinal; written once, under the mutex
static std::mutex g_bn_static_mutex;

static int bn_static_scratch(double** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) {
        set_error("bn_stats_static: cudaGetDevice: %s", cudaGetErrorString(e));
        return NPFK_ERR_CUDA;
    }
    if (dev < 0 || dev >= kBnStaticMaxDevices) {
        set_error("bn_stats_static: device ordinal %d outside [0, %d)", dev, kBnStaticMaxDevices);
        return NPFK_ERR_INVALID;
    }
    std::lock_guard<std::mutex> lock(g_bn_static_mutex);
    if (!g_bn_static_ptr[dev]) {
        double* ptr = nullptr;
        e = cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_bn_static_scratch);
        if (e != cudaSuccess || !ptr) {
            set_error("bn_stats_static: cudaGetSymbolAddress on device %d: %s", dev, cudaGetErrorString(e));
            return NPFK_ERR_CUDA;
        }
        g_bn_static_ptr[dev] = ptr;
    }
    *out = g_bn_static_ptr[dev];
    return NPFK_OK;
}

}  // namespace npfk

using namespace npfk;

extern "C" {

int npfk_bn_stats(const float* x, int64_t rows, int C, double* scratch, float* stats, float* running_mean,
                  float* running_var, float momentum, int64_t* num_batches_tracked, void* stream) {
    NPFK_REQUIRE(x && scratch && stats, "bn_stats: null pointer");
    NPFK_REQUIRE(rows > 0 && C > 0, "bn_stats: bad shape");
    NPFK_REQUIRE((running_mean == nullptr) == (running_var == nullptr), "bn_stats: running stats must come in pairs");
    cudaStream_t st = as_stream(stream);
    cudaError_t e = cudaMemsetAsync(scratch, 0, sizeof(double) * (2 * C + 1), st);      // sums + the blocks-done ticket
    if (e != cudaSuccess) {
        set_error("bn_stats: memset: %s", cudaGetErrorString(e));
        return NPFK_ERR_CUDA;
    }
    NPFK_REQUIRE(C <= 65535 * 128, "bn_stats: too many channels");
    dim3 pgrid((unsigned)ceil_div64(rows, kBnRows), (unsigned)((C + 127) / 128));
    NPFK_REQUIRE(pgrid.x <= 2147483647u / pgrid.y, "bn_stats: grid too large");
    NPFK_LAUNCH(bn_stats_kernel, pgrid, 256, 0, st, x, rows, C, scratch, stats, running_mean, running_var, momentum,
                reinterpret_cast<long long*>(num_batches_tracked));
    return check_launch("bn_stats");
}

int npfk_bn_stats_static(const float* x, int64_t rows, int C, float* stats, float* running_mean, float* running_var,
                         float momentum, int64_t* num_batches_tracked, void* stream) {
    NPFK_REQUIRE(x && stats, "bn_stats_static: null pointer");
    NPFK_REQUIRE(rows > 0 && C > 0, "bn_stats_static: bad shape");
    NPFK_REQUIRE(C <= kBnStaticMaxC,
                 "bn_stats_static: C = %d exceeds the built-in scratch (NPFK_BN_STATS_STATIC_MAX_C = %d channels); "
                 "use npfk_bn_stats with caller scratch", C, kBnStaticMaxC);
    NPFK_REQUIRE((running_mean == nullptr) == (running_var == nullptr),
                 "bn_stats_static: running stats must come in pairs");
    double* scratch = nullptr;
    const int rc = bn_static_scratch(&scratch);
    if (rc != NPFK_OK) return rc;
    cudaStream_t st = as_stream(stream);
    // the array keeps the previous call's sums and ticket: both must start at zero
    cudaError_t e = cudaMemsetAsync(scratch, 0, sizeof(double) * (2 * C + 1), st);
    if (e != cudaSuccess) {
        set_error("bn_stats_static: memset: %s", cudaGetErrorString(e));
        return NPFK_ERR_CUDA;
    }
    dim3 pgrid((unsigned)ceil_div64(rows, kBnRows), (unsigned)((C + 127) / 128));
    NPFK_REQUIRE(pgrid.x <= 2147483647u / pgrid.y, "bn_stats_static: grid too large");
    NPFK_LAUNCH(bn_stats_kernel, pgrid, 256, 0, st, x, rows, C, scratch, stats, running_mean, running_var, momentum,
                reinterpret_cast<long long*>(num_batches_tracked));
    return check_launch("bn_stats_static");
}

int npfk_dwconv_fwd(const float* x, const float* mean, const float* var, const float* gamma, const float* beta,
                    float eps, const float* w, const float* bias, const float* res, float* y, int B, int L, int C,
                    int ks, int relu, void* stream) {
    NPFK_REQUIRE(x && w && y, "dwconv_fwd: null pointer");
    NPFK_REQUIRE(B > 0 && L > 0 && C > 0, "dwconv_fwd: bad shape");
    NPFK_REQUIRE(ks > 0 && ks % 2 == 1 && ks <= kDwMaxKs, "dwconv_fwd: kernel size %d must be odd and <= %d", ks, kDwMaxKs);
    NPFK_REQUIRE(!mean || (var && gamma && beta), "dwconv_fwd: incomplete normalisation arguments");
    NormArgs n{mean, var, gamma, beta, eps};
    const int64_t total = (int64_t)B * L * C;
    cudaStream_t st = as_stream(stream);
    if (B <= 65535) {
        constexpr int kWinLC = 8;
        dim3 wgrid((C + 127) / 128, (L + kWinLC - 1) / kWinLC, B);
#define NPFK_DW_WIN(KS_)                                                                                              \
    case KS_:                                                                                                         \
        NPFK_LAUNCH((dwconv_fwd_win_kernel<KS_, kWinLC>), wgrid, 128, 0, st, x, n, w, bias, res, y, L, C, relu);      \
        return check_launch("dwconv_fwd_win");
        switch (ks) {
            NPFK_DW_WIN(3)
            NPFK_DW_WIN(5)
            NPFK_DW_WIN(7)
            NPFK_DW_WIN(11)
            NPFK_DW_WIN(19)
            default: break;
        }
#undef NPFK_DW_WIN
    }
    const unsigned grid = ew_grid2(total);
    switch (ks) {
        case 5: NPFK_LAUNCH(dwconv_fwd_kernel<5>, grid, 256, 0, st, x, n, w, bias, res, y, B, L, C, ks, relu); break;
        case 11: NPFK_LAUNCH(dwconv_fwd_kernel<11>, grid, 256, 0, st, x, n, w, bias, res, y, B, L, C, ks, relu); break;
        case 19: NPFK_LAUNCH(dwconv_fwd_kernel<19>, grid, 256, 0, st, x, n, w, bias, res, y, B, L, C, ks, relu); break;
        default: NPFK_LAUNCH(dwconv_fwd_kernel<0>, grid, 256, 0, st, x, n, w, bias, res, y, B, L, C, ks, relu); break;
    }
    return check_launch("dwconv_fwd");
}

int npfk_dwconv_bwd(const float* dy, const float* x, const float* mean, const float* var, const float* gamma,
                    const float* beta, float eps, const float* w, float* dh, float* dw, float* dbias, float* sums,
                    int B, int L, int C, int ks, int relu, void* stream) {
    NPFK_REQUIRE(dy && x && w && dh && dw, "dwconv_bwd: null pointer");
    NPFK_REQUIRE(B > 0 && B <= 65535 && L > 0 && C > 0, "dwconv_bwd: bad shape");
    NPFK_REQUIRE(ks > 0 && ks % 2 == 1 && ks <= kDwMaxKs, "dwconv_bwd: kernel size %d must be odd and <= %d", ks, kDwMaxKs);
    NPFK_REQUIRE(!mean || (var && gamma && beta), "dwconv_bwd: incomplete normalisation arguments");
    NormArgs n{mean, var, gamma, beta, eps};
    constexpr int kWinLC = 8;
    dim3 wgrid((C + 31) / 32, (unsigned)ceil_div64((int64_t)((L + kWinLC - 1) / kWinLC) * B, kDwSlices));
#define NPFK_DW_WIN(KS_)                                                                                              \
    case KS_:                                                                                                         \
        NPFK_LAUNCH((dwconv_bwd_win_kernel<KS_, kWinLC>), wgrid, 32 * kDwSlices, 0, as_stream(stream), dy, x, n, w,   \
                    dh, dw, dbias, sums, B, L, C, relu);                                                              \
        return check_launch("dwconv_bwd_win");
    switch (ks) {
        NPFK_DW_WIN(3)
        NPFK_DW_WIN(5)
        NPFK_DW_WIN(7)
        NPFK_DW_WIN(11)
        NPFK_DW_WIN(19)
        default: break;
    }
#undef NPFK_DW_WIN
    dim3 grid((C + 127) / 128, (L + kDwLChunk - 1) / kDwLChunk, B);
    NPFK_LAUNCH(dwconv_bwd_kernel, grid, 128, 0, as_stream(stream), dy, x, n, w, dh, dw, dbias, sums, L, C, ks, relu);
    return check_launch("dwconv_bwd");
}

int npfk_bn_bwd(const float* dh, const float* x, const float* mean, const float* var, const float* gamma, float eps,
                const float* sums, float* dx, float* dgamma, float* dbeta, int64_t rows, int C, int training,
                void* stream) {
    NPFK_REQUIRE(dh && x && mean && var && gamma && sums && dx && dgamma && dbeta, "bn_bwd: null pointer");
    if (rows * C <= 0) return NPFK_OK;
    NPFK_LAUNCH(bn_bwd_kernel, ew_grid2(rows * C), 256, 0, as_stream(stream), dh, x, mean, var, gamma, eps, sums, dx,
                dgamma, dbeta, rows, C, training);
    return check_launch("bn_bwd");
}

int npfk_avgpool2_fwd(const float* x, float* y, int B, int L, int C, void* stream) {
    NPFK_REQUIRE(x && y, "avgpool2_fwd: null pointer");
    const int64_t total = (int64_t)B * (L / 2) * C;
    if (total <= 0) return NPFK_OK;
    NPFK_LAUNCH(avgpool2_fwd_kernel, ew_grid2(total), 256, 0, as_stream(stream), x, y, B, L, C);
    return check_launch("avgpool2_fwd");
}

int npfk_avgpool2_bwd(const float* dy, float* dx, int B, int L, int C, void* stream) {
    NPFK_REQUIRE(dy && dx, "avgpool2_bwd: null pointer");
    const int64_t total = (int64_t)B * L * C;
    if (total <= 0) return NPFK_OK;
    NPFK_LAUNCH(avgpool2_bwd_kernel, ew_grid2(total), 256, 0, as_stream(stream), dy, dx, B, L, C);
    return check_launch("avgpool2_bwd");
}

int npfk_upsample2_fwd(const float* x, float* y, int64_t ldy, int B, int L, int C, void* stream) {
    NPFK_REQUIRE(x && y && ldy >= C, "upsample2_fwd: bad arguments");
    const int64_t total = (int64_t)B * 2 * L * C;
    if (total <= 0) return NPFK_OK;
    NPFK_LAUNCH(upsample2_fwd_kernel, ew_grid2(total), 256, 0, as_stream(stream), x, y, ldy, B, L, C);
    return check_launch("upsample2_fwd");
}

int npfk_upsample2_bwd(const float* dy, int64_t lddy, float* dx, int B, int L, int C, void* stream) {
    NPFK_REQUIRE(dy && dx && lddy >= C, "upsample2_bwd: bad arguments");
    const int64_t total = (int64_t)B * 2 * L * C;
    if (total <= 0) return NPFK_OK;
    NPFK_LAUNCH(upsample2_bwd_kernel, ew_grid2(total), 256, 0, as_stream(stream), dy, lddy, dx, B, L, C);
    return check_launch("upsample2_bwd");
}

int npfk_copy2d(const float* src, int64_t lds, float* dst, int64_t ldd, int64_t rows, int cols, void* stream) {
    NPFK_REQUIRE(src && dst && lds >= cols && ldd >= cols, "copy2d: bad arguments");
    if (rows * cols <= 0) return NPFK_OK;
    NPFK_LAUNCH(copy2d_kernel, ew_grid2(rows * cols), 256, 0, as_stream(stream), src, lds, dst, ldd, rows, cols);
    return check_launch("copy2d");
}

}  // extern "C"
