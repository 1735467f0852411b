// Predictive-distribution tail, Gaussian log-likelihood reductions, latent statistics, KL, input range flag
// and the flat Adam step.  All are HBM-bound streaming kernels.
//
// Reference: decode tail npf/neuralproc/base.py:347-353,116; sum_log_prob npf/losses.py:18-24;
// latent stats base.py:538-545; rsample base.py:512; KL npf/losses.py:146-148; range check base.py:241-247.
#include "common.cuh"

namespace npfk {

constexpr float kLogSqrt2Pi = 0.9189385332046727f;

inline unsigned ew_grid3(int64_t total, int block = 256) {
    int64_t g = ceil_div64(total, block);
    return (unsigned)max((int64_t)1, min(g, (int64_t)kNumSMs * 16));
}

// n rows of [2*ydim] -> loc[n*ydim], scale[n*ydim]
__global__ void loc_scale_fwd_kernel(const float* __restrict__ suff, float* __restrict__ loc, float* __restrict__ scale,
                                     int64_t n, int ydim) {
    const int64_t total = n * ydim;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = e / ydim;
        const int j = (int)(e - r * ydim);
        loc[e] = suff[r * 2 * ydim + j];
        scale[e] = 0.01f + 0.99f * softplus_f(suff[r * 2 * ydim + ydim + j]);
    }
}

__global__ void loc_scale_bwd_kernel(const float* __restrict__ suff, const float* __restrict__ dloc,
                                     const float* __restrict__ dscale, float* __restrict__ dsuff, int64_t n, int ydim) {
    const int64_t total = n * ydim;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = e / ydim;
        const int j = (int)(e - r * ydim);
        dsuff[r * 2 * ydim + j] = dloc ? dloc[e] : 0.f;
        const float s = suff[r * 2 * ydim + ydim + j];
        dsuff[r * 2 * ydim + ydim + j] = dscale ? dscale[e] * 0.99f * sigmoid_f(s) : 0.f;   // softplus' = sigmoid
    }
}

__device__ __forceinline__ float log_prob_from_suff(float mu, float sraw, float y) {
    const float sc = 0.01f + 0.99f * softplus_f(sraw);
    const float d = y - mu;
    return -(d * d) / (2.f * sc * sc) - logf(sc) - kLogSqrt2Pi;
}

// grid (T-chunks, B, Z): block reduces its chunk of targets, one atomicAdd per block
constexpr int kNllChunk = 2048;
__global__ void __launch_bounds__(256)
gauss_nll_fwd_kernel(const float* __restrict__ suff, const float* __restrict__ y, float* __restrict__ sum_lp, int B,
                     int64_t T, int ydim) {
    __shared__ float sm[8];
    const int z = blockIdx.z, b = blockIdx.y;
    const int64_t e0 = (int64_t)blockIdx.x * kNllChunk, e1 = min(T * ydim, e0 + kNllChunk);
    const float* sb = suff + ((int64_t)z * B + b) * T * 2 * ydim;
    const float* yb = y + (int64_t)b * T * ydim;
    float s = 0.f;
    for (int64_t e = e0 + threadIdx.x; e < e1; e += blockDim.x) {
        const int64_t t = e / ydim;
        const int j = (int)(e - t * ydim);
        s += log_prob_from_suff(sb[t * 2 * ydim + j], sb[t * 2 * ydim + ydim + j], yb[e]);
    }
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.f;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += sm[w];
        atomicAdd(sum_lp + (int64_t)z * B + b, t);
    }
}

__global__ void gauss_nll_bwd_kernel(const float* __restrict__ suff, const float* __restrict__ y,
                                     const float* __restrict__ dsum_lp, float* __restrict__ dsuff, int Z, int B,
                                     int64_t T, int ydim) {
    const int64_t per = T * ydim, total = (int64_t)Z * B * per;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const int64_t zb = e / per, r = e - zb * per;
        const int64_t b = zb % B;
        const int64_t t = r / ydim;
        const int j = (int)(r - t * ydim);
        const int64_t base = (zb * T + t) * 2 * ydim;
        const float mu = suff[base + j], sraw = suff[base + ydim + j];
        const float sc = 0.01f + 0.99f * softplus_f(sraw);
        const float d = y[b * per + r] - mu;
        const float g = dsum_lp[zb];
        const float inv = 1.f / sc;
        // d lp / d mu = d / s^2 ; d lp / d s = d^2/s^3 - 1/s
        dsuff[base + j] = g * d * inv * inv;
        dsuff[base + ydim + j] = g * (d * d * inv * inv * inv - inv) * 0.99f * sigmoid_f(sraw);
    }
}

__global__ void latent_stats_fwd_kernel(const float* __restrict__ suff, float* __restrict__ loc, float* __restrict__ scale,
                                        int64_t n, int zd) {
    const int64_t total = n * zd;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = e / zd;
        const int j = (int)(e - r * zd);
        loc[e] = suff[r * 2 * zd + j];
        scale[e] = 0.1f + 0.9f * sigmoid_f(suff[r * 2 * zd + zd + j]);
    }
}

__global__ void latent_stats_bwd_kernel(const float* __restrict__ suff, const float* __restrict__ dloc,
                                        const float* __restrict__ dscale, float* __restrict__ dsuff, int64_t n, int zd) {
    const int64_t total = n * zd;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = e / zd;
        const int j = (int)(e - r * zd);
        dsuff[r * 2 * zd + j] = dloc ? dloc[e] : 0.f;
        const float sg = sigmoid_f(suff[r * 2 * zd + zd + j]);
        dsuff[r * 2 * zd + zd + j] = dscale ? dscale[e] * 0.9f * sg * (1.f - sg) : 0.f;
    }
}

__global__ void rsample_fwd_kernel(const float* __restrict__ loc, const float* __restrict__ scale,
                                   const float* __restrict__ eps, float* __restrict__ z, int S, int64_t n) {
    const int64_t total = (int64_t)S * n;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i = e % n;
        z[e] = loc[i] + scale[i] * eps[e];
    }
}

__global__ void rsample_bwd_kernel(const float* __restrict__ dz, const float* __restrict__ eps, float* __restrict__ dloc,
                                   float* __restrict__ dscale, int S, int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        float a = 0.f, b = 0.f;
        for (int s = 0; s < S; ++s) {
            const float g = dz[(int64_t)s * n + i];
            a += g;
            b = fmaf(g, eps[(int64_t)s * n + i], b);
        }
        dloc[i] = a;
        dscale[i] = b;
    }
}

// one warp per row of zd latent dims
__global__ void kl_diag_fwd_kernel(const float* __restrict__ lq, const float* __restrict__ sq, const float* __restrict__ lp,
                                   const float* __restrict__ sp, float* __restrict__ kl, int64_t n, int zd) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t r = warp; r < n; r += nwarps) {
        float s = 0.f;
        for (int j = lane; j < zd; j += 32) {
            const int64_t i = r * zd + j;
            const float ratio = sq[i] / sp[i];
            const float vr = ratio * ratio;
            const float t = (lq[i] - lp[i]) / sp[i];
            s += 0.5f * (vr + t * t - 1.f - logf(vr));
        }
        s = warp_sum(s);
        if (lane == 0) kl[r] = s;
    }
}

__global__ void kl_diag_bwd_kernel(const float* __restrict__ lq, const float* __restrict__ sq, const float* __restrict__ lp,
                                   const float* __restrict__ sp, const float* __restrict__ dkl, float* __restrict__ dlq,
                                   float* __restrict__ dsq, float* __restrict__ dlp, float* __restrict__ dsp, int64_t n,
                                   int zd) {
    const int64_t total = n * zd;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const float g = dkl[i / zd];
        const float a = sq[i], b = sp[i], d = lq[i] - lp[i];
        const float ib2 = 1.f / (b * b);
        // KL = 0.5*(a^2/b^2 + d^2/b^2 - 1) - log a + log b
        dlq[i] = g * d * ib2;
        dlp[i] = -g * d * ib2;
        dsq[i] = g * (a * ib2 - 1.f / a);
        dsp[i] = g * (-(a * a + d * d) * ib2 / b + 1.f / b);
    }
}

__global__ void range_flag_kernel(const float* __restrict__ x, int64_t n, float lo, float hi, int* __restrict__ flag) {
    bool bad = false;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const float v = x[i];
        bad |= !(v >= lo && v <= hi);
    }
    if (__syncthreads_or(bad) && threadIdx.x == 0) atomicOr(flag, 1);
}

__global__ void adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                            int64_t n, float lr, float b1, float b2, float eps, float bc1, float bc2_sqrt,
                            float grad_scale) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const float gi = g[i] * grad_scale;
        const float mi = b1 * m[i] + (1.f - b1) * gi;
        const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
        m[i] = mi;
        v[i] = vi;
        // torch.optim.Adam: step_size = lr / bias_correction1 ; denom = sqrt(v)/sqrt(bias_correction2) + eps
        p[i] -= (lr / bc1) * mi / (sqrtf(vi) / bc2_sqrt + eps);
    }
}

// Adam with the step number kept on the device (CUDA-graph friendly: nothing in the launch depends on the host's
// step counter).  adam_step_inc_kernel advances the counter, adam_dev_kernel reads it and derives the bias corrections.
__global__ void adam_step_inc_kernel(int* step) { *step += 1; }

__global__ void adam_dev_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                                int64_t n, float lr, float b1, float b2, float eps, const int* __restrict__ step,
                                float grad_scale) {
    const float t = (float)__ldg(step);
    const float bc1 = 1.f - powf(b1, t);
    const float bc2_sqrt = sqrtf(1.f - powf(b2, t));
    const float step_size = lr / bc1;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const float gi = g[i] * grad_scale;
        const float mi = b1 * m[i] + (1.f - b1) * gi;
        const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
        m[i] = mi;
        v[i] = vi;
        p[i] -= step_size * mi / (sqrtf(vi) / bc2_sqrt + eps);
    }
}

}  // namespace npfk

using namespace npfk;

extern "C" {

int npfk_loc_scale_fwd(const float* suff, float* loc, float* scale, int64_t n, int ydim, void* stream) {
    NPFK_REQUIRE(suff && loc && scale && ydim > 0, "loc_scale_fwd: bad arguments");
    if (n <= 0) return NPFK_OK;
    NPFK_LAUNCH(loc_scale_fwd_kernel, ew_grid3(n * ydim), 256, 0, as_stream(stream), suff, loc, scale, n, ydim);
    return check_launch("loc_scale_fwd");
}

int npfk_loc_scale_bwd(const float* suff, const float* dloc, const float* dscale, float* dsuff, int64_t n, int ydim,
                       void* stream) {
    NPFK_REQUIRE(suff && dsuff && ydim > 0, "loc_scale_bwd: bad arguments");
    if (n <= 0) return NPFK_OK;
    NPFK_LAUNCH(loc_scale_bwd_kernel, ew_grid3(n * ydim), 256, 0, as_stream(stream), suff, dloc, dscale, dsuff, n, ydim);
    return check_launch("loc_scale_bwd");
}

int npfk_gauss_nll_fwd(const float* suff, const float* y, float* sum_lp, int Z, int B, int64_t T, int ydim,
                       void* stream) {
    NPFK_REQUIRE(suff && y && sum_lp, "gauss_nll_fwd: null pointer");
    NPFK_REQUIRE(Z > 0 && Z <= 65535 && B > 0 && B <= 65535 && T > 0 && ydim > 0, "gauss_nll_fwd: bad shape");
    dim3 grid((unsigned)ceil_div64(T * ydim, kNllChunk), B, Z);
    NPFK_LAUNCH(gauss_nll_fwd_kernel, grid, 256, 0, as_stream(stream), suff, y, sum_lp, B, T, ydim);
    return check_launch("gauss_nll_fwd");
}

int npfk_gauss_nll_bwd(const float* suff, const float* y, const float* dsum_lp, float* dsuff, int Z, int B, int64_t T,
                       int ydim, void* stream) {
    NPFK_REQUIRE(suff && y && dsum_lp && dsuff, "gauss_nll_bwd: null pointer");
    const int64_t total = (int64_t)Z * B * T * ydim;
    if (total <= 0) return NPFK_OK;
    NPFK_LAUNCH(gauss_nll_bwd_kernel, ew_grid3(total), 256, 0, as_stream(stream), suff, y, dsum_lp, dsuff, Z, B, T, ydim);
    return check_launch("gauss_nll_bwd");
}

int npfk_latent_stats_fwd(const float* suff, float* loc, float* scale, int64_t n, int zd, void* stream) {
    NPFK_REQUIRE(suff && loc && scale && zd > 0, "latent_stats_fwd: bad arguments");
    if (n <= 0) return NPFK_OK;
    NPFK_LAUNCH(latent_stats_fwd_kernel, ew_grid3(n * zd), 256, 0, as_stream(stream), suff, loc, scale, n, zd);
    return check_launch("latent_stats_fwd");
}

int npfk_latent_stats_bwd(const float* suff, const float* dloc, const float* dscale, float* dsuff, int64_t n, int zd,
                          void* stream) {
    NPFK_REQUIRE(suff && dsuff && zd > 0, "latent_stats_bwd: bad arguments");
    if (n <= 0) return NPFK_OK;
    NPFK_LAUNCH(latent_stats_bwd_kernel, ew_grid3(n * zd), 256, 0, as_stream(stream), suff, dloc, dscale, dsuff, n, zd);
    return check_launch("latent_stats_bwd");
}

int npfk_rsample_fwd(const float* loc, const float* scale, const float* eps, float* z, int S, int64_t n, void* stream) {
    NPFK_REQUIRE(loc && scale && eps && z && S > 0, "rsample_fwd: bad arguments");
    if (n <= 0) return NPFK_OK;
    NPFK_LAUNCH(rsample_fwd_kernel, ew_grid3((int64_t)S * n), 256, 0, as_stream(stream), loc, scale, eps, z, S, n);
    return check_launch("rsample_fwd");
}

int npfk_rsample_bwd(const float* dz, const float* eps, float* dloc, float* dscale, int S, int64_t n, void* stream) {
    NPFK_REQUIRE(dz && eps && dloc && dscale && S > 0, "rsample_bwd: bad arguments");
    if (n <= 0) return NPFK_OK;
    NPFK_LAUNCH(rsample_bwd_kernel, ew_grid3(n), 256, 0, as_stream(stream), dz, eps, dloc, dscale, S, n);
    return check_launch("rsample_bwd");
}

int npfk_kl_diag_fwd(const float* lq, const float* sq, const float* lp, const float* sp, float* kl, int64_t n, int zd,
                     void* stream) {
    NPFK_REQUIRE(lq && sq && lp && sp && kl && zd > 0, "kl_diag_fwd: bad arguments");
    if (n <= 0) return NPFK_OK;
    NPFK_LAUNCH(kl_diag_fwd_kernel, (unsigned)min(ceil_div64(n, 8), (int64_t)kNumSMs * 8), 256, 0, as_stream(stream), lq,
                sq, lp, sp, kl, n, zd);
    return check_launch("kl_diag_fwd");
}

int npfk_kl_diag_bwd(const float* lq, const float* sq, const float* lp, const float* sp, const float* dkl, float* dlq,
                     float* dsq, float* dlp, float* dsp, int64_t n, int zd, void* stream) {
    NPFK_REQUIRE(lq && sq && lp && sp && dkl && dlq && dsq && dlp && dsp && zd > 0, "kl_diag_bwd: bad arguments");
    if (n <= 0) return NPFK_OK;
    NPFK_LAUNCH(kl_diag_bwd_kernel, ew_grid3(n * zd), 256, 0, as_stream(stream), lq, sq, lp, sp, dkl, dlq, dsq, dlp, dsp,
                n, zd);
    return check_launch("kl_diag_bwd");
}

int npfk_range_flag(const float* x, int64_t n, float lo, float hi, int* flag, void* stream) {
    NPFK_REQUIRE(x && flag, "range_flag: null pointer");
    if (n <= 0) return NPFK_OK;
    NPFK_LAUNCH(range_flag_kernel, ew_grid3(n), 256, 0, as_stream(stream), x, n, lo, hi, flag);
    return check_launch("range_flag");
}

int npfk_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, float lr, float beta1,
                   float beta2, float eps, int step_count, float grad_scale, void* stream) {
    NPFK_REQUIRE(param && grad && exp_avg && exp_avg_sq && step_count >= 1, "adam_step: bad arguments");
    if (n <= 0) return NPFK_OK;
    const float bc1 = 1.f - powf(beta1, (float)step_count);
    const float bc2_sqrt = sqrtf(1.f - powf(beta2, (float)step_count));
    NPFK_LAUNCH(adam_kernel, ew_grid3(n), 256, 0, as_stream(stream), param, grad, exp_avg, exp_avg_sq, n, lr, beta1,
                beta2, eps, bc1, bc2_sqrt, grad_scale);
    return check_launch("adam_step");
}

int npfk_adam_step_dev(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, float lr, float beta1,
                       float beta2, float eps, int* step_count_dev, float grad_scale, void* stream) {
    NPFK_REQUIRE(param && grad && exp_avg && exp_avg_sq && step_count_dev, "adam_step_dev: bad arguments");
    NPFK_LAUNCH(adam_step_inc_kernel, 1, 1, 0, as_stream(stream), step_count_dev);
    if (n <= 0) return check_launch("adam_step_dev");
    NPFK_LAUNCH(adam_dev_kernel, ew_grid3(n), 256, 0, as_stream(stream), param, grad, exp_avg, exp_avg_sq, n, lr, beta1,
                beta2, eps, step_count_dev, grad_scale);
    return check_launch("adam_step_dev");
}

}  // extern "C"
