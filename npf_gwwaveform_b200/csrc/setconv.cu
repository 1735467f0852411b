// SetConv with the exponential RBF, x_dim = 1 (NPFK_MODE_FP32 parity path).
//
// Replaces the reference's broadcast materialisation [B,Nq,Nk,1] -> softmax/exp -> (weight*values).sum(2)
// (npf/architectures/setcnn.py:254-265 with ExpRBF :126-142 or the fork's UnsharedExpRBF :14-83).
// Weights live in registers / shared memory only; HBM sees keys, queries, values and the [B,Nq,C+1] output.
//
// CTA = 128 threads = one batch element x 64 queries.  Weight phase: thread (q, key-half) evaluates the RBF.
// Value phase: thread <-> channel, the 64 query accumulators live in registers.
#include <math_constants.h>

#include "common.cuh"

namespace npfk {

constexpr int kSCThreads = 128;
constexpr int kQT = 64;    // queries per CTA
constexpr int kKT = 32;    // keys per shared-memory chunk
constexpr int kCT = 128;   // channels per pass (one per thread)

struct SetConvArgs {
    const float* xk; int64_t sxk;
    const float* xq; int64_t sxq;
    const float* v;
    const float* theta;
    int64_t Nk, Nq; int C; float p; int unshared;
};

__device__ __forceinline__ float rbf_inp(float xk, float xq, float sigma, float p) {
    // reference op order: dist / sigma, then pow(p), then negate (setcnn.py:131-132)
    const float t = fabsf(xk - xq) / sigma;
    return (p == 2.f) ? -(t * t) : -powf(t, p);
}

// row maximum of inp over all keys (softmax stabiliser); identical value in both key-half threads after the merge
__device__ __forceinline__ void row_max_pass(const SetConvArgs& a, int b, float xq, float sigma, int half, int ql,
                                             float* sXk, float* sRed, float& m) {
    m = -CUDART_INF_F;
    for (int64_t k0 = 0; k0 < a.Nk; k0 += kKT) {
        const int nk = (int)min((int64_t)kKT, a.Nk - k0);
        __syncthreads();
        if (threadIdx.x < nk) sXk[threadIdx.x] = __ldg(a.xk + (int64_t)b * a.sxk + k0 + threadIdx.x);
        __syncthreads();
        for (int j = half; j < nk; j += 2) m = fmaxf(m, rbf_inp(sXk[j], xq, sigma, a.p));
    }
    __syncthreads();
    sRed[threadIdx.x] = m;
    __syncthreads();
    m = fmaxf(sRed[ql], sRed[ql + kQT]);
}

__global__ void __launch_bounds__(kSCThreads)
setconv_fwd_kernel(const SetConvArgs a, float* __restrict__ out, int64_t ldo) {
    __shared__ __align__(16) float sW[kKT][kQT];   // un-normalised weights of the current key chunk
    __shared__ float sXk[kKT];
    __shared__ float sRed[kSCThreads];
    __shared__ float sInvL[kQT];
    const int b = blockIdx.y;
    const int ql = threadIdx.x % kQT, half = threadIdx.x / kQT;
    const int64_t q0 = (int64_t)blockIdx.x * kQT;
    const int nq = (int)min((int64_t)kQT, a.Nq - q0);
    const float xq = (ql < nq) ? __ldg(a.xq + (int64_t)b * a.sxq + q0 + ql) : 0.f;
    const float sigma = 1e-5f + softplus_f(__ldg(a.theta));
    const float sigma_d = a.unshared ? 1e-5f + softplus_f(__ldg(a.theta + 1)) : sigma;

    float m = 0.f;
    if (!a.unshared) row_max_pass(a, b, xq, sigma, half, ql, sXk, sRed, m);

    for (int c0 = 0; c0 < max(a.C, 1); c0 += kCT) {
        const int c = c0 + threadIdx.x;
        float acc[kQT];
#pragma unroll
        for (int i = 0; i < kQT; ++i) acc[i] = 0.f;
        float l = 0.f;   // sum of un-normalised weights (shared) or density terms (unshared), per (q, half)
        for (int64_t k0 = 0; k0 < a.Nk; k0 += kKT) {
            const int nk = (int)min((int64_t)kKT, a.Nk - k0);
            __syncthreads();
            if (threadIdx.x < nk) sXk[threadIdx.x] = __ldg(a.xk + (int64_t)b * a.sxk + k0 + threadIdx.x);
            __syncthreads();
            for (int j = half; j < kKT; j += 2) {
                float e = 0.f;
                if (j < nk) {
                    if (a.unshared) {
                        e = expf(rbf_inp(sXk[j], xq, sigma, a.p));
                        l += expf(rbf_inp(sXk[j], xq, sigma_d, a.p));
                    } else {
                        e = expf(rbf_inp(sXk[j], xq, sigma, a.p) - m);
                        l += e;
                    }
                }
                sW[j][ql] = e;
            }
            __syncthreads();
            if (c < a.C) {
                for (int j = 0; j < nk; ++j) {
                    const float vv = __ldg(a.v + ((int64_t)b * a.Nk + k0 + j) * a.C + c);
#pragma unroll
                    for (int i = 0; i < kQT; i += 4) {
                        const float4 w4 = *reinterpret_cast<const float4*>(&sW[j][i]);
                        acc[i] = fmaf(w4.x, vv, acc[i]);
                        acc[i + 1] = fmaf(w4.y, vv, acc[i + 1]);
                        acc[i + 2] = fmaf(w4.z, vv, acc[i + 2]);
                        acc[i + 3] = fmaf(w4.w, vv, acc[i + 3]);
                    }
                }
            }
        }
        // merge the two key halves: l (and the density) per query
        __syncthreads();
        sRed[threadIdx.x] = l;
        __syncthreads();
        if (threadIdx.x < kQT) {
            const float lt = sRed[ql] + sRed[ql + kQT];
            float density, inv;
            if (a.unshared) {
                density = lt;
                inv = 1.f / (lt + 1e-8f);
            } else {
                density = lt * expf(m);
                inv = 1.f / lt;
            }
            sInvL[ql] = inv;
            if (c0 == 0 && ql < nq) out[((int64_t)b * a.Nq + q0 + ql) * ldo + a.C] = density;
        }
        __syncthreads();
        if (c < a.C) {
#pragma unroll
            for (int i = 0; i < kQT; ++i)
                if (i < nq) out[((int64_t)b * a.Nq + q0 + i) * ldo + c] = acc[i] * sInvL[i];
        }
    }
}

// dynamic shared memory of the backward kernel
struct SCBwdSmem {
    float dT[kCT][kQT + 1];   // dT chunk, channel-major (conflict-free for both access patterns)
    float V[kKT][kCT];        // value chunk
    float W[kKT][kQT];        // normalised weights
    float Xk[kKT];
    float Red[kSCThreads];
    float Dot[kQT];
    float InvL[kQT];
};

__global__ void __launch_bounds__(kSCThreads)
setconv_bwd_kernel(const SetConvArgs a, const float* __restrict__ out, const float* __restrict__ dout, int64_t ldo,
                   float* __restrict__ dv, float* __restrict__ dtheta) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SCBwdSmem& s = *reinterpret_cast<SCBwdSmem*>(smem_raw);
    const int b = blockIdx.y;
    const int ql = threadIdx.x % kQT, half = threadIdx.x / kQT;
    const int64_t q0 = (int64_t)blockIdx.x * kQT;
    const int nq = (int)min((int64_t)kQT, a.Nq - q0);
    const bool qv = ql < nq;
    const float xq = qv ? __ldg(a.xq + (int64_t)b * a.sxq + q0 + ql) : 0.f;
    const float th0 = __ldg(a.theta);
    const float sigma = 1e-5f + softplus_f(th0);
    const float th1 = a.unshared ? __ldg(a.theta + 1) : th0;
    const float sigma_d = a.unshared ? 1e-5f + softplus_f(th1) : sigma;
    const float* orow = out + ((int64_t)b * a.Nq + q0 + ql) * ldo;
    const float* drow = dout + ((int64_t)b * a.Nq + q0 + ql) * ldo;

    // ---- per-query scalars: softmax max m, normaliser l, dot = sum_c dT*T, dDens
    float m = 0.f;
    if (!a.unshared) row_max_pass(a, b, xq, sigma, half, ql, s.Xk, s.Red, m);
    float l = 0.f;
    for (int64_t k0 = 0; k0 < a.Nk; k0 += kKT) {
        const int nk = (int)min((int64_t)kKT, a.Nk - k0);
        __syncthreads();
        if (threadIdx.x < nk) s.Xk[threadIdx.x] = __ldg(a.xk + (int64_t)b * a.sxk + k0 + threadIdx.x);
        __syncthreads();
        for (int j = half; j < nk; j += 2)
            l += a.unshared ? expf(rbf_inp(s.Xk[j], xq, sigma_d, a.p)) : expf(rbf_inp(s.Xk[j], xq, sigma, a.p) - m);
    }
    float dot = 0.f;
    if (qv)
        for (int c = half; c < a.C; c += 2) dot = fmaf(__ldg(drow + c), __ldg(orow + c), dot);
    __syncthreads();
    s.Red[threadIdx.x] = l;
    __syncthreads();
    l = s.Red[ql] + s.Red[ql + kQT];
    __syncthreads();
    s.Red[threadIdx.x] = dot;
    __syncthreads();
    dot = s.Red[ql] + s.Red[ql + kQT];
    const float ddens = qv ? __ldg(drow + a.C) : 0.f;
    const float inv_l = a.unshared ? 1.f / (l + 1e-8f) : 1.f / l;
    const float emax = a.unshared ? 1.f : expf(m);
    // unshared: total gradient reaching the density sum (direct + through the weight normaliser)
    const float dD = a.unshared ? ddens - dot * inv_l : 0.f;

    float dsig0 = 0.f, dsig1 = 0.f;
    for (int64_t k0 = 0; k0 < a.Nk; k0 += kKT) {
        const int nk = (int)min((int64_t)kKT, a.Nk - k0);
        __syncthreads();
        if (threadIdx.x < nk) s.Xk[threadIdx.x] = __ldg(a.xk + (int64_t)b * a.sxk + k0 + threadIdx.x);
        __syncthreads();
        // normalised weights of this chunk
        for (int j = half; j < kKT; j += 2) {
            float w = 0.f;
            if (j < nk && qv) {
                const float inp = rbf_inp(s.Xk[j], xq, sigma, a.p);
                w = (a.unshared ? expf(inp) : expf(inp - m)) * inv_l;
            }
            s.W[j][ql] = w;
        }
        float dp[kKT / 2];   // dPraw[q][k] for this thread's keys j = half + 2*i
#pragma unroll
        for (int i = 0; i < kKT / 2; ++i) dp[i] = 0.f;

        for (int c0 = 0; c0 < a.C; c0 += kCT) {
            const int ncc = min(kCT, a.C - c0);
            __syncthreads();
            // stage dT[q][c0:c0+ncc] (transposed) and V[k][c0:c0+ncc]
            for (int e = threadIdx.x; e < kQT * kCT; e += kSCThreads) {
                const int qq = e / kCT, cc = e % kCT;
                s.dT[cc][qq] = (qq < nq && cc < ncc) ? __ldg(dout + ((int64_t)b * a.Nq + q0 + qq) * ldo + c0 + cc) : 0.f;
            }
            for (int e = threadIdx.x; e < kKT * kCT; e += kSCThreads) {
                const int kk = e / kCT, cc = e % kCT;
                s.V[kk][cc] = (kk < nk && cc < ncc) ? __ldg(a.v + ((int64_t)b * a.Nk + k0 + kk) * a.C + c0 + cc) : 0.f;
            }
            __syncthreads();
            // dPraw[q][k] += sum_c dT[q][c] V[k][c]   (thread = (q, key-half))
            for (int cc = 0; cc < ncc; ++cc) {
                const float t = s.dT[cc][ql];
#pragma unroll
                for (int i = 0; i < kKT / 2; ++i) dp[i] = fmaf(t, s.V[half + 2 * i][cc], dp[i]);
            }
            // dV[k][c] += sum_q W[k][q] dT[q][c]       (thread = channel)
            const int cc = threadIdx.x;
            if (cc < ncc) {
                for (int kb = 0; kb < nk; kb += 8) {
                    float accv[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) accv[i] = 0.f;
                    for (int qq = 0; qq < nq; ++qq) {
                        const float t = s.dT[cc][qq];
#pragma unroll
                        for (int i = 0; i < 8; ++i) accv[i] = fmaf(s.W[kb + i][qq], t, accv[i]);
                    }
#pragma unroll
                    for (int i = 0; i < 8; ++i)
                        if (kb + i < nk) atomicAdd(dv + ((int64_t)b * a.Nk + k0 + kb + i) * a.C + c0 + cc, accv[i]);
                }
            }
        }
        // d inp and d sigma for this thread's (q, keys)
        if (qv) {
#pragma unroll
            for (int i = 0; i < kKT / 2; ++i) {
                const int j = half + 2 * i;
                if (j < nk) {
                    const float inp = rbf_inp(s.Xk[j], xq, sigma, a.p);
                    if (a.unshared) {
                        const float w = expf(inp) * inv_l;
                        dsig0 += (dp[i] * w) * (-a.p * inp / sigma);
                        const float inp1 = rbf_inp(s.Xk[j], xq, sigma_d, a.p);
                        dsig1 += (dD * expf(inp1)) * (-a.p * inp1 / sigma_d);
                    } else {
                        // d inp = softmax part + density part.  The softmax part sums to zero over keys, so its
                        // contribution sum_k ds_k*inp_k equals sum_k ds_k*(inp_k - m): same value, no cancellation
                        // between terms of magnitude |m| (far queries have |m| ~ 1e4).
                        const float e = expf(inp - m);
                        const float ds = e * inv_l * (dp[i] - dot);
                        const float dd = ddens * e * emax;
                        dsig0 += (ds * (inp - m) + dd * inp) * (-a.p / sigma);
                    }
                }
            }
        }
    }
    // CTA reduction of d sigma -> d theta (sigma = 1e-5 + softplus(theta) => d sigma / d theta = sigmoid(theta))
    dsig0 = warp_sum(dsig0);
    dsig1 = warp_sum(dsig1);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) {
        s.Red[threadIdx.x >> 5] = dsig0;
        s.Red[4 + (threadIdx.x >> 5)] = dsig1;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const float t0 = s.Red[0] + s.Red[1] + s.Red[2] + s.Red[3];
        atomicAdd(dtheta, t0 * sigmoid_f(th0));
        if (a.unshared) {
            const float t1 = s.Red[4] + s.Red[5] + s.Red[6] + s.Red[7];
            atomicAdd(dtheta + 1, t1 * sigmoid_f(th1));
        }
    }
}

// ---- few value channels (context -> grid: C = y_dim, usually 1).  The tiled kernels above spend their time staging
// 128-channel tiles that are almost empty; here one warp owns one (task, query), its lanes stride over the keys
// (coalesced, L1-resident across the CTA's warps) and warp shuffles finish the sums.  No shared-memory tiles.
constexpr int kSmallC = 4;
constexpr int kSmallWarps = 8;

template <int C>
__global__ void __launch_bounds__(kSmallWarps * 32)
setconv_small_fwd_kernel(const SetConvArgs a, float* __restrict__ out, int64_t ldo) {
    const int b = blockIdx.y, lane = threadIdx.x & 31;
    const int64_t q = (int64_t)blockIdx.x * kSmallWarps + (threadIdx.x >> 5);
    if (q >= a.Nq) return;
    const float xq = __ldg(a.xq + (int64_t)b * a.sxq + q);
    const float sigma = 1e-5f + softplus_f(__ldg(a.theta));
    const float sigma_d = a.unshared ? 1e-5f + softplus_f(__ldg(a.theta + 1)) : sigma;
    const float* xk = a.xk + (int64_t)b * a.sxk;
    const float* v = a.v + (int64_t)b * a.Nk * C;
    float m = 0.f;
    if (!a.unshared) {
        m = -CUDART_INF_F;
        for (int64_t k = lane; k < a.Nk; k += 32) m = fmaxf(m, rbf_inp(__ldg(xk + k), xq, sigma, a.p));
        m = warp_max(m);
    }
    float l = 0.f, acc[C];
#pragma unroll
    for (int c = 0; c < C; ++c) acc[c] = 0.f;
    for (int64_t k = lane; k < a.Nk; k += 32) {
        const float x = __ldg(xk + k);
        const float e = expf(rbf_inp(x, xq, sigma, a.p) - m);
        l += a.unshared ? expf(rbf_inp(x, xq, sigma_d, a.p)) : e;
#pragma unroll
        for (int c = 0; c < C; ++c) acc[c] = fmaf(e, __ldg(v + k * C + c), acc[c]);
    }
    l = warp_sum(l);
#pragma unroll
    for (int c = 0; c < C; ++c) acc[c] = warp_sum(acc[c]);
    if (lane == 0) {
        float* o = out + ((int64_t)b * a.Nq + q) * ldo;
        const float inv = a.unshared ? 1.f / (l + 1e-8f) : 1.f / l;
#pragma unroll
        for (int c = 0; c < C; ++c) o[c] = acc[c] * inv;
        o[C] = a.unshared ? l : l * expf(m);
    }
}

// gradient of the length scale(s) only (the values are data: dv == nullptr)
template <int C>
__global__ void __launch_bounds__(kSmallWarps * 32)
setconv_small_bwd_kernel(const SetConvArgs a, const float* __restrict__ out, const float* __restrict__ dout,
                         int64_t ldo, float* __restrict__ dtheta) {
    __shared__ float sRed[2][kSmallWarps];
    const int b = blockIdx.y, lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int64_t q = (int64_t)blockIdx.x * kSmallWarps + wid;
    const float th0 = __ldg(a.theta);
    const float sigma = 1e-5f + softplus_f(th0);
    const float th1 = a.unshared ? __ldg(a.theta + 1) : th0;
    const float sigma_d = a.unshared ? 1e-5f + softplus_f(th1) : sigma;
    float dsig0 = 0.f, dsig1 = 0.f;
    if (q < a.Nq) {
        const float xq = __ldg(a.xq + (int64_t)b * a.sxq + q);
        const float* xk = a.xk + (int64_t)b * a.sxk;
        const float* v = a.v + (int64_t)b * a.Nk * C;
        const float* orow = out + ((int64_t)b * a.Nq + q) * ldo;
        const float* drow = dout + ((int64_t)b * a.Nq + q) * ldo;
        float dT[C], dot = 0.f;
#pragma unroll
        for (int c = 0; c < C; ++c) {
            dT[c] = __ldg(drow + c);
            dot = fmaf(dT[c], __ldg(orow + c), dot);
        }
        const float ddens = __ldg(drow + C);
        float m = 0.f, l = 0.f;
        if (!a.unshared) {
            m = -CUDART_INF_F;
            for (int64_t k = lane; k < a.Nk; k += 32) m = fmaxf(m, rbf_inp(__ldg(xk + k), xq, sigma, a.p));
            m = warp_max(m);
        }
        for (int64_t k = lane; k < a.Nk; k += 32)
            l += expf(rbf_inp(__ldg(xk + k), xq, a.unshared ? sigma_d : sigma, a.p) - m);
        l = warp_sum(l);
        const float inv_l = a.unshared ? 1.f / (l + 1e-8f) : 1.f / l;
        const float emax = a.unshared ? 1.f : expf(m);
        const float dD = a.unshared ? ddens - dot * inv_l : 0.f;
        for (int64_t k = lane; k < a.Nk; k += 32) {
            const float x = __ldg(xk + k);
            const float inp = rbf_inp(x, xq, sigma, a.p);
            float dp = 0.f;
#pragma unroll
            for (int c = 0; c < C; ++c) dp = fmaf(dT[c], __ldg(v + k * C + c), dp);
            if (a.unshared) {
                const float w = expf(inp) * inv_l;
                dsig0 += (dp * w) * (-a.p * inp / sigma);
                const float inp1 = rbf_inp(x, xq, sigma_d, a.p);
                dsig1 += (dD * expf(inp1)) * (-a.p * inp1 / sigma_d);
            } else {   // same split as setconv_bwd_kernel: softmax part against (inp - m), density part against inp
                const float e = expf(inp - m);
                const float ds = e * inv_l * (dp - dot);
                const float dd = ddens * e * emax;
                dsig0 += (ds * (inp - m) + dd * inp) * (-a.p / sigma);
            }
        }
    }
    dsig0 = warp_sum(dsig0);
    dsig1 = warp_sum(dsig1);
    if (lane == 0) {
        sRed[0][wid] = dsig0;
        sRed[1][wid] = dsig1;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        float t0 = 0.f, t1 = 0.f;
#pragma unroll
        for (int i = 0; i < kSmallWarps; ++i) {
            t0 += sRed[0][i];
            t1 += sRed[1][i];
        }
        atomicAdd(dtheta, t0 * sigmoid_f(th0));
        if (a.unshared) atomicAdd(dtheta + 1, t1 * sigmoid_f(th1));
    }
}

}  // namespace npfk

using namespace npfk;

extern "C" {

int npfk_setconv_fwd(const float* xk, int64_t sxk, const float* xq, int64_t sxq, const float* v, const float* theta,
                     float* out, int64_t ldo, int B, int64_t Nk, int64_t Nq, int C, float p, int unshared,
                     void* stream) {
    NPFK_REQUIRE(xk && xq && v && theta && out, "setconv_fwd: null pointer");
    NPFK_REQUIRE(B > 0 && B <= 65535 && Nk > 0 && Nq > 0 && C > 0, "setconv_fwd: bad shape B=%d Nk=%lld Nq=%lld C=%d", B,
                 (long long)Nk, (long long)Nq, C);
    NPFK_REQUIRE(ldo >= C + 1, "setconv_fwd: ldo=%lld < C+1", (long long)ldo);
    NPFK_REQUIRE(p > 0.f, "setconv_fwd: p must be positive");
    SetConvArgs a{xk, sxk, xq, sxq, v, theta, Nk, Nq, C, p, unshared};
    if (C <= kSmallC) {
        dim3 grid((unsigned)ceil_div64(Nq, kSmallWarps), B);
        switch (C) {
            case 1: NPFK_LAUNCH(setconv_small_fwd_kernel<1>, grid, kSmallWarps * 32, 0, as_stream(stream), a, out, ldo); break;
            case 2: NPFK_LAUNCH(setconv_small_fwd_kernel<2>, grid, kSmallWarps * 32, 0, as_stream(stream), a, out, ldo); break;
            case 3: NPFK_LAUNCH(setconv_small_fwd_kernel<3>, grid, kSmallWarps * 32, 0, as_stream(stream), a, out, ldo); break;
            default: NPFK_LAUNCH(setconv_small_fwd_kernel<4>, grid, kSmallWarps * 32, 0, as_stream(stream), a, out, ldo); break;
        }
        return check_launch("setconv_fwd");
    }
    dim3 grid((unsigned)ceil_div64(Nq, kQT), B);
    NPFK_LAUNCH(setconv_fwd_kernel, grid, kSCThreads, 0, as_stream(stream), a, out, ldo);
    return check_launch("setconv_fwd");
}

int npfk_setconv_bwd(const float* xk, int64_t sxk, const float* xq, int64_t sxq, const float* v, const float* theta,
                     const float* out, const float* dout, int64_t ldo, float* dv, float* dtheta, int B, int64_t Nk,
                     int64_t Nq, int C, float p, int unshared, void* stream) {
    NPFK_REQUIRE(xk && xq && v && theta && out && dout && dtheta, "setconv_bwd: null pointer");
    NPFK_REQUIRE(B > 0 && B <= 65535 && Nk > 0 && Nq > 0 && C > 0, "setconv_bwd: bad shape");
    NPFK_REQUIRE(ldo >= C + 1, "setconv_bwd: ldo=%lld < C+1", (long long)ldo);
    SetConvArgs a{xk, sxk, xq, sxq, v, theta, Nk, Nq, C, p, unshared};
    if (!dv && C <= kSmallC) {   // values are data (context -> grid): only the length scale has a gradient
        dim3 grid((unsigned)ceil_div64(Nq, kSmallWarps), B);
        switch (C) {
            case 1: NPFK_LAUNCH(setconv_small_bwd_kernel<1>, grid, kSmallWarps * 32, 0, as_stream(stream), a, out, dout, ldo, dtheta); break;
            case 2: NPFK_LAUNCH(setconv_small_bwd_kernel<2>, grid, kSmallWarps * 32, 0, as_stream(stream), a, out, dout, ldo, dtheta); break;
            case 3: NPFK_LAUNCH(setconv_small_bwd_kernel<3>, grid, kSmallWarps * 32, 0, as_stream(stream), a, out, dout, ldo, dtheta); break;
            default: NPFK_LAUNCH(setconv_small_bwd_kernel<4>, grid, kSmallWarps * 32, 0, as_stream(stream), a, out, dout, ldo, dtheta); break;
        }
        return check_launch("setconv_bwd");
    }
    NPFK_REQUIRE(dv, "setconv_bwd: dv may be null only for C <= %d", kSmallC);
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(setconv_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)sizeof(SCBwdSmem));
        if (e != cudaSuccess) {
            set_error("setconv_bwd: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
            return NPFK_ERR_CUDA;
        }
        attr_set = true;
    }
    dim3 grid((unsigned)ceil_div64(Nq, kQT), B);
    NPFK_LAUNCH(setconv_bwd_kernel, grid, kSCThreads, sizeof(SCBwdSmem), as_stream(stream), a, out, dout, ldo, dv,
                dtheta);
    return check_launch("setconv_bwd");
}

}  // extern "C"
