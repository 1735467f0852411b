// SetConv grid -> target interpolation with the resizer hoisted onto the keys (ConvCNP induced -> target).
//
// Reference: SetConv.forward (npf/architectures/setcnn.py:254-268) computes
//     out_q = W_T (sum_k w_qk v_k) + w_d * density_q + b          with  resizer.weight = [W_T | w_d]
// on every one of the B*Nq target rows (a [B*Nq, C+1] x [C+1, O] GEMM after a [B,Nq,Nk,C] broadcast product).
// The resizer is linear, so it is applied once per KEY instead (v'_k = W_T v_k, a [B*Nk, C] GEMM with Nk = 192
// grid points) and this kernel produces the final rows directly:
//     out_q = sum_k w_qk v'_k + density_q * w_d + b
// HBM traffic per target row: one 4*O-byte write forward, one 4*O-byte read backward (540 B
// per target is the algorithmic figure of the unhoisted form).
//
// The RBF weights decay like exp(-(d/sigma)^p): a key contributes less than 2^-40 of the row's largest weight
// once (d/sigma)^p exceeds the nearest key's by 28, which is below fp32 resolution of the sum.  Keys are visited
// in chunks of 32; a chunk is skipped when its [min,max] position interval is that far from every query of the
// CTA, and inside a visited chunk only keys with a non-negligible weight for some query reach the value phase.
// No assumption is made on key order: unsorted keys only make the skip test succeed less often.
#include <math_constants.h>
#include <stdlib.h>

#include "common.cuh"

namespace npfk {

constexpr int kIThreads = 128;
constexpr int kIQ = 64;     // queries per CTA
constexpr int kIK = 32;     // keys per chunk
constexpr int kIC = 128;    // channels per pass (one per thread)
constexpr float kCutShared = -28.f;     // exp(-28) = 6.9e-13 relative to the row maximum
constexpr float kCutUnshared = -46.f;   // exp(-46) = 1.1e-20 absolute (weights are divided by density + 1e-8)

struct InterpArgs {
    const float* xk; int64_t sxk;
    const float* xq; int64_t sxq;
    const float* v;        // [B,Nk,C] values with the resizer already applied
    const float* theta;    // [1] or [2]
    const float* wd;       // [C] density column of the resizer weight
    const float* bias;     // [C] or null
    int64_t Nk, Nq; int C; float p; int unshared;
};

__device__ __forceinline__ float interp_inp_d(float d, float sigma, float p) {
    // reference op order: dist / sigma, then pow(p), then negate (setcnn.py:131-132)
    const float t = d / sigma;
    return (p == 2.f) ? -(t * t) : -powf(t, p);
}

// distance to the nearest key of the batch element, per query (identical in both key-half threads afterwards)
__device__ __forceinline__ float nearest_key_dist(const InterpArgs& a, int b, float xq, int half, int ql, float* sK,
                                                  float* sRed) {
    float dmin = CUDART_INF_F;
    for (int64_t k0 = 0; k0 < a.Nk; k0 += kIThreads) {
        const int n = (int)min((int64_t)kIThreads, a.Nk - k0);
        __syncthreads();
        if (threadIdx.x < n) sK[threadIdx.x] = __ldg(a.xk + (int64_t)b * a.sxk + k0 + threadIdx.x);
        __syncthreads();
        for (int j = half; j < n; j += 2) dmin = fminf(dmin, fabsf(sK[j] - xq));
    }
    __syncthreads();
    sRed[threadIdx.x] = dmin;
    __syncthreads();
    return fminf(sRed[ql], sRed[ql + kIQ]);
}

// loads key chunk k0 into sXk, publishes its position interval in sBnd and clears the per-key flags;
// returns (CTA-uniform) whether any valid query of the CTA can receive a non-negligible weight from the chunk
__device__ __forceinline__ bool chunk_begin(const InterpArgs& a, int b, int64_t k0, int nk, float xq, bool qv, float m,
                                            float sigma, float sigma_d, float* sXk, float* sBnd, int* sAct) {
    __syncthreads();
    if (threadIdx.x < kIK) {
        const bool ok = threadIdx.x < nk;
        const float x = ok ? __ldg(a.xk + (int64_t)b * a.sxk + k0 + threadIdx.x) : 0.f;
        sXk[threadIdx.x] = x;
        sAct[threadIdx.x] = 0;
        float lo = ok ? x : CUDART_INF_F, hi = ok ? x : -CUDART_INF_F;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
            hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
        }
        if (threadIdx.x == 0) {
            sBnd[0] = lo;
            sBnd[1] = hi;
        }
    }
    __syncthreads();
    const float dq = fmaxf(0.f, fmaxf(sBnd[0] - xq, xq - sBnd[1]));
    bool hit;
    if (a.unshared)
        hit = fmaxf(interp_inp_d(dq, sigma, a.p), interp_inp_d(dq, sigma_d, a.p)) > kCutUnshared;
    else
        hit = interp_inp_d(dq, sigma, a.p) - m > kCutShared;
    return __syncthreads_or(qv && hit) != 0;
}

constexpr int kIPiece = 1024;   // keys staged in shared memory at a time (32 chunks)

// Forward.  Per 1024-key piece the key positions are staged ONCE in shared memory together with the [min,max] interval
// of each 32-key chunk; the chunk loop then costs one barrier per skipped chunk and touches global memory only for the
// value rows of keys that matter (fetched four at a time, so their latencies overlap).
__global__ void __launch_bounds__(kIThreads, 3)
setconv_interp_fwd_kernel(const InterpArgs a, float* __restrict__ out, float* __restrict__ stats) {
    __shared__ __align__(16) float sW[kIK][kIQ];   // un-normalised weights of the current key chunk
    __shared__ float sKeys[kIPiece];
    __shared__ float sLo[kIPiece / kIK], sHi[kIPiece / kIK];
    __shared__ unsigned sMask[2];                  // active keys of the chunk (bit j), double-buffered by chunk parity
    __shared__ float sRed[kIThreads];
    __shared__ float sInvL[kIQ];
    __shared__ float sDens[kIQ];
    const int b = blockIdx.y;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ql = threadIdx.x % kIQ, half = threadIdx.x / kIQ;
    const int64_t q0 = (int64_t)blockIdx.x * kIQ;
    const int nq = (int)min((int64_t)kIQ, a.Nq - q0);
    const bool qv = ql < nq;
    const float xq = qv ? __ldg(a.xq + (int64_t)b * a.sxq + q0 + ql) : 0.f;
    const float sigma = 1e-5f + softplus_f(__ldg(a.theta));
    const float sigma_d = a.unshared ? 1e-5f + softplus_f(__ldg(a.theta + 1)) : sigma;
    const float* xkb = a.xk + (int64_t)b * a.sxk;
    if (threadIdx.x < 2) sMask[threadIdx.x] = 0u;

    // softmax stabiliser: inp at the nearest key (bit-identical to the maximum of inp over the keys)
    float m = 0.f;
    if (!a.unshared) {
        float dmin = CUDART_INF_F;
        for (int64_t p0 = 0; p0 < a.Nk; p0 += kIPiece) {
            const int np = (int)min((int64_t)kIPiece, a.Nk - p0);
            __syncthreads();
            for (int k = threadIdx.x; k < np; k += kIThreads) sKeys[k] = __ldg(xkb + p0 + k);
            __syncthreads();
            for (int jj = half; jj < np; jj += 2) dmin = fminf(dmin, fabsf(sKeys[jj] - xq));
        }
        __syncthreads();
        sRed[threadIdx.x] = dmin;
        __syncthreads();
        m = interp_inp_d(fminf(sRed[ql], sRed[ql + kIQ]), sigma, a.p);
    }

    for (int c0 = 0; c0 < a.C; c0 += kIC) {
        const int c = c0 + threadIdx.x;
        float acc[kIQ];
#pragma unroll
        for (int i = 0; i < kIQ; ++i) acc[i] = 0.f;
        float l = 0.f;   // sum of un-normalised weights (shared) or of the density terms (unshared), per (q, half)
        unsigned gen = 0;
        for (int64_t p0 = 0; p0 < a.Nk; p0 += kIPiece) {
            const int np = (int)min((int64_t)kIPiece, a.Nk - p0);
            const int nch = (np + kIK - 1) / kIK;
            if (a.Nk > kIPiece || a.unshared || c0 > 0 || p0 > 0) {     // otherwise sKeys still holds this piece
                __syncthreads();
                for (int k = threadIdx.x; k < np; k += kIThreads) sKeys[k] = __ldg(xkb + p0 + k);
            }
            __syncthreads();
            for (int ch = warp; ch < nch; ch += kIThreads / 32) {       // [min,max] of every chunk of the piece
                const bool ok = ch * kIK + lane < np;
                const float x = ok ? sKeys[ch * kIK + lane] : 0.f;
                float lo = ok ? x : CUDART_INF_F, hi = ok ? x : -CUDART_INF_F;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
                    hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
                }
                if (lane == 0) {
                    sLo[ch] = lo;
                    sHi[ch] = hi;
                }
            }
            __syncthreads();
            for (int ch = 0; ch < nch; ++ch) {
                const float dq = fmaxf(0.f, fmaxf(sLo[ch] - xq, xq - sHi[ch]));
                bool hit;
                if (a.unshared) hit = fmaxf(interp_inp_d(dq, sigma, a.p), interp_inp_d(dq, sigma_d, a.p)) > kCutUnshared;
                else hit = interp_inp_d(dq, sigma, a.p) - m > kCutShared;
                // the barrier also orders the previous chunk's value phase before sW is rewritten
                if (!__syncthreads_or(qv && hit)) continue;
                const int nk = min(kIK, np - ch * kIK);
                const float* ck = sKeys + ch * kIK;
                unsigned bits = 0u;
                for (int jj = half; jj < kIK; jj += 2) {
                    float e = 0.f;
                    if (jj < nk && qv) {
                        const float d = fabsf(ck[jj] - xq);
                        if (a.unshared) {
                            const float i0 = interp_inp_d(d, sigma, a.p), i1 = interp_inp_d(d, sigma_d, a.p);
                            e = expf(i0);
                            l += expf(i1);
                            if (fmaxf(i0, i1) > kCutUnshared) bits |= 1u << jj;
                        } else {
                            const float i0 = interp_inp_d(d, sigma, a.p) - m;
                            e = expf(i0);
                            l += e;
                            if (i0 > kCutShared) bits |= 1u << jj;
                        }
                    }
                    sW[jj][ql] = e;
                }
                ++gen;
                bits = __reduce_or_sync(0xffffffffu, bits);
                if (lane == 0 && bits) atomicOr(&sMask[gen & 1], bits);
                __syncthreads();
                unsigned mask = sMask[gen & 1];
                if (threadIdx.x == 0) sMask[(gen + 1) & 1] = 0u;       // the other buffer: read before this barrier
                if (c < a.C) {
                    const float* vb = a.v + ((int64_t)b * a.Nk + p0 + ch * kIK) * a.C + c;
                    while (mask) {
                        int jx[4];
                        float vv[4];
#pragma unroll
                        for (int t = 0; t < 4; ++t) {
                            jx[t] = mask ? __ffs(mask) - 1 : -1;
                            if (mask) mask &= mask - 1;
                            vv[t] = jx[t] >= 0 ? __ldg(vb + (int64_t)jx[t] * a.C) : 0.f;
                        }
#pragma unroll
                        for (int t = 0; t < 4; ++t) {
                            if (jx[t] < 0) break;
#pragma unroll
                            for (int i = 0; i < kIQ; i += 4) {
                                const float4 w4 = *reinterpret_cast<const float4*>(&sW[jx[t]][i]);
                                acc[i] = fmaf(w4.x, vv[t], acc[i]);
                                acc[i + 1] = fmaf(w4.y, vv[t], acc[i + 1]);
                                acc[i + 2] = fmaf(w4.z, vv[t], acc[i + 2]);
                                acc[i + 3] = fmaf(w4.w, vv[t], acc[i + 3]);
                            }
                        }
                    }
                }
            }
        }
        // merge the two key halves: normaliser and density per query
        __syncthreads();
        sRed[threadIdx.x] = l;
        __syncthreads();
        if (threadIdx.x < kIQ) {
            const float lt = sRed[ql] + sRed[ql + kIQ];
            float density, inv;
            if (a.unshared) {
                density = lt;
                inv = 1.f / (lt + 1e-8f);
            } else {
                density = lt * expf(m);
                inv = 1.f / lt;
            }
            sInvL[ql] = inv;
            sDens[ql] = density;
            if (c0 == 0 && qv) {
                stats[((int64_t)b * a.Nq + q0 + ql) * 2] = density;
                stats[((int64_t)b * a.Nq + q0 + ql) * 2 + 1] = lt;
            }
        }
        __syncthreads();
        if (c < a.C) {
            const float wdc = __ldg(a.wd + c);
            const float bc = a.bias ? __ldg(a.bias + c) : 0.f;
            float* op = out + ((int64_t)b * a.Nq + q0) * a.C + c;
#pragma unroll
            for (int i = 0; i < kIQ; ++i)
                if (i < nq) __stcs(op + (int64_t)i * a.C, fmaf(acc[i], sInvL[i], fmaf(sDens[i], wdc, bc)));
        }
    }
}

// ------------------------------------------------------------------------------------------ backward
struct InterpBwdSmem {
    float dT[kIC][kIQ + 1];   // dout tile, channel-major (conflict-free for row and column readers)
    float V[kIK][kIC];        // value rows of the active keys of the chunk (compacted)
    float W[kIK][kIQ];        // normalised weights
    float Xk[kIK];
    float K[kIThreads];
    float Red[kIThreads];
    float Red4[4][kIThreads];
    int Act[kIK];
    int ActIdx[kIK];
    int nAct;
    float Bnd[2];
};

// dv:[B,Nk,C] "+=", dwd:[C] "+=", dbias:[C] "+=" (may be null), dtheta:[1 or 2] "+="
__global__ void __launch_bounds__(kIThreads, 3)
setconv_interp_bwd_kernel(const InterpArgs a, const float* __restrict__ stats, const float* __restrict__ dout,
                          float* __restrict__ dv, float* __restrict__ dwd, float* __restrict__ dbias,
                          float* __restrict__ dtheta) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    InterpBwdSmem& s = *reinterpret_cast<InterpBwdSmem*>(smem_raw);
    const int b = blockIdx.y;
    const int ql = threadIdx.x % kIQ, half = threadIdx.x / kIQ;
    const int64_t q0 = (int64_t)blockIdx.x * kIQ;
    const int nq = (int)min((int64_t)kIQ, a.Nq - q0);
    const bool qv = ql < nq;
    const float xq = qv ? __ldg(a.xq + (int64_t)b * a.sxq + q0 + ql) : 0.f;
    const float th0 = __ldg(a.theta);
    const float sigma = 1e-5f + softplus_f(th0);
    const float th1 = a.unshared ? __ldg(a.theta + 1) : th0;
    const float sigma_d = a.unshared ? 1e-5f + softplus_f(th1) : sigma;
    const bool single = a.C <= kIC;

    float m = 0.f;
    if (!a.unshared) m = interp_inp_d(nearest_key_dist(a, b, xq, half, ql, s.K, s.Red), sigma, a.p);
    const float dens = qv ? __ldg(stats + ((int64_t)b * a.Nq + q0 + ql) * 2) : 0.f;
    const float lsum = qv ? __ldg(stats + ((int64_t)b * a.Nq + q0 + ql) * 2 + 1) : 1.f;
    const float inv_l = a.unshared ? 1.f / (lsum + 1e-8f) : 1.f / lsum;
    const float emax = a.unshared ? 1.f : expf(m);

    // dout tile of channel pass c0 -> s.dT (transposed); rows beyond nq / channels beyond C are zero
    auto stage_dT = [&](int c0) {
        const int ncc = min(kIC, a.C - c0);
        for (int e = threadIdx.x; e < kIQ * kIC; e += kIThreads) {
            const int qq = e / kIC, cc = e % kIC;
            s.dT[cc][qq] = (qq < nq && cc < ncc) ? __ldg(dout + ((int64_t)b * a.Nq + q0 + qq) * a.C + c0 + cc) : 0.f;
        }
    };
    float r[kIQ];   // this thread's channel row of the dout tile (kept across the key loop when C <= 128)

    // ---- prologue over channel passes: d bias, d wd, and d density_q = dout_q . wd
    float ddens = 0.f;
    for (int c0 = 0; c0 < a.C; c0 += kIC) {
        const int ncc = min(kIC, a.C - c0);
        __syncthreads();
        stage_dT(c0);
        s.Red[threadIdx.x] = dens;     // Red[0..63] = density per query (both halves write the same value)
        __syncthreads();
        if ((int)threadIdx.x < ncc) {
            float sb = 0.f, sd = 0.f;
#pragma unroll
            for (int i = 0; i < kIQ; ++i) {
                r[i] = s.dT[threadIdx.x][i];
                sb += r[i];
                sd = fmaf(r[i], s.Red[i], sd);
            }
            if (dbias) atomicAdd(dbias + c0 + threadIdx.x, sb);
            atomicAdd(dwd + c0 + threadIdx.x, sd);
        }
        for (int cc = half; cc < ncc; cc += 2) ddens = fmaf(s.dT[cc][ql], __ldg(a.wd + c0 + cc), ddens);
    }
    __syncthreads();
    s.Red[threadIdx.x] = ddens;
    __syncthreads();
    ddens = s.Red[ql] + s.Red[ql + kIQ];

    // per-(query, key-half) partial sums over the active keys
    float sA = 0.f, sB = 0.f, sDot = 0.f, sD = 0.f;
    for (int64_t k0 = 0; k0 < a.Nk; k0 += kIK) {
        const int nk = (int)min((int64_t)kIK, a.Nk - k0);
        if (!chunk_begin(a, b, k0, nk, xq, qv, m, sigma, sigma_d, s.Xk, s.Bnd, s.Act)) continue;
        for (int j = half; j < kIK; j += 2) {
            float w = 0.f;
            if (j < nk && qv) {
                const float d = fabsf(s.Xk[j] - xq);
                if (a.unshared) {
                    const float i0 = interp_inp_d(d, sigma, a.p), i1 = interp_inp_d(d, sigma_d, a.p);
                    w = expf(i0) * inv_l;
                    if (fmaxf(i0, i1) > kCutUnshared) s.Act[j] = 1;
                } else {
                    const float i0 = interp_inp_d(d, sigma, a.p) - m;
                    w = expf(i0) * inv_l;
                    if (i0 > kCutShared) s.Act[j] = 1;
                }
            }
            s.W[j][ql] = w;
        }
        __syncthreads();
        if (threadIdx.x < kIK) {   // compact the active keys of the chunk
            const bool act = s.Act[threadIdx.x] != 0;
            const unsigned mask = __ballot_sync(0xffffffffu, act);
            if (act) s.ActIdx[__popc(mask & ((1u << threadIdx.x) - 1u))] = threadIdx.x;
            if (threadIdx.x == 0) s.nAct = __popc(mask);
        }
        __syncthreads();
        const int nAct = s.nAct;
        float dp[kIK / 2];   // dout_q . v_k for this thread's active keys a = half + 2*i
#pragma unroll
        for (int i = 0; i < kIK / 2; ++i) dp[i] = 0.f;

        for (int c0 = 0; c0 < a.C; c0 += kIC) {
            const int ncc = min(kIC, a.C - c0);
            if (!single) {
                __syncthreads();
                stage_dT(c0);
            }
            for (int e = threadIdx.x; e < nAct * kIC; e += kIThreads) {
                const int aa = e / kIC, cc = e % kIC;
                s.V[aa][cc] = cc < ncc ? __ldg(a.v + ((int64_t)b * a.Nk + k0 + s.ActIdx[aa]) * a.C + c0 + cc) : 0.f;
            }
            __syncthreads();
            if (!single && (int)threadIdx.x < ncc) {
#pragma unroll
                for (int i = 0; i < kIQ; ++i) r[i] = s.dT[threadIdx.x][i];
            }
            // dp[q][k] += sum_c dT[q][c] v[k][c]      (thread = (q, key-half))
            // four channels per step: one 16-byte broadcast load of v serves four FMAs (dT and V are zero beyond ncc)
            for (int cc = 0; cc < ncc; cc += 4) {
                const float t0 = s.dT[cc][ql], t1 = s.dT[cc + 1][ql], t2 = s.dT[cc + 2][ql], t3 = s.dT[cc + 3][ql];
#pragma unroll
                for (int i = 0; i < kIK / 2; ++i)
                    if (half + 2 * i < nAct) {
                        const float4 v4 = *reinterpret_cast<const float4*>(&s.V[half + 2 * i][cc]);
                        dp[i] = fmaf(t0, v4.x, dp[i]);
                        dp[i] = fmaf(t1, v4.y, dp[i]);
                        dp[i] = fmaf(t2, v4.z, dp[i]);
                        dp[i] = fmaf(t3, v4.w, dp[i]);
                    }
            }
            // dv[k][c] += sum_q w[k][q] dT[q][c]        (thread = channel, dout row in registers)
            if ((int)threadIdx.x < ncc) {
                for (int aa = 0; aa < nAct; ++aa) {
                    const int j = s.ActIdx[aa];
                    float accv = 0.f;
#pragma unroll
                    for (int i = 0; i < kIQ; i += 4) {
                        const float4 w4 = *reinterpret_cast<const float4*>(&s.W[j][i]);
                        accv = fmaf(w4.x, r[i], accv);
                        accv = fmaf(w4.y, r[i + 1], accv);
                        accv = fmaf(w4.z, r[i + 2], accv);
                        accv = fmaf(w4.w, r[i + 3], accv);
                    }
                    atomicAdd(dv + ((int64_t)b * a.Nk + k0 + j) * a.C + c0 + threadIdx.x, accv);
                }
            }
            if (!single) __syncthreads();
        }
        // length-scale terms of this thread's (query, active keys)
        if (qv) {
#pragma unroll
            for (int i = 0; i < kIK / 2; ++i) {
                if (half + 2 * i < nAct) {
                    const int j = s.ActIdx[half + 2 * i];
                    const float d = fabsf(s.Xk[j] - xq);
                    const float inp = interp_inp_d(d, sigma, a.p);
                    const float w = s.W[j][ql];
                    sDot = fmaf(w, dp[i], sDot);
                    if (a.unshared) {
                        sA = fmaf(w * dp[i], inp, sA);
                        const float i1 = interp_inp_d(d, sigma_d, a.p);
                        sD = fmaf(expf(i1), i1, sD);
                    } else {
                        // softmax part: sum_k w_k (dp_k - dot)(inp_k - m) = sA - dot * sB, with the row maximum
                        // subtracted so far queries (|m| ~ 1e4) do not cancel catastrophically
                        sA = fmaf(w * dp[i], inp - m, sA);
                        sB = fmaf(w, inp - m, sB);
                        sD = fmaf(expf(inp - m) * emax, inp, sD);
                    }
                }
            }
        }
    }
    // merge the key halves per query, then reduce d sigma over the CTA
    __syncthreads();
    s.Red4[0][threadIdx.x] = sA;
    s.Red4[1][threadIdx.x] = sB;
    s.Red4[2][threadIdx.x] = sDot;
    s.Red4[3][threadIdx.x] = sD;
    __syncthreads();
    float dsig0 = 0.f, dsig1 = 0.f;
    if (threadIdx.x < kIQ && qv) {
        const float A = s.Red4[0][ql] + s.Red4[0][ql + kIQ], Bq = s.Red4[1][ql] + s.Red4[1][ql + kIQ];
        const float dot = s.Red4[2][ql] + s.Red4[2][ql + kIQ], D = s.Red4[3][ql] + s.Red4[3][ql + kIQ];
        if (a.unshared) {
            dsig0 = A * (-a.p / sigma);
            dsig1 = (ddens - dot * inv_l) * D * (-a.p / sigma_d);
        } else {
            dsig0 = (A - dot * Bq + ddens * D) * (-a.p / sigma);
        }
    }
    dsig0 = warp_sum(dsig0);
    dsig1 = warp_sum(dsig1);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) {
        s.Red[threadIdx.x >> 5] = dsig0;
        s.Red[4 + (threadIdx.x >> 5)] = dsig1;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        atomicAdd(dtheta, (s.Red[0] + s.Red[1] + s.Red[2] + s.Red[3]) * sigmoid_f(th0));
        if (a.unshared) atomicAdd(dtheta + 1, (s.Red[4] + s.Red[5] + s.Red[6] + s.Red[7]) * sigmoid_f(th1));
    }
}


// ------------------------------------------------------------------------------------------ forward, resident values
// Grid -> target interpolation when one task's value rows fit in shared memory (C == 128, Nk <= 256: the ConvCNP case,
// 192 induced points).  The write of 512 B per target is the only large HBM stream, so the kernel is organised around
// it: a CTA stages the task's [Nk, 128] values once (L2 hits after the first CTA of a task), then each warp takes 32
// queries.
//   phase 1 (lane = query): nearest key by bisection when the keys are sorted (they are a linspace for ConvCNP; the
//     sortedness is checked on the device), the window of keys whose weight is not negligible (the same 2^-40 / e^-46
//     cut-offs as the tiled kernel above), un-normalised weights of the warp's window into a [slot][query] table;
//   phase 2 (lane = 4 channels): out[q, :] = sum_slot w[slot][q] * V[key(slot), :] with 16 queries accumulated in
//     registers per pass, so a value row is read from shared memory once per 16 queries, and one 512-byte streaming store
//     per target row.
// Unsorted keys take the same code with the window [0, Nk); a window wider than kV2Slots is processed in rounds.
constexpr int kV2Threads = 512;
constexpr int kV2Slots = 32;
constexpr int kV2MaxKeys = 256;

__global__ void __launch_bounds__(kV2Threads, 1)
setconv_interp_fwd_v2_kernel(const InterpArgs a, float* __restrict__ out, float* __restrict__ stats) {
    extern __shared__ __align__(16) float smem_v2[];
    float* sV = smem_v2;                                   // [Nk][128]
    float* sKeys = sV + (size_t)a.Nk * 128;                // [Nk]
    float* sWq = sKeys + kV2MaxKeys;                       // [16 warps][kV2Slots][32]
    float* sQ = sWq + 16 * kV2Slots * 32;                  // [16 warps][2][32]: 1 / normaliser, density
    const int b = blockIdx.y;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int Nk = (int)a.Nk;
    const float* xkb = a.xk + (int64_t)b * a.sxk;
    {
        const float4* src = reinterpret_cast<const float4*>(a.v + (int64_t)b * a.Nk * 128);
        float4* dst = reinterpret_cast<float4*>(sV);
        for (int i = threadIdx.x; i < Nk * 32; i += kV2Threads) dst[i] = __ldg(src + i);
        for (int i = threadIdx.x; i < Nk; i += kV2Threads) sKeys[i] = __ldg(xkb + i);
    }
    __syncthreads();
    bool ok = true;
    for (int i = threadIdx.x; i + 1 < Nk; i += kV2Threads) ok = ok && (sKeys[i] <= sKeys[i + 1]);
    const bool sorted = __syncthreads_and(ok) != 0;

    const float sigma = 1e-5f + softplus_f(__ldg(a.theta));
    const float sigma_d = a.unshared ? 1e-5f + softplus_f(__ldg(a.theta + 1)) : sigma;
    const int64_t q0 = (int64_t)blockIdx.x * (kV2Threads) + warp * 32;      // 32 queries per warp, 512 per CTA
    if (q0 >= a.Nq) return;
    const int nq = (int)min((int64_t)32, a.Nq - q0);
    const bool qv = lane < nq;
    const float xq = qv ? __ldg(a.xq + (int64_t)b * a.sxq + q0 + lane) : sKeys[0];
    float* wq = sWq + warp * kV2Slots * 32;
    float* sq = sQ + warp * 64;

    // ---- phase 1: stabiliser (inp at the nearest key) and the window of keys that matter for this query
    float m = 0.f;
    int kl = 0, kr = Nk - 1;
    if (sorted) {
        int lo = 0, hi = Nk;                    // first key >= xq
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (sKeys[mid] < xq) lo = mid + 1; else hi = mid;
        }
        int near = lo < Nk ? lo : Nk - 1;
        if (lo > 0 && (lo >= Nk || fabsf(sKeys[lo - 1] - xq) <= fabsf(sKeys[lo] - xq))) near = lo - 1;
        const float dmin = fabsf(sKeys[near] - xq);
        if (!a.unshared) m = interp_inp_d(dmin, sigma, a.p);
        kl = kr = near;
        if (a.unshared) {
            while (kl > 0 && fmaxf(interp_inp_d(fabsf(sKeys[kl - 1] - xq), sigma, a.p),
                                   interp_inp_d(fabsf(sKeys[kl - 1] - xq), sigma_d, a.p)) > kCutUnshared) --kl;
            while (kr < Nk - 1 && fmaxf(interp_inp_d(fabsf(sKeys[kr + 1] - xq), sigma, a.p),
                                        interp_inp_d(fabsf(sKeys[kr + 1] - xq), sigma_d, a.p)) > kCutUnshared) ++kr;
        } else {
            while (kl > 0 && interp_inp_d(fabsf(sKeys[kl - 1] - xq), sigma, a.p) - m > kCutShared) --kl;
            while (kr < Nk - 1 && interp_inp_d(fabsf(sKeys[kr + 1] - xq), sigma, a.p) - m > kCutShared) ++kr;
        }
    } else if (!a.unshared) {
        float dmin = CUDART_INF_F;
        for (int k = 0; k < Nk; ++k) dmin = fminf(dmin, fabsf(sKeys[k] - xq));
        m = interp_inp_d(dmin, sigma, a.p);
    }
    // the warp's window: union over its queries
    int wl = qv ? kl : Nk, wr = qv ? kr : -1;      // lanes past the last query do not widen the window
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        wl = min(wl, __shfl_xor_sync(0xffffffffu, wl, o));
        wr = max(wr, __shfl_xor_sync(0xffffffffu, wr, o));
    }
    const bool single = (wr - wl + 1) <= kV2Slots;
    float l = 0.f;
    // weights of window slice [k0, k0 + kV2Slots) into the table; `sum` adds them to the normaliser
    auto fill = [&](int k0, bool sum) {
        __syncwarp();
#pragma unroll 4
        for (int sidx = 0; sidx < kV2Slots; ++sidx) {
            const int k = k0 + sidx;
            float e = 0.f;
            if (k <= wr) {
                const float d = fabsf(sKeys[k] - xq);
                if (a.unshared) {
                    e = expf(interp_inp_d(d, sigma, a.p));
                    if (sum) l += expf(interp_inp_d(d, sigma_d, a.p));
                } else {
                    e = expf(interp_inp_d(d, sigma, a.p) - m);
                    if (sum) l += e;
                }
            }
            wq[sidx * 32 + lane] = e;
        }
        __syncwarp();
    };
    if (single) {
        fill(wl, true);
    } else {
        for (int k = wl; k <= wr; ++k) {          // normaliser over the whole window first
            const float d = fabsf(sKeys[k] - xq);
            l += a.unshared ? expf(interp_inp_d(d, sigma_d, a.p)) : expf(interp_inp_d(d, sigma, a.p) - m);
        }
    }
    {
        float density, inv;
        if (a.unshared) {
            density = l;
            inv = 1.f / (l + 1e-8f);
        } else {
            density = l * expf(m);
            inv = 1.f / l;
        }
        sq[lane] = inv;
        sq[32 + lane] = density;
        if (qv) *reinterpret_cast<float2*>(stats + ((int64_t)b * a.Nq + q0 + lane) * 2) = make_float2(density, l);
    }
    __syncwarp();

    // ---- phase 2: lane = channels 4 lane .. 4 lane + 3
    const float4 wd4 = __ldg(reinterpret_cast<const float4*>(a.wd) + lane);
    const float4 b4 = a.bias ? __ldg(reinterpret_cast<const float4*>(a.bias) + lane) : make_float4(0.f, 0.f, 0.f, 0.f);
    const float4* v4 = reinterpret_cast<const float4*>(sV) + lane;
#pragma unroll 1
    for (int h = 0; h < 2; ++h) {
        if (h * 16 >= nq) break;
        float4 acc[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int k0 = wl; k0 <= wr; k0 += kV2Slots) {
            if (!single) fill(k0, false);
            const int ns = min(kV2Slots, wr - k0 + 1);
#pragma unroll 2
            for (int sidx = 0; sidx < ns; ++sidx) {
                const float4 vv = v4[(size_t)(k0 + sidx) * 32];
                const float4* w4 = reinterpret_cast<const float4*>(wq + sidx * 32 + h * 16);
#pragma unroll
                for (int i4 = 0; i4 < 4; ++i4) {
                    const float4 w = w4[i4];
                    acc[4 * i4].x = fmaf(w.x, vv.x, acc[4 * i4].x); acc[4 * i4].y = fmaf(w.x, vv.y, acc[4 * i4].y);
                    acc[4 * i4].z = fmaf(w.x, vv.z, acc[4 * i4].z); acc[4 * i4].w = fmaf(w.x, vv.w, acc[4 * i4].w);
                    acc[4 * i4 + 1].x = fmaf(w.y, vv.x, acc[4 * i4 + 1].x); acc[4 * i4 + 1].y = fmaf(w.y, vv.y, acc[4 * i4 + 1].y);
                    acc[4 * i4 + 1].z = fmaf(w.y, vv.z, acc[4 * i4 + 1].z); acc[4 * i4 + 1].w = fmaf(w.y, vv.w, acc[4 * i4 + 1].w);
                    acc[4 * i4 + 2].x = fmaf(w.z, vv.x, acc[4 * i4 + 2].x); acc[4 * i4 + 2].y = fmaf(w.z, vv.y, acc[4 * i4 + 2].y);
                    acc[4 * i4 + 2].z = fmaf(w.z, vv.z, acc[4 * i4 + 2].z); acc[4 * i4 + 2].w = fmaf(w.z, vv.w, acc[4 * i4 + 2].w);
                    acc[4 * i4 + 3].x = fmaf(w.w, vv.x, acc[4 * i4 + 3].x); acc[4 * i4 + 3].y = fmaf(w.w, vv.y, acc[4 * i4 + 3].y);
                    acc[4 * i4 + 3].z = fmaf(w.w, vv.z, acc[4 * i4 + 3].z); acc[4 * i4 + 3].w = fmaf(w.w, vv.w, acc[4 * i4 + 3].w);
                }
            }
        }
        float* op = out + ((int64_t)b * a.Nq + q0 + h * 16) * 128 + 4 * lane;
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            if (h * 16 + i < nq) {
                const float inv = sq[h * 16 + i], dens = sq[32 + h * 16 + i];
                float4 o;
                o.x = fmaf(acc[i].x, inv, fmaf(dens, wd4.x, b4.x));
                o.y = fmaf(acc[i].y, inv, fmaf(dens, wd4.y, b4.y));
                o.z = fmaf(acc[i].z, inv, fmaf(dens, wd4.z, b4.z));
                o.w = fmaf(acc[i].w, inv, fmaf(dens, wd4.w, b4.w));
                __stcs(reinterpret_cast<float4*>(op + (int64_t)i * 128), o);
            }
        }
    }
}


// ------------------------------------------------------------------------------------------ backward, resident values
// Same organisation as setconv_interp_fwd_v2_kernel: one task's value rows in shared memory, a warp per 32 queries.
//   (a) lane = query: window and stabiliser as in the forward; dout_q . V_k for the keys of the warp's window, 32
//       channels at a time (dout row quads from global memory, value rows as shared-memory broadcasts); d density_q =
//       dout_q . wd; normalised weights into the [slot][query] table; the four length-scale sums of the tiled kernel;
//   (b) lane = 4 channels: dV[slot] += w[slot][q] * dout_q accumulated in registers over the warp's 32 queries, then one
//       16-byte red.global.add per (key, lane) -- a few hundred KB of reductions per task instead of one atomic per
//       (CTA, key, channel); d bias / d wd column sums in the same sweep over the dout rows.
// Windows wider than kB2Slots keys are walked in rounds (the dout rows are then re-read from L1 / L2).
constexpr int kB2Slots = 16;

__global__ void __launch_bounds__(kV2Threads, 1)
setconv_interp_bwd_v2_kernel(const InterpArgs a, const float* __restrict__ stats, const float* __restrict__ dout,
                             float* __restrict__ dv, float* __restrict__ dwd, float* __restrict__ dbias,
                             float* __restrict__ dtheta) {
    extern __shared__ __align__(16) float smem_v2[];
    float* sV = smem_v2;                                   // [Nk][128]
    float* sKeys = sV + (size_t)a.Nk * 128;                // [kV2MaxKeys]
    float* sWq = sKeys + kV2MaxKeys;                       // [16 warps][kB2Slots][32]
    float* sQ = sWq + 16 * kB2Slots * 32;                  // [16 warps][32]: density per query
    float* sRed = sQ + 16 * 32;                            // [16 warps][2 * 128 + 2]
    const int b = blockIdx.y;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int Nk = (int)a.Nk;
    const float* xkb = a.xk + (int64_t)b * a.sxk;
    {
        const float4* src = reinterpret_cast<const float4*>(a.v + (int64_t)b * a.Nk * 128);
        float4* dst = reinterpret_cast<float4*>(sV);
        for (int i = threadIdx.x; i < Nk * 32; i += kV2Threads) dst[i] = __ldg(src + i);
        for (int i = threadIdx.x; i < Nk; i += kV2Threads) sKeys[i] = __ldg(xkb + i);
    }
    __syncthreads();
    bool ok = true;
    for (int i = threadIdx.x; i + 1 < Nk; i += kV2Threads) ok = ok && (sKeys[i] <= sKeys[i + 1]);
    const bool sorted = __syncthreads_and(ok) != 0;

    const float th0 = __ldg(a.theta);
    const float sigma = 1e-5f + softplus_f(th0);
    const float th1 = a.unshared ? __ldg(a.theta + 1) : th0;
    const float sigma_d = a.unshared ? 1e-5f + softplus_f(th1) : sigma;
    const int64_t q0 = (int64_t)blockIdx.x * kV2Threads + warp * 32;
    const int nq = (int)max((int64_t)0, min((int64_t)32, a.Nq - q0));
    const bool qv = lane < nq;
    float4 db4 = make_float4(0.f, 0.f, 0.f, 0.f), dw4 = db4;      // column sums of this warp (lane = 4 channels)
    float dsig0 = 0.f, dsig1 = 0.f;
    if (nq > 0) {
        const float xq = qv ? __ldg(a.xq + (int64_t)b * a.sxq + q0 + lane) : sKeys[0];
        float* wq = sWq + warp * kB2Slots * 32;
        float* sq = sQ + warp * 32;
        const float* drow = dout + ((int64_t)b * a.Nq + q0 + (qv ? lane : 0)) * 128;      // this lane's query row

        // ---- window and stabiliser (identical to the forward)
        float m = 0.f;
        int kl = 0, kr = Nk - 1;
        if (sorted) {
            int lo = 0, hi = Nk;
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                if (sKeys[mid] < xq) lo = mid + 1; else hi = mid;
            }
            int near = lo < Nk ? lo : Nk - 1;
            if (lo > 0 && (lo >= Nk || fabsf(sKeys[lo - 1] - xq) <= fabsf(sKeys[lo] - xq))) near = lo - 1;
            const float dmin = fabsf(sKeys[near] - xq);
            if (!a.unshared) m = interp_inp_d(dmin, sigma, a.p);
            kl = kr = near;
            if (a.unshared) {
                while (kl > 0 && fmaxf(interp_inp_d(fabsf(sKeys[kl - 1] - xq), sigma, a.p),
                                       interp_inp_d(fabsf(sKeys[kl - 1] - xq), sigma_d, a.p)) > kCutUnshared) --kl;
                while (kr < Nk - 1 && fmaxf(interp_inp_d(fabsf(sKeys[kr + 1] - xq), sigma, a.p),
                                            interp_inp_d(fabsf(sKeys[kr + 1] - xq), sigma_d, a.p)) > kCutUnshared) ++kr;
            } else {
                while (kl > 0 && interp_inp_d(fabsf(sKeys[kl - 1] - xq), sigma, a.p) - m > kCutShared) --kl;
                while (kr < Nk - 1 && interp_inp_d(fabsf(sKeys[kr + 1] - xq), sigma, a.p) - m > kCutShared) ++kr;
            }
        } else if (!a.unshared) {
            float dmin = CUDART_INF_F;
            for (int k = 0; k < Nk; ++k) dmin = fminf(dmin, fabsf(sKeys[k] - xq));
            m = interp_inp_d(dmin, sigma, a.p);
        }
        int wl = qv ? kl : Nk, wr = qv ? kr : -1;      // lanes past the last query do not widen the window
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            wl = min(wl, __shfl_xor_sync(0xffffffffu, wl, o));
            wr = max(wr, __shfl_xor_sync(0xffffffffu, wr, o));
        }
        const float dens = qv ? __ldg(stats + ((int64_t)b * a.Nq + q0 + lane) * 2) : 0.f;
        const float lsum = qv ? __ldg(stats + ((int64_t)b * a.Nq + q0 + lane) * 2 + 1) : 1.f;
        const float inv_l = a.unshared ? 1.f / (lsum + 1e-8f) : 1.f / lsum;
        const float emax = a.unshared ? 1.f : expf(m);
        sq[lane] = dens;

        float ddens = 0.f, sA = 0.f, sB = 0.f, sDot = 0.f, sD = 0.f;
        const float4* v4 = reinterpret_cast<const float4*>(sV);
        for (int k0 = wl; k0 <= wr; k0 += kB2Slots) {
            const int ns = min(kB2Slots, wr - k0 + 1);
            const bool first = k0 == wl;
            // ---- (a) lane = query: dp[s] = dout_q . V[k0 + s]
            float dp[kB2Slots];
#pragma unroll
            for (int i = 0; i < kB2Slots; ++i) dp[i] = 0.f;
#pragma unroll 1
            for (int cb = 0; cb < 4; ++cb) {
                float4 d[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) d[j] = qv ? __ldg(reinterpret_cast<const float4*>(drow) + cb * 8 + j)
                                                       : make_float4(0.f, 0.f, 0.f, 0.f);
                if (first) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const float4 w4 = __ldg(reinterpret_cast<const float4*>(a.wd) + cb * 8 + j);
                        ddens = fmaf(d[j].x, w4.x, ddens); ddens = fmaf(d[j].y, w4.y, ddens);
                        ddens = fmaf(d[j].z, w4.z, ddens); ddens = fmaf(d[j].w, w4.w, ddens);
                    }
                }
#pragma unroll
                for (int sidx = 0; sidx < kB2Slots; ++sidx) {
                    if (sidx < ns) {
                        const float4* vr = v4 + (size_t)(k0 + sidx) * 32 + cb * 8;
                        float acc = dp[sidx];
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const float4 vv = vr[j];
                            acc = fmaf(d[j].x, vv.x, acc); acc = fmaf(d[j].y, vv.y, acc);
                            acc = fmaf(d[j].z, vv.z, acc); acc = fmaf(d[j].w, vv.w, acc);
                        }
                        dp[sidx] = acc;
                    }
                }
            }
            // weights of the slice, the table for (b), and the length-scale sums
            __syncwarp();
#pragma unroll
            for (int sidx = 0; sidx < kB2Slots; ++sidx) {
                float w = 0.f;
                if (sidx < ns && qv) {
                    const float d = fabsf(sKeys[k0 + sidx] - xq);
                    const float inp = interp_inp_d(d, sigma, a.p);
                    if (a.unshared) {
                        const float i1 = interp_inp_d(d, sigma_d, a.p);
                        w = expf(inp) * inv_l;
                        sA = fmaf(w * dp[sidx], inp, sA);
                        sD = fmaf(expf(i1), i1, sD);
                    } else {
                        const float e = expf(inp - m);
                        w = e * inv_l;
                        sA = fmaf(w * dp[sidx], inp - m, sA);
                        sB = fmaf(w, inp - m, sB);
                        sD = fmaf(e * emax, inp, sD);
                    }
                    sDot = fmaf(w, dp[sidx], sDot);
                }
                wq[sidx * 32 + lane] = w;
            }
            __syncwarp();
            // ---- (b) lane = channels 4 lane .. 4 lane + 3
            float4 dvacc[kB2Slots];
#pragma unroll
            for (int i = 0; i < kB2Slots; ++i) dvacc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            const float4* dq4 = reinterpret_cast<const float4*>(dout + ((int64_t)b * a.Nq + q0) * 128) + lane;
#pragma unroll 2
            for (int q = 0; q < nq; ++q) {
                const float4 d4 = __ldg(dq4 + (size_t)q * 32);
                if (first) {
                    const float dn = sq[q];
                    db4.x += d4.x; db4.y += d4.y; db4.z += d4.z; db4.w += d4.w;
                    dw4.x = fmaf(dn, d4.x, dw4.x); dw4.y = fmaf(dn, d4.y, dw4.y);
                    dw4.z = fmaf(dn, d4.z, dw4.z); dw4.w = fmaf(dn, d4.w, dw4.w);
                }
#pragma unroll
                for (int sidx = 0; sidx < kB2Slots; ++sidx) {
                    if (sidx < ns) {
                        const float w = wq[sidx * 32 + q];
                        dvacc[sidx].x = fmaf(w, d4.x, dvacc[sidx].x); dvacc[sidx].y = fmaf(w, d4.y, dvacc[sidx].y);
                        dvacc[sidx].z = fmaf(w, d4.z, dvacc[sidx].z); dvacc[sidx].w = fmaf(w, d4.w, dvacc[sidx].w);
                    }
                }
            }
            float* dvb = dv + ((int64_t)b * a.Nk + k0) * 128 + 4 * lane;
#pragma unroll
            for (int sidx = 0; sidx < kB2Slots; ++sidx)
                if (sidx < ns)
                    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dvb + (size_t)sidx * 128),
                                 "f"(dvacc[sidx].x), "f"(dvacc[sidx].y), "f"(dvacc[sidx].z), "f"(dvacc[sidx].w)
                                 : "memory");
        }
        if (qv) {
            if (a.unshared) {
                dsig0 = sA * (-a.p / sigma);
                dsig1 = (ddens - sDot * inv_l) * sD * (-a.p / sigma_d);
            } else {
                dsig0 = (sA - sDot * sB + ddens * sD) * (-a.p / sigma);
            }
        }
        dsig0 = warp_sum(dsig0);
        dsig1 = warp_sum(dsig1);
    }
    // ---- block reduction of d bias, d wd (lane = 4 channels) and d sigma, one atomic per (CTA, element)
    float* r = sRed + warp * 260;
    *reinterpret_cast<float4*>(r + 4 * lane) = db4;
    *reinterpret_cast<float4*>(r + 128 + 4 * lane) = dw4;
    if (lane == 0) { r[256] = dsig0; r[257] = dsig1; }
    __syncthreads();
    for (int i = threadIdx.x; i < 258; i += kV2Threads) {
        float sacc = 0.f;
#pragma unroll
        for (int w = 0; w < 16; ++w) sacc += sRed[w * 260 + i];
        if (i < 128) { if (dbias) atomicAdd(dbias + i, sacc); }
        else if (i < 256) atomicAdd(dwd + (i - 128), sacc);
        else if (i == 256) atomicAdd(dtheta, sacc * sigmoid_f(th0));
        else if (a.unshared) atomicAdd(dtheta + 1, sacc * sigmoid_f(th1));
    }
}

}  // namespace npfk

using namespace npfk;

extern "C" {

int npfk_setconv_interp_fwd(const float* xk, int64_t sxk, const float* xq, int64_t sxq, const float* v,
                            const float* theta, const float* wd, const float* bias, float* out, float* stats, int B,
                            int64_t Nk, int64_t Nq, int C, float p, int unshared, void* stream) {
    NPFK_REQUIRE(xk && xq && v && theta && wd && out && stats, "setconv_interp_fwd: null pointer");
    NPFK_REQUIRE(B > 0 && B <= 65535 && Nk > 0 && Nq > 0 && C > 0, "setconv_interp_fwd: bad shape B=%d Nk=%lld Nq=%lld C=%d",
                 B, (long long)Nk, (long long)Nq, C);
    NPFK_REQUIRE(p > 0.f, "setconv_interp_fwd: p must be positive");
    InterpArgs a{xk, sxk, xq, sxq, v, theta, wd, bias, Nk, Nq, C, p, unshared};
    static int v2_env = -1;
    if (v2_env < 0) {
        const char* e = getenv("NPFK_SETCONV_V2");
        v2_env = e ? atoi(e) : 1;
    }
    auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
    if ((v2_env == 1 || v2_env == 2) && C == 128 && Nk <= kV2MaxKeys && al16(v) && al16(wd) && al16(out) && al16(stats) && (!bias || al16(bias)) &&
        ceil_div64(Nq, kV2Threads) <= 65535) {
        // one task's value rows resident in shared memory, warp per 32 queries (see the kernel's header)
        const size_t smem = ((size_t)Nk * 128 + kV2MaxKeys + 16 * kV2Slots * 32 + 16 * 64) * sizeof(float);
        static bool attr_done = false;
        if (!attr_done) {
            cudaError_t e = cudaFuncSetAttribute(setconv_interp_fwd_v2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                 (int)((kV2MaxKeys * 128 + kV2MaxKeys + 16 * kV2Slots * 32 + 16 * 64) * sizeof(float)));
            if (e != cudaSuccess) {
                set_error("setconv_interp_fwd: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
                return NPFK_ERR_CUDA;
            }
            attr_done = true;
        }
        dim3 grid2((unsigned)ceil_div64(Nq, kV2Threads), B);
        NPFK_LAUNCH(setconv_interp_fwd_v2_kernel, grid2, kV2Threads, smem, as_stream(stream), a, out, stats);
        return check_launch("setconv_interp_fwd_v2");
    }
    dim3 grid((unsigned)ceil_div64(Nq, kIQ), B);
    NPFK_LAUNCH(setconv_interp_fwd_kernel, grid, kIThreads, 0, as_stream(stream), a, out, stats);
    return check_launch("setconv_interp_fwd");
}

int npfk_setconv_interp_bwd(const float* xk, int64_t sxk, const float* xq, int64_t sxq, const float* v,
                            const float* theta, const float* wd, const float* stats, const float* dout, float* dv,
                            float* dwd, float* dbias, float* dtheta, int B, int64_t Nk, int64_t Nq, int C, float p,
                            int unshared, void* stream) {
    NPFK_REQUIRE(xk && xq && v && theta && wd && stats && dout && dv && dwd && dtheta, "setconv_interp_bwd: null pointer");
    NPFK_REQUIRE(B > 0 && B <= 65535 && Nk > 0 && Nq > 0 && C > 0, "setconv_interp_bwd: bad shape");
    NPFK_REQUIRE(p > 0.f, "setconv_interp_bwd: p must be positive");
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(setconv_interp_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)sizeof(InterpBwdSmem));
        if (e != cudaSuccess) {
            set_error("setconv_interp_bwd: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
            return NPFK_ERR_CUDA;
        }
        attr_set = true;
    }
    InterpArgs a{xk, sxk, xq, sxq, v, theta, wd, nullptr, Nk, Nq, C, p, unshared};
    static int v2_env = -1;
    if (v2_env < 0) {
        const char* e = getenv("NPFK_SETCONV_V2");
        v2_env = e ? atoi(e) : 1;
    }
    auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
    if ((v2_env == 1 || v2_env == 3) && C == 128 && Nk <= kV2MaxKeys && al16(v) && al16(wd) && al16(dout) && al16(dv) && al16(stats) &&
        ceil_div64(Nq, kV2Threads) <= 65535) {
        const size_t fl = (size_t)kV2MaxKeys + 16 * kB2Slots * 32 + 16 * 32 + 16 * 260;
        const size_t smem = ((size_t)Nk * 128 + fl) * sizeof(float);
        static bool attr2 = false;
        if (!attr2) {
            cudaError_t e = cudaFuncSetAttribute(setconv_interp_bwd_v2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                 (int)(((size_t)kV2MaxKeys * 128 + fl) * sizeof(float)));
            if (e != cudaSuccess) {
                set_error("setconv_interp_bwd: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
                return NPFK_ERR_CUDA;
            }
            attr2 = true;
        }
        dim3 grid2((unsigned)ceil_div64(Nq, kV2Threads), B);
        NPFK_LAUNCH(setconv_interp_bwd_v2_kernel, grid2, kV2Threads, smem, as_stream(stream), a, stats, dout, dv, dwd, dbias,
                    dtheta);
        return check_launch("setconv_interp_bwd_v2");
    }
    dim3 grid((unsigned)ceil_div64(Nq, kIQ), B);
    NPFK_LAUNCH(setconv_interp_bwd_kernel, grid, kIThreads, sizeof(InterpBwdSmem), as_stream(stream), a, stats, dout, dv,
                dwd, dbias, dtheta);
    return check_launch("setconv_interp_bwd");
}

}  // extern "C"
