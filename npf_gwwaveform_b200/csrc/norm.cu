// Row-wise helpers of the attentive / latent paths: residual + LayerNorm, broadcast merge, set mean.
#include "common.cuh"

namespace npfk {

// ---------------------------------------------------------------- add + LayerNorm (one warp per row)
template <int PL>  // values per lane: C <= 32*PL
__global__ void __launch_bounds__(256)
add_layernorm_fwd_kernel(const float* __restrict__ a, const float* __restrict__ b, const float* __restrict__ gamma,
                         const float* __restrict__ beta, float* __restrict__ out, float* __restrict__ mean,
                         float* __restrict__ rstd, int64_t rows, int C, float eps) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t row = warp; row < rows; row += nwarps) {
        float v[PL];
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < PL; ++i) {
            const int c = lane + 32 * i;
            float t = 0.f;
            if (c < C) {
                t = a[row * C + c];
                if (b) t += b[row * C + c];
            }
            v[i] = t;
            s += t;
        }
        const float mu = warp_sum(s) / (float)C;
        float q = 0.f;
#pragma unroll
        for (int i = 0; i < PL; ++i) {
            const int c = lane + 32 * i;
            const float d = (c < C) ? v[i] - mu : 0.f;
            q = fmaf(d, d, q);
        }
        const float rs = rsqrtf(warp_sum(q) / (float)C + eps);
#pragma unroll
        for (int i = 0; i < PL; ++i) {
            const int c = lane + 32 * i;
            if (c < C) out[row * C + c] = (v[i] - mu) * rs * gamma[c] + beta[c];
        }
        if (lane == 0) {
            mean[row] = mu;
            rstd[row] = rs;
        }
    }
}

// C == 128 (r_dim of every benchmark config): a lane owns 4 consecutive columns (one 16-byte load per tensor per row)
// and a warp works on two rows at a time, so four 16-byte loads per lane are in flight; gamma / beta stay in registers.
__global__ void __launch_bounds__(256)
add_layernorm_fwd_c128_kernel(const float* __restrict__ a, const float* __restrict__ b, const float* __restrict__ gamma,
                              const float* __restrict__ beta, float* __restrict__ out, float* __restrict__ mean,
                              float* __restrict__ rstd, int64_t rows, float eps) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const float4 g4 = __ldg(reinterpret_cast<const float4*>(gamma) + lane);
    const float4 be4 = __ldg(reinterpret_cast<const float4*>(beta) + lane);
    for (int64_t row0 = warp * 2; row0 < rows; row0 += nwarps * 2) {
        const bool two = row0 + 1 < rows;
        const int64_t row1 = two ? row0 + 1 : row0;          // clamped: the loads stay unconditional
        const float4* a0 = reinterpret_cast<const float4*>(a + row0 * 128) + lane;
        const float4* a1 = reinterpret_cast<const float4*>(a + row1 * 128) + lane;
        float4 x0 = *a0, x1 = *a1;
        if (b) {
            const float4 y0 = *(reinterpret_cast<const float4*>(b + row0 * 128) + lane);
            const float4 y1 = *(reinterpret_cast<const float4*>(b + row1 * 128) + lane);
            x0.x += y0.x; x0.y += y0.y; x0.z += y0.z; x0.w += y0.w;
            x1.x += y1.x; x1.y += y1.y; x1.z += y1.z; x1.w += y1.w;
        }
        const float mu0 = warp_sum((x0.x + x0.y) + (x0.z + x0.w)) * (1.f / 128.f);
        const float mu1 = warp_sum((x1.x + x1.y) + (x1.z + x1.w)) * (1.f / 128.f);
        x0.x -= mu0; x0.y -= mu0; x0.z -= mu0; x0.w -= mu0;
        x1.x -= mu1; x1.y -= mu1; x1.z -= mu1; x1.w -= mu1;
        const float rs0 = rsqrtf(warp_sum(fmaf(x0.x, x0.x, fmaf(x0.y, x0.y, fmaf(x0.z, x0.z, x0.w * x0.w)))) * (1.f / 128.f) + eps);
        const float rs1 = rsqrtf(warp_sum(fmaf(x1.x, x1.x, fmaf(x1.y, x1.y, fmaf(x1.z, x1.z, x1.w * x1.w)))) * (1.f / 128.f) + eps);
        float4 o0, o1;
        o0.x = x0.x * rs0 * g4.x + be4.x; o0.y = x0.y * rs0 * g4.y + be4.y;
        o0.z = x0.z * rs0 * g4.z + be4.z; o0.w = x0.w * rs0 * g4.w + be4.w;
        o1.x = x1.x * rs1 * g4.x + be4.x; o1.y = x1.y * rs1 * g4.y + be4.y;
        o1.z = x1.z * rs1 * g4.z + be4.z; o1.w = x1.w * rs1 * g4.w + be4.w;
        *(reinterpret_cast<float4*>(out + row0 * 128) + lane) = o0;
        if (two) *(reinterpret_cast<float4*>(out + row1 * 128) + lane) = o1;
        if (lane == 0) {
            mean[row0] = mu0;
            rstd[row0] = rs0;
            if (two) {
                mean[row1] = mu1;
                rstd[row1] = rs1;
            }
        }
    }
}

struct LnDeltaArgs {
    float* delta;          // [B, H, Nqp] or null
    int H;
    int64_t Nq, Nqp;
};

template <int PL, bool HAS2, bool HASD>
__global__ void __launch_bounds__(256)
add_layernorm_bwd_kernel(const float* __restrict__ dout, const float* __restrict__ dout2, const float* __restrict__ a,
                         const float* __restrict__ b,
                         const float* __restrict__ gamma, const float* __restrict__ mean,
                         const float* __restrict__ rstd, float* __restrict__ ds, float* __restrict__ dgamma,
                         float* __restrict__ dbeta, int64_t rows, int C, const LnDeltaArgs da) {
    __shared__ float sg[8][33];
    __shared__ float sb[8][33];
    const int lane = threadIdx.x & 31, wl = threadIdx.x >> 5;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    float accg[PL], accb[PL], gm[PL];
#pragma unroll
    for (int i = 0; i < PL; ++i) {
        accg[i] = accb[i] = 0.f;
        const int c = lane + 32 * i;
        gm[i] = (c < C) ? gamma[c] : 0.f;
    }
    // delta output: the CTA's 8 warps own 8 consecutive rows of one task (Nq % 8 == 0); their 8 x H sums go through
    // shared memory so that each head's 8 values leave as one 32-byte store.  (d_b, d_q) = (task, query) of the
    // CTA's first row, advanced without a 64-bit division per row.
    __shared__ __align__(16) float sD[2][8][8];
    int64_t d_b = 0, d_q = 0;
    uint32_t d_it = 0;
    if (HASD) {
        d_b = (warp - wl) / da.Nq;
        d_q = (warp - wl) - d_b * da.Nq;
    }
    for (int64_t row = warp; row < rows; row += nwarps) {
        const float mu = mean[row], rs = rstd[row];
        float xh[PL], dxh[PL], av[HASD ? PL : 1];
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int i = 0; i < PL; ++i) {
            const int c = lane + 32 * i;
            float t = 0.f, g = 0.f;
            if (HASD) {
                // unconditional loads from a clamped column + selects: keeps all loads of the row in flight together
                // (with the extra live value the compiler otherwise turns `c < C` into branches between the chunks)
                const bool ok = c < C;
                const int64_t e = row * C + (ok ? c : 0);
                const float ta = a[e];
                const float tb = b ? b[e] : 0.f;
                float gg = dout[e];
                if (HAS2) gg += dout2[e];
                av[i] = ok ? ta : 0.f;
                t = ok ? (ta + tb - mu) * rs : 0.f;
                g = ok ? gg : 0.f;
            } else if (c < C) {
                t = a[row * C + c];
                if (b) t += b[row * C + c];
                g = dout[row * C + c];
                if (HAS2) g += dout2[row * C + c];    // second gradient of the output (fan-out summed here, not by autograd)
                t = (t - mu) * rs;
            }
            xh[i] = t;
            dxh[i] = g * gm[i];
            accg[i] = fmaf(g, t, accg[i]);
            accb[i] += g;
            s1 += dxh[i];
            s2 = fmaf(dxh[i], t, s2);
        }
        s1 = warp_sum(s1) / (float)C;
        s2 = warp_sum(s2) / (float)C;
#pragma unroll
        for (int i = 0; i < PL; ++i) {
            const int c = lane + 32 * i;
            const float dsv = rs * (dxh[i] - s1 - xh[i] * s2);
            if (c < C) ds[row * C + c] = dsv;
            if (HASD) {
                // heads of 16 columns: columns lane + 32 i of a half warp belong to head 2 i + lane / 16
                float pr = (c < C) ? dsv * av[i] : 0.f;
                pr += __shfl_xor_sync(0xffffffffu, pr, 8);
                pr += __shfl_xor_sync(0xffffffffu, pr, 4);
                pr += __shfl_xor_sync(0xffffffffu, pr, 2);
                pr += __shfl_xor_sync(0xffffffffu, pr, 1);
                const int h = 2 * i + (lane >> 4);
                if ((lane & 15) == 0 && h < da.H) sD[d_it & 1][h][wl] = pr;
            }
        }
        if (HASD) {
            __syncthreads();          // same trip count in all 8 warps: rows % 8 == 0
            if (wl == 0 && lane < da.H) {
                const float4* src = reinterpret_cast<const float4*>(sD[d_it & 1][lane]);
                float4* dst = reinterpret_cast<float4*>(da.delta + (d_b * da.H + lane) * da.Nqp + d_q);
                dst[0] = src[0];
                dst[1] = src[1];
            }
            ++d_it;
            d_q += nwarps;
            while (d_q >= da.Nq) {
                d_q -= da.Nq;
                ++d_b;
            }
        }
    }
#pragma unroll
    for (int i = 0; i < PL; ++i) {
        sg[wl][lane] = accg[i];
        sb[wl][lane] = accb[i];
        __syncthreads();
        const int c = lane + 32 * i;
        if (wl == 0 && c < C) {
            float tg = 0.f, tb = 0.f;
#pragma unroll
            for (int w = 0; w < 8; ++w) {
                tg += sg[w][lane];
                tb += sb[w][lane];
            }
            atomicAdd(dgamma + c, tg);
            atomicAdd(dbeta + c, tb);
        }
        __syncthreads();
    }
}

// ---------------------------------------------------------------- out[z,b,t,:] = relu(a[b,t,:] + r[z,b,:])
__global__ void bcast_add_relu_fwd_kernel(const float* __restrict__ a, const float* __restrict__ r,
                                          float* __restrict__ out, int Z, int B, int64_t T, int C) {
    const int64_t total = (int64_t)Z * B * T * C;
    const int64_t btc = (int64_t)B * T * C, tc = T * C;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const int64_t z = e / btc, rem = e - z * btc;
        const int64_t bb = rem / tc;
        const int c = (int)(rem % C);
        out[e] = fmaxf(a[rem] + r[(z * B + bb) * C + c], 0.f);
    }
}

__global__ void bcast_add_relu_bwd_da_kernel(const float* __restrict__ dout, const float* __restrict__ out,
                                             float* __restrict__ da, int Z, int64_t btc) {
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < btc; e += (int64_t)gridDim.x * blockDim.x) {
        float s = 0.f;
        for (int z = 0; z < Z; ++z) {
            const int64_t i = (int64_t)z * btc + e;
            s += out[i] > 0.f ? dout[i] : 0.f;
        }
        da[e] = s;
    }
}

constexpr int kBcastTChunk = 256;
// grid (T chunks, B, Z); threads stride over C; dr[z,b,c] += sum_t
__global__ void bcast_add_relu_bwd_dr_kernel(const float* __restrict__ dout, const float* __restrict__ out,
                                             float* __restrict__ dr, int B, int64_t T, int C) {
    const int z = blockIdx.z, b = blockIdx.y;
    const int64_t t0 = (int64_t)blockIdx.x * kBcastTChunk, t1 = min(T, t0 + kBcastTChunk);
    const int64_t base = ((int64_t)z * B + b) * T * C;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        float s = 0.f;
        for (int64_t t = t0; t < t1; ++t) {
            const int64_t i = base + t * C + c;
            s += out[i] > 0.f ? dout[i] : 0.f;
        }
        atomicAdd(dr + ((int64_t)z * B + b) * C + c, s);
    }
}

// Vectorised variants for C % 4 == 0 (r_dim = 128).  Forward: a thread owns one float4 of a[b,t,:] and writes its
// Z outputs (one index decomposition per thread instead of three 64-bit divisions per element).
__global__ void __launch_bounds__(256)
bcast_add_relu_fwd_v4_kernel(const float4* __restrict__ a, const float4* __restrict__ r, float4* __restrict__ out, int Z,
                             int B, int64_t T, int C4) {
    const int64_t btc4 = (int64_t)B * T * C4, tc4 = T * C4;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < btc4; e += (int64_t)gridDim.x * blockDim.x) {
        const int64_t bb = e / tc4;
        const int c4 = (int)(e % C4);
        const float4 av = __ldg(a + e);
        const float4* rp = r + bb * C4 + c4;
#pragma unroll 4
        for (int z = 0; z < Z; ++z) {
            const float4 rv = __ldg(rp + (int64_t)z * B * C4);
            float4 o;
            o.x = fmaxf(av.x + rv.x, 0.f);
            o.y = fmaxf(av.y + rv.y, 0.f);
            o.z = fmaxf(av.z + rv.z, 0.f);
            o.w = fmaxf(av.w + rv.w, 0.f);
            __stcs(out + (int64_t)z * btc4 + e, o);
        }
    }
}

// Backward in ONE pass over dout/out: CTA = (b, 64 consecutive t) x all z; thread = (t row of 8, float4 of channels).
// da[b,t,:] = sum_z g accumulates in registers, dr[z,b,:] += sum_t g through a shared-memory reduction per z.
constexpr int kBcTRows = 8, kBcTIter = 8;      // 64 positions per CTA
__global__ void __launch_bounds__(256)
bcast_add_relu_bwd_v4_kernel(const float4* __restrict__ dout, const float4* __restrict__ out, float4* __restrict__ da,
                             float* __restrict__ dr, int Z, int B, int64_t T, int C4) {
    __shared__ float4 red[kBcTRows][32];
    const int b = blockIdx.y;
    const int lane = threadIdx.x & 31, trow = threadIdx.x >> 5;
    const int64_t t0 = (int64_t)blockIdx.x * (kBcTRows * kBcTIter);
    const int64_t btc4 = (int64_t)B * T * C4;
    for (int cb = 0; cb < C4; cb += 32) {
        const int c4 = cb + lane;
        const bool cok = c4 < C4;
        float4 acc[kBcTIter];
#pragma unroll
        for (int i = 0; i < kBcTIter; ++i) acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int z = 0; z < Z; ++z) {
            float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int i = 0; i < kBcTIter; ++i) {
                const int64_t t = t0 + i * kBcTRows + trow;
                if (cok && t < T) {
                    const int64_t idx = (int64_t)z * btc4 + ((int64_t)b * T + t) * C4 + c4;
                    const float4 o = __ldcs(out + idx), g = __ldcs(dout + idx);
                    float4 m;
                    m.x = o.x > 0.f ? g.x : 0.f;
                    m.y = o.y > 0.f ? g.y : 0.f;
                    m.z = o.z > 0.f ? g.z : 0.f;
                    m.w = o.w > 0.f ? g.w : 0.f;
                    acc[i].x += m.x; acc[i].y += m.y; acc[i].z += m.z; acc[i].w += m.w;
                    s.x += m.x; s.y += m.y; s.z += m.z; s.w += m.w;
                }
            }
            if (dr) {
                red[trow][lane] = s;
                __syncthreads();
                if (trow == 0 && cok) {
                    float4 tsum = red[0][lane];
#pragma unroll
                    for (int w = 1; w < kBcTRows; ++w) {
                        const float4 v = red[w][lane];
                        tsum.x += v.x; tsum.y += v.y; tsum.z += v.z; tsum.w += v.w;
                    }
                    float* dst = dr + (((int64_t)z * B + b) * C4 + c4) * 4;
                    atomicAdd(dst, tsum.x);
                    atomicAdd(dst + 1, tsum.y);
                    atomicAdd(dst + 2, tsum.z);
                    atomicAdd(dst + 3, tsum.w);
                }
                __syncthreads();
            }
        }
        if (da && cok) {
#pragma unroll
            for (int i = 0; i < kBcTIter; ++i) {
                const int64_t t = t0 + i * kBcTRows + trow;
                if (t < T) da[((int64_t)b * T + t) * C4 + c4] = acc[i];
            }
        }
    }
}

// ---------------------------------------------------------------- mean over the set dimension
// grid (B); 256 threads = 32 channel lanes x 8 row lanes
__global__ void __launch_bounds__(256)
set_mean_fwd_kernel(const float* __restrict__ x, float* __restrict__ out, int64_t N, int C) {
    __shared__ float sm[8][33];
    const int b = blockIdx.x, lane = threadIdx.x & 31, wl = threadIdx.x >> 5;
    const float* xb = x + (int64_t)b * N * C;
    for (int c0 = 0; c0 < C; c0 += 32) {
        const int c = c0 + lane;
        float s = 0.f;
        if (c < C)
            for (int64_t n = wl; n < N; n += 8) s += xb[n * C + c];
        sm[wl][lane] = s;
        __syncthreads();
        if (wl == 0 && c < C) {
            float t = 0.f;
#pragma unroll
            for (int w = 0; w < 8; ++w) t += sm[w][lane];
            out[(int64_t)b * C + c] = t / (float)N;
        }
        __syncthreads();
    }
}

__global__ void set_mean_bwd_kernel(const float* __restrict__ dout, float* __restrict__ dx, int B, int64_t N, int C) {
    const int64_t total = (int64_t)B * N * C, nc = N * C;
    const float inv = 1.f / (float)N;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const int64_t b = e / nc;
        const int c = (int)(e % C);
        dx[e] = dout[b * C + c] * inv;
    }
}

inline unsigned ew_grid(int64_t total, int block = 256) {
    int64_t g = ceil_div64(total, block);
    return (unsigned)max((int64_t)1, min(g, (int64_t)kNumSMs * 16));
}

}  // namespace npfk

using namespace npfk;

extern "C" {

int npfk_add_layernorm_fwd(const float* a, const float* b, const float* gamma, const float* beta, float* out,
                           float* mean, float* rstd, int64_t rows, int C, float eps, void* stream) {
    NPFK_REQUIRE(a && gamma && beta && out && mean && rstd, "add_layernorm_fwd: null pointer");
    NPFK_REQUIRE(C > 0 && C <= 1024, "add_layernorm_fwd: C=%d outside (0,1024]", C);
    if (rows <= 0) return NPFK_OK;
    cudaStream_t st = as_stream(stream);
    const unsigned grid = (unsigned)min(ceil_div64(rows, 8), (int64_t)kNumSMs * 8);
    auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
    if (C == 128 && al16(a) && al16(b) && al16(gamma) && al16(beta) && al16(out)) {
        const unsigned g2 = (unsigned)min(ceil_div64(rows, 16), (int64_t)kNumSMs * 8);
        NPFK_LAUNCH(add_layernorm_fwd_c128_kernel, g2, 256, 0, st, a, b, gamma, beta, out, mean, rstd, rows, eps);
    } else if (C <= 128) NPFK_LAUNCH(add_layernorm_fwd_kernel<4>, grid, 256, 0, st, a, b, gamma, beta, out, mean, rstd, rows, C, eps);
    else if (C <= 256) NPFK_LAUNCH(add_layernorm_fwd_kernel<8>, grid, 256, 0, st, a, b, gamma, beta, out, mean, rstd, rows, C, eps);
    else NPFK_LAUNCH(add_layernorm_fwd_kernel<32>, grid, 256, 0, st, a, b, gamma, beta, out, mean, rstd, rows, C, eps);
    return check_launch("add_layernorm_fwd");
}

int npfk_add_layernorm_bwd(const float* dout, const float* dout2, const float* a, const float* b, const float* gamma,
                           const float* mean, const float* rstd, float* ds, float* dgamma, float* dbeta, int64_t rows,
                           int C, float* attn_delta, int H, int64_t Nq, int64_t Nqp, void* stream) {
    NPFK_REQUIRE(dout && a && gamma && mean && rstd && ds && dgamma && dbeta, "add_layernorm_bwd: null pointer");
    NPFK_REQUIRE(C > 0 && C <= 1024, "add_layernorm_bwd: C=%d outside (0,1024]", C);
    NPFK_REQUIRE(!attn_delta || (H > 0 && C == 16 * H && C <= 128 && Nq > 0 && Nq % 8 == 0 && Nqp >= Nq &&
                                 Nqp % 8 == 0 && rows % Nq == 0 && (reinterpret_cast<uintptr_t>(attn_delta) & 15) == 0),
                 "add_layernorm_bwd: attn_delta needs C == 16*H <= 128, Nq %% 8 == 0 and rows == B*Nq (C=%d H=%d "
                 "rows=%lld Nq=%lld)", C, H, (long long)rows, (long long)Nq);
    const LnDeltaArgs da{attn_delta, H, Nq, Nqp};
    if (rows <= 0) return NPFK_OK;
    cudaStream_t st = as_stream(stream);
    const unsigned grid = (unsigned)min(ceil_div64(rows, 8 * 16), (int64_t)kNumSMs * 4);
#define NPFK_LN_BWD(PL_, H2_, HD_)                                                                                   \
    NPFK_LAUNCH((add_layernorm_bwd_kernel<PL_, H2_, HD_>), grid, 256, 0, st, dout, dout2, a, b, gamma, mean, rstd, ds, \
                dgamma, dbeta, rows, C, da)
    if (attn_delta) {          // C == 16 H <= 128 checked above
        if (dout2) NPFK_LN_BWD(4, true, true);
        else NPFK_LN_BWD(4, false, true);
    } else if (dout2) {
        if (C <= 128) NPFK_LN_BWD(4, true, false);
        else if (C <= 256) NPFK_LN_BWD(8, true, false);
        else NPFK_LN_BWD(32, true, false);
    } else {
        if (C <= 128) NPFK_LN_BWD(4, false, false);
        else if (C <= 256) NPFK_LN_BWD(8, false, false);
        else NPFK_LN_BWD(32, false, false);
    }
#undef NPFK_LN_BWD
    return check_launch("add_layernorm_bwd");
}

int npfk_bcast_add_relu_fwd(const float* a, const float* r, float* out, int Z, int B, int64_t T, int C, void* stream) {
    NPFK_REQUIRE(a && r && out, "bcast_add_relu_fwd: null pointer");
    const int64_t total = (int64_t)Z * B * T * C;
    if (total <= 0) return NPFK_OK;
    const bool v4 = (C % 4 == 0) && ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(r) |
                                      reinterpret_cast<uintptr_t>(out)) & 15) == 0;
    if (v4) {
        NPFK_LAUNCH(bcast_add_relu_fwd_v4_kernel, ew_grid(total / Z / 4), 256, 0, as_stream(stream),
                    reinterpret_cast<const float4*>(a), reinterpret_cast<const float4*>(r),
                    reinterpret_cast<float4*>(out), Z, B, T, C / 4);
        return check_launch("bcast_add_relu_fwd_v4");
    }
    NPFK_LAUNCH(bcast_add_relu_fwd_kernel, ew_grid(total), 256, 0, as_stream(stream), a, r, out, Z, B, T, C);
    return check_launch("bcast_add_relu_fwd");
}

int npfk_bcast_add_relu_bwd(const float* dout, const float* out, float* da, float* dr, int Z, int B, int64_t T, int C,
                            void* stream) {
    NPFK_REQUIRE(dout && out, "bcast_add_relu_bwd: null pointer");
    const int64_t btc = (int64_t)B * T * C;
    if (btc * Z <= 0) return NPFK_OK;
    cudaStream_t st = as_stream(stream);
    int rc = NPFK_OK;
    const bool v4 = (C % 4 == 0) && B <= 65535 &&
                    ((reinterpret_cast<uintptr_t>(dout) | reinterpret_cast<uintptr_t>(out) |
                      reinterpret_cast<uintptr_t>(da) | reinterpret_cast<uintptr_t>(dr)) & 15) == 0;
    if (v4 && (da || dr)) {
        dim3 grid((unsigned)ceil_div64(T, kBcTRows * kBcTIter), B);
        NPFK_LAUNCH(bcast_add_relu_bwd_v4_kernel, grid, 256, 0, st, reinterpret_cast<const float4*>(dout),
                    reinterpret_cast<const float4*>(out), reinterpret_cast<float4*>(da), dr, Z, B, T, C / 4);
        return check_launch("bcast_add_relu_bwd_v4");
    }
    if (da) {
        NPFK_LAUNCH(bcast_add_relu_bwd_da_kernel, ew_grid(btc), 256, 0, st, dout, out, da, Z, btc);
        rc = check_launch("bcast_add_relu_bwd_da");
        if (rc != NPFK_OK) return rc;
    }
    if (dr) {
        NPFK_REQUIRE(B <= 65535 && Z <= 65535, "bcast_add_relu_bwd: B, Z must be <= 65535");
        dim3 grid((unsigned)ceil_div64(T, kBcastTChunk), B, Z);
        NPFK_LAUNCH(bcast_add_relu_bwd_dr_kernel, grid, min(C, 256), 0, st, dout, out, dr, B, T, C);
        rc = check_launch("bcast_add_relu_bwd_dr");
    }
    return rc;
}

int npfk_set_mean_fwd(const float* x, float* out, int B, int64_t N, int C, void* stream) {
    NPFK_REQUIRE(x && out, "set_mean_fwd: null pointer");
    NPFK_REQUIRE(B > 0 && N > 0 && C > 0, "set_mean_fwd: bad shape");
    NPFK_LAUNCH(set_mean_fwd_kernel, B, 256, 0, as_stream(stream), x, out, N, C);
    return check_launch("set_mean_fwd");
}

int npfk_set_mean_bwd(const float* dout, float* dx, int B, int64_t N, int C, void* stream) {
    NPFK_REQUIRE(dout && dx, "set_mean_bwd: null pointer");
    const int64_t total = (int64_t)B * N * C;
    if (total <= 0) return NPFK_OK;
    NPFK_LAUNCH(set_mean_bwd_kernel, ew_grid(total), 256, 0, as_stream(stream), dout, dx, B, N, C);
    return check_launch("set_mean_bwd");
}

}  // extern "C"
