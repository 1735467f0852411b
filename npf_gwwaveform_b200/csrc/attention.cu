// Multi-head cross attention, flash style, fp32 (NPFK_MODE_FP32 parity path).
//
// Replaces the reference's permute-copy / bmm / softmax / bmm chain
// (npf/architectures/attention.py:507-527, 204-220, 129-164).  Heads are addressed in place inside the
// feature dimension; logits [B*H, Nq, Nk] are never written to HBM.
//
// Work split: a group of LPQ lanes owns one query (forward, dq) or one key (dk/dv); each lane keeps
// D/LPQ features of its row in registers, the other side of the product is streamed through shared
// memory in tiles and broadcast to all lanes.
#include <math_constants.h>

#include "common.cuh"

namespace npfk {

constexpr int kAttnThreads = 128;
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

template <int LPQ>
__device__ __forceinline__ float group_sum(float v) {
#pragma unroll
    for (int o = LPQ / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

template <int D>
struct AttnTile {
    static constexpr int TS = (2048 / D) > 256 ? 256 : (2048 / D);  // streamed rows per shared-memory tile
};

// cooperative load of rows [r0, r0+n) x D features of head h into smem (row-major, D contiguous)
template <int D>
__device__ __forceinline__ void load_rows(float* __restrict__ dst, const float* __restrict__ src, int64_t r0, int n,
                                          int64_t ld) {
    for (int e = threadIdx.x; e < n * D; e += kAttnThreads) {
        const int r = e / D, d = e % D;
        dst[e] = __ldg(src + (r0 + r) * ld + d);
    }
}

template <int D, int LPQ>
__global__ void __launch_bounds__(kAttnThreads)
mha_fwd_kernel(const float* __restrict__ q, const float* __restrict__ k, const float* __restrict__ v,
               float* __restrict__ out, float* __restrict__ lse, int64_t Nq, int64_t Nk, int H, float scale) {
    constexpr int DL = D / LPQ, QPB = kAttnThreads / LPQ, TS = AttnTile<D>::TS, KC = 8;
    __shared__ __align__(16) float sK[TS * D];
    __shared__ __align__(16) float sV[TS * D];
    const int h = blockIdx.y, b = blockIdx.z;
    const int64_t ld = (int64_t)H * D;
    const int part = threadIdx.x % LPQ;
    const int64_t qi = (int64_t)blockIdx.x * QPB + threadIdx.x / LPQ;
    const bool valid = qi < Nq;
    const float* kb = k + (int64_t)b * Nk * ld + h * D;
    const float* vb = v + (int64_t)b * Nk * ld + h * D;

    float qr[DL], acc[DL];
    const float sl2 = scale * kLog2e;
#pragma unroll
    for (int d = 0; d < DL; ++d) {
        qr[d] = valid ? __ldg(q + ((int64_t)b * Nq + qi) * ld + h * D + part * DL + d) * sl2 : 0.f;
        acc[d] = 0.f;
    }
    float m = -CUDART_INF_F, l = 0.f;

    for (int64_t k0 = 0; k0 < Nk; k0 += TS) {
        const int nk = (int)min((int64_t)TS, Nk - k0);
        __syncthreads();
        load_rows<D>(sK, kb, k0, nk, ld);
        load_rows<D>(sV, vb, k0, nk, ld);
        __syncthreads();
        for (int j0 = 0; j0 < nk; j0 += KC) {
            float s[KC];
            float cm = -CUDART_INF_F;
#pragma unroll
            for (int c = 0; c < KC; ++c) {
                const int j = j0 + c;
                float t = 0.f;
                if (j < nk) {
#pragma unroll
                    for (int d = 0; d < DL; ++d) t = fmaf(qr[d], sK[j * D + part * DL + d], t);
                }
                t = group_sum<LPQ>(t);
                s[c] = (j < nk) ? t : -CUDART_INF_F;
                cm = fmaxf(cm, s[c]);
            }
            const float mn = fmaxf(m, cm);
            const float corr = exp2f(m - mn);
            l *= corr;
#pragma unroll
            for (int d = 0; d < DL; ++d) acc[d] *= corr;
#pragma unroll
            for (int c = 0; c < KC; ++c) {
                const int j = j0 + c;
                if (j < nk) {
                    const float pj = exp2f(s[c] - mn);
                    l += pj;
#pragma unroll
                    for (int d = 0; d < DL; ++d) acc[d] = fmaf(pj, sV[j * D + part * DL + d], acc[d]);
                }
            }
            m = mn;
        }
    }
    if (valid) {
        const float inv = 1.f / l;
        float* o = out + ((int64_t)b * Nq + qi) * ld + h * D + part * DL;
#pragma unroll
        for (int d = 0; d < DL; ++d) o[d] = acc[d] * inv;
        if (part == 0) lse[((int64_t)b * H + h) * Nq + qi] = (m + log2f(l)) * kLn2;
    }
}

// dq, plus delta[b,h,q] = sum_d dO*O for the dk/dv pass
template <int D, int LPQ>
__global__ void __launch_bounds__(kAttnThreads)
mha_bwd_dq_kernel(const float* __restrict__ q, const float* __restrict__ k, const float* __restrict__ v,
                  const float* __restrict__ out, const float* __restrict__ lse, const float* __restrict__ dout,
                  float* __restrict__ dq, float* __restrict__ delta, int64_t Nq, int64_t Nk, int H, float scale) {
    constexpr int DL = D / LPQ, QPB = kAttnThreads / LPQ, TS = AttnTile<D>::TS;
    __shared__ __align__(16) float sK[TS * D];
    __shared__ __align__(16) float sV[TS * D];
    const int h = blockIdx.y, b = blockIdx.z;
    const int64_t ld = (int64_t)H * D;
    const int part = threadIdx.x % LPQ;
    const int64_t qi = (int64_t)blockIdx.x * QPB + threadIdx.x / LPQ;
    const bool valid = qi < Nq;
    const float* kb = k + (int64_t)b * Nk * ld + h * D;
    const float* vb = v + (int64_t)b * Nk * ld + h * D;
    const int64_t roff = ((int64_t)b * Nq + qi) * ld + h * D + part * DL;

    float qr[DL], dor[DL], dqr[DL];
    float dl = 0.f;
#pragma unroll
    for (int d = 0; d < DL; ++d) {
        qr[d] = valid ? __ldg(q + roff + d) : 0.f;
        dor[d] = valid ? __ldg(dout + roff + d) : 0.f;
        dl = fmaf(dor[d], valid ? __ldg(out + roff + d) : 0.f, dl);
        dqr[d] = 0.f;
    }
    dl = group_sum<LPQ>(dl);
    const float my_lse = valid ? __ldg(lse + ((int64_t)b * H + h) * Nq + qi) : 0.f;

    for (int64_t k0 = 0; k0 < Nk; k0 += TS) {
        const int nk = (int)min((int64_t)TS, Nk - k0);
        __syncthreads();
        load_rows<D>(sK, kb, k0, nk, ld);
        load_rows<D>(sV, vb, k0, nk, ld);
        __syncthreads();
        for (int j = 0; j < nk; ++j) {
            float s = 0.f, dp = 0.f;
#pragma unroll
            for (int d = 0; d < DL; ++d) {
                s = fmaf(qr[d], sK[j * D + part * DL + d], s);
                dp = fmaf(dor[d], sV[j * D + part * DL + d], dp);
            }
            s = group_sum<LPQ>(s);
            dp = group_sum<LPQ>(dp);
            const float pj = expf(s * scale - my_lse);
            const float ds = pj * (dp - dl) * scale;
#pragma unroll
            for (int d = 0; d < DL; ++d) dqr[d] = fmaf(ds, sK[j * D + part * DL + d], dqr[d]);
        }
    }
    if (valid) {
#pragma unroll
        for (int d = 0; d < DL; ++d) dq[roff + d] = dqr[d];
        if (part == 0) delta[((int64_t)b * H + h) * Nq + qi] = dl;
    }
}

// dk, dv: one lane group per key, queries streamed through shared memory
template <int D, int LPQ>
__global__ void __launch_bounds__(kAttnThreads)
mha_bwd_dkv_kernel(const float* __restrict__ q, const float* __restrict__ k, const float* __restrict__ v,
                   const float* __restrict__ lse, const float* __restrict__ delta, const float* __restrict__ dout,
                   float* __restrict__ dk, float* __restrict__ dv, int64_t Nq, int64_t Nk, int H, float scale) {
    constexpr int DL = D / LPQ, KPB = kAttnThreads / LPQ, TS = AttnTile<D>::TS;
    __shared__ __align__(16) float sQ[TS * D];
    __shared__ __align__(16) float sdO[TS * D];
    __shared__ float sLse[TS];
    __shared__ float sDelta[TS];
    const int h = blockIdx.y, b = blockIdx.z;
    const int64_t ld = (int64_t)H * D;
    const int part = threadIdx.x % LPQ;
    const int64_t ki = (int64_t)blockIdx.x * KPB + threadIdx.x / LPQ;
    const bool valid = ki < Nk;
    const float* qb = q + (int64_t)b * Nq * ld + h * D;
    const float* dob = dout + (int64_t)b * Nq * ld + h * D;
    const int64_t roff = ((int64_t)b * Nk + ki) * ld + h * D + part * DL;

    float kr[DL], vr[DL], dkr[DL], dvr[DL];
#pragma unroll
    for (int d = 0; d < DL; ++d) {
        kr[d] = valid ? __ldg(k + roff + d) : 0.f;
        vr[d] = valid ? __ldg(v + roff + d) : 0.f;
        dkr[d] = 0.f;
        dvr[d] = 0.f;
    }
    for (int64_t q0 = 0; q0 < Nq; q0 += TS) {
        const int nq = (int)min((int64_t)TS, Nq - q0);
        __syncthreads();
        load_rows<D>(sQ, qb, q0, nq, ld);
        load_rows<D>(sdO, dob, q0, nq, ld);
        for (int e = threadIdx.x; e < nq; e += kAttnThreads) {
            sLse[e] = __ldg(lse + ((int64_t)b * H + h) * Nq + q0 + e);
            sDelta[e] = __ldg(delta + ((int64_t)b * H + h) * Nq + q0 + e);
        }
        __syncthreads();
        for (int i = 0; i < nq; ++i) {
            float s = 0.f, dp = 0.f;
#pragma unroll
            for (int d = 0; d < DL; ++d) {
                s = fmaf(kr[d], sQ[i * D + part * DL + d], s);
                dp = fmaf(vr[d], sdO[i * D + part * DL + d], dp);
            }
            s = group_sum<LPQ>(s);
            dp = group_sum<LPQ>(dp);
            const float pi = expf(s * scale - sLse[i]);
            const float ds = pi * (dp - sDelta[i]) * scale;
#pragma unroll
            for (int d = 0; d < DL; ++d) {
                dvr[d] = fmaf(pi, sdO[i * D + part * DL + d], dvr[d]);
                dkr[d] = fmaf(ds, sQ[i * D + part * DL + d], dkr[d]);
            }
        }
    }
    if (valid) {
#pragma unroll
        for (int d = 0; d < DL; ++d) {
            dk[roff + d] = dkr[d];
            dv[roff + d] = dvr[d];
        }
    }
}

template <int D, int LPQ>
static int launch_fwd(const float* q, const float* k, const float* v, float* out, float* lse, int B, int64_t Nq,
                      int64_t Nk, int H, float scale, cudaStream_t st) {
    constexpr int QPB = kAttnThreads / LPQ;
    dim3 grid((unsigned)ceil_div64(Nq, QPB), H, B);
    NPFK_LAUNCH((mha_fwd_kernel<D, LPQ>), grid, kAttnThreads, 0, st, q, k, v, out, lse, Nq, Nk, H, scale);
    return check_launch("mha_fwd");
}

template <int D, int LPQ>
static int launch_bwd(const float* q, const float* k, const float* v, const float* out, const float* lse,
                      const float* dout, float* dq, float* dk, float* dv, float* delta, int B, int64_t Nq, int64_t Nk,
                      int H, float scale, cudaStream_t st) {
    constexpr int PB = kAttnThreads / LPQ;
    dim3 gq((unsigned)ceil_div64(Nq, PB), H, B);
    NPFK_LAUNCH((mha_bwd_dq_kernel<D, LPQ>), gq, kAttnThreads, 0, st, q, k, v, out, lse, dout, dq, delta, Nq, Nk, H,
                scale);
    int rc = check_launch("mha_bwd_dq");
    if (rc != NPFK_OK) return rc;
    dim3 gk((unsigned)ceil_div64(Nk, PB), H, B);
    NPFK_LAUNCH((mha_bwd_dkv_kernel<D, LPQ>), gk, kAttnThreads, 0, st, q, k, v, lse, delta, dout, dk, dv, Nq, Nk, H,
                scale);
    return check_launch("mha_bwd_dkv");
}

namespace tc {
bool attn_tc_eligible(const float* q, const float* k, const float* v, const float* out, int B, int64_t Nq, int64_t Nk,
                      int H, int D);
int attn_fwd_tc(const float* q, const float* k, const float* v, float* out, float* lse, int B, int64_t Nq, int64_t Nk,
                int H, int D, float scale, cudaStream_t st);
int64_t attn_bwd_ws_floats(int B, int64_t Nq, int H);
int attn_bwd_tc(const float* q, const float* k, const float* v, const float* out, const float* lse, const float* dout,
                float* dq, float* dk, float* dv, float* ws, int B, int64_t Nq, int64_t Nk, int H, int D, float scale,
                int delta_ready, cudaStream_t st);
}  // namespace tc

}  // namespace npfk

using namespace npfk;

#define NPFK_ATTN_DISPATCH(FN, ...)                                   \
    switch (D) {                                                      \
        case 4: return FN<4, 1>(__VA_ARGS__);                         \
        case 8: return FN<8, 1>(__VA_ARGS__);                         \
        case 16: return FN<16, 1>(__VA_ARGS__);                       \
        case 32: return FN<32, 2>(__VA_ARGS__);                       \
        case 64: return FN<64, 4>(__VA_ARGS__);                       \
        case 128: return FN<128, 8>(__VA_ARGS__);                     \
        default:                                                      \
            set_error("mha: head size %d not in {4,8,16,32,64,128}", D); \
            return NPFK_ERR_UNSUPPORTED;                              \
    }

extern "C" {

int npfk_mha_fwd(const float* q, const float* k, const float* v, float* out, float* lse, int B, int64_t Nq, int64_t Nk,
                 int H, int D, float scale, int mode, void* stream) {
    NPFK_REQUIRE(q && k && v && out && lse, "mha_fwd: null pointer");
    NPFK_REQUIRE(B > 0 && Nq > 0 && Nk > 0 && H > 0, "mha_fwd: bad shape B=%d Nq=%lld Nk=%lld H=%d", B, (long long)Nq,
                 (long long)Nk, H);
    NPFK_REQUIRE(H <= 65535 && B <= 65535, "mha_fwd: H and B must be <= 65535");
    cudaStream_t st = as_stream(stream);
    if (mode == NPFK_MODE_FAST && tc::attn_tc_eligible(q, k, v, out, B, Nq, Nk, H, D))
        return tc::attn_fwd_tc(q, k, v, out, lse, B, Nq, Nk, H, D, scale, st);
    NPFK_ATTN_DISPATCH(launch_fwd, q, k, v, out, lse, B, Nq, Nk, H, scale, st)
}

int64_t npfk_mha_bwd_ws_floats(int B, int64_t Nq, int H) { return tc::attn_bwd_ws_floats(B, Nq, H); }

int npfk_mha_bwd(const float* q, const float* k, const float* v, const float* out, const float* lse, const float* dout,
                 float* dq, float* dk, float* dv, float* delta_ws, int B, int64_t Nq, int64_t Nk, int H, int D,
                 float scale, int delta_ready, int mode, void* stream) {
    NPFK_REQUIRE(q && k && v && out && lse && dout && dq && dk && dv && delta_ws, "mha_bwd: null pointer");
    NPFK_REQUIRE(B > 0 && Nq > 0 && Nk > 0 && H > 0, "mha_bwd: bad shape");
    NPFK_REQUIRE(H <= 65535 && B <= 65535, "mha_bwd: H and B must be <= 65535");
    cudaStream_t st = as_stream(stream);
    if (mode == NPFK_MODE_FAST && tc::attn_tc_eligible(q, k, v, out, B, Nq, Nk, H, D))
        return tc::attn_bwd_tc(q, k, v, out, lse, dout, dq, dk, dv, delta_ws, B, Nq, Nk, H, D, scale, delta_ready, st);
    if (delta_ready) {
        set_error("mha_bwd: delta_ready is only understood by the tensor-core path (fast mode, D in {16,32})");
        return NPFK_ERR_UNSUPPORTED;
    }
    NPFK_ATTN_DISPATCH(launch_bwd, q, k, v, out, lse, dout, dq, dk, dv, delta_ws, B, Nq, Nk, H, scale, st)
}

}  // extern "C"
