// C-ABI of the fused MLP chain (K1) and of the decoder tail's backward.
//
//   npfk_mlp_chain_fwd   -> tc_mlp_chain.cu (tcgen05, activations resident in TMEM)
//   npfk_nll_head_bwd    one pass over the last hidden activation: derivative of the Gaussian log-likelihood and of the
//                        predictive-distribution transform, the 128 -> 2 head's dgrad (with the ReLU mask of the layer
//                        below) and its weight / bias gradients.  Replaces autograd of base.py:347-353,116 + losses.py:18-24.
#include "tc_common.cuh"

namespace npfk {
namespace tc {
int tc_mlp_chain_fwd(const float* x, int64_t ldx, int64_t M, int L, const float* const* w, const float* const* bias,
                     int relu_mask, int res_layer, const float* res, int64_t ldres, int64_t res_rows, float* const* out,
                     const float* head_w, const float* head_b, float* suff, const float* y, int64_t y_rows, int64_t T,
                     float* sum_lp, cudaStream_t st);
bool chain_bwd2_eligible(const float* g, int64_t ldg, const float* h1, int64_t ldh1, const float* w1, int64_t ldw1,
                         const float* h0, int64_t ldh0, const float* w0, int64_t ldw0, const float* dx, int64_t lddx,
                         const float* dw1, int64_t lddw1, const float* dw0, int64_t lddw0, int64_t M);
int tc_chain_bwd2(const float* g, int64_t ldg, const float* h1, int64_t ldh1, const float* w1, int64_t ldw1,
                  const float* h0, int64_t ldh0, const float* w0, int64_t ldw0, int flags, float* dx, int64_t lddx,
                  float* dw1, int64_t lddw1, float* db1, float* dw0, int64_t lddw0, float* db0, int64_t M,
                  cudaStream_t st);
}

// A warp per 32 rows at a time, a lane owns 4 consecutive columns (16-byte loads / stores); the head's weight and bias
// gradients are accumulated in registers over all rows of the warp, reduced across the block's warps in shared memory
// and added to HBM with one atomic per (block, element).
constexpr int kHbWarps = 8;
__global__ void __launch_bounds__(kHbWarps * 32)
nll_head_bwd_kernel(const float* __restrict__ h, int64_t ldh, const float* __restrict__ head_w,
                    const float* __restrict__ suff, const float* __restrict__ y, int64_t y_rows, int64_t T,
                    const float* __restrict__ dsum_lp, const float* __restrict__ dsuff_in, int masked,
                    float* __restrict__ g, float* __restrict__ dw, float* __restrict__ db, int64_t M) {
    __shared__ __align__(16) float red[kHbWarps][2 * 128 + 4];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float4 w0 = __ldg(reinterpret_cast<const float4*>(head_w) + lane);
    const float4 w1 = __ldg(reinterpret_cast<const float4*>(head_w + 128) + lane);
    float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0;
    float b0 = 0.f, b1 = 0.f;
    // A warp takes 32 consecutive rows per iteration.  Phase 1, lane = row: the per-row scalars (d lp / d suff: softplus,
    // sigmoid, reciprocal, the task index) are computed ONCE per row by the lane that owns it, all lanes busy -- as a
    // "one lane computes, the others wait" branch per row they cost the warp ~100 issue slots per row and made the kernel
    // issue-bound at 60 % of HBM.  Phase 2, lane = 4 columns: the scalars of row r are broadcast with two shuffles, the rows'
    // 512-byte loads are issued four at a time.
    const int64_t nwarps = (int64_t)gridDim.x * kHbWarps;
    for (int64_t base = ((int64_t)blockIdx.x * kHbWarps + warp) * 32; base < M; base += nwarps * 32) {
        const int64_t row = base + lane;
        float my0 = 0.f, my1 = 0.f;
        if (row < M) {
            if (dsuff_in) {
                const float2 di = __ldg(reinterpret_cast<const float2*>(dsuff_in) + row);
                my0 = di.x;
                my1 = di.y;
            }
            if (dsum_lp) {
                // lp = -(y - mu)^2 / (2 s^2) - log s - c,  s = 0.01 + 0.99 softplus(raw):  d lp/d mu = d / s^2,
                // d lp/d raw = (d^2 / s^3 - 1 / s) * 0.99 sigmoid(raw).  M, T, y_rows < 2^31 (checked by the entry point)
                const float2 sf = __ldg(reinterpret_cast<const float2*>(suff) + row);
                const float gl = __ldg(dsum_lp + (uint32_t)row / (uint32_t)T);
                const float yv = __ldg(y + (uint32_t)row % (uint32_t)y_rows);
                const float sc = 0.01f + 0.99f * softplus_f(sf.y), inv = 1.f / sc;
                const float d = yv - sf.x;
                my0 = fmaf(gl * d, inv * inv, my0);
                my1 = fmaf(gl * (d * d * inv * inv * inv - inv), 0.99f * sigmoid_f(sf.y), my1);
            }
        }
        const int nr = (int)min((int64_t)32, M - base);
        constexpr int U = 4;
        for (int r0 = 0; r0 < nr; r0 += U) {
            float4 hv[U];
#pragma unroll
            for (int u = 0; u < U; ++u)
                if (r0 + u < nr) hv[u] = __ldg(reinterpret_cast<const float4*>(h + (base + r0 + u) * ldh) + lane);
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (r0 + u >= nr) break;
                const float d0 = __shfl_sync(0xffffffffu, my0, r0 + u), d1 = __shfl_sync(0xffffffffu, my1, r0 + u);
                float4 o;
                o.x = fmaf(d0, w0.x, d1 * w1.x); o.y = fmaf(d0, w0.y, d1 * w1.y);
                o.z = fmaf(d0, w0.z, d1 * w1.z); o.w = fmaf(d0, w0.w, d1 * w1.w);
                if (masked) {
                    if (!(hv[u].x > 0.f)) o.x = 0.f;
                    if (!(hv[u].y > 0.f)) o.y = 0.f;
                    if (!(hv[u].z > 0.f)) o.z = 0.f;
                    if (!(hv[u].w > 0.f)) o.w = 0.f;
                }
                if (g) __stcs(reinterpret_cast<float4*>(g + (base + r0 + u) * 128) + lane, o);
                a0.x = fmaf(d0, hv[u].x, a0.x); a0.y = fmaf(d0, hv[u].y, a0.y); a0.z = fmaf(d0, hv[u].z, a0.z); a0.w = fmaf(d0, hv[u].w, a0.w);
                a1.x = fmaf(d1, hv[u].x, a1.x); a1.y = fmaf(d1, hv[u].y, a1.y); a1.z = fmaf(d1, hv[u].z, a1.z); a1.w = fmaf(d1, hv[u].w, a1.w);
                b0 += d0;
                b1 += d1;
            }
        }
    }
    float* r = red[warp];
    *reinterpret_cast<float4*>(r + 4 * lane) = a0;
    *reinterpret_cast<float4*>(r + 128 + 4 * lane) = a1;
    if (lane == 0) { r[256] = b0; r[257] = b1; }
    __syncthreads();
    for (int i = threadIdx.x; i < 258; i += blockDim.x) {
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < kHbWarps; ++w) s += red[w][i];
        if (i < 256) { if (dw) atomicAdd(dw + i, s); }
        else if (db) atomicAdd(db + (i - 256), s);
    }
}
}  // namespace npfk

extern "C" {

int npfk_mlp_chain_fwd(const float* x, int64_t ldx, int64_t M, int L, const float* const* w, const float* const* bias,
                       int relu_mask, int res_layer, const float* res, int64_t ldres, int64_t res_rows, float* const* out,
                       const float* head_w, const float* head_b, float* suff, const float* y, int64_t y_rows, int64_t T,
                       float* sum_lp, int mode, void* stream) {
    using namespace npfk;
    NPFK_REQUIRE(x && w, "mlp_chain_fwd: null pointer");
    if (mode != NPFK_MODE_FAST) {
        set_error("mlp_chain_fwd: the fused chain is the tensor-core (NPFK_MODE_FAST) path; NPFK_MODE_FP32 runs layer by layer "
                  "through npfk_linear_fwd");
        return NPFK_ERR_UNSUPPORTED;
    }
    if (M <= 0) return NPFK_OK;
    return tc::tc_mlp_chain_fwd(x, ldx, M, L, w, bias, relu_mask, res_layer, res, ldres, res_rows, out, head_w, head_b, suff,
                                y, y_rows, T, sum_lp, as_stream(stream));
}

int npfk_nll_head_bwd(const float* h, int64_t ldh, const float* head_w, const float* suff, const float* y, int64_t y_rows,
                      int64_t T, const float* dsum_lp, const float* dsuff_in, int masked, float* g, float* dw, float* db,
                      int64_t M, void* stream) {
    using namespace npfk;
    NPFK_REQUIRE(h && head_w && suff, "nll_head_bwd: null pointer");
    NPFK_REQUIRE(!dsum_lp || (y && T > 0 && y_rows > 0), "nll_head_bwd: dsum_lp needs y, y_rows and T");
    NPFK_REQUIRE(M < 2147483647LL && T < 2147483647LL && y_rows < 2147483647LL, "nll_head_bwd: sizes must be below 2^31");
    NPFK_REQUIRE((ldh & 3) == 0 && ((reinterpret_cast<uintptr_t>(h) | reinterpret_cast<uintptr_t>(g) |
                                     reinterpret_cast<uintptr_t>(head_w)) & 15) == 0 &&
                     (reinterpret_cast<uintptr_t>(suff) & 7) == 0 && (reinterpret_cast<uintptr_t>(dsuff_in) & 7) == 0,
                 "nll_head_bwd: h / g / head_w must be 16-byte aligned (row pitch a multiple of 4 floats), suff 8-byte aligned");
    if (M <= 0) return NPFK_OK;
    int64_t blocks = ceil_div64(M, kHbWarps * 32);       // a warp takes 32 rows per iteration
    if (blocks > (int64_t)kNumSMs * 6) blocks = (int64_t)kNumSMs * 6;
    if (blocks < 1) blocks = 1;
    NPFK_LAUNCH(nll_head_bwd_kernel, (unsigned)blocks, kHbWarps * 32, 0, as_stream(stream), h, ldh, head_w, suff, y,
                y_rows > 0 ? y_rows : 1, T > 0 ? T : 1, dsum_lp, dsuff_in, masked, g, dw, db, M);
    return check_launch("nll_head_bwd");
}

/* bytes of caller-provided workspace an op needs for the given sizes (0 = none); SURVEY.md section 8b */
int npfk_linear_bwd_fused2(const float* g, int64_t ldg, const float* x1, int64_t ldx1, const float* w1, int64_t ldw1,
                           const float* x0, int64_t ldx0, const float* w0, int64_t ldw0, int flags, float* dx,
                           int64_t lddx, float* dw1, int64_t lddw1, float* db1, float* dw0, int64_t lddw0, float* db0,
                           int64_t M, void* stream) {
    using namespace npfk;
    NPFK_REQUIRE(g && x1 && w1 && x0 && w0 && dx && dw1 && dw0, "linear_bwd_fused2: null pointer");
    NPFK_REQUIRE((flags & ~3) == 0, "linear_bwd_fused2: unknown flag bits %d", flags);
    if (M <= 0) return NPFK_OK;
    if (!tc::chain_bwd2_eligible(g, ldg, x1, ldx1, w1, ldw1, x0, ldx0, w0, ldw0, dx, lddx, dw1, lddw1, dw0, lddw0, M)) {
        set_error("linear_bwd_fused2: needs M >= 128, 16-byte aligned operands and row pitches that are multiples of 4 "
                  "floats (got M=%lld); run the two layers through npfk_linear_bwd_fused", (long long)M);
        return NPFK_ERR_UNSUPPORTED;
    }
    return tc::tc_chain_bwd2(g, ldg, x1, ldx1, w1, ldw1, x0, ldx0, w0, ldw0, flags, dx, lddx, dw1, lddw1, db1, dw0, lddw0,
                             db0, M, as_stream(stream));
}

int64_t npfk_workspace_bytes(int op, int64_t d0, int64_t d1, int64_t d2) {
    switch (op) {
        case NPFK_OP_MHA_BWD: return 4 * npfk_mha_bwd_ws_floats((int)d0, d1, (int)d2);   // (B, Nq, H)
        case NPFK_OP_BN_STATS: return 16 * d0 + 8;                                       // (C): double sums [2C] + ticket
        default: return 0;
    }
}

}  // extern "C"
