// K1 -- a whole ReLU-MLP chain of 128 -> 128 layers in ONE kernel, activations resident in tensor memory.
//
//   h_0 = X ;  h_{l+1} = act_l(h_l . W_l^T + b_l [+ res at layer res_layer])        l = 0 .. L-1   (L <= 8)
//   optional tail on the last activation (the decoder of every 1-D model ends this way):
//     suff = h_L . W_h^T + b_h        (128 -> 2 head, FFMA in the epilogue registers)
//     lp   = log N(y ; suff_0, 0.01 + 0.99 softplus(suff_1)),  sum_lp[row / T] += lp   (Gaussian log-likelihood + sum over targets)
//
// Replaces the addmm + ReLU chain of MLP.forward (reference npf/architectures/mlp.py:95-109), the sum-merge
// relu(x1 + MLP(x2)) of MergeFlatInputs (npf/architectures/encoders.py:175-183), the decode tail
// (npf/neuralproc/base.py:347-353,116) and sum_log_prob (npf/losses.py:18-24).
//
// Layout.  A CTA works on two 128-row tiles at a time ("streams" 0 and 1).  Per stream, tensor memory holds the current
// activation tile as the A operand of tcgen05.mma.kind::tf32 (128 lanes x 128 columns, one 32-bit column per K element)
// and the fp32 accumulator (128 columns): 2 x 256 = all 512 columns.  Layer weights (tf32-rounded fp32, K-major, 64 KB)
// are streamed by TMA through two shared-memory slots shared by both streams, so layer l+1 loads while layer l runs.
// Between layers an activation never leaves the SM: the epilogue warps read the accumulator (tcgen05.ld), add the bias
// (+ residual), apply the ReLU and write the tile back into the A columns (tcgen05.st).  What goes to HBM is only what
// the caller asks for: the activations the backward pass needs (`store_mask`; TMA stores from a swizzled staging box),
// or nothing but [M, 2] sufficient statistics and the per-task log-likelihood sums in inference.
//
//   warp 0      TMA producer of the weight slots
//   warp 1      TMEM allocator + MMA issuer (one elected thread), alternating between the two streams
//   warps 2..5  stream 0: X copy-in (TMA box -> registers -> tcgen05.st), per-layer epilogues, head + NLL tail
//   warps 6..9  stream 1: the same
// While one stream's epilogue runs, the other stream's MMAs keep the tensor pipe busy.
#include <math.h>

#include "tc_common.cuh"

namespace npfk {
namespace tc {

constexpr int kCThreads = 320;
constexpr int kCMaxLayers = 8;
constexpr int kCWSlot = 65536;                 // one 128 x 128 fp32 layer = four K-major [128 rows x 32 floats] boxes
constexpr int kCBox = 4096;                    // [32 rows x 32 floats]
constexpr int kCOffBox = 2 * kCWSlot;          // 8 epilogue warps x 2 boxes
constexpr int kCOffBias = kCOffBox + 16 * kCBox;
constexpr int kCOffHead = kCOffBias + kCMaxLayers * 512;
constexpr int kCOffBars = kCOffHead + 1024 + 16;
constexpr int kCNumBars = 2 + 2 + 2 + 2 + 16;  // w_full, w_empty, acc_full, a_ready, box bars
constexpr int kCSmem = 1024 + kCOffBars + kCNumBars * 8 + 16;
constexpr float kLogSqrt2PiC = 0.9189385332046727f;

struct ChainMaps {
    CUtensorMap x;                   // [M, 128], box 32 x 32
    CUtensorMap w[kCMaxLayers];      // [128, 128], box 32 x 128 (K-major rows of W)
    CUtensorMap out[kCMaxLayers];    // [M, 128], box 32 x 32 (only the layers in store_mask are encoded)
};

struct ChainParams {
    int64_t M;
    int L;
    int relu_mask;                   // bit l: ReLU after layer l
    int store_mask;                  // bit l: the output of layer l is written to HBM (maps.out[l])
    int res_layer;                   // -1, or the layer whose pre-activation gets + res[row % res_rows]
    const float* res;
    int64_t ldres, res_rows;
    const float* bias[kCMaxLayers];  // [128] or null
    const float* head_w;             // [2, 128] fp32 or null
    const float* head_b;             // [2] or null
    float* suff;                     // [M, 2] or null
    const float* y;                  // [y_rows] or null: targets of the Gaussian log-likelihood
    int64_t y_rows;                  // y index = row % y_rows (z-samples share the targets)
    int64_t T;                       // rows per task
    float* sum_lp;                   // [M / T], "+="
    int dbg;                         // timing experiments only (NPFK_CHAIN_DBG): 1 = epilogues do nothing but the barrier hand-shake
    int* g_abort;
};

__device__ __forceinline__ float softplus_c(float x) { return x > 20.f ? x : log1pf(expf(x)); }
// max(x, 0) rounded to the nearest tf32, ties away from zero: what cvt.rna.tf32.f32 gives for a non-negative input, as two
// integer instructions (the compiler's cvt sequence carries an extra exponent test per element)
__device__ __forceinline__ float relu_round_tf32(float x) {
    return __uint_as_float((__float_as_uint(fmaxf(x, 0.f)) + 0x1000u) & 0xFFFFE000u);
}

__global__ void __launch_bounds__(kCThreads, 1)
tc_mlp_chain_fwd_kernel(const __grid_constant__ ChainMaps maps, const ChainParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* sW = base;
    float* sBias = reinterpret_cast<float*>(base + kCOffBias);
    float* sHead = reinterpret_cast<float*>(base + kCOffHead);     // [2][128] then b[2]
    uint64_t* w_full = reinterpret_cast<uint64_t*>(base + kCOffBars);
    uint64_t* w_empty = w_full + 2;
    uint64_t* acc_full = w_empty + 2;
    uint64_t* a_ready = acc_full + 2;
    uint64_t* box_bar = a_ready + 2;          // [8 warps][2]
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(box_bar + 16);
    volatile int* cta_abort = reinterpret_cast<volatile int*>(tmem_ptr + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t tiles = (p.M + 127) / 128;
    const int64_t pairs = (tiles + 1) / 2;
    const int L = p.L;

    if (threadIdx.x == 0) {
        prefetch_tmap(&maps.x);
        for (int l = 0; l < L; ++l) {
            prefetch_tmap(&maps.w[l]);
            if ((p.store_mask >> l) & 1) prefetch_tmap(&maps.out[l]);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&w_full[i], 1);
            mbar_init(&w_empty[i], 1);
            mbar_init(&acc_full[i], 1);
            mbar_init(&a_ready[i], 4);
        }
        for (int i = 0; i < 16; ++i) mbar_init(&box_bar[i], 1);
        *cta_abort = 0;
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_ptr, 512);
    if (warp >= 2) {
        const int t = threadIdx.x - 64;       // 0 .. 255
        for (int i = t; i < L * 128; i += 256) {
            const float* b = p.bias[i >> 7];
            sBias[i] = b ? __ldg(b + (i & 127)) : 0.f;
        }
        if (p.head_w) {
            sHead[t] = __ldg(p.head_w + t);
            if (t < 2) sHead[256 + t] = p.head_b ? __ldg(p.head_b + t) : 0.f;
        }
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp == 0) {
        if (lane == 0) {   // ------------------------------------------------ TMA producer: weight slots
            uint32_t cnt = 0;
            bool alive = true;
            for (int64_t pr = blockIdx.x; pr < pairs && alive; pr += gridDim.x) {
                for (int l = 0; l < L; ++l, ++cnt) {
                    const uint32_t slot = cnt & 1, ph = (cnt >> 1) & 1;
                    if (!mbar_wait(&w_empty[slot], ph ^ 1, cta_abort, p.g_abort, 61)) { alive = false; break; }
                    mbar_arrive_expect_tx(&w_full[slot], (uint32_t)kCWSlot);
                    for (int kb = 0; kb < 4; ++kb)
                        tma_load_2d(&maps.w[l], &w_full[slot], sW + slot * kCWSlot + kb * 16384, kb * 32, 0);
                }
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {   // ---------------------------------------------- MMA issuer
            const uint32_t idesc = make_idesc_tf32(128, false, false);
            const uint64_t d_w0 = make_sdesc_sw128(smem_u32(sW), 16, 1024);
            uint32_t cnt = 0, ar[2] = {0, 0};
            bool alive = true;
            for (int64_t pr = blockIdx.x; pr < pairs && alive; pr += gridDim.x) {
                const bool act1 = 2 * pr + 1 < tiles;
                for (int l = 0; l < L && alive; ++l, ++cnt) {
                    const uint32_t slot = cnt & 1, ph = (cnt >> 1) & 1;
                    if (!mbar_wait(&w_full[slot], ph, cta_abort, p.g_abort, 62)) { alive = false; break; }
                    for (int s = 0; s < 2; ++s) {
                        if (s == 1 && !act1) break;
                        if (!mbar_wait(&a_ready[s], ar[s] & 1, cta_abort, p.g_abort, 63)) { alive = false; break; }
                        ++ar[s];
                        tc_fence_after();
                        const uint32_t tA = tmem_base + (uint32_t)s * 256u, tD = tA + 128u;
                        const uint64_t bw = sdesc_advance(d_w0, slot * (uint32_t)kCWSlot);
#pragma unroll
                        for (int kb = 0; kb < 4; ++kb)
#pragma unroll
                            for (int k = 0; k < 4; ++k)
                                umma_tf32_ts(tD, tA + (uint32_t)(kb * 32 + k * 8), sdesc_advance(bw, kb * 16384 + k * 32), idesc,
                                             (kb | k) != 0);
                        umma_commit(&acc_full[s]);
                    }
                    if (alive) umma_commit(&w_empty[slot]);
                }
            }
        }
    } else {   // -------------------------------------------------------------- epilogue warps
        const int s = (warp - 2) >> 2;           // stream
        const int q = warp & 3;                  // TMEM lane quarter this warp may access
        uint8_t* boxes = base + kCOffBox + (warp - 2) * 2 * kCBox;
        uint64_t* bbar = box_bar + (warp - 2) * 2;
        const uint32_t tA = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)s * 256u, tD = tA + 128u;
        uint32_t af = 0, ld_cnt[2] = {0, 0}, st_chunk = 0;
        const uint32_t sw = (uint32_t)(lane & 7);
        bool alive = true;
        auto wait_all = [&](uint64_t* bb, uint32_t ph, int code) -> bool {
            bool ok = mbar_wait(bb, ph, cta_abort, p.g_abort, code);
            return __all_sync(0xffffffffu, ok);
        };
        for (int64_t pr = blockIdx.x; pr < pairs && alive; pr += gridDim.x) {
            const int64_t tile = 2 * pr + s;
            if (tile >= tiles) break;
            const int64_t row0 = tile * 128 + q * 32, row = row0 + lane;
            const bool rows_ok = row0 < p.M, row_ok = row < p.M;
            // ---- X copy-in: four [32 x 32] boxes -> registers -> A columns of this stream
            if (lane == 0) tma_store_wait_read<0>();     // staging boxes of the previous tile have been read by their stores
            __syncwarp();
            if (rows_ok && lane == 0) {
                for (int c = 0; c < 2; ++c) {
                    mbar_arrive_expect_tx(&bbar[c], (uint32_t)kCBox);
                    tma_load_2d(&maps.x, &bbar[c], boxes + c * kCBox, c * 32, (int)row0);
                }
            }
            for (int c = 0; c < 4; ++c) {
                uint32_t v[32];
                if (rows_ok) {
                    if (!wait_all(&bbar[c & 1], ld_cnt[c & 1] & 1, 64)) { alive = false; break; }
                    ++ld_cnt[c & 1];
                    const uint8_t* bx = boxes + (c & 1) * kCBox + lane * 128;
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const float4 x4 = *reinterpret_cast<const float4*>(bx + (((uint32_t)j ^ sw) << 4));
                        v[4 * j] = __float_as_uint(x4.x); v[4 * j + 1] = __float_as_uint(x4.y);
                        v[4 * j + 2] = __float_as_uint(x4.z); v[4 * j + 3] = __float_as_uint(x4.w);
                    }
                    __syncwarp();
                    if (lane == 0 && c + 2 < 4) {
                        mbar_arrive_expect_tx(&bbar[c & 1], (uint32_t)kCBox);
                        tma_load_2d(&maps.x, &bbar[c & 1], boxes + (c & 1) * kCBox, (c + 2) * 32, (int)row0);
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = 0u;
                }
                tmem_st_32x32(tA + (uint32_t)(c * 32), v);
            }
            if (!alive) break;
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&a_ready[s]);
            const int64_t rrow = (p.res_layer >= 0 && row_ok) ? row % p.res_rows : 0;
            // ---- layers
            for (int l = 0; l < L; ++l) {
                if (!wait_all(&acc_full[s], af & 1, 65)) { alive = false; break; }
                ++af;
                tc_fence_after();
                const bool last = l == L - 1;
                const bool relu = (p.relu_mask >> l) & 1, store = (p.store_mask >> l) & 1;
                const bool has_res = l == p.res_layer;
                const float* bl = sBias + l * 128;
                float h0 = 0.f, h1 = 0.f;
                // accumulator chunks are double-buffered in registers: the tcgen05.ld of chunk c + 1 is in flight while
                // chunk c is processed (the loop is fully unrolled so that both buffers keep static register names)
                uint32_t v[2][32];
                if (!(p.dbg & 1)) tmem_ld_32x32(tD, v[0]);
#pragma unroll
                for (int c = 0; c < ((p.dbg & 1) ? 0 : 4); ++c) {
                    float4 r4[8];
                    if (has_res && row_ok) {     // issued before the accumulator wait so that the two latencies overlap
                        const float4* rp = reinterpret_cast<const float4*>(p.res + rrow * p.ldres + c * 32);
#pragma unroll
                        for (int j = 0; j < 8; ++j) r4[j] = __ldg(rp + j);
                    }
                    tmem_ld_wait();
                    if (c + 1 < 4) tmem_ld_32x32(tD + (uint32_t)((c + 1) * 32), v[(c + 1) & 1]);
                    float f[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[c & 1][j]) + bl[c * 32 + j];
                    if (has_res && row_ok) {
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            f[4 * j] += r4[j].x; f[4 * j + 1] += r4[j].y; f[4 * j + 2] += r4[j].z; f[4 * j + 3] += r4[j].w;
                        }
                    }
                    if (relu) {      // hidden activations are rounded to the nearest tf32 (as npfk_linear_fwd does), so the
#pragma unroll               // truncation by the next layer's MMA is exact
                        for (int j = 0; j < 32; ++j) f[j] = relu_round_tf32(f[j]);
                    }
                    if (!last) {
                        uint32_t o[32];
#pragma unroll
                        for (int j = 0; j < 32; ++j) o[j] = __float_as_uint(f[j]);
                        tmem_st_32x32(tA + (uint32_t)(c * 32), o);
                    } else if (p.head_w) {
                        const float4* w0 = reinterpret_cast<const float4*>(sHead + c * 32);
                        const float4* w1 = reinterpret_cast<const float4*>(sHead + 128 + c * 32);
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const float4 a = w0[j], b = w1[j];
                            h0 = fmaf(f[4 * j], a.x, h0); h0 = fmaf(f[4 * j + 1], a.y, h0);
                            h0 = fmaf(f[4 * j + 2], a.z, h0); h0 = fmaf(f[4 * j + 3], a.w, h0);
                            h1 = fmaf(f[4 * j], b.x, h1); h1 = fmaf(f[4 * j + 1], b.y, h1);
                            h1 = fmaf(f[4 * j + 2], b.z, h1); h1 = fmaf(f[4 * j + 3], b.w, h1);
                        }
                    }
                    if (store) {
                        uint8_t* sb = boxes + (st_chunk & 1) * kCBox;
                        if (lane == 0) tma_store_wait_read<1>();   // the box stored two chunks ago has been read
                        __syncwarp();
#pragma unroll
                        for (int j = 0; j < 8; ++j)
                            *reinterpret_cast<float4*>(sb + lane * 128 + (((uint32_t)j ^ sw) << 4)) =
                                make_float4(f[4 * j], f[4 * j + 1], f[4 * j + 2], f[4 * j + 3]);
                        fence_proxy_async();
                        __syncwarp();
                        if (lane == 0 && rows_ok) {
                            tma_store_2d(&maps.out[l], sb, c * 32, (int)row0);
                            tma_store_commit();
                        }
                        ++st_chunk;
                    }
                }
                if (!last) {
                    tmem_st_wait();
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&a_ready[s]);
                } else if (p.head_w) {
                    // ---- 128 -> 2 head, predictive-distribution transform and Gaussian log-likelihood of this row
                    h0 += sHead[256];
                    h1 += sHead[257];
                    if (p.suff && row_ok) *reinterpret_cast<float2*>(p.suff + row * 2) = make_float2(h0, h1);
                    if (p.y) {
                        float lp = 0.f;
                        if (row_ok) {
                            const float sc = 0.01f + 0.99f * softplus_c(h1);
                            const float d = __ldg(p.y + row % p.y_rows) - h0;
                            lp = -(d * d) / (2.f * sc * sc) - logf(sc) - kLogSqrt2PiC;
                        }
                        if (rows_ok) {
                            const int64_t r_last = min(row0 + 31, p.M - 1);
                            const int64_t t0 = row0 / p.T;
                            if (t0 == r_last / p.T) {          // the warp's 32 rows belong to one task
                                lp = warp_sum(lp);
                                if (lane == 0) atomicAdd(p.sum_lp + t0, lp);
                            } else if (row_ok) {
                                atomicAdd(p.sum_lp + row / p.T, lp);
                            }
                        }
                    }
                }
            }
        }
        if (lane == 0) tma_store_wait_all<0>();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// ------------------------------------------------------------------------------------------ host
static bool c_aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

int tc_mlp_chain_fwd(const float* x, int64_t ldx, int64_t M, int L, const float* const* w, const float* const* bias,
                     int relu_mask, int res_layer, const float* res, int64_t ldres, int64_t res_rows, float* const* out,
                     const float* head_w, const float* head_b, float* suff, const float* y, int64_t y_rows, int64_t T,
                     float* sum_lp, cudaStream_t st) {
    if (L < 1 || L > kCMaxLayers || M < 1 || M > 2147483647LL) {
        set_error("mlp_chain: 1 <= L <= %d layers and 1 <= M < 2^31 rows (got L=%d M=%lld)", kCMaxLayers, L, (long long)M);
        return NPFK_ERR_INVALID;
    }
    if (!c_aligned16(x) || (ldx & 3) || (res && (!c_aligned16(res) || (ldres & 3) || res_rows <= 0))) {
        set_error("mlp_chain: x / res must be 16-byte aligned with row pitches that are multiples of 4 floats");
        return NPFK_ERR_INVALID;
    }
    if (res_layer >= L || (res_layer >= 0 && !res)) {
        set_error("mlp_chain: res_layer %d without a residual tensor, or past the last layer", res_layer);
        return NPFK_ERR_INVALID;
    }
    if (y && (!head_w || !sum_lp || T <= 0 || y_rows <= 0)) {
        set_error("mlp_chain: the log-likelihood tail needs the 128 -> 2 head, sum_lp, T and y_rows");
        return NPFK_ERR_INVALID;
    }
    if (suff && !c_aligned16(suff)) {
        set_error("mlp_chain: suff must be 16-byte aligned");
        return NPFK_ERR_INVALID;
    }
    ChainMaps maps;
    ChainParams p{};
    int rc = encode_tmap_2d(&maps.x, x, 128, (uint64_t)M, (uint64_t)ldx, 32);
    if (rc) return rc;
    for (int l = 0; l < L; ++l) {
        if (!w[l] || !c_aligned16(w[l])) {
            set_error("mlp_chain: weight %d is null or not 16-byte aligned", l);
            return NPFK_ERR_INVALID;
        }
        rc = encode_tmap_2d(&maps.w[l], w[l], 128, 128, 128, 128);
        if (rc) return rc;
        p.bias[l] = bias ? bias[l] : nullptr;
        if (out && out[l]) {
            if (!c_aligned16(out[l])) {
                set_error("mlp_chain: output %d is not 16-byte aligned", l);
                return NPFK_ERR_INVALID;
            }
            rc = encode_tmap_2d(&maps.out[l], out[l], 128, (uint64_t)M, 128, 32);
            if (rc) return rc;
            p.store_mask |= 1 << l;
        }
    }
    p.M = M; p.L = L;
    p.relu_mask = relu_mask;
    p.res_layer = res ? res_layer : -1;
    p.res = res; p.ldres = ldres; p.res_rows = res_rows;
    p.head_w = head_w; p.head_b = head_b; p.suff = suff;
    p.y = y; p.y_rows = y_rows; p.T = T > 0 ? T : 1; p.sum_lp = sum_lp;
    p.g_abort = abort_flag_ptr();
    static int dbg_env = -1;
    if (dbg_env < 0) {
        const char* e = getenv("NPFK_CHAIN_DBG");
        dbg_env = e ? atoi(e) : 0;
    }
    p.dbg = dbg_env;
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(tc_mlp_chain_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kCSmem);
        if (e != cudaSuccess) {
            set_error("mlp_chain: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
            return NPFK_ERR_CUDA;
        }
        attr_done = true;
    }
    const int64_t tiles = (M + 127) / 128, pairs = (tiles + 1) / 2;
    const int gx = (int)(pairs < kNumSMs ? pairs : kNumSMs);
    tc_mlp_chain_fwd_kernel<<<gx, kCThreads, kCSmem, st>>>(maps, p);
    count_launch();
    return check_launch("tc_mlp_chain_fwd");
}

}  // namespace tc
}  // namespace npfk
