// K1b -- backward of TWO consecutive 128 -> 128 layers of a ReLU MLP in ONE pass (NPFK_MODE_FAST).
//
//   upper layer (weights W1, saved input H1):   dW1 += G^T . H1 ;  db1 += sum_m G ;  Gm = (G . W1) * (H1 > 0)
//   lower layer (weights W0, saved input H0):   dW0 += Gm^T . H0 ; db0 += sum_m Gm ; dX = (Gm . W0) [* (H0 > 0)]
//
// The layer-wise kernel (tc_gemm_bwd.cu) moves three activation-sized tensors per layer (G, H in; dX out).  Here the
// gradient between the two layers, Gm, never leaves the SM: per 128-row tile the kernel reads G, H1, H0 and writes dX,
// four tensors for two layers instead of six.  Replaces autograd of two nn.Linear + ReLU of MLP.forward (reference
// npf/architectures/mlp.py:95-109).
//
// Per tile, shared memory holds ONE copy of the current gradient tile (G, later overwritten by Gm), one saved-input tile
// (H1, then H0) and one weight matrix (W1, then W0), all as four [128 rows x 32 floats] boxes in the MN-major tf32
// layout (128B swizzle, 32B atom).  That layout serves every smem operand: wgrad reads the gradient and H with the row
// index as K, dgrad reads W with its row index as K.  The dgrad A operand (gradient rows, K-major) comes from tensor
// memory instead, so no second copy of G in the other swizzle is needed: the epilogue warps copy G into TMEM for the upper
// layer and write Gm to TMEM (and, in the MN-major layout, to shared memory for the lower wgrad) themselves.
//
//   warp 0      TMA producer (G, H1, W1, H0, W0 per tile; weights come from L2)
//   warp 1      TMEM allocator + MMA issuer: wgrad1, dgrad1, wgrad0, dgrad0 per tile
//   warps 2..5  mid epilogue (mask, tf32 rounding, Gm -> TMEM + smem), dX epilogue (TMA store);
//               dW1 / dW0 at the end of the CTA (red.global.add.v4.f32)
//   warps 6..9  gradient readers: G copy-in (smem -> TMEM), column sums of G / Gm (bias gradients)
//   warps 10..13 mask readers: ReLU mask bits of H1 / H0
// Every buffer is single, so the order of the MMAs is what creates overlap: per tile wgrad1, dgrad1, wgrad0, dgrad0 --
// after wgrad0 the gradient and H buffers are free and the next tile's G / H1 stream in behind dgrad0 and the dX epilogue,
// and wgrad1 of the next tile runs while that epilogue still reads the accumulator.
// TMEM (512 columns): A operand 128 | dX accumulator 128 | dW1 128 | dW0 128.
#include <stdlib.h>

#include "tc_common.cuh"

namespace npfk {
namespace tc {

constexpr int kBThreads = 448;
constexpr int kBTile = 65536;                 // [128 x 128] fp32 as four [128 rows x 32 floats] boxes
constexpr int kBBox = 16384;
constexpr int kBOffG = 0;
constexpr int kBOffH = kBTile;
constexpr int kBOffW = 2 * kBTile;
constexpr int kBOffEpi = 3 * kBTile;          // per epilogue warp: one [32 x 32] staging box for the dX store
constexpr int kBOffMask = kBOffEpi + 4 * 4096;
constexpr int kBOffBars = kBOffMask + 2 * 128 * 16;
constexpr int kBNumBars = 16;
constexpr int kBSmem = 1024 + kBOffBars + kBNumBars * 8 + 16;
constexpr uint32_t kBTmA = 0, kBTmAcc = 128, kBTmDw1 = 256, kBTmDw0 = 384;

struct ChainBwdParams {
    int64_t M;
    float* dw1; int64_t lddw1; float* db1;
    float* dw0; int64_t lddw0; float* db0;
    int masked0;          // dX is zeroed where H0 <= 0
    int accumulate;       // dX tiles are added to the destination (TMA reduce) instead of stored
    int dbg;              // timing experiments only (NPFK_CHAIN_BWD_DBG): 1 epilogues idle, 2 readers idle, 4 no wgrad, 8 no
                          // dgrad, 16 no G/H loads, 32 no W loads, 64 no G copy-in
    int* g_abort;
};

// 32 floats of row `r` (0..127) of box `c` of a tile in the MN-major 32B-atom layout: chunk k8 (8 floats) sits at
// r * 128 + ((k8 ^ (r & 3)) << 5)
__device__ __forceinline__ uint8_t* tile_chunk(uint8_t* tile, int c, int r, int k8) {
    return tile + c * kBBox + r * 128 + ((k8 ^ (r & 3)) << 5);
}

__device__ __forceinline__ uint32_t relu_bits8(const float4 a, const float4 b) {
    return (a.x > 0.f ? 1u : 0u) | (a.y > 0.f ? 2u : 0u) | (a.z > 0.f ? 4u : 0u) | (a.w > 0.f ? 8u : 0u) |
           (b.x > 0.f ? 16u : 0u) | (b.y > 0.f ? 32u : 0u) | (b.z > 0.f ? 64u : 0u) | (b.w > 0.f ? 128u : 0u);
}

__global__ void __launch_bounds__(kBThreads, 1)
tc_chain_bwd2_kernel(const __grid_constant__ CUtensorMap tmG, const __grid_constant__ CUtensorMap tmH1,
                     const __grid_constant__ CUtensorMap tmH0, const __grid_constant__ CUtensorMap tmW1,
                     const __grid_constant__ CUtensorMap tmW0, const __grid_constant__ CUtensorMap tmC,
                     const ChainBwdParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* sG = base + kBOffG;
    uint8_t* sH = base + kBOffH;
    uint8_t* sW = base + kBOffW;
    uint8_t* epi = base + kBOffEpi;
    uint4* maskbuf = reinterpret_cast<uint4*>(base + kBOffMask);
    uint64_t* bars = reinterpret_cast<uint64_t*>(base + kBOffBars);
    uint64_t* g_full = bars + 0;       // TMA: G tile landed
    uint64_t* g_free = bars + 1;       // lower wgrad done + readers done with Gm: the G buffer may be reloaded
    uint64_t* h_full = bars + 2;
    uint64_t* h_free = bars + 3;       // wgrad done + readers done with the H tile
    uint64_t* w_full = bars + 4;
    uint64_t* w_free = bars + 5;       // dgrad done
    uint64_t* a_ready = bars + 6;      // epilogue warps: A operand in TMEM (and Gm in smem) written
    uint64_t* gmid_full = bars + 7;    // epilogue warps: Gm in smem written (for the readers)
    uint64_t* acc_full = bars + 8;     // dgrad accumulator complete
    uint64_t* mask_full = bars + 9;    // [2]: per layer
    uint64_t* mask_empty = bars + 11;  // [2]
    uint64_t* dw_full = bars + 13;
    uint64_t* acc_empty = bars + 14;   // dX epilogue has read the dgrad accumulator
    uint64_t* a_free = bars + 15;      // lower dgrad done: the A columns may take the next tile's G
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + kBNumBars);
    volatile int* cta_abort = reinterpret_cast<volatile int*>(tmem_ptr + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t num_tiles = (p.M + 127) / 128;

    if (threadIdx.x == 0) {
        prefetch_tmap(&tmG); prefetch_tmap(&tmH1); prefetch_tmap(&tmH0);
        prefetch_tmap(&tmW1); prefetch_tmap(&tmW0); prefetch_tmap(&tmC);
        mbar_init(g_full, 1);
        mbar_init(g_free, 5);
        mbar_init(h_full, 1);
        mbar_init(h_free, 5);
        mbar_init(w_full, 1);
        mbar_init(w_free, 1);
        mbar_init(a_ready, 4);
        mbar_init(gmid_full, 4);
        mbar_init(acc_full, 1);
        for (int b = 0; b < 2; ++b) {
            mbar_init(&mask_full[b], b == 0 ? 8 : 4);     // upper: mask readers + gradient readers (G summed)
            mbar_init(&mask_empty[b], 4);
        }
        mbar_init(dw_full, 1);
        mbar_init(acc_empty, 4);
        mbar_init(a_free, 1);
        *cta_abort = 0;
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_ptr, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp == 0) {
        if (lane == 0) {   // ---------------- TMA producer
            uint32_t gc = 0, hc = 0, wc = 0;
            bool alive = true;
            for (int64_t tile = blockIdx.x; tile < num_tiles && alive; tile += gridDim.x) {
                const int row0 = (int)(tile * 128);
                if (!mbar_wait(g_free, (gc & 1) ^ 1, cta_abort, p.g_abort, 71)) break;
                ++gc;
                if (p.dbg & 16) mbar_arrive(g_full);
                else {
                    mbar_arrive_expect_tx(g_full, (uint32_t)kBTile);
                    for (int j = 0; j < 4; ++j) tma_load_2d(&tmG, g_full, sG + j * kBBox, j * 32, row0);
                }
                for (int step = 0; step < 2; ++step) {
                    if (!mbar_wait(h_free, (hc & 1) ^ 1, cta_abort, p.g_abort, 72)) { alive = false; break; }
                    ++hc;
                    if (p.dbg & 16) mbar_arrive(h_full);
                    else {
                        mbar_arrive_expect_tx(h_full, (uint32_t)kBTile);
                        for (int j = 0; j < 4; ++j)
                            tma_load_2d(step == 0 ? &tmH1 : &tmH0, h_full, sH + j * kBBox, j * 32, row0);
                    }
                    if (!mbar_wait(w_free, (wc & 1) ^ 1, cta_abort, p.g_abort, 73)) { alive = false; break; }
                    ++wc;
                    if (p.dbg & 32) mbar_arrive(w_full);
                    else {
                        mbar_arrive_expect_tx(w_full, (uint32_t)kBTile);
                        for (int j = 0; j < 4; ++j)
                            tma_load_2d(step == 0 ? &tmW1 : &tmW0, w_full, sW + j * kBBox, j * 32, 0);
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {   // ---------------- MMA issuer
            const uint32_t idesc_w = make_idesc_tf32(128, true, true);     // wgrad: A = grad^T, B = H, both MN-major
            const uint32_t idesc_d = make_idesc_tf32(128, false, true);    // dgrad: A from TMEM, B = W MN-major
            const uint64_t dG = make_sdesc_mn32(smem_u32(sG), kBBox, 512);
            const uint64_t dH = make_sdesc_mn32(smem_u32(sH), kBBox, 512);
            const uint64_t dW = make_sdesc_mn32(smem_u32(sW), kBBox, 512);
            const uint32_t tA = tmem_base + kBTmA, tAcc = tmem_base + kBTmAcc;
            uint32_t gc = 0, hc = 0, wc = 0, ac = 0, n = 0;
            bool first = true;
            for (int64_t tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++n) {
                // ---- upper layer.  The wgrad needs only the TMA tiles and its own accumulator: it runs while the
                // previous tile's dX epilogue is still reading the dgrad accumulator.
                if (!mbar_wait(g_full, gc & 1, cta_abort, p.g_abort, 74)) break;
                ++gc;
                if (!mbar_wait(h_full, hc & 1, cta_abort, p.g_abort, 75)) break;
                ++hc;
                tc_fence_after();
#pragma unroll
                for (int s = 0; s < ((p.dbg & 4) ? 0 : 16); ++s)
                    umma_tf32(tmem_base + kBTmDw1, sdesc_advance(dG, s * 1024), sdesc_advance(dH, s * 1024), idesc_w,
                              (first && s == 0) ? 0u : 1u);
                umma_commit(h_free);
                if (!mbar_wait(a_ready, ac & 1, cta_abort, p.g_abort, 76)) break;
                ++ac;
                if (!mbar_wait(w_full, wc & 1, cta_abort, p.g_abort, 77)) break;
                ++wc;
                if (!mbar_wait(acc_empty, (n & 1) ^ 1, cta_abort, p.g_abort, 93)) break;
                tc_fence_after();
#pragma unroll
                for (int s = 0; s < ((p.dbg & 8) ? 0 : 16); ++s)
                    umma_tf32_ts(tAcc, tA + (uint32_t)(s * 8), sdesc_advance(dW, s * 1024), idesc_d, s != 0 ? 1u : 0u);
                umma_commit(acc_full);
                umma_commit(w_free);
                // ---- lower layer.  The wgrad goes first: when it is done the gradient buffer is free, and the next tile's
                // G (and H1) stream in behind the lower dgrad and the dX epilogue.
                if (!mbar_wait(a_ready, ac & 1, cta_abort, p.g_abort, 78)) break;
                ++ac;
                if (!mbar_wait(h_full, hc & 1, cta_abort, p.g_abort, 80)) break;
                ++hc;
                tc_fence_after();
#pragma unroll
                for (int s = 0; s < ((p.dbg & 4) ? 0 : 16); ++s)
                    umma_tf32(tmem_base + kBTmDw0, sdesc_advance(dG, s * 1024), sdesc_advance(dH, s * 1024), idesc_w,
                              (first && s == 0) ? 0u : 1u);
                umma_commit(h_free);
                umma_commit(g_free);
                if (!mbar_wait(w_full, wc & 1, cta_abort, p.g_abort, 79)) break;
                ++wc;
                tc_fence_after();
#pragma unroll
                for (int s = 0; s < ((p.dbg & 8) ? 0 : 16); ++s)
                    umma_tf32_ts(tAcc, tA + (uint32_t)(s * 8), sdesc_advance(dW, s * 1024), idesc_d, s != 0 ? 1u : 0u);
                umma_commit(acc_full);
                umma_commit(w_free);
                umma_commit(a_free);
                first = false;
            }
            umma_commit(dw_full);
        }
    } else if (warp >= 10) {   // ---------------- mask readers: ReLU mask bits of H1 / H0
        const int r = (warp - 10) * 32 + lane;   // row whose mask bits this thread extracts
        uint32_t hc = 0, n = 0;
        auto wait_all = [&](uint64_t* bb, uint32_t ph, int code) -> bool {
            bool ok = mbar_wait(bb, ph, cta_abort, p.g_abort, code);
            return __all_sync(0xffffffffu, ok);
        };
        auto extract_mask = [&](int buf) {
            uint32_t wds[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                uint32_t bits = 0;
#pragma unroll
                for (int k8 = 0; k8 < 4; ++k8) {
                    const float4* ptr = reinterpret_cast<const float4*>(tile_chunk(sH, j, r, k8));
                    bits |= relu_bits8(ptr[0], ptr[1]) << (k8 * 8);
                }
                wds[j] = bits;
            }
            maskbuf[buf * 128 + r] = make_uint4(wds[0], wds[1], wds[2], wds[3]);
        };
        for (int64_t tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++n) {
            for (int layer = 0; layer < 2; ++layer) {      // 0: upper (H1), 1: lower (H0)
                if (!wait_all(h_full, hc & 1, 82)) goto mask_done;
                ++hc;
                if (!wait_all(&mask_empty[layer], (n & 1) ^ 1, 83)) goto mask_done;
                if ((layer == 0 || p.masked0) && !(p.dbg & 2)) extract_mask(layer);
                __syncwarp();
                if (lane == 0) {
                    mbar_arrive(&mask_full[layer]);
                    mbar_arrive(h_free);
                }
            }
        }
    mask_done:;
    } else if (warp >= 6) {   // ---------------- gradient readers: G copy-in, bias-gradient column sums of G / Gm
        const int q = warp & 3;                  // TMEM lane quarter this warp may access
        const int r = q * 32 + lane;             // row copied into TMEM
        const int w4 = warp - 6;                 // column sums: this warp's 32 rows, 4 columns per lane
        const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16);
        float4 db1s = make_float4(0.f, 0.f, 0.f, 0.f), db0s = db1s;
        uint32_t gc = 0, gm = 0, n = 0;
        auto wait_all = [&](uint64_t* bb, uint32_t ph, int code) -> bool {
            bool ok = mbar_wait(bb, ph, cta_abort, p.g_abort, code);
            return __all_sync(0xffffffffu, ok);
        };
        // columns 4 lane .. 4 lane + 3 (box lane / 8, 32-byte chunk (lane % 8) / 2, its half lane % 2) over the warp's 32 rows
        auto colsum = [&](float4& acc) {
            const uint8_t* sA = sG + (lane >> 3) * kBBox + (lane & 1) * 16;
            const int k8 = (lane & 7) >> 1;
#pragma unroll 8
            for (int i = 0; i < 32; ++i) {
                const int rr = w4 * 32 + i;
                const float4 x = *reinterpret_cast<const float4*>(sA + rr * 128 + ((k8 ^ (rr & 3)) << 5));
                acc.x += x.x; acc.y += x.y; acc.z += x.z; acc.w += x.w;
            }
        };
        for (int64_t tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++n) {
            // ---- G copy-in: smem (MN-major layout) -> registers -> A columns, once the previous lower dgrad has read them
            if (!wait_all(g_full, gc & 1, 81)) break;
            ++gc;
            if (!wait_all(a_free, (n & 1) ^ 1, 94)) break;
            tc_fence_after();
#pragma unroll
            for (int c = 0; c < ((p.dbg & 64) ? 0 : 4); ++c) {
                uint32_t v[32];
#pragma unroll
                for (int k8 = 0; k8 < 4; ++k8) {
                    const float4* ptr = reinterpret_cast<const float4*>(tile_chunk(sG, c, r, k8));
                    const float4 x0 = ptr[0], x1 = ptr[1];
                    v[k8 * 8 + 0] = __float_as_uint(x0.x); v[k8 * 8 + 1] = __float_as_uint(x0.y);
                    v[k8 * 8 + 2] = __float_as_uint(x0.z); v[k8 * 8 + 3] = __float_as_uint(x0.w);
                    v[k8 * 8 + 4] = __float_as_uint(x1.x); v[k8 * 8 + 5] = __float_as_uint(x1.y);
                    v[k8 * 8 + 6] = __float_as_uint(x1.z); v[k8 * 8 + 7] = __float_as_uint(x1.w);
                }
                tmem_st_32x32(lane_base + kBTmA + (uint32_t)(c * 32), v);
            }
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(a_ready);
            if (p.db1 && !(p.dbg & 2)) colsum(db1s);
            __syncwarp();
            if (lane == 0) mbar_arrive(&mask_full[0]);      // the column sums of G are taken: Gm may overwrite it
            // ---- column sums of Gm, then the gradient buffer is free for the next tile
            if (!wait_all(gmid_full, gm & 1, 86)) break;
            ++gm;
            if (p.db0 && !(p.dbg & 2)) colsum(db0s);
            __syncwarp();
            if (lane == 0) mbar_arrive(g_free);
        }
        if (*cta_abort == 0) {
            if (p.db1) {
                atomicAdd(p.db1 + 4 * lane + 0, db1s.x); atomicAdd(p.db1 + 4 * lane + 1, db1s.y);
                atomicAdd(p.db1 + 4 * lane + 2, db1s.z); atomicAdd(p.db1 + 4 * lane + 3, db1s.w);
            }
            if (p.db0) {
                atomicAdd(p.db0 + 4 * lane + 0, db0s.x); atomicAdd(p.db0 + 4 * lane + 1, db0s.y);
                atomicAdd(p.db0 + 4 * lane + 2, db0s.z); atomicAdd(p.db0 + 4 * lane + 3, db0s.w);
            }
        }
    } else {   // ---------------- epilogue warps 2..5: TMEM lane quarter = warp % 4
        const int q = warp & 3;
        const int r = q * 32 + lane;             // row of the tile owned by this thread
        uint8_t* stg = epi + (warp - 2) * 4096;
        const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16);
        uint32_t af = 0, n = 0;
        auto wait_all = [&](uint64_t* bb, uint32_t ph, int code) -> bool {
            bool ok = mbar_wait(bb, ph, cta_abort, p.g_abort, code);
            return __all_sync(0xffffffffu, ok);
        };
        // masked value rounded to the nearest tf32 (ties away from zero, what cvt.rna.tf32.f32 gives) as integer ops
        auto mask_round = [](uint32_t v, uint32_t bits, int j) -> uint32_t {
            return ((bits >> j) & 1u) ? ((v + 0x1000u) & 0xFFFFE000u) : 0u;
        };
        for (int64_t tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++n) {
            const int64_t row0 = tile * 128 + q * 32;
            // ---- mid epilogue: Gm = round_tf32((G . W1) * (H1 > 0)) -> A columns + the G buffer
            if (!wait_all(&mask_full[0], n & 1, 88)) break;
            uint4 mk = maskbuf[r];
            __syncwarp();
            if (lane == 0) mbar_arrive(&mask_empty[0]);
            if (!wait_all(acc_full, af & 1, 89)) break;
            ++af;
            tc_fence_after();
            if (!(p.dbg & 1)) {
                const uint32_t mw[4] = {mk.x, mk.y, mk.z, mk.w};
                uint32_t v[2][32];
                tmem_ld_32x32(lane_base + kBTmAcc, v[0]);
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    tmem_ld_wait();
                    if (c + 1 < 4) tmem_ld_32x32(lane_base + kBTmAcc + (uint32_t)((c + 1) * 32), v[(c + 1) & 1]);
                    const uint32_t bits = mw[c];
                    uint32_t o[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) o[j] = mask_round(v[c & 1][j], bits, j);
                    tmem_st_32x32(lane_base + kBTmA + (uint32_t)(c * 32), o);
#pragma unroll
                    for (int k8 = 0; k8 < 4; ++k8) {
                        uint4* ptr = reinterpret_cast<uint4*>(tile_chunk(sG, c, r, k8));
                        ptr[0] = make_uint4(o[k8 * 8 + 0], o[k8 * 8 + 1], o[k8 * 8 + 2], o[k8 * 8 + 3]);
                        ptr[1] = make_uint4(o[k8 * 8 + 4], o[k8 * 8 + 5], o[k8 * 8 + 6], o[k8 * 8 + 7]);
                    }
                }
            }
            tmem_st_wait();
            fence_proxy_async();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(a_ready);
                mbar_arrive(gmid_full);
            }
            // ---- dX epilogue: (Gm . W0) [* (H0 > 0)], tf32 rounding, swizzled staging box, TMA store
            if (!wait_all(&mask_full[1], n & 1, 90)) break;
            mk = p.masked0 ? maskbuf[128 + r] : make_uint4(~0u, ~0u, ~0u, ~0u);
            __syncwarp();
            if (lane == 0) mbar_arrive(&mask_empty[1]);
            if (!wait_all(acc_full, af & 1, 91)) break;
            ++af;
            tc_fence_after();
            if (!(p.dbg & 1)) {
                const uint32_t mw[4] = {mk.x, mk.y, mk.z, mk.w};
                uint32_t v[2][32];
                tmem_ld_32x32(lane_base + kBTmAcc, v[0]);
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    tmem_ld_wait();
                    if (c + 1 < 4) tmem_ld_32x32(lane_base + kBTmAcc + (uint32_t)((c + 1) * 32), v[(c + 1) & 1]);
                    else {           // accumulator fully read: the next tile's upper dgrad may overwrite it
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(acc_empty);
                    }
                    const uint32_t bits = mw[c];
                    uint32_t o[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) o[j] = mask_round(v[c & 1][j], bits, j);
                    if (lane == 0) tma_store_wait_read<0>();   // the box stored one chunk ago has been read
                    __syncwarp();
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        *reinterpret_cast<uint4*>(stg + lane * 128 + ((j ^ (lane & 7)) << 4)) =
                            make_uint4(o[4 * j], o[4 * j + 1], o[4 * j + 2], o[4 * j + 3]);
                    fence_proxy_async();
                    __syncwarp();
                    if (lane == 0 && row0 < p.M) {
                        if (p.accumulate) tma_reduce_add_2d(&tmC, stg, c * 32, (int)row0);
                        else tma_store_2d(&tmC, stg, c * 32, (int)row0);
                        tma_store_commit();
                    }
                }
            } else {
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(acc_empty);
            }
        }
        if (lane == 0) tma_store_wait_all<0>();
        // dW1, dW0: accumulated over all tiles of this CTA
        if (*cta_abort == 0 && (int64_t)blockIdx.x < num_tiles) {
            if (wait_all(dw_full, 0, 92)) {
                tc_fence_after();
#pragma unroll
                for (int which = 0; which < 2; ++which) {
                    float* out = which == 0 ? p.dw1 + (int64_t)r * p.lddw1 : p.dw0 + (int64_t)r * p.lddw0;
                    const uint32_t tcol = which == 0 ? kBTmDw1 : kBTmDw0;
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        uint32_t v[32];
                        tmem_ld_32x32(lane_base + tcol + (uint32_t)(c * 32), v);
                        tmem_ld_wait();
#pragma unroll
                        for (int j = 0; j < 32; j += 4)
                            red_add_v4(out + c * 32 + j, __uint_as_float(v[j]), __uint_as_float(v[j + 1]),
                                       __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3]));
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// ------------------------------------------------------------------------------------------ host
static bool b_aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

bool chain_bwd2_eligible(const float* g, int64_t ldg, const float* h1, int64_t ldh1, const float* w1, int64_t ldw1,
                         const float* h0, int64_t ldh0, const float* w0, int64_t ldw0, const float* dx, int64_t lddx,
                         const float* dw1, int64_t lddw1, const float* dw0, int64_t lddw0, int64_t M) {
    if (M < 128 || M > 2147483647LL) return false;
    if (!b_aligned16(g) || !b_aligned16(h1) || !b_aligned16(w1) || !b_aligned16(h0) || !b_aligned16(w0) ||
        !b_aligned16(dx) || !b_aligned16(dw1) || !b_aligned16(dw0))
        return false;
    if ((ldg & 3) || (ldh1 & 3) || (ldw1 & 3) || (ldh0 & 3) || (ldw0 & 3) || (lddx & 3) || (lddw1 & 3) || (lddw0 & 3))
        return false;
    return true;
}

int tc_chain_bwd2(const float* g, int64_t ldg, const float* h1, int64_t ldh1, const float* w1, int64_t ldw1,
                  const float* h0, int64_t ldh0, const float* w0, int64_t ldw0, int flags, float* dx, int64_t lddx,
                  float* dw1, int64_t lddw1, float* db1, float* dw0, int64_t lddw0, float* db0, int64_t M,
                  cudaStream_t st) {
    CUtensorMap tmG, tmH1, tmH0, tmW1, tmW0, tmC;
    int rc = encode_tmap_2d(&tmG, g, 128, (uint64_t)M, (uint64_t)ldg, 128, true);
    if (rc) return rc;
    rc = encode_tmap_2d(&tmH1, h1, 128, (uint64_t)M, (uint64_t)ldh1, 128, true);
    if (rc) return rc;
    rc = encode_tmap_2d(&tmH0, h0, 128, (uint64_t)M, (uint64_t)ldh0, 128, true);
    if (rc) return rc;
    rc = encode_tmap_2d(&tmW1, w1, 128, 128, (uint64_t)ldw1, 128, true);
    if (rc) return rc;
    rc = encode_tmap_2d(&tmW0, w0, 128, 128, (uint64_t)ldw0, 128, true);
    if (rc) return rc;
    rc = encode_tmap_2d(&tmC, dx, 128, (uint64_t)M, (uint64_t)lddx, 32);
    if (rc) return rc;
    ChainBwdParams p{};
    p.M = M;
    p.dw1 = dw1; p.lddw1 = lddw1; p.db1 = db1;
    p.dw0 = dw0; p.lddw0 = lddw0; p.db0 = db0;
    p.masked0 = flags & 1;
    p.accumulate = (flags >> 1) & 1;
    p.g_abort = abort_flag_ptr();
    {
        static int dbg = -1;
        if (dbg < 0) {
            const char* e = getenv("NPFK_CHAIN_BWD_DBG");
            dbg = e ? atoi(e) : 0;
        }
        p.dbg = dbg;
    }
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(tc_chain_bwd2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kBSmem);
        if (e != cudaSuccess) {
            set_error("tc chain bwd: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
            return NPFK_ERR_CUDA;
        }
        attr_done = true;
    }
    const int64_t tiles = (M + 127) / 128;
    const int gx = (int)(tiles < kNumSMs ? tiles : kNumSMs);
    tc_chain_bwd2_kernel<<<gx, kBThreads, kBSmem, st>>>(tmG, tmH1, tmH0, tmW1, tmW0, tmC, p);
    count_launch();
    return check_launch("tc_chain_bwd2");
}

}  // namespace tc
}  // namespace npfk
