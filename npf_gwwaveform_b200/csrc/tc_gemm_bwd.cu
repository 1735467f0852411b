// Fused Linear backward for NPFK_MODE_FAST: dgrad and wgrad of one 128 -> 128 layer in ONE pass over G and H.
//
//   dX[M,128] = (G[M,128] . W[128,128]) * (H > 0)        tcgen05 kind::tf32, A = G K-major, B = W MN-major (resident)
//   dW[128,128] += G^T . H ; db[128] += sum_m G          A = G MN-major, B = H MN-major, accumulated in TMEM over M
//
// The separate kernels (tc_gemm.cu) move 5 activation-sized tensors per layer (dgrad: G, mask, dX; wgrad: G, H);
// here G and H are read from HBM once and dX written once (3 tensors): the second copy of G (other swizzle) comes
// from L2, and the ReLU mask is extracted as bits from the H tile that the wgrad consumes anyway.
// Replaces autograd of nn.Linear + ReLU inside MLP.forward (reference npf/architectures/mlp.py:95-109).
//
// Persistent CTAs of 10 warps (1 CTA / SM):
//   warp 0      TMA producer: W once, then per 128-row tile 4 "W-stages" (32 rows of G and H, MN-major boxes)
//               and 2 "D-stages" (two 128x32 K-major boxes of G)
//   warp 1      TMEM allocator + MMA issuer
//   warps 2..5  dX epilogue: TMEM -> registers -> mask bits, tf32 rounding -> swizzled smem -> TMA store;
//               dW at the end of the CTA (red.global.add.v4.f32)
//   warps 6..9  stage readers: column sums of G (db) and ReLU mask bits of H from the W-stages in shared memory
#include "tc_common.cuh"

namespace npfk {
namespace tc {

constexpr int kFThreads = 320;
constexpr int kFDim = 128;                    // N == K == 128
constexpr int kFStageBytes = 32768;
constexpr int kFStages = 4;
constexpr int kFWBytes = kFDim * kFDim * 4;   // resident W, MN-major boxes
constexpr int kFEpiBytes = 4 * 4096;          // per epilogue warp: one 32x32 fp32 staging box (smem goes to the ring)
constexpr int kFMaskBytes = 2 * 128 * 16;     // two tiles x 128 rows x 128 bits
constexpr int kFOffRing = kFWBytes;
constexpr int kFOffEpi = kFOffRing + kFStages * kFStageBytes;
constexpr int kFOffMask = kFOffEpi + kFEpiBytes;
constexpr int kFOffBars = kFOffMask + kFMaskBytes;
constexpr int kFNumBars = 2 * kFStages + 10;
constexpr int kFSmem = 1024 + kFOffBars + kFNumBars * 8 + 16;
constexpr uint32_t kFTmemCols = 512;          // dX 2 x 128 columns, dW 128 columns
constexpr uint32_t kFTmemDw = 256;

struct FusedBwdParams {
    int64_t M;
    float* dw; int64_t lddw;
    float* db;            // [128] "+=" or null
    int masked;           // dX is zeroed where H <= 0
    int accumulate;       // dX tiles are added to the destination (TMA reduce) instead of stored
    int* g_abort;
};

__global__ void __launch_bounds__(kFThreads, 1)
tc_linear_bwd_fused_kernel(const __grid_constant__ CUtensorMap tmGk, const __grid_constant__ CUtensorMap tmGm,
                           const __grid_constant__ CUtensorMap tmH, const __grid_constant__ CUtensorMap tmW,
                           const __grid_constant__ CUtensorMap tmC, const FusedBwdParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* sW = base;
    uint8_t* ring = base + kFOffRing;
    uint8_t* epi = base + kFOffEpi;
    uint4* maskbuf = reinterpret_cast<uint4*>(base + kFOffMask);
    uint64_t* full = reinterpret_cast<uint64_t*>(base + kFOffBars);
    uint64_t* empty = full + kFStages;
    uint64_t* tmem_full = empty + kFStages;   // [2]
    uint64_t* tmem_empty = tmem_full + 2;     // [2]
    uint64_t* mask_full = tmem_empty + 2;     // [2]
    uint64_t* mask_empty = mask_full + 2;     // [2]
    uint64_t* w_full = mask_empty + 2;
    uint64_t* dw_full = w_full + 1;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(dw_full + 1);
    volatile int* cta_abort = reinterpret_cast<volatile int*>(tmem_ptr + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t num_tiles = (p.M + 127) / 128;

    if (threadIdx.x == 0) {
        prefetch_tmap(&tmGk);
        prefetch_tmap(&tmGm);
        prefetch_tmap(&tmH);
        prefetch_tmap(&tmW);
        prefetch_tmap(&tmC);
        for (int s = 0; s < kFStages; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], 5);          // MMA commit + 4 reader warps
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&tmem_full[b], 1);
            mbar_init(&tmem_empty[b], 4);
            mbar_init(&mask_full[b], 4);
            mbar_init(&mask_empty[b], 4);
        }
        mbar_init(w_full, 1);
        mbar_init(dw_full, 1);
        *cta_abort = 0;
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_ptr, kFTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp == 0) {
        if (lane == 0) {   // ---------------- TMA producer
            mbar_arrive_expect_tx(w_full, (uint32_t)kFWBytes);
            for (int kb = 0; kb < 4; ++kb)
                for (int j = 0; j < 4; ++j) tma_load_2d(&tmW, w_full, sW + (kb * 4 + j) * 4096, j * 32, kb * 32);
            int stage = 0;
            uint32_t phase = 0;
            bool alive = true;
            for (int64_t tile = blockIdx.x; tile < num_tiles && alive; tile += gridDim.x) {
                const int row0 = (int)(tile * 128);
                for (int i = 0; i < 6; ++i) {
                    if (!mbar_wait(&empty[stage], phase ^ 1, cta_abort, p.g_abort, 21)) { alive = false; break; }
                    uint8_t* st = ring + (size_t)stage * kFStageBytes;
                    mbar_arrive_expect_tx(&full[stage], (uint32_t)kFStageBytes);
                    if (i < 4) {        // W-stage: rows row0 + 32 i .. + 31 of G and H, four 32x32 boxes each
                        for (int j = 0; j < 4; ++j) tma_load_2d(&tmGm, &full[stage], st + j * 4096, j * 32, row0 + i * 32);
                        for (int j = 0; j < 4; ++j)
                            tma_load_2d(&tmH, &full[stage], st + 16384 + j * 4096, j * 32, row0 + i * 32);
                    } else {            // D-stage: columns 64 (i-4) .. + 63 of G as two 128 x 32 K-major boxes
                        for (int j = 0; j < 2; ++j)
                            tma_load_2d(&tmGk, &full[stage], st + j * 16384, ((i - 4) * 2 + j) * 32, row0);
                    }
                    if (++stage == kFStages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {   // ---------------- MMA issuer
            const uint32_t idesc_w = make_idesc_tf32(kFDim, true, true);
            const uint32_t idesc_d = make_idesc_tf32(kFDim, false, true);
            const uint32_t ring0 = smem_u32(ring);
            const uint64_t dA_w = make_sdesc_mn32(ring0, 4096, 512), dB_w = make_sdesc_mn32(ring0 + 16384, 4096, 512);
            const uint64_t dA_d = make_sdesc_sw128(ring0, 16, 1024);
            const uint64_t dB_d = make_sdesc_mn32(smem_u32(sW), 4096, 512);
            const uint32_t d_dw = tmem_base + kFTmemDw;
            int stage = 0;
            uint32_t phase = 0, it = 0;
            bool alive = mbar_wait(w_full, 0, cta_abort, p.g_abort, 22);
            bool first = true;
            for (int64_t tile = blockIdx.x; tile < num_tiles && alive; tile += gridDim.x, ++it) {
                const uint32_t buf = it & 1, bphase = (it >> 1) & 1;
                for (int i = 0; i < 4 && alive; ++i) {      // wgrad: dW += G_rows^T . H_rows
                    if (!mbar_wait(&full[stage], phase, cta_abort, p.g_abort, 23)) { alive = false; break; }
                    tc_fence_after();
                    const uint32_t soff = (uint32_t)stage * (uint32_t)kFStageBytes;
                    const uint64_t a0 = sdesc_advance(dA_w, soff), b0 = sdesc_advance(dB_w, soff);
#pragma unroll
                    for (int s = 0; s < 4; ++s)
                        umma_tf32(d_dw, sdesc_advance(a0, s * 1024), sdesc_advance(b0, s * 1024), idesc_w,
                                  (first && i == 0 && s == 0) ? 0u : 1u);
                    umma_commit(&empty[stage]);
                    if (++stage == kFStages) { stage = 0; phase ^= 1; }
                }
                first = false;
                if (!alive) break;
                if (!mbar_wait(&tmem_empty[buf], bphase ^ 1, cta_abort, p.g_abort, 24)) break;
                tc_fence_after();
                const uint32_t d_dx = tmem_base + buf * (uint32_t)kFDim;
                for (int i = 0; i < 2 && alive; ++i) {      // dgrad: dX = G . W
                    if (!mbar_wait(&full[stage], phase, cta_abort, p.g_abort, 25)) { alive = false; break; }
                    tc_fence_after();
                    const uint32_t soff = (uint32_t)stage * (uint32_t)kFStageBytes;
#pragma unroll
                    for (int j = 0; j < 2; ++j) {
                        const int kb = i * 2 + j;
                        const uint64_t a0 = sdesc_advance(dA_d, soff + j * 16384);
                        const uint64_t b0 = sdesc_advance(dB_d, (uint32_t)kb * 16384u);
#pragma unroll
                        for (int s = 0; s < 4; ++s)
                            umma_tf32(d_dx, sdesc_advance(a0, s * 32), sdesc_advance(b0, s * 1024), idesc_d,
                                      (kb | s) != 0 ? 1u : 0u);
                    }
                    umma_commit(&empty[stage]);
                    if (++stage == kFStages) { stage = 0; phase ^= 1; }
                }
                if (alive) umma_commit(&tmem_full[buf]);
            }
            if (alive) umma_commit(dw_full);
        }
    } else if (warp >= 6) {   // ---------------- stage readers: db column sums and ReLU mask bits
        const int rq = warp - 6;
        const int t = rq * 32 + lane;           // column of G summed by this thread
        const int box = t >> 5, col = t & 31;
        float dbsum = 0.f;
        int stage = 0;
        uint32_t phase = 0, it = 0;
        bool alive = true;
        for (int64_t tile = blockIdx.x; tile < num_tiles && alive; tile += gridDim.x, ++it) {
            const uint32_t buf = it & 1, bphase = (it >> 1) & 1;
            for (int i = 0; i < 6; ++i) {
                bool ok = mbar_wait(&full[stage], phase, cta_abort, p.g_abort, 26);
                ok = __all_sync(0xffffffffu, ok);
                if (!ok) { alive = false; break; }
                if (i < 4) {
                    const uint8_t* st = ring + (size_t)stage * kFStageBytes;
                    if (p.db) {
                        const uint8_t* sA = st + box * 4096;
#pragma unroll 8
                        for (int r = 0; r < 32; ++r)
                            dbsum += *reinterpret_cast<const float*>(sA + r * 128 + (((col >> 3) ^ (r & 3)) << 5) + (col & 7) * 4);
                    }
                    if (p.masked && i == rq) {      // this warp's 32 rows of H: lane = row, 128 columns -> 4 words
                        bool mok = mbar_wait(&mask_empty[buf], bphase ^ 1, cta_abort, p.g_abort, 27);
                        mok = __all_sync(0xffffffffu, mok);
                        if (!mok) { alive = false; break; }
                        const uint8_t* sH = st + 16384 + lane * 128;
                        uint32_t wds[4];
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            uint32_t bits = 0;
#pragma unroll
                            for (int c8 = 0; c8 < 4; ++c8) {
                                const float4* ptr = reinterpret_cast<const float4*>(sH + j * 4096 + ((c8 ^ (lane & 3)) << 5));
                                const float4 x0 = ptr[0], x1 = ptr[1];
                                bits |= (x0.x > 0.f ? 1u : 0u) << (c8 * 8 + 0);
                                bits |= (x0.y > 0.f ? 1u : 0u) << (c8 * 8 + 1);
                                bits |= (x0.z > 0.f ? 1u : 0u) << (c8 * 8 + 2);
                                bits |= (x0.w > 0.f ? 1u : 0u) << (c8 * 8 + 3);
                                bits |= (x1.x > 0.f ? 1u : 0u) << (c8 * 8 + 4);
                                bits |= (x1.y > 0.f ? 1u : 0u) << (c8 * 8 + 5);
                                bits |= (x1.z > 0.f ? 1u : 0u) << (c8 * 8 + 6);
                                bits |= (x1.w > 0.f ? 1u : 0u) << (c8 * 8 + 7);
                            }
                            wds[j] = bits;
                        }
                        maskbuf[buf * 128 + rq * 32 + lane] = make_uint4(wds[0], wds[1], wds[2], wds[3]);
                        __syncwarp();
                        if (lane == 0) mbar_arrive(&mask_full[buf]);
                    }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&empty[stage]);
                if (++stage == kFStages) { stage = 0; phase ^= 1; }
            }
        }
        if (alive && p.db) atomicAdd(p.db + t, dbsum);
    } else {   // ---------------- epilogue warps 2..5: TMEM lane quarter = warp % 4
        const int q = warp & 3;
        uint8_t* stg = epi + (warp - 2) * 4096;
        uint32_t it = 0, chunk = 0;
        bool alive = true;
        for (int64_t tile = blockIdx.x; tile < num_tiles && alive; tile += gridDim.x, ++it) {
            const uint32_t buf = it & 1, bphase = (it >> 1) & 1;
            uint4 mk = make_uint4(~0u, ~0u, ~0u, ~0u);
            if (p.masked) {
                bool ok = mbar_wait(&mask_full[buf], bphase, cta_abort, p.g_abort, 28);
                ok = __all_sync(0xffffffffu, ok);
                if (!ok) break;
                mk = maskbuf[buf * 128 + q * 32 + lane];
                __syncwarp();
                if (lane == 0) mbar_arrive(&mask_empty[buf]);
            }
            bool ok = mbar_wait(&tmem_full[buf], bphase, cta_abort, p.g_abort, 29);
            ok = __all_sync(0xffffffffu, ok);
            if (!ok) break;
            tc_fence_after();
            const int64_t row0 = tile * 128 + q * 32;
            const uint32_t mw[4] = {mk.x, mk.y, mk.z, mk.w};
#pragma unroll
            for (int cc = 0; cc < 4; ++cc, ++chunk) {
                uint32_t v[32];
                tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + buf * (uint32_t)kFDim + (uint32_t)(cc * 32), v);
                tmem_ld_wait();
                if (cc == 3) {   // accumulator buffer fully read: hand it back to the MMA warp
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&tmem_empty[buf]);
                }
                float f[32];
                const uint32_t bits = mw[cc];
#pragma unroll
                for (int j = 0; j < 32; ++j) f[j] = ((bits >> j) & 1u) ? round_tf32(__uint_as_float(v[j])) : 0.f;
                uint8_t* sb = stg;
                if (lane == 0) tma_store_wait_read<0>();   // the box stored one chunk ago has been read
                __syncwarp();
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    *reinterpret_cast<float4*>(sb + lane * 128 + ((j ^ (lane & 7)) << 4)) =
                        make_float4(f[4 * j], f[4 * j + 1], f[4 * j + 2], f[4 * j + 3]);
                fence_proxy_async();
                __syncwarp();
                if (lane == 0 && row0 < p.M) {
                    if (p.accumulate) tma_reduce_add_2d(&tmC, sb, cc * 32, (int)row0);
                    else tma_store_2d(&tmC, sb, cc * 32, (int)row0);
                    tma_store_commit();
                }
            }
        }
        if (lane == 0) tma_store_wait_all<0>();
        // dW: accumulated over all tiles of this CTA
        if (alive && (int64_t)blockIdx.x < num_tiles) {
            bool ok = mbar_wait(dw_full, 0, cta_abort, p.g_abort, 30);
            ok = __all_sync(0xffffffffu, ok);
            if (ok) {
                tc_fence_after();
                float* out = p.dw + (int64_t)(q * 32 + lane) * p.lddw;
#pragma unroll
                for (int cc = 0; cc < 4; ++cc) {
                    uint32_t v[32];
                    tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + kFTmemDw + (uint32_t)(cc * 32), v);
                    tmem_ld_wait();
#pragma unroll
                    for (int j = 0; j < 32; j += 4)
                        red_add_v4(out + cc * 32 + j, __uint_as_float(v[j]), __uint_as_float(v[j + 1]),
                                   __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3]));
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kFTmemCols);
    }
}

// ------------------------------------------------------------------------------------------ host
static bool f_aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

bool fused_bwd_eligible(const float* g, int64_t ldg, const float* h, int64_t ldh, const float* w, int64_t ldw,
                        const float* dx, int64_t lddx, const float* dw, int64_t lddw, int64_t M, int N, int K) {
    if (N != kFDim || K != kFDim || M < 128 || M > 2147483647LL) return false;
    if (!f_aligned16(g) || !f_aligned16(h) || !f_aligned16(w) || !f_aligned16(dx) || !f_aligned16(dw)) return false;
    if ((ldg & 3) || (ldh & 3) || (ldw & 3) || (lddx & 3) || (lddw & 3)) return false;
    return true;
}

int tc_linear_bwd_fused(const float* g, int64_t ldg, const float* h, int64_t ldh, const float* w, int64_t ldw, int masked,
                        float* dx, int64_t lddx, float* dw, int64_t lddw, float* db, int64_t M, cudaStream_t st) {
    CUtensorMap tmGk, tmGm, tmH, tmW, tmC;
    int rc = encode_tmap_2d(&tmGk, g, kFDim, (uint64_t)M, (uint64_t)ldg, 128);
    if (rc) return rc;
    rc = encode_tmap_2d(&tmGm, g, kFDim, (uint64_t)M, (uint64_t)ldg, 32, true);
    if (rc) return rc;
    rc = encode_tmap_2d(&tmH, h, kFDim, (uint64_t)M, (uint64_t)ldh, 32, true);
    if (rc) return rc;
    rc = encode_tmap_2d(&tmW, w, kFDim, kFDim, (uint64_t)ldw, 32, true);
    if (rc) return rc;
    rc = encode_tmap_2d(&tmC, dx, kFDim, (uint64_t)M, (uint64_t)lddx, 32);
    if (rc) return rc;
    FusedBwdParams p{};
    p.M = M;
    p.dw = dw; p.lddw = lddw; p.db = db;
    p.masked = masked & 1;
    p.accumulate = (masked >> 1) & 1;
    p.g_abort = abort_flag_ptr();
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(tc_linear_bwd_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kFSmem);
        if (e != cudaSuccess) {
            set_error("tc fused bwd: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
            return NPFK_ERR_CUDA;
        }
        attr_done = true;
    }
    const int64_t tiles = (M + 127) / 128;
    const int gx = (int)(tiles < kNumSMs ? tiles : kNumSMs);
    tc_linear_bwd_fused_kernel<<<gx, kFThreads, kFSmem, st>>>(tmGk, tmGm, tmH, tmW, tmC, p);
    count_launch();
    return check_launch("tc_linear_bwd_fused");
}

}  // namespace tc
}  // namespace npfk
