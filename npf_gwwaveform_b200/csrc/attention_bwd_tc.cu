// tcgen05 / TMEM / TMA backward of the multi-head cross attention (NPFK_MODE_FAST).
//
// Autograd of softmax(Q K^T * scale) V of the reference (npf/architectures/attention.py:129-164, 204-220 with the head
// split/merge of :507-527), flash style: logits are recomputed from q/k and the saved log-sum-exp, never stored.
//
// One CTA owns (batch item, 32-float column box = 32/D heads, tile of 128 keys) and streams the queries in half
// tiles of 64.  The transposed problem is used so that KEYS sit on the TMEM lanes:
//     S^T  = K Q^T          (A = K tile, B = Q half tile, both K-major)           128 x 64 in TMEM
//     dP^T = V dO^T         (A = V tile, B = dO half tile)                        128 x 64 in TMEM
//   thread = key row:  P^T = exp2(S^T*scale*log2e - lse2[q]),  dS^T = P^T * (dP^T - delta[q])
//     P^T and dS^T are written back IN PLACE into TMEM (tcgen05.st) and become the A operands of
//     dV += P^T dO          (A from TMEM, B = dO half tile MN-major)              accumulated over all queries
//     dK += dS^T Q          (A from TMEM, B = Q half tile MN-major)               accumulated over all queries
//     dS^T also goes to shared memory (MN-major, 32-byte-atom swizzle) as the A operand of
//     dQ  = dS K            (M = queries, B = K tile MN-major), reduced into HBM with red.global.add.v4.f32
//                           because the other key tiles of the same task add to the same rows.
// scale is folded into the dQ / dK drains.  K, Q and dO are each needed K-major and MN-major; tf32 MN-major
// operands only exist in the 32-byte-atom swizzle, so TMA loads them twice with two tensor maps (L2 hits).
//   warp 0  TMA producer    warp 1  TMEM allocator + MMA issuer    warps 2.. : NSETS softmax warpgroups (unit u -> warpgroup u mod NSETS)
// Default: two units in flight, four softmax warps per unit, dQ product with M = 128.  Measured alternatives kept behind
// developer switches (C2 shape, same box): NPFK_ATTN_BWD_SETS=3 (three units in flight on a two-stage
// ring: 4.62 vs 3.98 ms), NPFK_ATTN_BWD_WPU=8 (two warps per TMEM lane quarter: 4.32 vs 4.17 ms), NPFK_ATTN_BWD_DQ_M64=1
// (dQ with UMMA M = 64, accumulator rows 16 j.. in lanes 0..15 of quarter j: 4.40 vs 4.19 ms).
#include <math.h>

#include "tc_common.cuh"

namespace npfk {
namespace tc {

constexpr int kBK = 128;                       // keys per CTA (TMEM lanes)
constexpr int kBQ = 64;                        // queries per half tile (UMMA N of S^T)
constexpr int kBStgBytes = kBK * kBQ * 4;      // 32 KB: dS^T staging, 2 blocks of [128 keys x 32 queries]
constexpr int kBStgBlock = kBK * 128;          // 16 KB
constexpr int kBTileQ = kBQ * 128;             // 8 KB: one half tile of a 32-float box
constexpr int kBStageBytes = 4 * kBTileQ + 1024;   // Q(K-major) Q(MN) dO(K-major) dO(MN) + lse2/delta
constexpr int kBTileK = kBK * 128;             // 16 KB
constexpr uint32_t kBTmemCols = 512;

// NSETS units (half tile x head) are in flight at once: one TMEM set (S^T | dP^T, 128 columns), one dS^T staging
// buffer, one dQ accumulator and one softmax warpgroup per set.  Head dim 16 runs three sets on a two-stage Q/dO ring
// (3 x 128 + 64 + 3 x 16 = 496 TMEM columns, 210 KB of shared memory); head dim 32 needs 32-column dQ buffers and
// keeps two sets on a three-stage ring.
// WPU = softmax warps per unit: 4 (a warp owns 32 keys x all 64 queries of the unit) or 8 (two warps share a TMEM lane
// quarter and take 32 queries each, halving the serial softmax time of a unit: the unit chain is latency-bound).
template <int NSETS, int WPU = 4>
struct BwdCfg {
    static constexpr int kStages = NSETS == 3 ? 2 : 3;
    static constexpr int kThreads = 64 + 32 * WPU * NSETS;    // TMA warp, MMA warp, NSETS softmax groups of WPU warps
    static constexpr int kStgOff = 0;
    static constexpr int kStageOff = NSETS * kBStgBytes;
    static constexpr int kKkmOff = kStageOff + kStages * kBStageBytes;
    static constexpr int kKmnOff = kKkmOff + kBTileK;
    static constexpr int kVOff = kKmnOff + kBTileK;
    static constexpr int kBarOff = kVOff + kBTileK;
    static constexpr int kNumBars = 1 + 2 * kStages + 4 * NSETS + 1;
    static constexpr int kSmem = 1024 + kBarOff + kNumBars * 8 + 16;
    static constexpr uint32_t kTmemAcc = NSETS * 128;           // dK / dV accumulators (64 columns)
    static constexpr uint32_t kTmemDq = NSETS * 128 + 64;       // NSETS dQ accumulators of D columns
};

struct AttnBwdParams {
    float* dq;            // [B, Nq, E]  "+=" (zero-filled by the host wrapper)
    float* dk;            // [B, Nk, E]
    float* dv;
    const float* ws;      // lse2_pad [B,H,Nqp] then delta_pad [B,H,Nqp]
    int B, H, E;
    int64_t Nq, Nk, Nqp;
    float scale;
    int k_tiles, half_tiles, nsplit, tiles_per_split;
    int dq_m64;           // dQ product with UMMA M = 64 (1: rows in lanes 0-15 of each quarter, 2: rows in lanes 0-63)
    int dbg;              // timing experiments only (NPFK_ATTN_DBG): 1 no softmax math, 2 no dQ MMAs, 4 no dV/dK MMAs, 8 no dq red, 16 no staging stores
    int* g_abort;
};

// delta[b,h,q] = sum_d dO*O, lse2 = lse * log2(e); both padded to a multiple of 64 queries (+inf / 0) so that the
// main kernel can fetch them with aligned 1-D bulk copies and padded queries get P = 0.
__global__ void attn_bwd_prep_kernel(const float* __restrict__ out, const float* __restrict__ dout,
                                     const float* __restrict__ lse, float* __restrict__ ws, int B, int H, int D,
                                     int64_t Nq, int64_t Nqp, int delta_ready) {
    const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int bh = blockIdx.y;
    if (q >= Nqp) return;
    const int b = bh / H, h = bh % H;
    float delta = 0.f, l2 = INFINITY;
    if (delta_ready) {
        // delta of the real queries was written by the producer of dO (npfk_add_layernorm_bwd's attn_delta output):
        // only the log2-domain lse and the zero padding are left to do, without reading O and dO again
        if (q < Nq) l2 = __ldg(lse + (int64_t)bh * Nq + q) * 1.4426950408889634f;
        ws[(int64_t)bh * Nqp + q] = l2;
        if (q >= Nq) ws[((int64_t)B * H + bh) * Nqp + q] = 0.f;
        return;
    }
    if (q < Nq) {
        const int64_t off = ((int64_t)b * Nq + q) * (H * D) + h * D;
        const float4* o4 = reinterpret_cast<const float4*>(out + off);
        const float4* d4 = reinterpret_cast<const float4*>(dout + off);
        for (int i = 0; i < D / 4; ++i) {
            const float4 a = __ldg(o4 + i), g = __ldg(d4 + i);
            delta = fmaf(a.x, g.x, delta);
            delta = fmaf(a.y, g.y, delta);
            delta = fmaf(a.z, g.z, delta);
            delta = fmaf(a.w, g.w, delta);
        }
        l2 = __ldg(lse + (int64_t)bh * Nq + q) * 1.4426950408889634f;
    }
    ws[(int64_t)bh * Nqp + q] = l2;
    ws[((int64_t)B * H + bh) * Nqp + q] = delta;
}

struct BwdBars {
    uint64_t* kv_full;
    uint64_t* st_full;
    uint64_t* st_empty;
    uint64_t* s_full;
    uint64_t* p_ready;
    uint64_t* dq_full;
    uint64_t* dq_empty;
    uint64_t* fin;
};

__device__ __forceinline__ float ex2_approx_b(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

template <int D, int NSETS, int WPU>
__global__ void __launch_bounds__(BwdCfg<NSETS, WPU>::kThreads, 1)
attn_bwd_tc_kernel(const __grid_constant__ CUtensorMap tmQk, const __grid_constant__ CUtensorMap tmQm,
                   const __grid_constant__ CUtensorMap tmOk, const __grid_constant__ CUtensorMap tmOm,
                   const __grid_constant__ CUtensorMap tmKk, const __grid_constant__ CUtensorMap tmKm,
                   const __grid_constant__ CUtensorMap tmVk, const AttnBwdParams p) {
    constexpr int HB = 32 / D;     // heads per 32-float box
    using Cfg = BwdCfg<NSETS, WPU>;
    constexpr int kBStages = Cfg::kStages, kBStgOff = Cfg::kStgOff, kBStageOff = Cfg::kStageOff, kBKkmOff = Cfg::kKkmOff,
                  kBKmnOff = Cfg::kKmnOff, kBVOff = Cfg::kVOff, kBBarOff = Cfg::kBarOff, kBNumBars = Cfg::kNumBars;
    constexpr uint32_t kBTmemAcc = Cfg::kTmemAcc, kBTmemDq = Cfg::kTmemDq;
    constexpr int NH = D;          // UMMA N of the dV / dK / dQ products: one head
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint64_t* bars = reinterpret_cast<uint64_t*>(base + kBBarOff);
    BwdBars bar;
    bar.kv_full = bars;
    bar.st_full = bars + 1;
    bar.st_empty = bar.st_full + kBStages;
    bar.s_full = bar.st_empty + kBStages;
    bar.p_ready = bar.s_full + NSETS;
    bar.dq_full = bar.s_full + 2 * NSETS;
    bar.dq_empty = bar.s_full + 3 * NSETS;
    bar.fin = bar.s_full + 4 * NSETS;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + kBNumBars);
    volatile int* cta_abort = reinterpret_cast<volatile int*>(tmem_ptr + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int boxes = p.E / 32;
    int idx = blockIdx.x;
    const int split = idx % p.nsplit;  idx /= p.nsplit;
    const int kt = idx % p.k_tiles;    idx /= p.k_tiles;
    const int box = idx % boxes;
    const int b = idx / boxes;
    const int ht_begin = split * p.tiles_per_split;
    const int ht_end = min(p.half_tiles, ht_begin + p.tiles_per_split);
    const int U = (ht_end - ht_begin) * HB;                 // units = (half tile, head)
    constexpr uint32_t head_mn_off = 64u;                   // byte offset of head 1 inside an MN-major row (D = 16)

    if (threadIdx.x == 0) {
        prefetch_tmap(&tmQk); prefetch_tmap(&tmQm); prefetch_tmap(&tmOk); prefetch_tmap(&tmOm);
        prefetch_tmap(&tmKk); prefetch_tmap(&tmKm); prefetch_tmap(&tmVk);
        mbar_init(bar.kv_full, 1);
        for (int i = 0; i < kBStages; ++i) {
            mbar_init(&bar.st_full[i], 1);
            mbar_init(&bar.st_empty[i], 1);
        }
        for (int i = 0; i < NSETS; ++i) {
            mbar_init(&bar.s_full[i], 1);
            mbar_init(&bar.p_ready[i], WPU);
            mbar_init(&bar.dq_full[i], 1);
            mbar_init(&bar.dq_empty[i], p.dq_m64 == 1 ? 4 : 2);
        }
        mbar_init(bar.fin, 1);
        *cta_abort = 0;
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_ptr, kBTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp == 0) {
        if (lane == 0 && U > 0) {   // ---------------------------------------------------------- TMA producer
            mbar_arrive_expect_tx(bar.kv_full, 3u * kBTileK);
            tma_load_3d(&tmKk, bar.kv_full, base + kBKkmOff, box * 32, kt * kBK, b);
            tma_load_3d(&tmKm, bar.kv_full, base + kBKmnOff, box * 32, kt * kBK, b);
            tma_load_3d(&tmVk, bar.kv_full, base + kBVOff, box * 32, kt * kBK, b);
            const int64_t bh0 = (int64_t)b * p.H + box * HB;
            for (int ht = ht_begin; ht < ht_end; ++ht) {
                const uint32_t i = (uint32_t)(ht - ht_begin);
                const uint32_t st = i % kBStages, ph = (i / kBStages) & 1;
                if (!mbar_wait(&bar.st_empty[st], ph ^ 1, cta_abort, p.g_abort, 41)) break;
                uint8_t* s = base + kBStageOff + st * kBStageBytes;
                const bool skip_mn = (p.dbg & 32) != 0;   // timing experiments only
                mbar_arrive_expect_tx(&bar.st_full[st], (skip_mn ? 2u : 4u) * kBTileQ + (uint32_t)HB * 512u);
                tma_load_3d(&tmQk, &bar.st_full[st], s, box * 32, ht * kBQ, b);
                if (!skip_mn) tma_load_3d(&tmQm, &bar.st_full[st], s + kBTileQ, box * 32, ht * kBQ, b);
                tma_load_3d(&tmOk, &bar.st_full[st], s + 2 * kBTileQ, box * 32, ht * kBQ, b);
                if (!skip_mn) tma_load_3d(&tmOm, &bar.st_full[st], s + 3 * kBTileQ, box * 32, ht * kBQ, b);
                for (int hb = 0; hb < HB; ++hb) {
                    const float* l2 = p.ws + (bh0 + hb) * p.Nqp + (int64_t)ht * kBQ;
                    bulk_load_1d(s + 4 * kBTileQ + hb * 512, l2, 256, &bar.st_full[st]);
                    bulk_load_1d(s + 4 * kBTileQ + hb * 512 + 256, l2 + (int64_t)p.B * p.H * p.Nqp, 256, &bar.st_full[st]);
                }
            }
        }
    } else if (warp == 1) {
        if (U > 0 && elect_one()) {   // ------------------------------------------------------- MMA issuer
            const uint32_t idesc_s = make_idesc_tf32(kBQ, false, false);
            const uint32_t idesc_kv = make_idesc_tf32(NH, false, true);    // A from TMEM, B MN-major
            // A, B MN-major; M = 64 reads only the 64 staged query rows (an M = 128 product fetches 64 more rows of A
            // from shared memory that nobody uses: the kernel is shared-memory-bandwidth bound)
            const uint32_t idesc_dq = make_idesc_tf32(NH, true, true, p.dq_m64 ? 64 : 128);
            const uint32_t kkm = smem_u32(base + kBKkmOff), kmn = smem_u32(base + kBKmnOff), vkm = smem_u32(base + kBVOff);
            const uint32_t stg0 = smem_u32(base + kBStgOff), stage0 = smem_u32(base + kBStageOff);
            // base descriptors; per-MMA operands are these plus a byte offset (sdesc_advance)
            const uint64_t d_kkm = make_sdesc_sw128(kkm, 16, 1024), d_vkm = make_sdesc_sw128(vkm, 16, 1024);
            const uint64_t d_kmn = make_sdesc_mn32(kmn, 16384, 512);
            const uint64_t d_stage_k = make_sdesc_sw128(stage0, 16, 1024);      // Q / dO half tiles, K-major
            const uint64_t d_stage_m = make_sdesc_mn32(stage0, 16384, 512);     // Q / dO half tiles, MN-major
            const uint64_t d_stg = make_sdesc_mn32(stg0, kBStgBlock, 512);
            auto issue_s = [&](int u) -> bool {     // S^T and dP^T of unit u
                const int i = u / HB, hb = u % HB;
                const uint32_t st = (uint32_t)i % kBStages;
                if (hb == 0) {
                    if (!mbar_wait(&bar.st_full[st], ((uint32_t)i / kBStages) & 1, cta_abort, p.g_abort, 42)) return false;
                    tc_fence_after();
                }
                const uint32_t koff = (uint32_t)(hb * D * 4);
                const uint32_t tS = tmem_base + (uint32_t)(u % NSETS) * 128u;
                const uint64_t dk = sdesc_advance(d_kkm, koff), dv = sdesc_advance(d_vkm, koff);
                const uint64_t dq = sdesc_advance(d_stage_k, st * kBStageBytes + koff);
                const uint64_t dd = sdesc_advance(dq, 2 * kBTileQ);
#pragma unroll
                for (int s = 0; s < D / 8; ++s)
                    umma_tf32(tS, sdesc_advance(dk, s * 32), sdesc_advance(dq, s * 32), idesc_s, s > 0);
#pragma unroll
                for (int s = 0; s < D / 8; ++s)
                    umma_tf32(tS + 64, sdesc_advance(dv, s * 32), sdesc_advance(dd, s * 32), idesc_s, s > 0);
                umma_commit(&bar.s_full[u % NSETS]);
                return true;
            };
            auto role = [&]() {
                if (!mbar_wait(bar.kv_full, 0, cta_abort, p.g_abort, 43)) return;
                tc_fence_after();
                // S^T / dP^T run NSETS - 1 units ahead of the products that consume P^T / dS^T
                for (int u = 0; u < NSETS - 1 && u < U; ++u)
                    if (!issue_s(u)) return;
                for (int u = 0; u < U; ++u) {
                    if (u + NSETS - 1 < U && !issue_s(u + NSETS - 1)) return;
                    const int i = u / HB, hb = u % HB;
                    const uint32_t st = (uint32_t)i % kBStages, set = (uint32_t)(u % NSETS), ph = ((uint32_t)(u / NSETS)) & 1;
                    if (!mbar_wait(&bar.p_ready[set], ph, cta_abort, p.g_abort, 44)) return;
                    if (!mbar_wait(&bar.dq_empty[set], ph ^ 1, cta_abort, p.g_abort, 45)) return;
                    tc_fence_after();
                    const uint32_t tS = tmem_base + set * 128u;
                    const uint32_t hoff = (uint32_t)hb * head_mn_off;
                    const uint64_t d_do = sdesc_advance(d_stage_m, st * kBStageBytes + 3 * kBTileQ + hoff);
                    const uint64_t d_q = sdesc_advance(d_stage_m, st * kBStageBytes + kBTileQ + hoff);
                    const uint64_t d_ds = sdesc_advance(d_stg, set * kBStgBytes);
                    const uint64_t d_k = sdesc_advance(d_kmn, hoff);
                    const uint32_t acc = tmem_base + kBTmemAcc + (uint32_t)hb * 2u * (uint32_t)NH;
                    const uint32_t first = (i == 0) ? 0u : 1u;
                    if (!(p.dbg & 4))
#pragma unroll
                    for (int j = 0; j < kBQ / 8; ++j)      // dV += P^T dO
                        umma_tf32_ts(acc + NH, tS + 8 * j, sdesc_advance(d_do, j * 1024),
                                     idesc_kv, first | (uint32_t)(j > 0));
                    if (!(p.dbg & 4))
#pragma unroll
                    for (int j = 0; j < kBQ / 8; ++j)      // dK += dS^T Q
                        umma_tf32_ts(acc, tS + 64 + 8 * j, sdesc_advance(d_q, j * 1024),
                                     idesc_kv, first | (uint32_t)(j > 0));
                    if (!(p.dbg & 2))
#pragma unroll
                    for (int j = 0; j < kBK / 8; ++j)      // dQ = dS K   (rows 64..127 of the accumulator are unused)
                        umma_tf32(tmem_base + kBTmemDq + set * (uint32_t)D, sdesc_advance(d_ds, j * 1024), sdesc_advance(d_k, j * 1024),
                                  idesc_dq, j > 0);
                    umma_commit(&bar.dq_full[set]);
                    if (hb == HB - 1) umma_commit(&bar.st_empty[st]);
                }
                umma_commit(bar.fin);
            };
            role();
        }
    } else {
        // ------------------------------------------------------------------------------ softmax / drain warps
        const int qd = warp & 3;                       // TMEM lane quarter this warp may access
        const int wg = (warp - 2) / WPU;               // softmax group: units u = wg mod NSETS (= TMEM set, staging buffer, dQ buffer)
        const int half = ((warp - 2) % WPU) >> 2;      // WPU == 8: which 32 queries of the unit this warp takes
        constexpr int NC = 8 / WPU;                    // 32-query chunks per warp
        const int kk = qd * 32 + lane;                 // key row inside the tile
        const uint32_t lane_base = tmem_base + ((uint32_t)(qd * 32) << 16);
        const float k2 = p.scale * 1.4426950408889634f;
        const bool key_ok = (int64_t)kt * kBK + kk < p.Nk;
        const uint32_t sw = (uint32_t)(kk & 3);
        const uint32_t hsel = (uint32_t)((kk >> 2) & 1);   // alternate the 16-byte halves: conflict-free staging stores
        auto wait_all = [&](uint64_t* bb, uint32_t ph, int code) -> bool {
            bool ok = mbar_wait(bb, ph, cta_abort, p.g_abort, code);
            return __all_sync(0xffffffffu, ok);
        };
        // dQ of unit u: rows 0..63 of the accumulator = the 64 queries of the half tile (lane quarters 0 and 1)
        auto drain_dq = [&](int u) -> bool {
            const uint32_t set = (uint32_t)(u % NSETS);
            if (!wait_all(&bar.dq_full[set], ((uint32_t)(u / NSETS)) & 1, 51)) return false;
            // accumulator rows -> TMEM lanes: M = 128 / M = 64 "flat": query row r in lane r (quarters 0 and 1);
            // M = 64 datapath layout: rows 16 j .. 16 j + 15 in lanes 0..15 of quarter j (all four quarters take part)
            const bool quarters = p.dq_m64 == 1;
            if ((quarters || qd < 2) && half == WPU / 4 - 1) {
                tc_fence_after();
                const int i = u / HB, hb = u % HB;
                uint32_t r[D];
                const uint32_t ta = lane_base + kBTmemDq + set * (uint32_t)D;
                if constexpr (D == 16) tmem_ld_32x16(ta, r);
                else tmem_ld_32x32(ta, r);
                tmem_ld_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&bar.dq_empty[set]);
                const int64_t q = (int64_t)(ht_begin + i) * kBQ + (quarters ? qd * 16 + lane : kk);
                if (q < p.Nq && !(p.dbg & 8) && (!quarters || lane < 16)) {
                    float* dst = p.dq + ((int64_t)b * p.Nq + q) * p.E + box * 32 + hb * D;
#pragma unroll
                    for (int d4 = 0; d4 < D / 4; ++d4)
                        red_add_v4(dst + 4 * d4, __uint_as_float(r[4 * d4]) * p.scale, __uint_as_float(r[4 * d4 + 1]) * p.scale,
                                   __uint_as_float(r[4 * d4 + 2]) * p.scale, __uint_as_float(r[4 * d4 + 3]) * p.scale);
                }
            }
            return true;
        };
        auto role = [&]() {
            int last = -1;
            for (int u = wg; u < U; u += NSETS) {
                const int i = u / HB, hb = u % HB;
                const uint32_t set = (uint32_t)wg, st = (uint32_t)i % kBStages;
                if (u >= NSETS && !drain_dq(u - NSETS)) return;          // also frees staging[set]
                last = u;
                if (!wait_all(&bar.st_full[st], ((uint32_t)i / kBStages) & 1, 52)) return;
                if (!wait_all(&bar.s_full[set], ((uint32_t)(u / NSETS)) & 1, 53)) return;
                tc_fence_after();
                const uint8_t* stats = base + kBStageOff + st * kBStageBytes + 4 * kBTileQ + hb * 512;
                uint8_t* stg = base + kBStgOff + set * kBStgBytes + kk * 128;
                const uint32_t tS = lane_base + set * 128u;
#pragma unroll
                for (int c = half; c < ((p.dbg & 1) ? 0 : half + NC); ++c) {          // 32 queries at a time
                    uint32_t s[32], g[32];
                    tmem_ld_32x32(tS + c * 32, s);
                    tmem_ld_32x32(tS + 64 + c * 32, g);
                    tmem_ld_wait();
#pragma unroll
                    for (int j4 = 0; j4 < 8; ++j4) {
                        const float4 l2 = *reinterpret_cast<const float4*>(stats + (c * 32 + j4 * 4) * 4);
                        const float4 dl = *reinterpret_cast<const float4*>(stats + 256 + (c * 32 + j4 * 4) * 4);
                        const float l2a[4] = {l2.x, l2.y, l2.z, l2.w}, dla[4] = {dl.x, dl.y, dl.z, dl.w};
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const int j = j4 * 4 + e;
                            float pv = ex2_approx_b(fmaf(__uint_as_float(s[j]), k2, -l2a[e]));
                            pv = key_ok ? pv : 0.f;
                            const float ds = pv * (__uint_as_float(g[j]) - dla[e]);
                            s[j] = __float_as_uint(pv);
                            g[j] = __float_as_uint(ds);
                        }
                    }
                    tmem_st_32x32(tS + c * 32, s);
                    tmem_st_32x32(tS + 64 + c * 32, g);
                    // dS^T row of this key into the MN-major staging block c: 32-byte chunk ch at (ch ^ (kk & 3))
                    uint8_t* blk = stg + c * kBStgBlock;
                    if (!(p.dbg & 16))
#pragma unroll
                    for (int ch = 0; ch < 4; ++ch) {
                        // both 16-byte halves of chunk ch; lanes with hsel store the upper half first (conflict-free),
                        // chosen with selects so that g[] keeps static register indices
                        const float4 lo = make_float4(__uint_as_float(g[8 * ch]), __uint_as_float(g[8 * ch + 1]),
                                                      __uint_as_float(g[8 * ch + 2]), __uint_as_float(g[8 * ch + 3]));
                        const float4 hi = make_float4(__uint_as_float(g[8 * ch + 4]), __uint_as_float(g[8 * ch + 5]),
                                                      __uint_as_float(g[8 * ch + 6]), __uint_as_float(g[8 * ch + 7]));
                        uint8_t* dst = blk + (((uint32_t)ch ^ sw) << 5);
                        *reinterpret_cast<float4*>(dst + (hsel << 4)) = hsel ? hi : lo;
                        *reinterpret_cast<float4*>(dst + ((hsel ^ 1u) << 4)) = hsel ? lo : hi;
                    }
                }
                tmem_st_wait();
                fence_proxy_async();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&bar.p_ready[set]);
            }
            if (last >= 0 && !drain_dq(last)) return;
            // ---- dK / dV of this key tile
            if (!wait_all(bar.fin, 0, 54)) return;
            tc_fence_after();
            const int64_t key = (int64_t)kt * kBK + kk;
#pragma unroll
            for (int hb = wg; hb < HB; hb += NSETS) {       // warpgroup g drains head g of the box
                uint32_t rk[D], rv[D];
                const uint32_t acc = lane_base + kBTmemAcc + (uint32_t)hb * 2u * (uint32_t)NH;
                const bool do_k = WPU == 4 || half == 0, do_v = WPU == 4 || half == 1;   // two warps per lane quarter: one each
                if constexpr (D == 16) {
                    tmem_ld_32x16(acc, rk);
                    tmem_ld_32x16(acc + NH, rv);
                } else {
                    tmem_ld_32x32(acc, rk);
                    tmem_ld_32x32(acc + NH, rv);
                }
                tmem_ld_wait();
                if (key < p.Nk) {
                    const int64_t off = ((int64_t)b * p.Nk + key) * p.E + box * 32 + hb * D;
#pragma unroll
                    for (int d4 = 0; d4 < D / 4; ++d4) {
                        const float4 a = make_float4(__uint_as_float(rk[4 * d4]) * p.scale, __uint_as_float(rk[4 * d4 + 1]) * p.scale,
                                                     __uint_as_float(rk[4 * d4 + 2]) * p.scale, __uint_as_float(rk[4 * d4 + 3]) * p.scale);
                        const float4 c = make_float4(__uint_as_float(rv[4 * d4]), __uint_as_float(rv[4 * d4 + 1]),
                                                     __uint_as_float(rv[4 * d4 + 2]), __uint_as_float(rv[4 * d4 + 3]));
                        if (p.nsplit == 1) {
                            if (do_k) *reinterpret_cast<float4*>(p.dk + off + 4 * d4) = a;
                            if (do_v) *reinterpret_cast<float4*>(p.dv + off + 4 * d4) = c;
                        } else {
                            if (do_k) red_add_v4(p.dk + off + 4 * d4, a.x, a.y, a.z, a.w);
                            if (do_v) red_add_v4(p.dv + off + 4 * d4, c.x, c.y, c.z, c.w);
                        }
                    }
                }
            }
            tc_fence_before();
        };
        if (U > 0) role();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kBTmemCols);
    }
}

int64_t attn_bwd_ws_floats(int B, int64_t Nq, int H) {
    const int64_t nqp = (Nq + kBQ - 1) / kBQ * kBQ;
    return 2 * (int64_t)B * H * nqp;
}

int attn_bwd_tc(const float* q, const float* k, const float* v, const float* out, const float* lse, const float* dout,
                float* dq, float* dk, float* dv, float* ws, int B, int64_t Nq, int64_t Nk, int H, int D, float scale,
                int delta_ready, cudaStream_t st) {
    const int E = H * D;
    if ((reinterpret_cast<uintptr_t>(ws) & 15) || (reinterpret_cast<uintptr_t>(dq) & 15) ||
        (reinterpret_cast<uintptr_t>(dk) & 15) || (reinterpret_cast<uintptr_t>(dv) & 15) ||
        (reinterpret_cast<uintptr_t>(dout) & 15)) {
        set_error("attn_bwd_tc: pointers must be 16-byte aligned");
        return NPFK_ERR_INVALID;
    }
    CUtensorMap tmQk, tmQm, tmOk, tmOm, tmKk, tmKm, tmVk;
    int rc = encode_tmap_3d(&tmQk, q, E, Nq, B, E, (uint64_t)Nq * E, kBQ, false);
    if (!rc) rc = encode_tmap_3d(&tmQm, q, E, Nq, B, E, (uint64_t)Nq * E, kBQ, true);
    if (!rc) rc = encode_tmap_3d(&tmOk, dout, E, Nq, B, E, (uint64_t)Nq * E, kBQ, false);
    if (!rc) rc = encode_tmap_3d(&tmOm, dout, E, Nq, B, E, (uint64_t)Nq * E, kBQ, true);
    if (!rc) rc = encode_tmap_3d(&tmKk, k, E, Nk, B, E, (uint64_t)Nk * E, kBK, false);
    if (!rc) rc = encode_tmap_3d(&tmKm, k, E, Nk, B, E, (uint64_t)Nk * E, kBK, true);
    if (!rc) rc = encode_tmap_3d(&tmVk, v, E, Nk, B, E, (uint64_t)Nk * E, kBK, false);
    if (rc) return rc;

    AttnBwdParams p{};
    p.dq = dq; p.dk = dk; p.dv = dv; p.ws = ws;
    p.B = B; p.H = H; p.E = E;
    p.Nq = Nq; p.Nk = Nk;
    p.Nqp = (Nq + kBQ - 1) / kBQ * kBQ;
    p.scale = scale;
    p.k_tiles = (int)((Nk + kBK - 1) / kBK);
    p.half_tiles = (int)(p.Nqp / kBQ);
    const int64_t items = (int64_t)B * (E / 32) * p.k_tiles;
    int64_t ns = (2 * kNumSMs + items - 1) / items;
    if (ns < 1) ns = 1;
    if (ns > p.half_tiles) ns = p.half_tiles;
    p.tiles_per_split = (int)((p.half_tiles + ns - 1) / ns);
    p.nsplit = (p.half_tiles + p.tiles_per_split - 1) / p.tiles_per_split;
    static int dbg_env = -1;
    if (dbg_env < 0) {
        const char* e = getenv("NPFK_ATTN_DBG");
        dbg_env = e ? atoi(e) : 0;
    }
    p.dbg = dbg_env;
    static int m64_env = -1;
    if (m64_env < 0) {
        const char* e = getenv("NPFK_ATTN_BWD_DQ_M64");
        m64_env = e ? atoi(e) : 0;
    }
    p.dq_m64 = m64_env;

    p.g_abort = abort_flag_ptr();

    // prep: delta and lse*log2e into the padded workspace; dq (and dk/dv when the queries are split) are "+="
    {
        dim3 grid((unsigned)((p.Nqp + 255) / 256), (unsigned)(B * H));
        attn_bwd_prep_kernel<<<grid, 256, 0, st>>>(out, dout, lse, ws, B, H, D, Nq, p.Nqp, delta_ready);
        count_launch();
    }
    cudaMemsetAsync(dq, 0, sizeof(float) * (size_t)B * Nq * E, st);
    if (p.nsplit > 1) {
        cudaMemsetAsync(dk, 0, sizeof(float) * (size_t)B * Nk * E, st);
        cudaMemsetAsync(dv, 0, sizeof(float) * (size_t)B * Nk * E, st);
    }
    static int sets_env = -1, wpu_env = -1;
    if (sets_env < 0) {
        const char* e = getenv("NPFK_ATTN_BWD_SETS");
        sets_env = (e && atoi(e) == 3) ? 3 : 2;
        e = getenv("NPFK_ATTN_BWD_WPU");
        wpu_env = (e && atoi(e) == 8) ? 8 : 4;
    }
    const int64_t grid = items * p.nsplit;
    if (grid > 2147483647LL) {
        set_error("attn_bwd_tc: grid too large");
        return NPFK_ERR_INVALID;
    }
    auto launch = [&](auto kernel, int threads, int smem, bool& attr_done) -> int {
        if (!attr_done) {
            cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
            if (e != cudaSuccess) {
                set_error("attn_bwd_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
                return NPFK_ERR_CUDA;
            }
            attr_done = true;
        }
        kernel<<<(unsigned)grid, threads, smem, st>>>(tmQk, tmQm, tmOk, tmOm, tmKk, tmKm, tmVk, p);
        return NPFK_OK;
    };
    static bool done[5] = {false, false, false, false, false};
    if (D == 16 && sets_env == 3) rc = launch(attn_bwd_tc_kernel<16, 3, 4>, BwdCfg<3, 4>::kThreads, BwdCfg<3, 4>::kSmem, done[0]);
    else if (D == 16 && wpu_env == 4) rc = launch(attn_bwd_tc_kernel<16, 2, 4>, BwdCfg<2, 4>::kThreads, BwdCfg<2, 4>::kSmem, done[1]);
    else if (D == 16) rc = launch(attn_bwd_tc_kernel<16, 2, 8>, BwdCfg<2, 8>::kThreads, BwdCfg<2, 8>::kSmem, done[2]);
    else if (wpu_env == 4) rc = launch(attn_bwd_tc_kernel<32, 2, 4>, BwdCfg<2, 4>::kThreads, BwdCfg<2, 4>::kSmem, done[3]);
    else rc = launch(attn_bwd_tc_kernel<32, 2, 8>, BwdCfg<2, 8>::kThreads, BwdCfg<2, 8>::kSmem, done[4]);
    if (rc) return rc;
    count_launch();
    return check_launch("attn_bwd_tc");
}

}  // namespace tc
}  // namespace npfk
