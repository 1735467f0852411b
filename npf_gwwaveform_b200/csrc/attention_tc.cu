// tcgen05 / TMEM / TMA multi-head cross attention (targets -> context) for NPFK_MODE_FAST.
//
// Replaces _make_multiheaded / DotAttender.score / softmax / bmm / _concatenate_multiheads of the reference
// (npf/architectures/attention.py:507-527, 204-220, 129-164).  Heads are slices of the feature dimension and are
// read in place by TMA: one 32-float (128-byte) column box of q/k/v holds 32/D heads; a head is addressed by a
// byte offset inside the swizzled row, so no permute copies and no [B*H, Nq, Nk] logits ever reach HBM.
//
// One CTA owns (batch item, column box) and walks query tiles of 128 rows.  K/V of the box stay resident in
// shared memory when the context fits the ring (<= 512 points), otherwise they are streamed per pass.
//   warp 0      TMA producer (Q tiles, K/V chunks of 64 keys)
//   warp 1      TMEM allocator + tcgen05.mma issuer (one thread)
//   warps 2..5  softmax: thread = query row (TMEM lane), tcgen05.ld of the logits, exp2, P to swizzled smem
// Forward per (query tile, head):  pass 1  S = Q K^T per chunk -> running row maximum (no exponentials);
//                                  pass 2  S again -> P = exp2((S - max) * scale*log2e) -> O += P V in TMEM.
// Recomputing S (2 tiny MMAs per chunk) is cheaper than rescaling O in TMEM when the maximum moves.
#include <math.h>

#include "tc_common.cuh"

namespace npfk {
namespace tc {

constexpr int kAQT = 128;                    // queries per tile (UMMA M, TMEM lanes)
constexpr int kAKC = 64;                     // keys per chunk (UMMA N of S, K of PV)
constexpr int kARing = 8;                    // K/V ring slots
constexpr int kAThreads = 192;
constexpr int kAKBytes = kAKC * 128;         // one chunk of a 32-float box
constexpr int kASlotBytes = 2 * kAKBytes;    // K chunk + V chunk
constexpr int kAQBytes = kAQT * 128;
constexpr int kAPBytes = kAQT * kAKC * 4;
constexpr int kARingOff = 0;
constexpr int kAQOff = kARing * kASlotBytes;             // 131072
constexpr int kAPOff = kAQOff + 2 * kAQBytes;            // 163840
constexpr int kABarOff = kAPOff + 2 * kAPBytes;          // 229376
constexpr int kANumBars = 2 * kARing + 20;
constexpr int kAFwdSmem = 1024 + kABarOff + kANumBars * 8 + 16;
constexpr uint32_t kATmemCols = 256;         // S: 2 x 64 columns, O: 2 x 32 columns
constexpr uint32_t kATmemO = 128;

struct AttnParams {
    float* out;          // [B, Nq, E]
    float* lse;          // [B, H, Nq]
    int B, H, E;
    int64_t Nq, Nk;
    float scale;
    int n_chunks, q_tiles, nsplit, tiles_per_split;
    int* g_abort;
};

struct AttnBars {
    uint64_t* kv_full;
    uint64_t* kv_empty;
    uint64_t* q_full;
    uint64_t* q_empty;
    uint64_t* s_full;
    uint64_t* s_empty;
    uint64_t* p_full;
    uint64_t* p_empty;
    uint64_t* o_full;
    uint64_t* o_empty;
};

__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

template <int D>
__global__ void __launch_bounds__(kAThreads, 1)
attn_fwd_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                   const __grid_constant__ CUtensorMap tmV, const AttnParams p) {
    constexpr int HB = 32 / D;     // heads per 32-float box
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint64_t* bars = reinterpret_cast<uint64_t*>(base + kABarOff);
    AttnBars bar;
    bar.kv_full = bars;
    bar.kv_empty = bars + kARing;
    bar.q_full = bars + 2 * kARing;
    bar.q_empty = bar.q_full + 2;
    bar.s_full = bar.q_full + 4;
    bar.s_empty = bar.q_full + 6;
    bar.p_full = bar.q_full + 8;
    bar.p_empty = bar.q_full + 10;
    bar.o_full = bar.q_full + 12;
    bar.o_empty = bar.q_full + 14;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + kANumBars);
    volatile int* cta_abort = reinterpret_cast<volatile int*>(tmem_ptr + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int boxes = p.E / 32;
    const int item = blockIdx.x / p.nsplit, split = blockIdx.x % p.nsplit;
    const int b = item / boxes, box = item % boxes;
    const int qt_begin = split * p.tiles_per_split;
    const int qt_end = min(p.q_tiles, qt_begin + p.tiles_per_split);
    const int nc = p.n_chunks;
    const bool resident = nc <= kARing;

    if (threadIdx.x == 0) {
        prefetch_tmap(&tmQ);
        prefetch_tmap(&tmK);
        prefetch_tmap(&tmV);
        for (int i = 0; i < kARing; ++i) {
            mbar_init(&bar.kv_full[i], 1);
            mbar_init(&bar.kv_empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&bar.q_full[i], 1);
            mbar_init(&bar.q_empty[i], 1);
            mbar_init(&bar.s_full[i], 1);
            mbar_init(&bar.s_empty[i], 4);
            mbar_init(&bar.p_full[i], 4);
            mbar_init(&bar.p_empty[i], 1);
            mbar_init(&bar.o_full[i], 1);
            mbar_init(&bar.o_empty[i], 4);
        }
        *cta_abort = 0;
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_ptr, kATmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp == 0) {
        if (lane == 0 && qt_begin < qt_end) {   // ------------------------------------------------ TMA producer
            auto load_kv = [&](int slot, int c) {
                uint8_t* sK = base + kARingOff + slot * kASlotBytes;
                mbar_arrive_expect_tx(&bar.kv_full[slot], (uint32_t)kASlotBytes);
                tma_load_3d(&tmK, &bar.kv_full[slot], sK, box * 32, c * kAKC, b);
                tma_load_3d(&tmV, &bar.kv_full[slot], sK + kAKBytes, box * 32, c * kAKC, b);
            };
            if (resident)
                for (int c = 0; c < nc; ++c) load_kv(c, c);
            uint32_t kv_it = 0;
            bool alive = true;
            for (int qt = qt_begin; qt < qt_end && alive; ++qt) {
                const uint32_t qi = (uint32_t)(qt - qt_begin);
                const uint32_t qb = qi & 1, qph = (qi >> 1) & 1;
                if (!mbar_wait(&bar.q_empty[qb], qph ^ 1, cta_abort, p.g_abort, 21)) break;
                mbar_arrive_expect_tx(&bar.q_full[qb], (uint32_t)kAQBytes);
                tma_load_3d(&tmQ, &bar.q_full[qb], base + kAQOff + qb * kAQBytes, box * 32, qt * kAQT, b);
                if (!resident) {
                    for (int rep = 0; rep < 2 * HB && alive; ++rep)
                        for (int c = 0; c < nc; ++c, ++kv_it) {
                            const uint32_t slot = kv_it % kARing, ph = (kv_it / kARing) & 1;
                            if (!mbar_wait(&bar.kv_empty[slot], ph ^ 1, cta_abort, p.g_abort, 22)) { alive = false; break; }
                            load_kv((int)slot, c);
                        }
                }
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {   // ------------------------------------------------------------------ MMA issuer
            const uint32_t idesc_s = make_idesc_tf32(kAKC, false, false);
            const uint32_t idesc_pv = make_idesc_tf32(32, false, true);
            const uint32_t ring_addr = smem_u32(base + kARingOff);
            const uint32_t p_addr0 = smem_u32(base + kAPOff);
            // base descriptors: per-MMA operands are these plus a byte offset (one add each)
            const uint64_t d_ring_k = make_sdesc_sw128(ring_addr, 16, 1024);
            const uint64_t d_ring_v = make_sdesc_mn32(ring_addr + kAKBytes, 4096, 512);
            const uint64_t d_q0 = make_sdesc_sw128(smem_u32(base + kAQOff), 16, 1024);
            const uint64_t d_p0 = make_sdesc_sw128(p_addr0, 16, 1024);
            uint32_t s_it = 0, p_it = 0, o_it = 0, kv_it = 0;
            auto role = [&]() {
                for (int qt = qt_begin; qt < qt_end; ++qt) {
                    const uint32_t qi = (uint32_t)(qt - qt_begin);
                    const uint32_t qb = qi & 1, qph = (qi >> 1) & 1;
                    if (!mbar_wait(&bar.q_full[qb], qph, cta_abort, p.g_abort, 23)) return;
                    tc_fence_after();
                    const uint64_t d_q = sdesc_advance(d_q0, qb * kAQBytes);
                    for (int hb = 0; hb < HB; ++hb) {
                        const uint32_t koff = (uint32_t)(hb * D * 4);
                        // S = Q_h K_h^T for chunk c into S buffer (s_it & 1); returns the ring slot used
                        auto issue_s = [&](int c, uint32_t kvi) -> bool {
                            const uint32_t slot = resident ? (uint32_t)c : kvi % kARing;
                            const uint32_t ph = resident ? 0u : (kvi / kARing) & 1;
                            if (!mbar_wait(&bar.kv_full[slot], ph, cta_abort, p.g_abort, 24)) return false;
                            const uint32_t sb = s_it & 1;
                            if (!mbar_wait(&bar.s_empty[sb], ((s_it >> 1) & 1) ^ 1, cta_abort, p.g_abort, 25)) return false;
                            tc_fence_after();
                            const uint64_t dq = sdesc_advance(d_q, koff), dk = sdesc_advance(d_ring_k, slot * kASlotBytes + koff);
#pragma unroll
                            for (int s = 0; s < D / 8; ++s)
                                umma_tf32(tmem_base + sb * kAKC, sdesc_advance(dq, s * 32), sdesc_advance(dk, s * 32), idesc_s, s > 0);
                            umma_commit(&bar.s_full[sb]);
                            ++s_it;
                            return true;
                        };
                        // pass 1: row maxima
                        for (int c = 0; c < nc; ++c) {
                            if (!issue_s(c, kv_it)) return;
                            if (!resident) {
                                umma_commit(&bar.kv_empty[kv_it % kARing]);
                                ++kv_it;
                            }
                        }
                        // pass 2: P V
                        const uint32_t ob = o_it & 1;
                        if (!mbar_wait(&bar.o_empty[ob], ((o_it >> 1) & 1) ^ 1, cta_abort, p.g_abort, 26)) return;
                        tc_fence_after();
                        if (!issue_s(0, kv_it)) return;
                        for (int c = 0; c < nc; ++c) {
                            if (c + 1 < nc && !issue_s(c + 1, kv_it + 1)) return;
                            const uint32_t pb = p_it & 1;
                            if (!mbar_wait(&bar.p_full[pb], (p_it >> 1) & 1, cta_abort, p.g_abort, 27)) return;
                            tc_fence_after();
                            const uint32_t slot = resident ? (uint32_t)c : kv_it % kARing;
                            const uint64_t dv = sdesc_advance(d_ring_v, slot * kASlotBytes);
                            const uint64_t dp = sdesc_advance(d_p0, pb * kAPBytes);
#pragma unroll
                            for (int j = 0; j < kAKC / 8; ++j)
                                umma_tf32(tmem_base + kATmemO + ob * 32,
                                          sdesc_advance(dp, (j >> 2) * (kAQT * 128) + (j & 3) * 32),
                                          sdesc_advance(dv, j * 1024), idesc_pv, (c | j) != 0);
                            umma_commit(&bar.p_empty[pb]);
                            ++p_it;
                            if (!resident) {
                                umma_commit(&bar.kv_empty[slot]);
                                ++kv_it;
                            }
                        }
                        umma_commit(&bar.o_full[ob]);
                        ++o_it;
                    }
                    umma_commit(&bar.q_empty[qb]);
                }
            };
            role();
        }
    } else {   // ---------------------------------------------------------------------------- softmax warps
        const int qd = warp & 3;
        const int row = qd * 32 + lane;
        const uint32_t lane_base = tmem_base + ((uint32_t)(qd * 32) << 16);
        const float k2 = p.scale * 1.4426950408889634f;
        uint32_t s_it = 0, p_it = 0, o_it = 0;
        auto wait_all = [&](uint64_t* bb, uint32_t ph, int code) -> bool {
            bool ok = mbar_wait(bb, ph, cta_abort, p.g_abort, code);
            return __all_sync(0xffffffffu, ok);
        };
        auto role = [&]() {
            for (int qt = qt_begin; qt < qt_end; ++qt) {
                const int64_t qrow = (int64_t)qt * kAQT + row;
                for (int hb = 0; hb < HB; ++hb) {
                    // ---- pass 1: running maximum of the raw logits
                    float m = -INFINITY;
                    for (int c = 0; c < nc; ++c) {
                        const uint32_t sb = s_it & 1;
                        if (!wait_all(&bar.s_full[sb], (s_it >> 1) & 1, 31)) return;
                        tc_fence_after();
                        const int64_t key0 = (int64_t)c * kAKC;
#pragma unroll
                        for (int half = 0; half < 2; ++half) {
                            uint32_t v[32];
                            tmem_ld_32x32(lane_base + sb * kAKC + half * 32, v);
                            tmem_ld_wait();
                            if (key0 + half * 32 + 32 <= p.Nk) {
#pragma unroll
                                for (int j = 0; j < 32; ++j) m = fmaxf(m, __uint_as_float(v[j]));
                            } else {
#pragma unroll
                                for (int j = 0; j < 32; ++j)
                                    if (key0 + half * 32 + j < p.Nk) m = fmaxf(m, __uint_as_float(v[j]));
                            }
                        }
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(&bar.s_empty[sb]);
                        ++s_it;
                    }
                    // ---- pass 2: P = exp2((S - m) * scale*log2e), row sums, P -> smem (K-major, 128-byte swizzle)
                    const float mk = m * k2;
                    float l = 0.f;
                    for (int c = 0; c < nc; ++c) {
                        const uint32_t sb = s_it & 1;
                        if (!wait_all(&bar.s_full[sb], (s_it >> 1) & 1, 32)) return;
                        tc_fence_after();
                        uint32_t v0[32], v1[32];
                        tmem_ld_32x32(lane_base + sb * kAKC, v0);
                        tmem_ld_32x32(lane_base + sb * kAKC + 32, v1);
                        tmem_ld_wait();
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(&bar.s_empty[sb]);
                        ++s_it;
                        const int64_t key0 = (int64_t)c * kAKC;
                        float pr[64];
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            pr[j] = ex2_approx(fmaf(__uint_as_float(v0[j]), k2, -mk));
                            pr[32 + j] = ex2_approx(fmaf(__uint_as_float(v1[j]), k2, -mk));
                        }
                        if (key0 + kAKC > p.Nk) {
#pragma unroll
                            for (int j = 0; j < 64; ++j)
                                if (key0 + j >= p.Nk) pr[j] = 0.f;
                        }
#pragma unroll
                        for (int j = 0; j < 64; ++j) l += pr[j];
                        const uint32_t pb = p_it & 1;
                        if (!wait_all(&bar.p_empty[pb], ((p_it >> 1) & 1) ^ 1, 33)) return;
                        uint8_t* pbuf = base + kAPOff + pb * kAPBytes + row * 128;
#pragma unroll
                        for (int ch = 0; ch < 16; ++ch) {
                            const float4 o = make_float4(pr[4 * ch], pr[4 * ch + 1], pr[4 * ch + 2], pr[4 * ch + 3]);
                            *reinterpret_cast<float4*>(pbuf + (ch >> 3) * (kAQT * 128) + (((ch & 7) ^ (row & 7)) << 4)) = o;
                        }
                        fence_proxy_async();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(&bar.p_full[pb]);
                        ++p_it;
                    }
                    // ---- epilogue of this head: O / l
                    const uint32_t ob = o_it & 1;
                    if (!wait_all(&bar.o_full[ob], (o_it >> 1) & 1, 34)) return;
                    tc_fence_after();
                    uint32_t o[D];
                    if constexpr (D == 16) tmem_ld_32x16(lane_base + kATmemO + ob * 32 + hb * D, o);
                    else tmem_ld_32x32(lane_base + kATmemO + ob * 32, o);
                    tmem_ld_wait();
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&bar.o_empty[ob]);
                    ++o_it;
                    if (qrow < p.Nq) {
                        const float inv = 1.f / l;
                        const int h = box * HB + hb;
                        float* op = p.out + ((int64_t)b * p.Nq + qrow) * p.E + h * D;
#pragma unroll
                        for (int d4 = 0; d4 < D / 4; ++d4) {
                            float4 r;
                            r.x = __uint_as_float(o[4 * d4]) * inv;
                            r.y = __uint_as_float(o[4 * d4 + 1]) * inv;
                            r.z = __uint_as_float(o[4 * d4 + 2]) * inv;
                            r.w = __uint_as_float(o[4 * d4 + 3]) * inv;
                            reinterpret_cast<float4*>(op)[d4] = r;
                        }
                        p.lse[((int64_t)b * p.H + h) * p.Nq + qrow] = m * p.scale + logf(l);
                    }
                }
            }
        };
        role();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kATmemCols);
    }
}

bool attn_tc_eligible(const float* q, const float* k, const float* v, const float* out, int B, int64_t Nq, int64_t Nk,
                      int H, int D) {
    if (D != 16 && D != 32) return false;
    if ((H * D) % 32 != 0) return false;
    if (Nq < 1 || Nk < 1 || Nq > 2147483647LL || Nk > 2147483647LL) return false;
    auto al = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
    return al(q) && al(k) && al(v) && al(out);
}

static void attn_split(AttnParams& p, int B) {
    const int items = B * (p.E / 32);
    int ns = (2 * kNumSMs + items - 1) / items;
    if (ns < 1) ns = 1;
    if (ns > p.q_tiles) ns = p.q_tiles;
    p.tiles_per_split = (p.q_tiles + ns - 1) / ns;
    p.nsplit = (p.q_tiles + p.tiles_per_split - 1) / p.tiles_per_split;
}

int attn_fwd_tc(const float* q, const float* k, const float* v, float* out, float* lse, int B, int64_t Nq, int64_t Nk,
                int H, int D, float scale, cudaStream_t st) {
    const int E = H * D;
    CUtensorMap tmQ, tmK, tmV;
    int rc = encode_tmap_3d(&tmQ, q, E, Nq, B, E, (uint64_t)Nq * E, kAQT, false);
    if (rc) return rc;
    rc = encode_tmap_3d(&tmK, k, E, Nk, B, E, (uint64_t)Nk * E, kAKC, false);
    if (rc) return rc;
    rc = encode_tmap_3d(&tmV, v, E, Nk, B, E, (uint64_t)Nk * E, kAKC, true);
    if (rc) return rc;
    AttnParams p{};
    p.out = out; p.lse = lse;
    p.B = B; p.H = H; p.E = E;
    p.Nq = Nq; p.Nk = Nk;
    p.scale = scale;
    p.n_chunks = (int)((Nk + kAKC - 1) / kAKC);
    p.q_tiles = (int)((Nq + kAQT - 1) / kAQT);
    attn_split(p, B);
    p.g_abort = abort_flag_ptr();
    static bool attr_done[2] = {false, false};
    const int di = D == 16 ? 0 : 1;
    if (!attr_done[di]) {
        cudaError_t e = D == 16 ? cudaFuncSetAttribute(attn_fwd_tc_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, kAFwdSmem)
                                : cudaFuncSetAttribute(attn_fwd_tc_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, kAFwdSmem);
        if (e != cudaSuccess) {
            set_error("attn_fwd_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
            return NPFK_ERR_CUDA;
        }
        attr_done[di] = true;
    }
    const int64_t grid = (int64_t)B * (E / 32) * p.nsplit;
    if (grid > 2147483647LL) {
        set_error("attn_fwd_tc: grid too large");
        return NPFK_ERR_INVALID;
    }
    if (D == 16) attn_fwd_tc_kernel<16><<<(unsigned)grid, kAThreads, kAFwdSmem, st>>>(tmQ, tmK, tmV, p);
    else attn_fwd_tc_kernel<32><<<(unsigned)grid, kAThreads, kAFwdSmem, st>>>(tmQ, tmK, tmV, p);
    count_launch();
    return check_launch("attn_fwd_tc");
}

}  // namespace tc
}  // namespace npfk
