// tcgen05 / TMEM / TMA multi-head cross attention (targets -> context) for NPFK_MODE_FAST.
//
// Replaces _make_multiheaded / DotAttender.score / softmax / bmm / _concatenate_multiheads of the reference
// (npf/architectures/attention.py:507-527, 204-220, 129-164).  Heads are slices of the feature dimension and are
// read in place by TMA: one 32-float (128-byte) column box of q/k/v holds 32/D heads; a head is addressed by a
// byte offset inside the swizzled row, so no permute copies and no [B*H, Nq, Nk] logits ever reach HBM.
//
// One CTA owns (batch item, column box) and walks query tiles of 128 rows.  K/V of the box stay resident in
// shared memory when the context fits the ring (<= 512 points), otherwise they are streamed once per unit pair.
// A unit is (query tile, head); units 2i and 2i+1 run concurrently on two softmax warpgroups:
//   warp 0            TMA producer (Q tiles, K/V chunks of 64 keys)
//   warp 1            TMEM allocator + tcgen05.mma issuer (one elected thread)
//   warps 2..5, 6..9  softmax warpgroups: thread = query row (TMEM lane)
// Per (unit, chunk):  S = Q_h K_c^T (TMEM)  ->  online softmax in registers: running max m, P = exp2((S-m)*scale*log2e)
// truncated to tf32 and written back IN PLACE over S (tcgen05.st)  ->  O_c = P V_c with A = P read from TMEM  ->
// the warpgroup folds O_c into its register accumulator, rescaled by exp2((m_old-m_new)*scale*log2e).
// Single pass: every logit is computed and read once; nothing but q, k, v, out and lse touches HBM.
#include <math.h>
#include <stdlib.h>

#include "tc_common.cuh"

namespace npfk {
namespace tc {

constexpr int kAQT = 128;                    // queries per tile (UMMA M, TMEM lanes)
constexpr int kAKC = 64;                     // keys per chunk (UMMA N of S, K of PV)
constexpr int kARing = 8;                    // K/V ring slots
constexpr int kAQSlots = 4;                  // Q tile slots
constexpr int kAThreads = 448;                // TMA + MMA warps and up to three softmax warpgroups
constexpr int kAKBytes = kAKC * 128;         // one chunk of a 32-float box
constexpr int kASlotBytes = 2 * kAKBytes;    // K chunk + V chunk
constexpr int kAQBytes = kAQT * 128;
constexpr int kARingOff = 0;
constexpr int kAQOff = kARing * kASlotBytes;             // 131072
constexpr int kABarOff = kAQOff + kAQSlots * kAQBytes;   // 196608
constexpr int kANumBars = 2 * kARing + 2 * kAQSlots + 18;
constexpr int kAFwdSmem = 1024 + kABarOff + kANumBars * 8 + 16;
constexpr uint32_t kATmemCols = 512;         // per warpgroup: S 2 x 64 columns, O 2 x 32 columns (192), two warpgroups
constexpr uint32_t kATmemWg = 192;
constexpr uint32_t kATmemO = 128;

struct AttnParams {
    float* out;          // [B, Nq, E]
    float* lse;          // [B, H, Nq]
    int B, H, E;
    int64_t Nq, Nk;
    float scale;
    int n_chunks, q_tiles, nsplit, tiles_per_split;
    int stagger_ns;      // start-up delay of warpgroup 1 (de-phases the two softmax warpgroups on the MUFU pipe)
    int* g_abort;
};

struct AttnBars {
    uint64_t* kv_full;    // [kARing]
    uint64_t* kv_empty;   // [kARing]
    uint64_t* q_full;     // [kAQSlots]
    uint64_t* q_empty;    // [kAQSlots]
    uint64_t* s_full;     // [2 warpgroups][2 buffers]
    uint64_t* p_ready;    // [2][2]
    uint64_t* o_full;     // [2][2]
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
    bar.q_empty = bar.q_full + kAQSlots;
    bar.s_full = bar.q_empty + kAQSlots;
    bar.p_ready = bar.s_full + 6;
    bar.o_full = bar.s_full + 12;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + kANumBars);
    volatile int* cta_abort = reinterpret_cast<volatile int*>(tmem_ptr + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int boxes = p.E / 32;
    const int item = blockIdx.x / p.nsplit, split = blockIdx.x % p.nsplit;
    const int b = item / boxes, box = item % boxes;
    const int qt_begin = split * p.tiles_per_split;
    const int qt_end = min(p.q_tiles, qt_begin + p.tiles_per_split);
    const int n_qt = max(qt_end - qt_begin, 0);
    const int U = n_qt * HB;                 // units = (query tile, head); unit u -> tile u / HB, head u % HB
    const int nc = p.n_chunks;
    const bool resident = nc <= kARing;
    // Resident K/V with head dim 16: THREE softmax warpgroups (S 2 x 64 + O 2 x 16 = 160 TMEM columns each) so that
    // three warps per scheduler share the MUFU pipe; otherwise two warpgroups of 192 columns.
    const int nwg = (resident && D == 16) ? 3 : 2;
    const uint32_t tm_wg = nwg == 3 ? 160u : kATmemWg;
    const uint32_t o_str = nwg == 3 ? 16u : 32u;

    if (threadIdx.x == 0) {
        prefetch_tmap(&tmQ);
        prefetch_tmap(&tmK);
        prefetch_tmap(&tmV);
        for (int i = 0; i < kARing; ++i) {
            mbar_init(&bar.kv_full[i], 1);
            mbar_init(&bar.kv_empty[i], 1);
        }
        for (int i = 0; i < kAQSlots; ++i) {
            mbar_init(&bar.q_full[i], 1);
            mbar_init(&bar.q_empty[i], HB);     // one commit per unit (head) that reads the tile
        }
        for (int i = 0; i < 6; ++i) {
            mbar_init(&bar.s_full[i], 1);
            mbar_init(&bar.p_ready[i], 4);
            mbar_init(&bar.o_full[i], 1);
        }
        *cta_abort = 0;
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_ptr, kATmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;
    const int n_pairs = (U + 1) / 2;

    if (warp == 0) {
        if (lane == 0 && U > 0) {   // ------------------------------------------------------------ TMA producer
            auto load_kv = [&](int slot, int c) {
                uint8_t* sK = base + kARingOff + slot * kASlotBytes;
                mbar_arrive_expect_tx(&bar.kv_full[slot], (uint32_t)kASlotBytes);
                tma_load_3d(&tmK, &bar.kv_full[slot], sK, box * 32, c * kAKC, b);
                tma_load_3d(&tmV, &bar.kv_full[slot], sK + kAKBytes, box * 32, c * kAKC, b);
            };
            if (resident)
                for (int c = 0; c < nc; ++c) load_kv(c, c);
            uint32_t kv_it = 0;
            int next_tile = 0;      // local index of the next Q tile to load
            bool alive = true;
            for (int pi = 0; pi < n_pairs && alive; ++pi) {
                const int last_tile = min(2 * pi + 1, U - 1) / HB;      // tiles used by units 2pi, 2pi+1
                for (; next_tile <= last_tile; ++next_tile) {
                    const uint32_t qs = (uint32_t)next_tile % kAQSlots, qph = ((uint32_t)next_tile / kAQSlots) & 1;
                    if (!mbar_wait(&bar.q_empty[qs], qph ^ 1, cta_abort, p.g_abort, 21)) { alive = false; break; }
                    mbar_arrive_expect_tx(&bar.q_full[qs], (uint32_t)kAQBytes);
                    tma_load_3d(&tmQ, &bar.q_full[qs], base + kAQOff + qs * kAQBytes, box * 32, (qt_begin + next_tile) * kAQT, b);
                }
                if (!resident && alive) {
                    for (int c = 0; c < nc; ++c, ++kv_it) {
                        const uint32_t slot = kv_it % kARing, ph = (kv_it / kARing) & 1;
                        if (!mbar_wait(&bar.kv_empty[slot], ph ^ 1, cta_abort, p.g_abort, 22)) { alive = false; break; }
                        load_kv((int)slot, c);
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (U > 0 && elect_one()) {   // ---------------------------------------------------------- MMA issuer
            const uint32_t idesc_s = make_idesc_tf32(kAKC, false, false);
            const uint32_t idesc_pv = make_idesc_tf32(D, false, true);          // A = P from TMEM, B = V MN-major
            const uint32_t ring_addr = smem_u32(base + kARingOff);
            const uint64_t d_ring_k = make_sdesc_sw128(ring_addr, 16, 1024);
            const uint64_t d_ring_v = make_sdesc_mn32(ring_addr + kAKBytes, 4096, 512);
            const uint64_t d_q0 = make_sdesc_sw128(smem_u32(base + kAQOff), 16, 1024);
            uint32_t cnt[2] = {0, 0};       // S / P / O buffer use counters per warpgroup
            uint32_t kv_it = 0;             // streamed K/V chunks consumed so far
            // Resident K/V: the two warpgroups are served as two INDEPENDENT streams of (unit, chunk) steps.  Stream g
            // always has S of its next two steps issued; when its P(n) is ready the issuer queues P.V(n) and then
            // S(n+2) into the buffer P.V(n) has just read (the tensor pipe executes in order).  Nothing makes one
            // warpgroup wait for the other, so their exp phases need not coincide on the MUFU pipe.
            auto role_streams = [&]() {
                for (int c = 0; c < nc; ++c)
                    if (!mbar_wait(&bar.kv_full[c], 0, cta_abort, p.g_abort, 24)) return;
                tc_fence_after();
                // Per-stream state.  Every loop over the streams is fully unrolled, so these arrays live in registers and
                // the (unit, chunk) of the next step is carried along instead of being divided out of a step counter: the
                // instruction latency of this one thread is on the critical path of all three warpgroups.
                uint32_t n_steps[3], s_n[3], p_n[3];     // steps in all, S issued, P.V issued
                int s_u[3], s_c[3], p_u[3], p_c[3];      // unit / chunk of the next S and of the next P.V
#pragma unroll
                for (int g = 0; g < 3; ++g) {
                    n_steps[g] = g < nwg ? (uint32_t)((U - g + nwg - 1) / nwg) * (uint32_t)nc : 0u;
                    s_n[g] = p_n[g] = 0u;
                    s_u[g] = p_u[g] = g;
                    s_c[g] = p_c[g] = 0;
                }
                // S of the next step of stream g.  Returns 1 = issued, 0 = not now, -1 = the CTA is aborting.  When the unit's
                // Q tile has not landed, waiting for it is safe only if its load no longer depends on this thread: the
                // slot's previous tile (t - kAQSlots) is released by the last P.V of BOTH of its heads, which belong to other
                // streams -- a stream that has run ahead must keep serving the others instead of blocking (deadlock).
                auto issue_s = [&](const int g) -> int {
                    const int u = s_u[g], c = s_c[g], t = u / HB, hb = u % HB;
                    const uint32_t qs = (uint32_t)t % kAQSlots, qph = ((uint32_t)t / kAQSlots) & 1;
                    if (c == 0) {
                        if (t >= kAQSlots && !mbar_test_wait(smem_u32(&bar.q_full[qs]), qph)) {
                            const int u_last = (t - kAQSlots) * HB + HB - 1;      // every unit up to here must be finished
                            bool released = true;
#pragma unroll
                            for (int g2 = 0; g2 < 3; ++g2) released &= (g2 >= nwg) | (p_u[g2] > u_last);
                            if (!released) return 0;
                        }
                        if (!mbar_wait(&bar.q_full[qs], qph, cta_abort, p.g_abort, 23)) return -1;
                        tc_fence_after();
                    }
                    const uint32_t ko = (uint32_t)(hb * D * 4), sb = s_n[g] & 1;
                    const uint64_t dq = sdesc_advance(d_q0, qs * kAQBytes + ko);
                    const uint64_t dk = sdesc_advance(d_ring_k, (uint32_t)c * kASlotBytes + ko);
                    const uint32_t tS = tmem_base + (uint32_t)g * tm_wg + sb * kAKC;
#pragma unroll
                    for (int s = 0; s < D / 8; ++s)
                        umma_tf32(tS, sdesc_advance(dq, s * 32), sdesc_advance(dk, s * 32), idesc_s, s > 0);
                    umma_commit(&bar.s_full[g * 2 + sb]);
                    ++s_n[g];
                    if (++s_c[g] == nc) {
                        s_c[g] = 0;
                        s_u[g] += nwg;
                    }
                    return 1;
                };
#pragma unroll
                for (int g = 0; g < 3; ++g)
#pragma unroll
                    for (int i = 0; i < 2; ++i)
                        if (s_n[g] < n_steps[g] && issue_s(g) < 0) return;       // the first tiles never wait for a slot
                uint32_t spin = 0;
                while (p_n[0] < n_steps[0] || p_n[1] < n_steps[1] || p_n[2] < n_steps[2]) {
                    bool progressed = false;
#pragma unroll
                    for (int g = 0; g < 3; ++g) {
                        const uint32_t n = p_n[g];
                        if (n >= n_steps[g]) continue;
                        if (s_n[g] < n_steps[g] && s_n[g] < n + 2) {      // an S issue was put off: retry
                            const int r = issue_s(g);
                            if (r < 0) return;
                            progressed |= r > 0;
                        }
                        if (n >= s_n[g]) continue;
                        const uint32_t sb = n & 1;
                        if (!mbar_try_wait(smem_u32(&bar.p_ready[g * 2 + sb]), (n >> 1) & 1)) continue;
                        tc_fence_after();
                        const int u = p_u[g], c = p_c[g], t = u / HB, hb = u % HB;
                        const uint32_t tP = tmem_base + (uint32_t)g * tm_wg + sb * kAKC;
                        const uint32_t tO = tmem_base + (uint32_t)g * tm_wg + kATmemO + sb * o_str;
                        const uint64_t dv = sdesc_advance(d_ring_v, (uint32_t)c * kASlotBytes + ((D == 16) ? (uint32_t)hb * 64u : 0u));
#pragma unroll
                        for (int j = 0; j < kAKC / 8; ++j)
                            umma_tf32_ts(tO, tP + 8 * j, sdesc_advance(dv, j * 1024), idesc_pv, j > 0);
                        umma_commit(&bar.o_full[g * 2 + sb]);
                        ++p_n[g];
                        if (++p_c[g] == nc) {
                            umma_commit(&bar.q_empty[(uint32_t)t % kAQSlots]);
                            p_c[g] = 0;
                            p_u[g] += nwg;
                        }
                        if (s_n[g] < n_steps[g] && issue_s(g) < 0) return;
                        progressed = true;
                    }
                    if (progressed) {
                        spin = 0;
                    } else if ((++spin & 255u) == 255u) {
                        if (*cta_abort) return;
                        if (spin > kSpinLimit) {
                            *cta_abort = 27;
                            atomicCAS(p.g_abort, 0, 27);
                            return;
                        }
                    }
                }
            };
            auto role = [&]() {
                if (resident) {
                    role_streams();
                    return;
                }
                for (int pi = 0; pi < n_pairs; ++pi) {
                    const int ng = min(2, U - 2 * pi);
                    uint64_t dq[2];
                    uint32_t koff[2], voff[2];
                    for (int g = 0; g < ng; ++g) {
                        const int u = 2 * pi + g, t = u / HB, hb = u % HB;
                        const uint32_t qs = (uint32_t)t % kAQSlots;
                        if (HB == 1 || g == 0) {
                            if (!mbar_wait(&bar.q_full[qs], ((uint32_t)t / kAQSlots) & 1, cta_abort, p.g_abort, 23)) return;
                        }
                        koff[g] = (uint32_t)(hb * D * 4);
                        voff[g] = (D == 16) ? (uint32_t)hb * 64u : 0u;
                        dq[g] = sdesc_advance(d_q0, qs * kAQBytes + koff[g]);
                    }
                    tc_fence_after();
                    // S(g, c) = Q_g K_c^T into S buffer (cnt[g] + c) & 1 of warpgroup g
                    auto issue_s = [&](int c) -> bool {
                        const uint32_t slot = resident ? (uint32_t)c : (kv_it + (uint32_t)c) % kARing;
                        const uint32_t ph = resident ? 0u : ((kv_it + (uint32_t)c) / kARing) & 1;
                        if (!mbar_wait(&bar.kv_full[slot], ph, cta_abort, p.g_abort, 24)) return false;
                        tc_fence_after();
                        for (int g = 0; g < ng; ++g) {
                            const uint32_t n = cnt[g] + (uint32_t)c, sb = n & 1;
                            const uint64_t dk = sdesc_advance(d_ring_k, slot * kASlotBytes + koff[g]);
                            const uint32_t tS = tmem_base + (uint32_t)g * tm_wg + sb * kAKC;
#pragma unroll
                            for (int s = 0; s < D / 8; ++s)
                                umma_tf32(tS, sdesc_advance(dq[g], s * 32), sdesc_advance(dk, s * 32), idesc_s, s > 0);
                            umma_commit(&bar.s_full[g * 2 + sb]);
                        }
                        return true;
                    };
                    if (!issue_s(0)) return;
                    for (int c = 0; c < nc; ++c) {
                        if (c + 1 < nc && !issue_s(c + 1)) return;
                        const uint32_t slot = resident ? (uint32_t)c : (kv_it + (uint32_t)c) % kARing;
                        for (int g = 0; g < ng; ++g) {
                            const uint32_t n = cnt[g] + (uint32_t)c, sb = n & 1;
                            if (!mbar_wait(&bar.p_ready[g * 2 + sb], (n >> 1) & 1, cta_abort, p.g_abort, 27)) return;
                            tc_fence_after();
                            const uint32_t tP = tmem_base + (uint32_t)g * tm_wg + sb * kAKC;
                            const uint32_t tO = tmem_base + (uint32_t)g * tm_wg + kATmemO + sb * o_str;
                            const uint64_t dv = sdesc_advance(d_ring_v, slot * kASlotBytes + voff[g]);
#pragma unroll
                            for (int j = 0; j < kAKC / 8; ++j)
                                umma_tf32_ts(tO, tP + 8 * j, sdesc_advance(dv, j * 1024), idesc_pv, j > 0);
                            umma_commit(&bar.o_full[g * 2 + sb]);
                        }
                        if (!resident) umma_commit(&bar.kv_empty[slot]);
                    }
                    if (!resident) kv_it += (uint32_t)nc;
                    for (int g = 0; g < ng; ++g) {
                        cnt[g] += (uint32_t)nc;
                        const int u = 2 * pi + g, t = u / HB;
                        umma_commit(&bar.q_empty[(uint32_t)t % kAQSlots]);
                    }
                }
            };
            role();
        }
    } else {
        // ------------------------------------------------------------------------------ softmax warpgroups
        const int qd = warp & 3;                       // TMEM lane quarter of this warp
        const int wg = (warp - 2) >> 2;                // warpgroup = unit parity
        const int row = qd * 32 + lane;
        const uint32_t wg_base = tmem_base + ((uint32_t)(qd * 32) << 16) + (uint32_t)wg * tm_wg;
        const float k2 = p.scale * 1.4426950408889634f;
        const int tail_valid = (int)(p.Nk - (int64_t)(nc - 1) * kAKC);      // valid keys in the last chunk (1..64)
        uint32_t cnt = 0;
        auto wait_all = [&](uint64_t* bb, uint32_t ph, int code) -> bool {
            bool ok = mbar_wait(bb, ph, cta_abort, p.g_abort, code);
            return __all_sync(0xffffffffu, ok);
        };
        auto role = [&]() {
            if (wg == 1 && p.stagger_ns > 0) __nanosleep((unsigned)p.stagger_ns);
            for (int u = wg; u < U; u += nwg) {
                const int t = u / HB, hb = u % HB;
                const int64_t qrow = (int64_t)(qt_begin + t) * kAQT + row;
                float m = -INFINITY, l = 0.f, alpha_prev = 1.f;
                float o[D];
#pragma unroll
                for (int d = 0; d < D; ++d) o[d] = 0.f;
                // fold O_c of buffer use n into the accumulator: o = o * alpha + O_c
                auto fold = [&](uint32_t n, float alpha) -> bool {
                    if (!wait_all(&bar.o_full[wg * 2 + (n & 1)], (n >> 1) & 1, 34)) return false;
                    tc_fence_after();
                    uint32_t r[D];
                    if constexpr (D == 16) tmem_ld_32x16(wg_base + kATmemO + (n & 1) * o_str, r);
                    else tmem_ld_32x32(wg_base + kATmemO + (n & 1) * o_str, r);
                    tmem_ld_wait();
#pragma unroll
                    for (int d = 0; d < D; ++d) o[d] = fmaf(o[d], alpha, __uint_as_float(r[d]));
                    return true;
                };
                for (int c = 0; c < nc; ++c, ++cnt) {
                    const uint32_t sb = cnt & 1;
                    if (!wait_all(&bar.s_full[wg * 2 + sb], (cnt >> 1) & 1, 31)) return;
                    tc_fence_after();
                    uint32_t v0[32], v1[32];
                    tmem_ld_32x32(wg_base + sb * kAKC, v0);
                    tmem_ld_32x32(wg_base + sb * kAKC + 32, v1);
                    tmem_ld_wait();
                    const bool tail = (c == nc - 1) && tail_valid < kAKC;
                    float mx = -INFINITY;
                    if (!tail) {
                        // four independent chains (a single running max is a 32-deep dependent FMNMX chain)
                        float mq[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
                        for (int j = 0; j < 32; ++j)
                            mq[j & 3] = fmaxf(mq[j & 3], fmaxf(__uint_as_float(v0[j]), __uint_as_float(v1[j])));
                        mx = fmaxf(fmaxf(mq[0], mq[1]), fmaxf(mq[2], mq[3]));
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            if (j < tail_valid) mx = fmaxf(mx, __uint_as_float(v0[j]));
                            if (32 + j < tail_valid) mx = fmaxf(mx, __uint_as_float(v1[j]));
                        }
                    }
                    const float m_new = fmaxf(m, mx);
                    const float mk = m_new * k2;
                    const float alpha = ex2_approx((m - m_new) * k2);      // m = -inf on the first chunk -> 0
                    float lq[4] = {0.f, 0.f, 0.f, 0.f};      // independent partial row sums (no 32-deep FADD chain)
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        // P is truncated to tf32 here (what the MMA would do anyway) so that the row sum l matches
                        // the values that enter P.V exactly
                        uint32_t a = __float_as_uint(ex2_approx(fmaf(__uint_as_float(v0[j]), k2, -mk))) & 0xffffe000u;
                        uint32_t bq = __float_as_uint(ex2_approx(fmaf(__uint_as_float(v1[j]), k2, -mk))) & 0xffffe000u;
                        if (tail) {
                            if (j >= tail_valid) a = 0u;
                            if (32 + j >= tail_valid) bq = 0u;
                        }
                        v0[j] = a;
                        v1[j] = bq;
                        lq[j & 3] += __uint_as_float(a) + __uint_as_float(bq);
                    }
                    l = fmaf(l, alpha, (lq[0] + lq[1]) + (lq[2] + lq[3]));
                    m = m_new;
                    tmem_st_32x32(wg_base + sb * kAKC, v0);
                    tmem_st_32x32(wg_base + sb * kAKC + 32, v1);
                    tmem_st_wait();
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&bar.p_ready[wg * 2 + sb]);
                    if (c > 0 && !fold(cnt - 1, alpha_prev)) return;
                    alpha_prev = alpha;
                }
                if (!fold(cnt - 1, alpha_prev)) return;
                tc_fence_before();
                if (qrow < p.Nq) {
                    const float inv = 1.f / l;
                    const int h = box * HB + hb;
                    float* op = p.out + ((int64_t)b * p.Nq + qrow) * p.E + h * D;
#pragma unroll
                    for (int d4 = 0; d4 < D / 4; ++d4)
                        reinterpret_cast<float4*>(op)[d4] =
                            make_float4(o[4 * d4] * inv, o[4 * d4 + 1] * inv, o[4 * d4 + 2] * inv, o[4 * d4 + 3] * inv);
                    p.lse[((int64_t)b * p.H + h) * p.Nq + qrow] = m * p.scale + logf(l);
                }
            }
        };
        if (U > 0 && wg < nwg) role();
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
    {
        static int stagger = -1;     // NPFK_ATTN_STAGGER_NS overrides the default (developer knob)
        if (stagger < 0) {
            const char* e = getenv("NPFK_ATTN_STAGGER_NS");
            stagger = e ? atoi(e) : 600;
        }
        p.stagger_ns = stagger;
    }
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
