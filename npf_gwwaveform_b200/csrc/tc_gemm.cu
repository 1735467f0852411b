// tcgen05 / TMEM / TMA GEMMs behind npfk_linear_* in NPFK_MODE_FAST (tf32 operands, fp32 accumulate).
//
//   forward  Y[M,N]  = act(X[M,K] . W[N,K]^T + bias + res)      A K-major, B K-major
//   dgrad    dX[M,K] = G[M,N] . W[N,K]                          A K-major, B MN-major (W read in place)
//   wgrad    dW[N,K] += G[M,N]^T . X[M,K] ; db[N] += sum_m G    A, B MN-major, accumulated in TMEM over M
//
// Persistent CTAs of 6 warps: warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer (one elected
// thread), warps 2..5 = epilogue (TMEM -> registers -> bias/residual/ReLU -> swizzled smem -> TMA store).
// fp32 tensors are loaded by TMA with the 128-byte swizzle and consumed by tcgen05.mma.kind::tf32 as they
// are; accumulators are double-buffered in TMEM so the epilogue of tile i overlaps the MMAs of tile i+1.
// Replaces the cuBLAS addmm + elementwise kernels of MLP.forward (reference npf/architectures/mlp.py:95-109).
#include "tc_common.cuh"

namespace npfk {
namespace tc {

__device__ int g_tc_abort = 0;

// ------------------------------------------------------------------------------------------ host helpers
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
        else
            (void)cudaGetLastError();
    }
    return fn;
}

int encode_tmap_2d(CUtensorMap* out, const float* base, uint64_t inner, uint64_t outer, uint64_t ld_elems,
                   uint32_t box_rows, bool atom32) {
    // The driver call needs a current context; autograd's backward threads may not have bound the primary
    // context yet (the runtime binds it lazily), so touch the runtime once per thread.
    static thread_local bool ctx_bound = false;
    if (!ctx_bound) {
        (void)cudaFree(nullptr);
        ctx_bound = true;
    }
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) {
        set_error("cuTensorMapEncodeTiled is not available from the driver");
        return NPFK_ERR_CUDA;
    }
    if ((reinterpret_cast<uintptr_t>(base) & 15) || (ld_elems & 3) || inner == 0 || outer == 0) {
        set_error("tensor map: base must be 16-byte aligned and the row pitch a multiple of 4 floats");
        return NPFK_ERR_INVALID;
    }
    cuuint64_t dims[2] = {inner, outer};
    cuuint64_t strides[1] = {ld_elems * sizeof(float)};
    cuuint32_t box[2] = {32, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE,
                    atom32 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed with CUresult %d (inner=%llu outer=%llu ld=%llu box_rows=%u)", (int)r,
                  (unsigned long long)inner, (unsigned long long)outer, (unsigned long long)ld_elems, box_rows);
        return NPFK_ERR_CUDA;
    }
    return NPFK_OK;
}

int encode_tmap_3d(CUtensorMap* out, const float* base, uint64_t cols, uint64_t rows, uint64_t batch, uint64_t ld_elems,
                   uint64_t batch_elems, uint32_t box_rows, bool atom32) {
    static thread_local bool ctx_bound = false;
    if (!ctx_bound) {
        (void)cudaFree(nullptr);
        ctx_bound = true;
    }
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) {
        set_error("cuTensorMapEncodeTiled is not available from the driver");
        return NPFK_ERR_CUDA;
    }
    if ((reinterpret_cast<uintptr_t>(base) & 15) || (ld_elems & 3) || (batch_elems & 3) || cols == 0 || rows == 0 ||
        batch == 0) {
        set_error("tensor map: base must be 16-byte aligned and the pitches multiples of 4 floats");
        return NPFK_ERR_INVALID;
    }
    cuuint64_t dims[3] = {cols, rows, batch};
    cuuint64_t strides[2] = {ld_elems * sizeof(float), batch_elems * sizeof(float)};
    cuuint32_t box[3] = {32, box_rows, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE,
                    atom32 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled(3d) failed with CUresult %d (cols=%llu rows=%llu batch=%llu ld=%llu)", (int)r,
                  (unsigned long long)cols, (unsigned long long)rows, (unsigned long long)batch,
                  (unsigned long long)ld_elems);
        return NPFK_ERR_CUDA;
    }
    return NPFK_OK;
}

constexpr int kThreads = 192;
constexpr int kBlockM = 128;        // rows per output tile = UMMA M
constexpr int kBlockK = 32;         // fp32 elements per 128-byte swizzle row
constexpr int kABytes = kBlockM * 128;
constexpr int kStageBytesMax = 227 * 1024;
constexpr int kEpiBytes = 4 * 4 * 4096;   // per epilogue warp: two 32x32 fp32 staging boxes + two side-input boxes

struct NtParams {
    int64_t M;            // output rows
    int N;                // output columns (all slices)
    int K;                // reduction length
    const float* bias;    // [N] or null
    const float* res;     // residual rows or null
    int64_t ldres, res_rows;
    const float* mask;    // dgrad: output is zeroed where mask[m, n] <= 0 (ReLU backward of the layer below)
    int64_t ldmask;
    int relu, round_out;
    int side_mode;        // 0: none / generic loads, 1: residual via TMA (tmE), 2: mask via TMA (tmE)
    int num_stages, b_bytes;   // smem ring depth, bytes of the B operand per stage
    int slice_n;          // output columns per CTA column slice (blockIdx.y): 256, or narrower when M alone cannot fill the SMs
    int* g_abort;
};

struct SmemLayout {
    uint8_t* stages;
    uint8_t* epi;
    float* bias;
    uint64_t* full;
    uint64_t* empty;
    uint64_t* tmem_full;
    uint64_t* tmem_empty;
    uint64_t* side_bar;   // [4 warps][2]
    uint32_t* tmem_ptr;
    volatile int* cta_abort;
};

__device__ __forceinline__ SmemLayout carve_smem(uint8_t* raw, int num_stages, int stage_bytes, int epi_bytes) {
    const uint32_t a = smem_u32(raw);
    uint8_t* base = raw + ((1024u - (a & 1023u)) & 1023u);
    SmemLayout s;
    s.stages = base;
    s.epi = base + (size_t)num_stages * stage_bytes;
    s.bias = reinterpret_cast<float*>(s.epi + epi_bytes);
    s.full = reinterpret_cast<uint64_t*>(s.epi + epi_bytes + 1024);
    s.empty = s.full + num_stages;
    s.tmem_full = s.empty + num_stages;
    s.tmem_empty = s.tmem_full + 2;
    s.side_bar = s.tmem_empty + 2;
    s.tmem_ptr = reinterpret_cast<uint32_t*>(s.side_bar + 8);
    s.cta_abort = reinterpret_cast<volatile int*>(s.tmem_ptr + 1);
    return s;
}
// bytes of dynamic shared memory for the layout above (plus alignment slack)
static inline int smem_bytes_for(int num_stages, int stage_bytes, int epi_bytes) {
    return 1024 + num_stages * stage_bytes + epi_bytes + 1024 + (2 * num_stages + 12) * 8 + 16;
}
__host__ __device__ inline uint32_t tmem_cols_for(int need) {
    uint32_t c = 32;
    while ((int)c < need) c <<= 1;
    return c;
}

// Y tile [128 x Nc] = A[128 x K] . B^T ; one CTA column slice of at most p.slice_n <= 256 columns (blockIdx.y)
template <bool B_MN>
__global__ void __launch_bounds__(kThreads, 1)
tc_gemm_nt_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                  const __grid_constant__ CUtensorMap tmC, const __grid_constant__ CUtensorMap tmE, const NtParams p) {
    extern __shared__ uint8_t smem_raw[];
    const int stage_bytes = kABytes + p.b_bytes;
    const SmemLayout sm = carve_smem(smem_raw, p.num_stages, stage_bytes, kEpiBytes);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n0 = blockIdx.y * p.slice_n;
    const int Nc = min(p.slice_n, p.N - n0);
    const int Npad = (Nc + 31) & ~31;
    const uint32_t tmem_cols = tmem_cols_for(2 * Npad);
    const int64_t num_tiles = (p.M + kBlockM - 1) / kBlockM;
    const int KB = (p.K + kBlockK - 1) / kBlockK;

    if (threadIdx.x == 0) {
        prefetch_tmap(&tmA);
        prefetch_tmap(&tmB);
        prefetch_tmap(&tmC);
        for (int s = 0; s < p.num_stages; ++s) {
            mbar_init(&sm.full[s], 1);
            mbar_init(&sm.empty[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&sm.tmem_full[b], 1);
            mbar_init(&sm.tmem_empty[b], 4);
        }
        for (int b = 0; b < 8; ++b) mbar_init(&sm.side_bar[b], 1);
        if (p.side_mode) prefetch_tmap(&tmE);
        *sm.cta_abort = 0;
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(sm.tmem_ptr, tmem_cols);
    if (warp >= 2)
        for (int i = threadIdx.x - 64; i < 256; i += 128) sm.bias[i] = (p.bias && i < Nc) ? __ldg(p.bias + n0 + i) : 0.f;
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *sm.tmem_ptr;

    if (warp == 0) {
        if (lane == 0) {   // ---------------- TMA producer
            int stage = 0;
            uint32_t phase = 0;
            const uint32_t b_tx = B_MN ? (uint32_t)(Npad / 32) * 4096u : (uint32_t)p.b_bytes;
            bool alive = true;
            for (int64_t tile = blockIdx.x; tile < num_tiles && alive; tile += gridDim.x) {
                for (int kb = 0; kb < KB; ++kb) {
                    if (!mbar_wait(&sm.empty[stage], phase ^ 1, sm.cta_abort, p.g_abort, 1)) { alive = false; break; }
                    uint8_t* sA = sm.stages + (size_t)stage * stage_bytes;
                    uint8_t* sB = sA + kABytes;
                    mbar_arrive_expect_tx(&sm.full[stage], (uint32_t)kABytes + b_tx);
                    tma_load_2d(&tmA, &sm.full[stage], sA, kb * kBlockK, (int)(tile * kBlockM));
                    if (!B_MN) {
                        tma_load_2d(&tmB, &sm.full[stage], sB, kb * kBlockK, n0);
                    } else {
                        for (int j = 0; j < Npad / 32; ++j)
                            tma_load_2d(&tmB, &sm.full[stage], sB + j * 4096, n0 + j * 32, kb * kBlockK);
                    }
                    if (++stage == p.num_stages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {   // ---------------- MMA issuer
            const uint32_t idesc = make_idesc_tf32(Nc, false, B_MN);
            const uint32_t stage0 = smem_u32(sm.stages);
            const uint64_t d_a0 = make_sdesc_sw128(stage0, 16, 1024);
            const uint64_t d_b0 = B_MN ? make_sdesc_mn32(stage0 + kABytes, 4096, 512) : make_sdesc_sw128(stage0 + kABytes, 16, 1024);
            int stage = 0;
            uint32_t phase = 0;
            uint32_t it = 0;
            bool alive = true;
            for (int64_t tile = blockIdx.x; tile < num_tiles && alive; tile += gridDim.x, ++it) {
                const uint32_t buf = it & 1, bphase = (it >> 1) & 1;
                if (!mbar_wait(&sm.tmem_empty[buf], bphase ^ 1, sm.cta_abort, p.g_abort, 2)) break;
                tc_fence_after();
                const uint32_t d = tmem_base + buf * (uint32_t)Npad;
                for (int kb = 0; kb < KB; ++kb) {
                    if (!mbar_wait(&sm.full[stage], phase, sm.cta_abort, p.g_abort, 3)) { alive = false; break; }
                    tc_fence_after();
                    const uint32_t soff = (uint32_t)stage * (uint32_t)stage_bytes;
                    const uint64_t ad0 = sdesc_advance(d_a0, soff), bd0 = sdesc_advance(d_b0, soff);
#pragma unroll
                    for (int s = 0; s < 4; ++s)
                        umma_tf32(d, sdesc_advance(ad0, s * 32), sdesc_advance(bd0, B_MN ? s * 1024 : s * 32), idesc,
                                  (kb | s) != 0);
                    umma_commit(&sm.empty[stage]);
                    if (++stage == p.num_stages) { stage = 0; phase ^= 1; }
                }
                if (alive) umma_commit(&sm.tmem_full[buf]);
            }
        }
    } else {   // ---------------- epilogue warps: TMEM lane quarter = warp % 4
        const int q = warp & 3;
        uint8_t* stg = sm.epi + q * 8192;
        uint8_t* side = sm.epi + 32768 + q * 8192;          // two 32x32 boxes of the mask / residual, TMA-prefetched
        uint64_t* sbar = sm.side_bar + q * 2;
        const int64_t res_rows = p.res_rows > 0 ? p.res_rows : p.M;
        uint32_t it = 0, chunk = 0;
        // side-input box of (tile, column c) -> buffer (chunk index & 1); issued one chunk ahead by lane 0
        auto side_load = [&](int64_t tile_, int c_, uint32_t chunk_) {
            const int64_t r0 = tile_ * kBlockM + q * 32;
            if (r0 < p.M) {
                mbar_arrive_expect_tx(&sbar[chunk_ & 1], 4096u);
                tma_load_2d(&tmE, &sbar[chunk_ & 1], side + (chunk_ & 1) * 4096, n0 + c_, (int)r0);
            }
        };
        if (p.side_mode && lane == 0 && (int64_t)blockIdx.x < num_tiles) side_load(blockIdx.x, 0, 0);
        for (int64_t tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
            const uint32_t buf = it & 1, bphase = (it >> 1) & 1;
            bool ok = mbar_wait(&sm.tmem_full[buf], bphase, sm.cta_abort, p.g_abort, 4);
            ok = __all_sync(0xffffffffu, ok);
            if (!ok) break;
            tc_fence_after();
            const int64_t row0 = tile * kBlockM + q * 32;
            const int64_t row = row0 + lane;
            const bool row_ok = row < p.M;
            for (int c = 0; c < Npad; c += 32, ++chunk) {
                if (p.side_mode && lane == 0) {      // prefetch the next chunk's box (its buffer was read one chunk ago)
                    if (c + 32 < Npad) side_load(tile, c + 32, chunk + 1);
                    else if (tile + gridDim.x < num_tiles) side_load(tile + gridDim.x, 0, chunk + 1);
                }
                uint32_t v[32];
                tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + buf * (uint32_t)Npad + (uint32_t)c, v);
                tmem_ld_wait();
                if (c + 32 >= Npad) {   // accumulator buffer fully read: hand it back to the MMA warp
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&sm.tmem_empty[buf]);
                }
                float f[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]) + sm.bias[c + j];
                const uint8_t* sbox = side + (chunk & 1) * 4096 + lane * 128;
                if (p.side_mode && row0 < p.M) {
                    bool sok = mbar_wait(&sbar[chunk & 1], (chunk >> 1) & 1, sm.cta_abort, p.g_abort, 5);
                    sok = __all_sync(0xffffffffu, sok);
                    if (!sok) break;
                }
                if (p.side_mode == 1 && row0 < p.M) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const float4 r4 = *reinterpret_cast<const float4*>(sbox + ((j ^ (lane & 7)) << 4));
                        f[4 * j] += r4.x; f[4 * j + 1] += r4.y; f[4 * j + 2] += r4.z; f[4 * j + 3] += r4.w;
                    }
                } else if (p.side_mode == 0 && p.res != nullptr && row_ok) {
                    const float* rp = p.res + (row % res_rows) * p.ldres + n0 + c;
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        if (c + 4 * j < Nc) {
                            const float4 r4 = __ldg(reinterpret_cast<const float4*>(rp) + j);
                            f[4 * j] += r4.x; f[4 * j + 1] += r4.y; f[4 * j + 2] += r4.z; f[4 * j + 3] += r4.w;
                        }
                }
                if (p.relu) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.f);
                }
                if (p.side_mode == 2 && row0 < p.M) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const float4 m4 = *reinterpret_cast<const float4*>(sbox + ((j ^ (lane & 7)) << 4));
                        if (!(m4.x > 0.f)) f[4 * j] = 0.f;
                        if (!(m4.y > 0.f)) f[4 * j + 1] = 0.f;
                        if (!(m4.z > 0.f)) f[4 * j + 2] = 0.f;
                        if (!(m4.w > 0.f)) f[4 * j + 3] = 0.f;
                    }
                } else if (p.side_mode == 0 && p.mask != nullptr && row_ok) {
                    const float* mp = p.mask + row * p.ldmask + n0 + c;
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        if (c + 4 * j < Nc) {
                            const float4 m4 = __ldg(reinterpret_cast<const float4*>(mp) + j);
                            if (!(m4.x > 0.f)) f[4 * j] = 0.f;
                            if (!(m4.y > 0.f)) f[4 * j + 1] = 0.f;
                            if (!(m4.z > 0.f)) f[4 * j + 2] = 0.f;
                            if (!(m4.w > 0.f)) f[4 * j + 3] = 0.f;
                        }
                }
                if (p.round_out) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) f[j] = round_tf32(f[j]);
                }
                uint8_t* sb = stg + (chunk & 1) * 4096;
                if (lane == 0) tma_store_wait_read<1>();   // the box stored two chunks ago has been read
                __syncwarp();
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    float4 o = make_float4(f[4 * j], f[4 * j + 1], f[4 * j + 2], f[4 * j + 3]);
                    *reinterpret_cast<float4*>(sb + lane * 128 + ((j ^ (lane & 7)) << 4)) = o;
                }
                fence_proxy_async();
                __syncwarp();
                if (lane == 0 && row0 < p.M) {
                    tma_store_2d(&tmC, sb, n0 + c, (int)row0);
                    tma_store_commit();
                }
            }
        }
        if (lane == 0) tma_store_wait_all<0>();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, tmem_cols);
    }
}

// ------------------------------------------------------------------------------------------ wgrad
struct TnParams {
    int64_t M;
    int N, K;
    float* dw;
    int64_t lddw;
    float* db;            // [N] "+=" or null
    int64_t kb_per_cta;   // 32-row blocks of M handled by one CTA
    int n_tiles;          // ceil(N / 128)
    int num_stages, b_bytes;
    int vec_red;          // dw rows are 16-byte aligned: red.global.add.v4.f32 in the epilogue
    int* g_abort;
};

// dW[n0:n0+128, k0:k0+Kc] += G[:, n-slice]^T . X[:, k-slice] over this CTA's range of rows.
// Both operands are MN-major: a stage holds 32 rows of G (4 boxes of 32 columns) and of X (Kc/32 boxes).
__global__ void __launch_bounds__(kThreads, 1)
tc_gemm_tn_kernel(const __grid_constant__ CUtensorMap tmG, const __grid_constant__ CUtensorMap tmX, const TnParams p) {
    extern __shared__ uint8_t smem_raw[];
    const int stage_bytes = kABytes + p.b_bytes;
    const SmemLayout sm = carve_smem(smem_raw, p.num_stages, stage_bytes, 0);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n0 = (blockIdx.y % p.n_tiles) * 128;
    const int kc_idx = blockIdx.y / p.n_tiles;
    const int k0 = kc_idx * 256;
    const int Kc = min(256, p.K - k0);
    const int Kn = (Kc + 15) & ~15;       // UMMA N
    const int Kpad = (Kc + 31) & ~31;     // loaded columns / TMEM columns
    const uint32_t tmem_cols = tmem_cols_for(Kpad);
    const int64_t kb_total = (p.M + kBlockK - 1) / kBlockK;
    const int64_t kb_begin = (int64_t)blockIdx.x * p.kb_per_cta;
    const int64_t kb_end = min(kb_total, kb_begin + p.kb_per_cta);
    const bool do_db = (p.db != nullptr) && kc_idx == 0;

    if (threadIdx.x == 0) {
        prefetch_tmap(&tmG);
        prefetch_tmap(&tmX);
        for (int s = 0; s < p.num_stages; ++s) {
            mbar_init(&sm.full[s], 1);
            mbar_init(&sm.empty[s], do_db ? 5 : 1);
        }
        mbar_init(&sm.tmem_full[0], 1);
        *sm.cta_abort = 0;
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(sm.tmem_ptr, tmem_cols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *sm.tmem_ptr;

    if (warp == 0) {
        if (lane == 0) {   // ---------------- TMA producer
            int stage = 0;
            uint32_t phase = 0;
            const uint32_t tx = (uint32_t)kABytes + (uint32_t)(Kpad / 32) * 4096u;
            for (int64_t kb = kb_begin; kb < kb_end; ++kb) {
                if (!mbar_wait(&sm.empty[stage], phase ^ 1, sm.cta_abort, p.g_abort, 11)) break;
                uint8_t* sA = sm.stages + (size_t)stage * stage_bytes;
                uint8_t* sB = sA + kABytes;
                mbar_arrive_expect_tx(&sm.full[stage], tx);
                for (int j = 0; j < 4; ++j) tma_load_2d(&tmG, &sm.full[stage], sA + j * 4096, n0 + j * 32, (int)(kb * kBlockK));
                for (int j = 0; j < Kpad / 32; ++j)
                    tma_load_2d(&tmX, &sm.full[stage], sB + j * 4096, k0 + j * 32, (int)(kb * kBlockK));
                if (++stage == p.num_stages) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {   // ---------------- MMA issuer
            const uint32_t idesc = make_idesc_tf32(Kn, true, true);
            const uint32_t stage0 = smem_u32(sm.stages);
            const uint64_t d_a0 = make_sdesc_mn32(stage0, 4096, 512), d_b0 = make_sdesc_mn32(stage0 + kABytes, 4096, 512);
            int stage = 0;
            uint32_t phase = 0;
            bool alive = true;
            for (int64_t kb = kb_begin; kb < kb_end; ++kb) {
                if (!mbar_wait(&sm.full[stage], phase, sm.cta_abort, p.g_abort, 13)) { alive = false; break; }
                tc_fence_after();
                const uint32_t soff = (uint32_t)stage * (uint32_t)stage_bytes;
                const uint64_t ad0 = sdesc_advance(d_a0, soff), bd0 = sdesc_advance(d_b0, soff);
#pragma unroll
                for (int s = 0; s < 4; ++s)
                    umma_tf32(tmem_base, sdesc_advance(ad0, s * 1024), sdesc_advance(bd0, s * 1024), idesc,
                              (kb > kb_begin || s > 0) ? 1u : 0u);
                umma_commit(&sm.empty[stage]);
                if (++stage == p.num_stages) { stage = 0; phase ^= 1; }
            }
            if (alive) umma_commit(&sm.tmem_full[0]);
        }
    } else {   // ---------------- epilogue warps
        const int q = warp & 3;
        const int t = (warp - 2) * 32 + lane;     // column of the G slice summed by this thread (db)
        bool alive = true;
        if (do_db) {
            float dbsum = 0.f;
            int stage = 0;
            uint32_t phase = 0;
            const int box = t >> 5, col = t & 31;
            for (int64_t kb = kb_begin; kb < kb_end; ++kb) {
                bool ok = mbar_wait(&sm.full[stage], phase, sm.cta_abort, p.g_abort, 15);
                ok = __all_sync(0xffffffffu, ok);
                if (!ok) { alive = false; break; }
                const uint8_t* sA = sm.stages + (size_t)stage * stage_bytes + box * 4096;
#pragma unroll 8
                for (int r = 0; r < 32; ++r)
                    dbsum += *reinterpret_cast<const float*>(sA + r * 128 + (((col >> 3) ^ (r & 3)) << 5) + (col & 7) * 4);
                __syncwarp();
                if (lane == 0) mbar_arrive(&sm.empty[stage]);
                if (++stage == p.num_stages) { stage = 0; phase ^= 1; }
            }
            if (alive && n0 + t < p.N) atomicAdd(p.db + n0 + t, dbsum);
        }
        if (alive) {
            bool ok = mbar_wait(&sm.tmem_full[0], 0, sm.cta_abort, p.g_abort, 14);
            ok = __all_sync(0xffffffffu, ok);
            if (ok) {
                tc_fence_after();
                const int n = n0 + q * 32 + lane;
                for (int c = 0; c < Kpad; c += 32) {
                    uint32_t v[32];
                    tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c, v);
                    tmem_ld_wait();
                    if (n < p.N) {
                        float* out = p.dw + (int64_t)n * p.lddw + k0 + c;
                        if (p.vec_red && c + 32 <= Kc) {     // 16-byte vector reductions: 4x fewer L2 atomics
#pragma unroll
                            for (int j = 0; j < 32; j += 4)
                                red_add_v4(out + j, __uint_as_float(v[j]), __uint_as_float(v[j + 1]),
                                           __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3]));
                        } else {
#pragma unroll
                            for (int j = 0; j < 32; ++j)
                                if (c + j < Kc) atomicAdd(out + j, __uint_as_float(v[j]));
                        }
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, tmem_cols);
    }
}

// ------------------------------------------------------------------------------------------ host launchers
int* abort_flag_ptr() {
    static int* p = nullptr;
    if (!p) {
        void* q = nullptr;
        if (cudaGetSymbolAddress(&q, g_tc_abort) == cudaSuccess) p = static_cast<int*>(q);
        else (void)cudaGetLastError();
    }
    return p;
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// shapes the tensor-core kernels take; everything else stays on the FFMA kernels
bool nt_eligible(const float* a, int64_t lda, const float* b, int64_t ldb, const float* c, int64_t ldc, int64_t M, int N,
                 int K, const float* res, int64_t ldres, const float* mask, int64_t ldmask) {
    if (M < 128 || N < 16 || (N % 16) != 0 || K < 8 || (K % 8) != 0) return false;
    if (M > 2147483647LL) return false;
    if (!aligned16(a) || !aligned16(b) || !aligned16(c) || (lda & 3) || (ldb & 3) || (ldc & 3)) return false;
    if (res && (!aligned16(res) || (ldres & 3))) return false;
    if (mask && (!aligned16(mask) || (ldmask & 3))) return false;
    return true;
}

template <bool B_MN>
static int launch_nt(const float* A, int64_t lda, const float* B, int64_t ldb, float* C, int64_t ldc, int64_t M, int N,
                     int K, const float* bias, const float* res, int64_t ldres, int64_t res_rows, const float* mask,
                     int64_t ldmask, int relu, int round_out, cudaStream_t st) {
    // column slices of 256; when the row tiles alone leave SMs idle (the UNet's 6144-row pointwise convolutions),
    // halve the slice while the finer split still fits one wave: more CTAs share the operand traffic and the epilogue
    const int64_t tiles = (M + kBlockM - 1) / kBlockM;
    int slice_n = 256;
    while (slice_n > 32 && tiles * ((N + slice_n / 2 - 1) / (slice_n / 2)) <= kNumSMs) slice_n >>= 1;
    const int n_slices = (N + slice_n - 1) / slice_n;
    const int Nc_max = N < slice_n ? N : slice_n;
    const int Npad = (Nc_max + 31) & ~31;
    CUtensorMap tmA, tmB, tmC;
    int rc = encode_tmap_2d(&tmA, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda, kBlockM);
    if (rc) return rc;
    if (!B_MN) rc = encode_tmap_2d(&tmB, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb, (uint32_t)Npad);   // W[N,K], box K32 x Npad
    else rc = encode_tmap_2d(&tmB, B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb, 32, true);                // W[R,N], box N32 x R32
    if (rc) return rc;
    rc = encode_tmap_2d(&tmC, C, (uint64_t)N, (uint64_t)M, (uint64_t)ldc, 32);
    if (rc) return rc;
    // mask, or a residual with one row per output row, is prefetched by TMA into the epilogue warps' boxes
    CUtensorMap tmE = tmC;
    int side_mode = 0;
    if (mask && !res) {
        rc = encode_tmap_2d(&tmE, mask, (uint64_t)N, (uint64_t)M, (uint64_t)ldmask, 32);
        side_mode = 2;
    } else if (res && !mask && (res_rows <= 0 || res_rows == M)) {
        rc = encode_tmap_2d(&tmE, res, (uint64_t)N, (uint64_t)M, (uint64_t)ldres, 32);
        side_mode = 1;
    }
    if (rc) return rc;
    NtParams p{};
    p.M = M; p.N = N; p.K = K;
    p.bias = bias; p.res = res; p.ldres = ldres; p.res_rows = res_rows;
    p.mask = mask; p.ldmask = ldmask;
    p.relu = relu; p.round_out = round_out;
    p.side_mode = side_mode;
    p.b_bytes = Npad * 128;
    p.slice_n = slice_n;
    const int stage_bytes = kABytes + p.b_bytes;
    int stages = (kStageBytesMax - 4096 - kEpiBytes) / stage_bytes;
    if (stages > 8) stages = 8;
    if (stages < 2) {
        set_error("tc gemm: tile does not fit shared memory (N=%d)", N);
        return NPFK_ERR_UNSUPPORTED;
    }
    p.num_stages = stages;
    p.g_abort = abort_flag_ptr();
    const int smem = smem_bytes_for(stages, stage_bytes, kEpiBytes);
    static bool attr_done[2] = {false, false};
    if (!attr_done[B_MN]) {
        cudaError_t e = cudaFuncSetAttribute(tc_gemm_nt_kernel<B_MN>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (e != cudaSuccess) {
            set_error("tc gemm: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
            return NPFK_ERR_CUDA;
        }
        attr_done[B_MN] = true;
    }
    int gx = kNumSMs / n_slices;
    if (gx < 1) gx = 1;
    if (tiles < gx) gx = (int)tiles;
    dim3 grid((unsigned)gx, (unsigned)n_slices);
    tc_gemm_nt_kernel<B_MN><<<grid, kThreads, smem, st>>>(tmA, tmB, tmC, tmE, p);
    count_launch();
    return check_launch("tc_gemm_nt");
}

int tc_linear_fwd(const float* x, int64_t ldx, const float* w, int64_t ldw, const float* bias, const float* res,
                  int64_t ldres, int64_t res_rows, float* y, int64_t ldy, int64_t M, int N, int K, int relu,
                  int round_out, cudaStream_t st) {
    return launch_nt<false>(x, ldx, w, ldw, y, ldy, M, N, K, bias, res, ldres, res_rows, nullptr, 0, relu, round_out, st);
}

// dx[M,K] = g[M,N] . w[N,K], zeroed where mask <= 0
int tc_linear_bwd_dx(const float* g, int64_t ldg, const float* w, int64_t ldw, const float* mask, int64_t ldmask,
                     float* dx, int64_t lddx, int64_t M, int N, int K, cudaStream_t st) {
    return launch_nt<true>(g, ldg, w, ldw, dx, lddx, M, /*out cols*/ K, /*reduction*/ N, nullptr, nullptr, 0, 0, mask,
                           ldmask, 0, 1, st);
}

bool tn_eligible(const float* g, int64_t ldg, const float* x, int64_t ldx, int64_t M, int N, int K) {
    if (M < 32 || N < 8 || K < 8 || (N & 3) || (K & 3)) return false;
    if (M > 2147483647LL) return false;
    if (!aligned16(g) || !aligned16(x) || (ldg & 3) || (ldx & 3)) return false;
    return true;
}

int tc_linear_bwd_dw(const float* g, int64_t ldg, const float* x, int64_t ldx, float* dw, int64_t lddw, float* db,
                     int64_t M, int N, int K, cudaStream_t st) {
    CUtensorMap tmG, tmX;
    int rc = encode_tmap_2d(&tmG, g, (uint64_t)N, (uint64_t)M, (uint64_t)ldg, 32, true);
    if (rc) return rc;
    rc = encode_tmap_2d(&tmX, x, (uint64_t)K, (uint64_t)M, (uint64_t)ldx, 32, true);
    if (rc) return rc;
    TnParams p{};
    p.M = M; p.N = N; p.K = K;
    p.dw = dw; p.lddw = lddw; p.db = db;
    p.n_tiles = (N + 127) / 128;
    const int k_chunks = (K + 255) / 256;
    const int Kc_max = K < 256 ? K : 256;
    const int Kpad = (Kc_max + 31) & ~31;
    p.b_bytes = (Kpad / 32) * 4096;
    const int stage_bytes = kABytes + p.b_bytes;
    int stages = (kStageBytesMax - 4096) / stage_bytes;
    if (stages > 8) stages = 8;
    p.num_stages = stages;
    p.g_abort = abort_flag_ptr();
    const int gy = p.n_tiles * k_chunks;
    const int64_t kb_total = (M + kBlockK - 1) / kBlockK;
    p.vec_red = (aligned16(dw) && (lddw & 3) == 0) ? 1 : 0;
    int64_t gx = kNumSMs / gy;             // one wave of CTAs (1 CTA/SM): bounds the reduction traffic on dW
    if (gx < 1) gx = 1;
    if (gx > kb_total) gx = kb_total;
    p.kb_per_cta = (kb_total + gx - 1) / gx;
    gx = (kb_total + p.kb_per_cta - 1) / p.kb_per_cta;
    const int smem = smem_bytes_for(stages, stage_bytes, 0);
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(tc_gemm_tn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (e != cudaSuccess) {
            set_error("tc wgrad: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
            return NPFK_ERR_CUDA;
        }
        attr_done = true;
    }
    dim3 grid((unsigned)gx, (unsigned)gy);
    tc_gemm_tn_kernel<<<grid, kThreads, smem, st>>>(tmG, tmX, p);
    count_launch();
    return check_launch("tc_gemm_tn");
}

}  // namespace tc
}  // namespace npfk

extern "C" int npfk_tc_abort_code(void) {
    int v = 0;
    if (cudaMemcpyFromSymbol(&v, npfk::tc::g_tc_abort, sizeof(int)) != cudaSuccess) {
        (void)cudaGetLastError();
        return -1;
    }
    if (v != 0) {
        const int zero = 0;
        (void)cudaMemcpyToSymbol(npfk::tc::g_tc_abort, &zero, sizeof(int));
    }
    return v;
}
