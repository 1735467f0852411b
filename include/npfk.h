/* npfk -- C-ABI of the B200 (sm_100a) kernels behind the npf neural-process hot path.
 *
 * This is the drop-in boundary.  Every entry point replaces a stretch of eager ATen calls
 * made by the reference's Python modules (citations are file:line in the reference checkout).
 * The reference is pure Python, so the binding a maintainer adds is ctypes.
 *
 * Conventions
 *   - plain pointers and sizes only; all pointers are DEVICE pointers unless stated otherwise;
 *   - fp32 tensors, row-major; "ld*" arguments are row strides in elements;
 *   - `stream` is a cudaStream_t passed as void*; every call only enqueues work (no device sync);
 *   - the library never allocates, frees or retains caller memory;
 *   - return value: NPFK_OK (0) or a negative NPFK_ERR_*; text via npfk_last_error() (thread-local);
 *   - outputs documented as "+=" are accumulated with atomics: the caller zero-fills them first.
 */
#ifndef NPFK_H_
#define NPFK_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NPFK_OK 0
#define NPFK_ERR_INVALID (-1)     /* bad shape / argument */
#define NPFK_ERR_UNSUPPORTED (-2) /* configuration outside what the kernels implement */
#define NPFK_ERR_CUDA (-3)        /* CUDA runtime error at launch */

#define NPFK_MODE_FP32 0 /* fp32 FFMA arithmetic: the <=1e-4 parity mode */
#define NPFK_MODE_FAST 1 /* tcgen05 tensor-core arithmetic: tf32 operands read from the fp32 tensors, fp32 accumulate in TMEM */

/* ---- library ---------------------------------------------------------------------------- */
int npfk_version(void);
const char* npfk_last_error(void);
/* number of kernels launched by this library since process start (bench.py's gpu_launches) */
int64_t npfk_launch_count(void);

/* ---- Linear / MLP layers ------------------------------------------------------------------
 * Replaces nn.Linear + ReLU chains of MLP.forward (npf/architectures/mlp.py:95-109), the sum-merge
 * relu(x1 + MLP(x2)) (npf/architectures/encoders.py:175-183), the SetConv resizer
 * (npf/architectures/setcnn.py:266-268), 1x1 convolutions (npf/architectures/cnn.py:214) and the
 * split-weight r_z_merger (npf/neuralproc/base.py:570-575).
 *
 *   y[m,n] = act( sum_k x[m,k] w[n,k] + bias[n] + res[(m % res_rows), n] )      act = relu or identity
 * bias/res may be NULL.  res_rows <= 0 means res_rows = M.
 */
int npfk_linear_fwd(const float* x, int64_t ldx, const float* w, int64_t ldw, const float* bias,
                    const float* res, int64_t ldres, int64_t res_rows, float* y, int64_t ldy,
                    int64_t M, int N, int K, int relu, int mode, void* stream);
/* g[m,n] = dy[m,n] * (y[m,n] > 0)  (ReLU backward; g may alias dy) */
int npfk_relu_bwd(const float* dy, int64_t lddy, const float* y, int64_t ldy, float* g, int64_t ldg,
                  int64_t M, int N, void* stream);
/* dx[m,k] = (mask == NULL || mask[m,k] > 0) ? sum_n g[m,n] w[n,k] : 0
 * mask = saved output of the ReLU layer below: its backward is fused into this epilogue. */
int npfk_linear_bwd_dx(const float* g, int64_t ldg, const float* w, int64_t ldw, const float* mask, int64_t ldmask,
                       float* dx, int64_t lddx, int64_t M, int N, int K, int mode, void* stream);
/* dw[n,k] += sum_m g[m,n] x[m,k] ; db[n] += sum_m g[m,n]  (db may be NULL) */
int npfk_linear_bwd_dw(const float* g, int64_t ldg, const float* x, int64_t ldx, float* dw, int64_t lddw,
                       float* db, int64_t M, int N, int K, int mode, void* stream);
/* dgrad + wgrad of one layer in one call: dx = (g . w) masked by (x > 0) when `masked`, dw += g^T . x,
 * db += sum_m g.  In NPFK_MODE_FAST with N == K == 128 this is ONE tcgen05 kernel that reads g and x from
 * HBM once (the ReLU mask is taken from x, the layer's saved input, which is also the wgrad operand);
 * every other shape/mode runs npfk_linear_bwd_dw followed by npfk_linear_bwd_dx.
 * `masked` is a bit set: NPFK_BWD_MASKED (1) as above; NPFK_BWD_DX_ACCUMULATE (2): dx += instead of dx = (the dX
 * tiles leave the kernel through a TMA reduce-add, so a gradient that already sits in dx -- the other branch of a
 * fan-out -- is summed without a separate pass); only the fused tensor-core path supports it, every other
 * shape/mode returns NPFK_ERR_UNSUPPORTED for that bit. */
#define NPFK_BWD_MASKED 1
#define NPFK_BWD_DX_ACCUMULATE 2
int npfk_linear_bwd_fused(const float* g, int64_t ldg, const float* x, int64_t ldx, const float* w, int64_t ldw,
                          int masked, float* dx, int64_t lddx, float* dw, int64_t lddw, float* db, int64_t M, int N,
                          int K, int mode, void* stream);
/* K1b: npfk_linear_bwd_fused for TWO consecutive 128 -> 128 layers in one kernel (NPFK_MODE_FAST arithmetic only).
 * Upper layer (w1, saved input x1 = a ReLU output), lower layer (w0, saved input x0):
 *   dw1 += g^T . x1 ; db1 += sum_m g ; gm = (g . w1) masked by (x1 > 0), rounded to tf32, kept on chip
 *   dw0 += gm^T . x0 ; db0 += sum_m gm ; dx = (gm . w0) [masked by (x0 > 0) when flags & NPFK_BWD_MASKED]
 * flags & NPFK_BWD_DX_ACCUMULATE: dx += (TMA reduce-add).  db1 / db0 may be NULL.  Per 128 rows the kernel reads g, x1,
 * x0 and writes dx: four activation-sized tensors for two layers instead of six (autograd of mlp.py:95-109).
 * All matrices [*, 128] fp32; returns NPFK_ERR_UNSUPPORTED for M < 128 or unaligned operands. */
int npfk_linear_bwd_fused2(const float* g, int64_t ldg, const float* x1, int64_t ldx1, const float* w1, int64_t ldw1,
                           const float* x0, int64_t ldx0, const float* w0, int64_t ldw0, int flags, float* dx,
                           int64_t lddx, float* dw1, int64_t lddw1, float* db1, float* dw0, int64_t lddw0, float* db0,
                           int64_t M, void* stream);
/* ---- K1: a ReLU-MLP chain of L <= 8 layers of 128 -> 128 as ONE kernel (NPFK_MODE_FAST only) ---------------
 * Replaces MLP.forward (npf/architectures/mlp.py:95-109), the sum-merge of MergeFlatInputs
 * (npf/architectures/encoders.py:175-183) and, with the tail, decode + sum_log_prob
 * (npf/neuralproc/base.py:347-353,116; npf/losses.py:18-24).
 *
 *   h_0 = x ; h_{l+1} = act_l( h_l . w[l]^T + bias[l] + (l == res_layer ? res[m % res_rows] : 0) ),  act_l = relu when bit l
 *   of relu_mask is set.  The activations stay in tensor memory between layers; out[l] != NULL asks for h_{l+1} in HBM
 *   ([M,128] contiguous; what the backward pass needs, or the chain's result).  w[l]: [128,128] row-major, tf32-rounded
 *   (npfk_round_tf32), contiguous.  w / bias / out are HOST arrays of L device pointers (bias, out and their entries may be
 *   NULL).
 *   Tail (head_w != NULL): suff[m,:] = h_L[m,:] . head_w^T + head_b  (head_w [2,128] fp32; suff [M,2] may be NULL), and with
 *   y != NULL also sum_lp[m / T] += log N(y[m % y_rows]; suff[m,0], 0.01 + 0.99 softplus(suff[m,1]))  (sum_lp zeroed by
 *   the caller).  Returns NPFK_ERR_UNSUPPORTED in NPFK_MODE_FP32 (run the layers through npfk_linear_fwd). */
int npfk_mlp_chain_fwd(const float* x, int64_t ldx, int64_t M, int L, const float* const* w, const float* const* bias,
                       int relu_mask, int res_layer, const float* res, int64_t ldres, int64_t res_rows, float* const* out,
                       const float* head_w, const float* head_b, float* suff, const float* y, int64_t y_rows, int64_t T,
                       float* sum_lp, int mode, void* stream);
/* Backward of the tail above in one pass over h = h_L [M,128]: per row dsuff = dsuff_in (may be NULL) + dsum_lp[m / T] *
 * d lp / d suff (dsum_lp may be NULL); g[m,:] = dsuff . head_w, zeroed where h <= 0 when `masked` (g may be NULL);
 * dw[2,128] += dsuff^T . h ; db[2] += sum_m dsuff (either may be NULL). */
int npfk_nll_head_bwd(const float* h, int64_t ldh, const float* head_w, const float* suff, const float* y, int64_t y_rows,
                      int64_t T, const float* dsum_lp, const float* dsuff_in, int masked, float* g, float* dw, float* db,
                      int64_t M, void* stream);
/* Bytes of caller-provided workspace `op` needs for the given sizes (the library never allocates): NPFK_OP_MHA_BWD
 * (B, Nq, H) -> delta_ws of npfk_mha_bwd; NPFK_OP_BN_STATS (C, -, -) -> scratch of npfk_bn_stats; 0 for every other op. */
#define NPFK_OP_MHA_BWD 1
#define NPFK_OP_BN_STATS 2
int64_t npfk_workspace_bytes(int op, int64_t d0, int64_t d1, int64_t d2);
/* dst[i] = round-to-nearest tf32(src[i]) (weights, before they meet tcgen05.mma.kind::tf32 which truncates) */
int npfk_round_tf32(const float* src, float* dst, int64_t n, void* stream);
/* != 0 when a tcgen05 kernel gave up on a bounded mbarrier wait since the last call (synchronises; tests only) */
int npfk_tc_abort_code(void);
/* out[r, n] = sum_{j} g[(j*rows + r), n] for j < reps  (gradient of a row-broadcast residual / bias) */
int npfk_rows_fold_sum(const float* g, int64_t ldg, float* out, int64_t ldo, int64_t rows, int64_t reps, int N,
                       void* stream);

/* ---- broadcast merge ------------------------------------------------------------------------
 * out[z,b,t,:] = relu(a[b,t,:] + r[z,b,:])     a:[B,T,C]  r:[Z,B,C]  out:[Z,B,T,C]
 * CNP/LNP decoder merge with the per-task representation hoisted out of the target loop
 * (npf/neuralproc/np.py:103-110 + encoders.py:175-183) and merge_r_z with W split into [W_r | W_z]
 * (npf/neuralproc/base.py:554-575).
 */
int npfk_bcast_add_relu_fwd(const float* a, const float* r, float* out, int Z, int B, int64_t T, int C, void* stream);
/* da[b,t,:] = sum_z dout*mask ; dr[z,b,:] += sum_t dout*mask   (mask = out > 0); da/dr may be NULL */
int npfk_bcast_add_relu_bwd(const float* dout, const float* out, float* da, float* dr, int Z, int B, int64_t T,
                            int C, void* stream);

/* ---- mean over the set dimension ------------------------------------------------------------
 * torch.mean(R, dim=1, keepdim=True) in CNP.encode_globally (np.py:95) / AttnLNP.rep_to_lat_input
 * (attnnp.py:181).  x:[B,N,C] -> out:[B,C];  bwd: dx[b,n,c] = dout[b,c]/N
 */
int npfk_set_mean_fwd(const float* x, float* out, int B, int64_t N, int C, void* stream);
int npfk_set_mean_bwd(const float* dout, float* dx, int B, int64_t N, int C, void* stream);

/* ---- multi-head cross attention --------------------------------------------------------------
 * Replaces _make_multiheaded / DotAttender.score / softmax / bmm / _concatenate_multiheads
 * (npf/architectures/attention.py:507-527, 204-220, 129-164): heads are read in place from the
 * feature dimension (no permute copies), logits are never materialised (online softmax).
 *   q:[B,Nq,H*D] k,v:[B,Nk,H*D] -> out:[B,Nq,H*D], lse:[B,H,Nq] (log-sum-exp of scaled logits)
 * D in {4,8,16,32,64,128}.  H=1, D=kq_size gives DotAttender("scaledot").
 * mode NPFK_MODE_FAST with D in {16,32} and H*D a multiple of 32 runs the tcgen05/TMEM kernels (tf32 operands).
 */
int npfk_mha_fwd(const float* q, const float* k, const float* v, float* out, float* lse, int B, int64_t Nq,
                 int64_t Nk, int H, int D, float scale, int mode, void* stream);
/* dq,dk,dv have the layouts of q,k,v (overwritten).  delta_ws: caller workspace of npfk_mha_bwd_ws_floats(B,Nq,H)
 * floats, 16-byte aligned: [B,H,Nqp] log2-domain lse, then [B,H,Nqp] delta = sum_d dO*O, Nqp = Nq rounded up to 64.
 * delta_ready != 0 (tensor-core path only): the delta half already holds the values of the real queries -- written by
 * the kernel that produced dO (npfk_add_layernorm_bwd, attn_delta) -- so O and dO are not read a second time. */
int64_t npfk_mha_bwd_ws_floats(int B, int64_t Nq, int H);
int npfk_mha_bwd(const float* q, const float* k, const float* v, const float* out, const float* lse,
                 const float* dout, float* dq, float* dk, float* dv, float* delta_ws, int B, int64_t Nq, int64_t Nk,
                 int H, int D, float scale, int delta_ready, int mode, void* stream);

/* ---- residual + LayerNorm ---------------------------------------------------------------------
 * LN(a + b) of TransformerAttender.forward (attention.py:583-588).  b may be NULL.
 * rows x C, C <= 1024.  mean/rstd:[rows] are saved for backward.
 */
int npfk_add_layernorm_fwd(const float* a, const float* b, const float* gamma, const float* beta, float* out,
                           float* mean, float* rstd, int64_t rows, int C, float eps, void* stream);
/* ds = d(a+b) ; dgamma, dbeta are "+=".  dout2 (may be NULL) is a second gradient of the output, added to dout while
 * it is read: where the LayerNorm output feeds two consumers (c and MLP(c) in the transformer block) the fan-out sum
 * costs one extra read here instead of a separate three-tensor add. */
int npfk_add_layernorm_bwd(const float* dout, const float* dout2, const float* a, const float* b, const float* gamma,
                           const float* mean, const float* rstd, float* ds, float* dgamma, float* dbeta,
                           int64_t rows, int C, float* attn_delta, int H, int64_t Nq, int64_t Nqp, void* stream);
/* attn_delta (may be NULL; needs C == 16*H <= 128, Nq % 8 == 0, 16-byte aligned): `a` is the output O of npfk_mha_fwd with H heads of 16 and
 * rows = B*Nq, so ds is the dO of npfk_mha_bwd: attn_delta[(b*H + h)*Nqp + q] = sum_d ds[row, 16h+d] * a[row, 16h+d]
 * is written while both are in registers (the delta half of npfk_mha_bwd's workspace, see delta_ready). */

/* ---- SetConv with the exponential RBF ----------------------------------------------------------
 * Replaces SetConv.forward + ExpRBF.forward up to (not including) the resizer Linear
 * (npf/architectures/setcnn.py:254-265, 126-142); x_dim = 1.
 *   xk:[B,Nk] (batch stride sxk, 0 = shared grid)  xq:[B,Nq] (batch stride sxq)  v:[B,Nk,C]
 *   out:[B,Nq,ldo], ldo >= C+1: columns [0,C) = sum_k softmax_k(-(|xk-xq|/sigma)^p) v_k, column C =
 *   density = sum_k exp(-(|xk-xq|/sigma)^p).  sigma = 1e-5 + softplus(*theta), theta a device scalar.
 * `unshared` != 0 selects the fork's UnsharedExpRBF (setcnn.py:14-83): theta has 2 entries (weights,
 * density) and weights = exp0 / (density + 1e-8) instead of a softmax.
 */
int npfk_setconv_fwd(const float* xk, int64_t sxk, const float* xq, int64_t sxq, const float* v,
                     const float* theta, float* out, int64_t ldo, int B, int64_t Nk, int64_t Nq, int C, float p,
                     int unshared, void* stream);
/* dv:[B,Nk,C] "+=" ; dtheta:[1 or 2] "+=" ; positions get no gradient (they are data).
 * dv may be NULL when C <= 4 (the values are data too, as for the context set): only dtheta is computed. */
int npfk_setconv_bwd(const float* xk, int64_t sxk, const float* xq, int64_t sxq, const float* v,
                     const float* theta, const float* out, const float* dout, int64_t ldo, float* dv,
                     float* dtheta, int B, int64_t Nk, int64_t Nq, int C, float p, int unshared, void* stream);

/* SetConv with the resizer Linear hoisted onto the keys (grid -> many targets; setcnn.py:254-268):
 *   out[b,q,:] = sum_k w_qk v[b,k,:] + density_q * wd + bias      v:[B,Nk,C] = values already multiplied by
 * resizer.weight[:, :C_in]^T, wd:[C] = resizer.weight[:, C_in], bias:[C] or NULL.  out:[B,Nq,C] (row pitch C).
 * stats:[B,Nq,2] = (density, softmax normaliser) saved for the backward pass.  Same RBF conventions as above.
 * Keys whose weight is below 2^-40 of the row maximum (UnsharedExpRBF: below 1e-20) are skipped. */
int npfk_setconv_interp_fwd(const float* xk, int64_t sxk, const float* xq, int64_t sxq, const float* v,
                            const float* theta, const float* wd, const float* bias, float* out, float* stats, int B,
                            int64_t Nk, int64_t Nq, int C, float p, int unshared, void* stream);
/* dv:[B,Nk,C] "+=" ; dwd:[C] "+=" ; dbias:[C] "+=" or NULL ; dtheta:[1 or 2] "+=". */
int npfk_setconv_interp_bwd(const float* xk, int64_t sxk, const float* xq, int64_t sxq, const float* v,
                            const float* theta, const float* wd, const float* stats, const float* dout, float* dv,
                            float* dwd, float* dbias, float* dtheta, int B, int64_t Nk, int64_t Nq, int C, float p,
                            int unshared, void* stream);

/* ---- context / target split on the device (npf/utils/datasplit.py:60-255) ------------------------
 * idx:[B,n] int64 <- n distinct indices of [lo, lo+N) per task, uniform over subsets and orders (what
 * GetRandomIndcs.__call__ draws with numpy on the host, datasplit.py:130-139); `shared` != 0: one draw for all
 * tasks (is_batch_share).  Deterministic in (seed, task). N <= 57344. */
int npfk_random_subset(int64_t* idx, int B, int N, int n, int64_t lo, int64_t seed, int shared, void* stream);
/* dst[b,i,:] = src[b, idx[b,i], :]  (CntxtTrgtGetter.select, datasplit.py:246-255); src:[B,N,D] dst:[B,n,D].
 * An index outside [0,N) writes zeros and sets *bad (may be NULL) instead of faulting. */
int npfk_gather_rows(const float* src, const int64_t* idx, float* dst, int B, int64_t N, int64_t n, int D, int* bad,
                     void* stream);

/* ---- 1-D depth-separable residual conv blocks, channel-last [B,L,C] ------------------------------
 * Replaces BatchNorm1d -> ReLU -> Conv1d(groups=C) of ResConvBlock.forward (npf/architectures/cnn.py:
 * 204-215, npf/utils/helpers.py:395-398) without the NLC<->NCL permutes (cnn.py:363-375).
 */
/* per-channel batch statistics of x:[rows,C]: stats[0:C] = mean, stats[C:2C] = biased variance.
 * If running_mean != NULL: running <- (1-momentum) running + momentum * (mean, unbiased var).
 * If num_batches_tracked != NULL: *num_batches_tracked += 1 (BatchNorm1d's int64 buffer). */
int npfk_bn_stats(const float* x, int64_t rows, int C, double* scratch /* [2C + 1], see npfk_workspace_bytes */, float* stats,
                  float* running_mean, float* running_var, float momentum, int64_t* num_batches_tracked,
                  void* stream);
/* npfk_bn_stats without caller scratch, for callers that cannot allocate.  No counterpart in the reference.
 * Same arguments and results as npfk_bn_stats minus `scratch`; the [2C + 1]
 * doubles live in a static __device__ array inside the library (64 KB + 256 B per GPU that loads it, held until
 * exit).  C <= NPFK_BN_STATS_STATIC_MAX_C, else NPFK_ERR_INVALID with the limit in npfk_last_error().  Calls no
 * cudaMalloc / cudaFree / allocating routine; every call clears and then shares the one array per GPU, so two calls
 * that overlap on the same GPU (different streams, or graphs replayed concurrently) race: serialise them. */
#define NPFK_BN_STATS_STATIC_MAX_C 4096
int npfk_bn_stats_static(const float* x, int64_t rows, int C, float* stats, float* running_mean, float* running_var,
                         float momentum, int64_t* num_batches_tracked, void* stream);
/* y[b,l,c] = sum_j w[c,j] * h[b,l+j-ks/2,c] + bias[c] + res[b,l,c]   (zero padding of h outside [0,L))
 *   h = act((x - mean[c]) * gamma[c]*rsqrt(var[c]+eps) + beta[c]),  act = relu if `relu` else identity
 * mean == NULL: no normalisation (Normalization=nn.Identity).  mean/var are the batch statistics from
 * npfk_bn_stats (training) or the running statistics (eval).  bias/res may be NULL.  ks odd, <= 31. */
int npfk_dwconv_fwd(const float* x, const float* mean, const float* var, const float* gamma, const float* beta,
                    float eps, const float* w, const float* bias, const float* res, float* y, int B, int L, int C,
                    int ks, int relu, void* stream);
/* dh[b,l,c] = act'(.) * sum_j w[c,j] dy[b,l-j+ks/2,c]  (gradient w.r.t. the normalised pre-activation);
 * dw:[C,ks] "+=", dbias:[C] "+=" (may be NULL), sums:[2C] "+=" (may be NULL): sum dh, sum dh*xhat. */
int npfk_dwconv_bwd(const float* dy, const float* x, const float* mean, const float* var, const float* gamma,
                    const float* beta, float eps, const float* w, float* dh, float* dw, float* dbias, float* sums,
                    int B, int L, int C, int ks, int relu, void* stream);
/* BatchNorm backward given dh and sums from npfk_dwconv_bwd.
 * training: dx = gamma*invstd*(dh - s1/n - xhat*s2/n) ; eval (training=0): dx = gamma*invstd*dh.
 * dgamma += s2, dbeta += s1.  dx may alias dh. */
int npfk_bn_bwd(const float* dh, const float* x, const float* mean, const float* var, const float* gamma,
                float eps, const float* sums, float* dx, float* dgamma, float* dbeta, int64_t rows, int C,
                int training, void* stream);
/* AvgPool1d(2) / its gradient;  x:[B,L,C] -> [B,L/2,C] */
int npfk_avgpool2_fwd(const float* x, float* y, int B, int L, int C, void* stream);
int npfk_avgpool2_bwd(const float* dy, float* dx, int B, int L, int C, void* stream);
/* F.interpolate(mode="linear", scale_factor=2, align_corners=True); y row stride ldy (writes into a concat buffer) */
int npfk_upsample2_fwd(const float* x, float* y, int64_t ldy, int B, int L, int C, void* stream);
/* dx:[B,L,C] "+=" from dy:[B,2L,*] with row stride lddy */
int npfk_upsample2_bwd(const float* dy, int64_t lddy, float* dx, int B, int L, int C, void* stream);
/* strided 2-D copy: dst[r, 0:cols] = src[r, 0:cols] */
int npfk_copy2d(const float* src, int64_t lds, float* dst, int64_t ldd, int64_t rows, int cols, void* stream);

/* ---- predictive distribution and losses -------------------------------------------------------------
 * decode tail (npf/neuralproc/base.py:347-353,116) + Normal.log_prob + sum over targets
 * (npf/losses.py:18-24, npf/utils/helpers.py:15-17).
 *   suff:[Z,B,T,2*ydim] -> loc, scale:[Z,B,T,ydim] with scale = 0.01 + 0.99*softplus(.)
 */
int npfk_loc_scale_fwd(const float* suff, float* loc, float* scale, int64_t n, int ydim, void* stream);
int npfk_loc_scale_bwd(const float* suff, const float* dloc, const float* dscale, float* dsuff, int64_t n,
                       int ydim, void* stream);
/* sum_lp[z,b] += sum_{t,j} log N(y[b,t,j]; loc, scale) computed straight from suff; y:[B,T,ydim] */
int npfk_gauss_nll_fwd(const float* suff, const float* y, float* sum_lp, int Z, int B, int64_t T, int ydim,
                       void* stream);
/* dsuff = d(sum_lp[z,b]) * d log p / d suff */
int npfk_gauss_nll_bwd(const float* suff, const float* y, const float* dsum_lp, float* dsuff, int Z, int B,
                       int64_t T, int ydim, void* stream);
/* latent sufficient statistics (base.py:538-545): loc = suff[..., :zd], scale = 0.1 + 0.9*sigmoid(suff[..., zd:]) */
int npfk_latent_stats_fwd(const float* suff, float* loc, float* scale, int64_t n, int zd, void* stream);
int npfk_latent_stats_bwd(const float* suff, const float* dloc, const float* dscale, float* dsuff, int64_t n,
                          int zd, void* stream);
/* z[s,i] = loc[i] + scale[i]*eps[s,i]  (rsample, base.py:512) and its gradients (dloc, dscale "=") */
int npfk_rsample_fwd(const float* loc, const float* scale, const float* eps, float* z, int S, int64_t n, void* stream);
int npfk_rsample_bwd(const float* dz, const float* eps, float* dloc, float* dscale, int S, int64_t n, void* stream);
/* kl[i] = sum_j KL(N(lq,sq) || N(lp,sp)) over the last dim zd (npf/losses.py:146-148) */
int npfk_kl_diag_fwd(const float* lq, const float* sq, const float* lp, const float* sp, float* kl, int64_t n,
                     int zd, void* stream);
int npfk_kl_diag_bwd(const float* lq, const float* sq, const float* lp, const float* sp, const float* dkl,
                     float* dlq, float* dsq, float* dlp, float* dsp, int64_t n, int zd, void* stream);

/* ---- input validation (npf/neuralproc/base.py:241-247, npf/utils/helpers.py:55-57) ------------------------
 * *flag |= 1 if any x outside [lo,hi] (or NaN).  flag is a device int32 the caller zero-fills and reads when
 * it chooses to (the reference forces a device->host sync every training forward). */
int npfk_range_flag(const float* x, int64_t n, float lo, float hi, int* flag, void* stream);

/* ---- optimiser on the flat parameter buffer (data-parallel training step) ------------------------------------
 * torch.optim.Adam semantics (no amsgrad, no weight decay): utils/train.py:243-260 uses Adam(lr=1e-3).
 * step_count is the 1-based step index. grad_scale multiplies the gradient first (e.g. 1/world_size). */
int npfk_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, float lr,
                   float beta1, float beta2, float eps, int step_count, float grad_scale, void* stream);
/* The same with the step index kept in device memory: *step_count_dev is incremented first (one-thread kernel), then
 * used as the 1-based step index.  Nothing in the launch depends on host state, so the optimiser step can be recorded
 * into a CUDA graph together with forward, backward and the gradient all-reduce (training.FlatAdamTrainer.capture). */
int npfk_adam_step_dev(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, float lr,
                       float beta1, float beta2, float eps, int* step_count_dev, float grad_scale, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* NPFK_H_ */
