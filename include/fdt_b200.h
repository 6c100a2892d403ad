/* C ABI of libfdt_b200.so — the hand-written sm_100a kernels of faster-distributed-training_b200.
 *
 * Every entry point takes raw device pointers plus the CUDA stream to launch on, allocates
 * nothing unless stated, and returns 0 on success, a cudaError_t value if a launch failed, or a
 * negative value for an argument it rejects.  Activations are bf16, NHWC viewed as [P, C]
 * (P = N*H*W); statistics, coefficients, parameters, gradients and optimizer state are fp32.
 * The reference project (SuperbTUM/Faster-Distributed-Training) ships no native code at all;
 * the "ref:" notes name the Python it replaces.
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* activation codes */
enum { FDT_ACT_NONE = 0, FDT_ACT_RELU = 1, FDT_ACT_CELU = 2, FDT_ACT_GELU = 3, FDT_ACT_TANH = 4 };

/* ---- channel-wise (batch-norm family) kernels, csrc/chan.cu ------------------------------------
 * ref: FusedConvBN2DFunction normalisation + batch_norm_backward (resnet.py:37-100), nn.BatchNorm2d */

/* sums[0:C] = sum_p a'[p,c], sums[C:2C] = sum_p a'[p,c]*b[p,c];  a' = a, or a*act'(out_act) when
 * out_act != NULL (then a' is also stored to g_store if non-NULL).  sums is cleared by the call. */
int fdt_chan_reduce2(const void* a, const void* b, const void* out_act, void* g_store, float* sums,
                     int64_t P, int C, int act, cudaStream_t st);
/* out = act(A[c]*x1 + B[c]*x2 + Cc[c] + res);  coef = A|B|Cc (3*C floats); x2, res may be NULL */
int fdt_chan_affine_act(const void* x1, const void* x2, const void* res, void* out, const float* coef,
                        int64_t P, int C, int act, cudaStream_t st);
/* form 0: reference batch-stat form, 1: BatchNorm2d train (updates running stats), 2: BatchNorm2d eval.
 * ctx (5*C floats: mean|inv|scale|k|gamma) is what the backward finalise reads. */
int fdt_bn_finalize_fwd(const float* sums, float* coef, float* ctx, const float* gamma, const float* beta,
                        float* running_mean, float* running_var, int C, float N, float eps, float momentum,
                        int form, cudaStream_t st);
int fdt_bn_finalize_bwd(const float* sums, const float* ctx, float* coef, float* dgamma, float* dbeta,
                        int C, float N, int form, cudaStream_t st);
/* statistics -> coefficients -> normalise [+res] + act behind one call.  scratch: 5*C floats of
 * caller-provided working memory (contents irrelevant on entry, garbage afterwards). */
int fdt_bn_act_fwd(const void* y, const void* res, void* out, float* ctx, float* scratch,
                   const float* gamma, const float* beta, float* running_mean, float* running_var,
                   int64_t P, int C, float eps, float momentum, int form, int act, cudaStream_t st);
/* one-launch variants with the per-channel coefficients derived inside the normalisation kernel (no separate finalize
 * launch): forward from sums = (Σy, Σy²) — writes ctx [5,C] and the running statistics; backward from sums = (Σg, Σg·y)
 * and the forward ctx — writes dgamma / dbeta (may be NULL).  fdt_chan_reduce2_acc: fdt_chan_reduce2 without clearing sums. */
int fdt_bn_act_fwd_fused(const void* y, const void* res, void* out, float* ctx, const float* sums, const float* gamma,
                         const float* beta, float* running_mean, float* running_var, int64_t P, int C, float eps,
                         float momentum, int form, int act, cudaStream_t st);
int fdt_bn_bwd_apply(const void* g, const void* y, void* dy, const float* sums, const float* ctx, float* dgamma,
                     float* dbeta, int64_t P, int C, int form, cudaStream_t st);
/* Backward of a ReLU unit without a residual input, mask re-derived from y with the forward's coefficients (ctx, beta or
   NULL) instead of read from the activated output: pass 1 accumulates (accumulate != 0) or writes sums[2C] = sum g*m,
   sum g*m*y; pass 2 writes dy and dgamma / dbeta.  C / 8 must divide 256. */
int fdt_chan_reduce2_masky(const void* g, const void* y, const float* ctx, const float* beta, float* sums, int accumulate,
                           int64_t P, int C, cudaStream_t st);
int fdt_bn_bwd_apply_masky(const void* g, const void* y, void* dy, const float* sums, const float* ctx, const float* beta,
                           float* dgamma, float* dbeta, int64_t P, int C, int form, cudaStream_t st);
int fdt_chan_reduce2_acc(const void* a, const void* b, const void* out_act, void* g_store, float* sums, int64_t P, int C,
                         int act, cudaStream_t st);
/* same, when scratch[0 .. 2C) already holds Σy, Σy² per channel (written by the conv GEMM epilogue: `sums` of
 * fdt_conv_fwd / fdt_gemm_nt) — skips the statistics pass over y; forms 0 and 1 only */
int fdt_bn_act_fwd_presummed(const void* y, const void* res, void* out, float* ctx, float* scratch,
                   const float* gamma, const float* beta, float* running_mean, float* running_var,
                   int64_t P, int C, float eps, float momentum, int form, int act, cudaStream_t st);
/* Same operation as fdt_bn_act_fwd for callers that cannot supply working memory and must not
 * allocate any (a host that gave the whole GPU to its own pool, a bare C program, a loop that has to
 * keep clear of cudaMalloc/cudaFree): the 5*C floats live in a fixed array inside the library, so
 * C is limited to 4096 channels (80 KiB per GPU, held until exit); a larger C returns -2 and prints
 * why.  The first call on a device resolves the array's address and should be made outside stream
 * capture.  All calls on a device share the array: calls ordered on one stream are fine, calls that
 * may overlap on the GPU race and must be serialised by the caller.  Allocates nothing, ever.
 * This entry point has no counterpart in the reference (which ships no native code). */
int fdt_bn_act_fwd_noalloc(const void* y, const void* res, void* out, float* ctx,
                           const float* gamma, const float* beta, float* running_mean, float* running_var,
                           int64_t P, int C, float eps, float momentum, int form, int act, cudaStream_t st);

/* ---- row-wise kernels, csrc/rowwise.cu ----------------------------------------------------------
 * ref: LayerNorm (transformer.py:230-242), Embeddings (:132-156), sub-layer dropout+residual (:245-275),
 *      mixup_criterion(_meta) (resnet50_test.py:451-457) */
/* Linear backward preparation: gz = gout ∘ act'(saved) (bf16 GEMM operand) and db += Σ_rows gz in one pass.
 * gout [T,N] bf16|fp32; saved [T,N] bf16 (z for GELU, the output for ReLU/tanh) or NULL; gz, db may be NULL. */
int fdt_linear_bwd_prep(const void* gout, int gout_is_f32, const void* saved, void* gz, float* db, int64_t T, int N,
                        int act, cudaStream_t st);
int fdt_linear_bwd_prep_dropout(const void* gout, int gout_is_f32, const void* saved, void* gz, float* db, int64_t T, int N,
                                int act, float p, const int64_t* seed_off, cudaStream_t st);
/* out = dropout(act(z), p) in one pass (bf16, n % 8 == 0); mask = Philox(seed_off[0], seed_off[1]) indexed by element/4 */
int fdt_act_dropout_fwd(const void* z, void* out, int64_t n, int act, float p, const int64_t* seed_off, cudaStream_t st);
int fdt_ln_fwd(const float* x, const float* a, const float* b, void* y_bf16, float* stats, int M, int D,
               float eps, cudaStream_t st);
int fdt_ln_bwd(const float* x, const float* a, const float* stats, const void* g_bf16, float* dx_acc,
               float* da, float* db, int M, int D, float eps, cudaStream_t st);
int fdt_embed_fwd(const int64_t* tokens, const int64_t* types, const float* tok_w, const float* pos_w,
                  const float* seg_w, float* out, int rows, int L, int D, float scale, cudaStream_t st);
int fdt_embed_bwd(const int64_t* tokens, const int64_t* types, const float* g, float* dtok, float* dpos,
                  float* dseg, int rows, int L, int D, float scale, cudaStream_t st);
int fdt_dropout_add(const void* x, int x_is_bf16, const float* res, float* out, int64_t n, float p,
                    const int64_t* seed_off, cudaStream_t st);
int fdt_dropout_bwd(const float* g, void* dx, int dx_is_bf16, int64_t n, float p, const int64_t* seed_off,
                    cudaStream_t st);
/* out[0] += mean loss, out[1] += lambda-weighted correct count (both must be 0 on entry) */
int fdt_mixup_ce(const float* logits, const int64_t* ya, const int64_t* yb, const float* lam, int lam_n,
                 float* out, float* dlogits, int B, int C, cudaStream_t st);

/* + dlam [B] = d loss / d lam_i for per-sample (learnable, meta-mixup) mixing weights */
int fdt_mixup_ce_ex(const float* logits, const int64_t* ya, const int64_t* yb, const float* lam, int lam_n,
                    float* out, float* dlogits, float* dlam, int B, int C, cudaStream_t st);
/* pooled classification head (transformer.py:61-84, 94-101): cls = bf16(x[:, 0, :]) and its gradient (zero elsewhere);
 * out = lam * drop(z) + (1 - lam) * drop(z[perm]) (pooled-vector dropout + manifold mixup; perm NULL = no mixing), dz fp32. */
int fdt_cls_gather(const float* x, void* cls_bf16, int B, int L, int D, cudaStream_t st);
int fdt_cls_scatter(const float* g, float* dx, int B, int L, int D, cudaStream_t st);
int fdt_head_mix_fwd(const void* z, const int64_t* perm, const float* lam, void* out, int B, int D, float p,
                     const int64_t* seed_off, cudaStream_t st);
int fdt_head_mix_bwd(const void* g, const int64_t* perm, const float* lam, float* dz, int B, int D, float p,
                     const int64_t* seed_off, cudaStream_t st);

/* ---- NGD Fisher-factor refresh, csrc/ngd_eig.cu -------------------------------------------------
 * ref: OnlineNaturalGradient._precondition_directions3 (ngd_optimizer.py:225-317) */
int fdt_ngd_eig_update(const float* K, const float* L, const float* d, const float* rho, const float* trXXt,
                       float* A, float* w_coeff, float* d1, float* rho1, int n, int R, int D, float N,
                       float alpha, float eta, int max_sweeps, cudaStream_t st);
/* descs: `count` 96-byte records {K,L,d,rho,tr,A,wc,d1,rho1 pointers; int R,D; float N,alpha,eta; int pad} */
int fdt_ngd_eig_update_multi(const void* descs, int count, int max_sweeps, cudaStream_t st);

/* ---- flat-buffer optimizer kernels, csrc/optim_multi.cu -----------------------------------------
 * ref: clip_grad_norm_ / GradScaler / NGD momentum+update loops (resnet50_test.py:544-548,
 *      ngd_optimizer.py:478-506), external madgrad package */
int fdt_sqnorm(const float* g, int64_t n, float* out, cudaStream_t st);           /* *out += sum g^2 */
int fdt_clip_scale(float* g, int64_t n, const float* sq, float max_norm, float* skipped, cudaStream_t st);
int fdt_sgd_momentum(float* p, const float* d, float* buf, int64_t n, const float* lr_ptr, float lr_val,
                     float momentum, float dampening, int nesterov, int first, cudaStream_t st);
int fdt_madgrad(float* p, const float* g, float* v, float* s, float* x0, int64_t n, float lamb, float ck,
                float decay, float eps, int mirror, cudaStream_t st);
int fdt_cast_bf16(const float* x, void* y_bf16, int64_t n, cudaStream_t st);
/* CUDA-graph-safe variants: the schedule-dependent hyper-parameters live in device memory.
 * fdt_sgd_momentum_dev: hyper = {lr, momentum, dampening}; buf may be NULL (no momentum).
 * fdt_clip_wd: g = finite(sq) ? g*min(1, max_norm/(sqrt(sq)+1e-6)) + wd*p : 0  (clip_grad_norm_ + the NGD/SGD weight-decay
 *   add of ngd_optimizer.py:478-479 in one pass; a non-finite norm zeroes the step and bumps *skipped).
 * fdt_madgrad_dev: hyper = {lr + eps, k, momentum}; lamb = hyper[0]*sqrt(hyper[1]+1); tick != 0 advances k afterwards. */
int fdt_sgd_momentum_dev(float* p, const float* d, float* buf, int64_t n, const float* hyper, int nesterov, int first,
                         void* shadow_bf16 /* optional: bf16 copy of the updated parameters, written in the same pass */,
                         cudaStream_t st);
int fdt_clip_wd(float* g, const float* p, int64_t n, const float* sq, float max_norm, float wd, float* skipped,
                cudaStream_t st);
int fdt_madgrad_dev(float* p, const float* g, float* v, float* s, float* x0, int64_t n, float* hyper, float decay,
                    float eps, int mirror, int tick, cudaStream_t st);

/* ---- tcgen05 / TMEM / TMA GEMM and convolution, csrc/gemm_sm100.cu, wgrad_sm100.cu, im2col.cu ------------
 * ref: F.conv2d / F.linear / convolution_backward call sites (resnet.py:21-34,83,108,204,221; transformer.py:168-176,213-226) */

/* out[M,N] (bf16 and/or fp32) = act(A[M,K] · B[N,K]^T + bias); sums (optional, [2N], zero on entry) += per-column sum, sum^2.
 * K % 64 == 0, N % 64 == 0.  Persistent kernel, 128 x {64,128} tiles, double-buffered TMEM accumulator, TMA-store epilogue. */
int fdt_gemm_nt(const void* A, const void* B, void* out_bf16, float* out_f32, const float* bias, float* sums,
                int M, int N, int K, int act, cudaStream_t st);
/* y[Nb,H,W,Cout] = conv(x[Nb,H,W,Cin], w[Cout,k,k,Cin]), stride 1, pad = k/2, NHWC bf16 (implicit GEMM: shifted 4-D TMA boxes). */
int fdt_conv_fwd(const void* x, const void* w, void* y, float* sums, int Nb, int H, int W, int Cin, int Cout,
                 int ksize, int pad, cudaStream_t st);
/* 2-CTA form (csrc/gemm2_sm100.cu): thread-block clusters of two CTAs, tcgen05.mma.cta_group::2 on 256 x 256 tiles, each CTA
 * staging half of A and half of B; N % 256 == 0, K % 64 == 0.  out bf16 = act(A B^T + bias) [+ addend]. */
int fdt_gemm2_nt(const void* A, const void* B, void* out_bf16, const float* bias, const void* addend, int M, int N, int K,
                 int act, cudaStream_t st);
/* generalised forms: `addend` (optional bf16 tensor indexed like the output, may alias it) is added in the epilogue — the
 * other branch's gradient at a residual fork; fdt_conv_ex also takes stride 2 (TMA element strides, no im2col);
 * fdt_conv_dgrad_s2 is the data gradient of a stride-2 convolution as four parity-class implicit GEMMs (no col2im):
 * dy [Nb,H/2,W/2,Cout], wt = tap-rotated transposed filter [Cin][k][k][Cout], dx [Nb,H,W,Cin]; accumulate != 0 adds into dx. */
int fdt_gemm_nt_ex(const void* A, const void* B, void* out_bf16, float* out_f32, const float* bias, float* sums,
                   const void* addend, int M, int N, int K, int act, cudaStream_t st);
int fdt_conv_ex(const void* x, const void* w, void* y, float* sums, const void* addend, int Nb, int H, int W, int Cin,
                int Cout, int ksize, int pad, int stride, cudaStream_t st);
int fdt_conv_dgrad_s2(const void* dy, const void* wt, void* dx, int accumulate, int Nb, int H, int W, int Cin, int Cout,
                      int ksize, int pad, cudaStream_t st);
/* dw[Cout][k][k][Cin] (fp32) += alpha * dY^T (*) X  (MN-major operands, split-K).  multicast != 0: dw is an NVLS multicast
 * address and the accumulation is an in-switch all-reduce (multimem.red) of finished tiles. */
int fdt_conv_wgrad(const void* dy, const void* x, float* dw, int Nb, int H, int W, int Cin, int Cout, int ksize,
                   int pad, float alpha, int multicast, cudaStream_t st);
/* same with stride 1 or 2 (H, W are the INPUT extents; the X tensor map walks every second pixel for stride 2) */
int fdt_conv_wgrad_ex(const void* dy, const void* x, float* dw, int Nb, int H, int W, int Cin, int Cout, int ksize,
                      int pad, int stride, float alpha, int multicast, cudaStream_t st);
/* dw[N][K] (fp32) += alpha * dY[M,N]^T · X[M,K] */
int fdt_linear_wgrad(const void* dy, const void* x, float* dw, int M, int N, int K, float alpha, int multicast,
                     cudaStream_t st);
/* per-step operand layouts of the data-gradient GEMMs (transposed / tap-rotated bf16 weights) for every layer in one launch.
 * descs: `count` 32-byte device records {const bf16* src; bf16* dst; int O, T, I, mode}.  csrc/weight_prep.cu */
int fdt_dgrad_weights(const void* descs, int count, cudaStream_t st);
/* same, one 64x64 tile per block over all weights (total_tiles = sum of T*ceil(O/64)*ceil(I/64)): balanced */
/* one block per 64 x 64 tile, with a caller-built table of total_tiles x {descriptor index, tile index in that weight} */
int fdt_dgrad_weights_map(const void* descs, const void* tile_map, int total_tiles, cudaStream_t st);
/* explicit im2col / col2im (strided convs, the 3-channel stem): A[N*Ho*Wo, Kp] <-> x[N,H,W,C] */
int fdt_im2col(const void* x, void* A, int N, int H, int W, int C, int k, int stride, int pad, int Ho, int Wo, int Kp,
               cudaStream_t st);
int fdt_col2im(const void* dA, void* dx, int N, int H, int W, int C, int k, int stride, int pad, int Ho, int Wo, int Kp,
               cudaStream_t st);

/* ---- fused attention, csrc/attention_sm100.cu (L <= 512, d_head = 64; KV-tile loop with online softmax) ---------
 * ref: ScaledDotProduct / MultiheadAttention (transformer.py:180-227).  q,k,v: bf16 [B*L, H*64] with row stride ldq (column
 * slices of a packed QKV matrix are fine); o, dO: row stride ldo; dq,dk,dv: row stride ldg; dq_acc: fp32 [B*L, H*64] work
 * buffer, needed when L > 128. */
int fdt_attn_fwd(const void* q, const void* k, const void* v, void* o, float* lse, const int64_t* key_mask,
                 int B, int H, int L, int ldq, int ldo, float scale, float p_drop, const int64_t* seed_off, int mask_compat,
                 cudaStream_t st);
int fdt_attn_bwd(const void* q, const void* k, const void* v, const void* o, const void* dO, const float* lse,
                 void* dq, void* dk, void* dv, float* dq_acc, const int64_t* key_mask, int B, int H, int L, int ldq, int ldo,
                 int ldg, float scale, float p_drop, const int64_t* seed_off, int mask_compat, cudaStream_t st);

/* ---- input / head kernels, csrc/input_head.cu ------------------------------------------------------------------
 * ref: loader transforms + mixup_data(_meta) (resnet50_test.py:301-318,355-401); AdaptiveAvgPool2d + Linear (resnet.py:279-281) */
int fdt_input_aug_mixup(const void* x_u8, const int64_t* perm, const float* lam, int lam_n, const int64_t* seed_off,
                        void* out_bf16, int B, int H, int W, int pad, int train, const float* mean3, const float* std3,
                        cudaStream_t st);
/* meta-mixup (resnet50_test.py:388-401): lam_mode 1 = `lam` holds theta, lam_i = sigmoid(theta_i) computed in the kernel;
 * _bwd: dlam[b] (+=) = sum_pixels g * (partner_a - partner_b) [* sigmoid'(theta_b)], partners recomputed from the uint8 images
 * with the forward's (seed, offset) */
int fdt_input_aug_mixup_ex(const void* x_u8, const int64_t* perm, const float* lam, int lam_n, int lam_mode,
                           const int64_t* seed_off, void* out_bf16, int B, int H, int W, int pad, int train,
                           const float* mean3, const float* std3, cudaStream_t st);
int fdt_input_aug_mixup_bwd(const void* x_u8, const int64_t* perm, const float* lamsrc, int lam_mode, const int64_t* seed_off,
                            const void* g_bf16, float* dlam, int B, int H, int W, int pad, int train, const float* mean3,
                            const float* std3, cudaStream_t st);
/* lam [1] ~ Beta(alpha, alpha) and perm [B] = uniform random permutation in one launch (either may be NULL);
 * ref: mixup_data's host-side Beta + torch.randperm (resnet50_test.py:355-362) */
int fdt_mixup_sample(const int64_t* seed_off, float alpha, int B, float* lam, int64_t* perm, cudaStream_t st);
int fdt_pool_fc_fwd(const void* x, const float* W, const float* bias, float* pooled, float* logits, int N, int HW,
                    int C, int K, cudaStream_t st);
int fdt_pool_fc_bwd(const float* dlogits, const float* W, const float* pooled, void* dx, float* dW, float* db,
                    int N, int HW, int C, int K, cudaStream_t st);

/* ---- NGD element-wise tails and opt-in tf32 tensor-core products, csrc/ngd_eig.cu, ngd_gemm_sm100.cu ----------
 * ref: ngd_optimizer.py:151-168,195-218,322-324 */
int fdt_ngd_sumsq2(const float* X, const float* Xh, float* sums, int n, int64_t M, cudaStream_t st);
int fdt_ngd_rescale(const float* X, const float* Xh, const float* sums, float* out, int n, int64_t M, cudaStream_t st);
/* same with the two sums given separately ([n] each): the end of a chain of factors rescales once with the first link's input
   norm and the last link's output norm */
int fdt_ngd_rescale2(const float* X, const float* Xh, const float* p_in, const float* p_out, float* out, int n, int64_t M,
                     cudaStream_t st);
int fdt_ngd_commit_multi(const void* descs, int count, int* flags, int max_elems, cudaStream_t st);
/* W1 = A (J + wc ∘ W) for every factor (80-byte descriptors: W1,d1,rho1,W,d,rho,A,J,wc pointers + int R, D), fused
 * with the finiteness check, then the guarded hand-over: 3 launches for the whole model */
int fdt_ngd_finish_multi(const void* descs, int count, int* flags, int max_D, int max_elems, cudaStream_t st);
/* fp32-accurate tensor-core products (3xTF32 split in shared memory; M/N-major operands transposed in shared memory):
 * X [n,N,D] (x_trans=0) or [n,D,N] (x_trans=1); W [n,R,D]; H [n,N,R] = X W^T; Xhat = X - H W (layout of X);
 * sums [2n] zero on entry -> sum X^2, sum Xhat^2.  J [n,R,D] = H^T X (zeroed inside). */
int fdt_ngd_project(const float* X, const float* W, float* H, float* Xhat, float* sums, int n, int N, int D, int R,
                    int x_trans, cudaStream_t st);
int fdt_ngd_stats_j(const float* H, const float* X, float* J, int n, int N, int D, int R, int x_trans, cudaStream_t st);
/* padded-pitch forms: W and J are [n, R, Dw] with Dw = D rounded up to 4 and zero padding columns, so that an axis length that
 * is not a multiple of 4 (x_trans only: the 30522-row token embedding) runs on the tensor cores too; Dw == 0 means D.
 * fdt_ngd_stats_hh: L [n,R,R] = H^T H (the N <= D case of ngd_optimizer.py:215-217). */
int fdt_ngd_project_ex(const float* X, const float* W, float* H, float* Xhat, float* sums, int n, int N, int D, int Dw,
                       int R, int x_trans, cudaStream_t st);
int fdt_ngd_stats_j_ex(const float* H, const float* X, float* J, int n, int N, int D, int Dw, int R, int x_trans,
                       cudaStream_t st);
int fdt_ngd_stats_hh(const float* H, float* L, int n, int N, int R, cudaStream_t st);
/* K [n,R,R] = J J^T and (L != NULL) L [n,R,R] = J W^T on the tensor cores (zeroed inside; 16 <= R <= 128, R % 4 == 0, D % 4 == 0) */
int fdt_ngd_stats_kl(const float* J, const float* W, float* K, float* L, int n, int D, int R, cudaStream_t st);
/* short axis (2 <= D <= 12, R == (D+1)/2) of the [n, A, D, B] view of a gradient: x̂ = x − Wᵀ(W x) in place of two
 * inner-dimension-3 SGEMMs and a transpose; sums [2n] = Σx², Σx̂²; J [n, R, D] = Σ h xᵀ or NULL.  csrc/ngd_axis.cu */
int fdt_ngd_small_axis(const float* X, const float* W, float* Xhat, float* sums, float* J, int n, int64_t A, int D,
                       int64_t B, int R, cudaStream_t st);

/* complete preconditioning step of SMALL factors (N*D + R*D + N*R <= 24576 floats) held in shared memory, one CTA per factor:
 * out = rescaled (X - X W^T W) in the layout of X, p0 [n] = sum X^2, J [n,R,ldj] = H^T X or NULL; W, J with row pitch ldw / ldj.
 * For gradient views whose rows are not 16-byte aligned (3-channel stem, [B,1,1,1] theta) — no cuBLAS for them either. */
int fdt_ngd_small_matrix(const float* X, const float* W, int ldw, float* out, float* p0, float* J, int ldj, int n,
                         int N, int D, int R, int x_trans, cudaStream_t st);
/* K [n,R,R] = J J^T and L [n,R,R] = J W^T of the short-axis factors (R <= 16, D <= 32; J, W [n,R,D] contiguous) */
int fdt_ngd_small_kl(const float* J, const float* W, float* K, float* L, int n, int R, int D, cudaStream_t st);

/* ---- in-kernel NVLink collectives over a symmetric heap, csrc/comm.cu -------------------------------------------
 * ref: DDP bucket all-reduce (resnet50_test.py:716), FSDP all-gather / reduce-scatter (transformer_test.py:387-392).
 * ctx: host pointer to {float* buf[8]; uint32_t* sig[8]; float* scratch[8]; float* mc; int rank, world, sig_off, pad1;} */
int fdt_comm_all_reduce_mean(const void* ctx, int64_t n, int use_multicast, cudaStream_t st);
int fdt_comm_reduce_scatter_mean(const void* ctx, int64_t lo, int64_t hi, int use_multicast, cudaStream_t st);
int fdt_comm_all_gather_cast(const void* ctx, int64_t region, void* shadow_bf16, int use_multicast, cudaStream_t st);
/* bucketed forms for overlap with backward / forward: per-rank [lo, hi) (may be empty) with a grid size common to all ranks;
 * per-owner range table `sub` (device, 2*world int64 {lo_r, hi_r}, absolute element offsets; regions may differ in length).  Σg² accumulates across
 * fdt_comm_reduce_scatter_range calls until fdt_comm_take_scalar. */
int fdt_comm_reduce_scatter_range(const void* ctx, int64_t lo, int64_t hi, int blocks, int use_multicast, cudaStream_t st);
int fdt_comm_all_gather_range(const void* ctx, const int64_t* sub, void* shadow_bf16, int blocks, cudaStream_t st);
/* broadcast of one owner's large range in two phases (scatter slices, then all-gather them): owner egress = 1x the range */
int fdt_comm_bcast_range(const void* ctx, int owner, int64_t lo, int64_t hi, void* shadow_bf16, int blocks, cudaStream_t st);
int fdt_comm_barrier(const void* ctx, cudaStream_t st);
int fdt_comm_take_scalar(float* slot, float* out, cudaStream_t st);
int fdt_comm_mc_red_add(float* dst_mc, const float* src, int64_t n, float alpha, cudaStream_t st);

#ifdef __cplusplus
}
#endif
