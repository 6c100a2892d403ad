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

#ifdef __cplusplus
}
#endif
