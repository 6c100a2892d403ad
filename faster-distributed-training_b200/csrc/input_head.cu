// Input-side and head-side fused kernels of the CIFAR ResNet step.
//
//  * fdt_input_aug_mixup : uint8 NHWC batch → random crop (pad 4) + horizontal flip + normalise, per sample
//    (Philox draws made in the kernel), blended with the same transform of the permuted partner sample:
//    out_b = λ_b·T_b(x_b) + (1−λ_b)·T_{π(b)}(x_{π(b)}), written as bf16 NHWC.  One pass over 3 KB/image.
//  * fdt_pool_fc_fwd / _bwd : global average pool + Linear(C → K ≤ 32) and its gradients.
//
// Parity: the loader-side transforms + mixup_data / mixup_data_meta of the reference (resnet50_test.py:
// 301-318, 355-401: per-sample CPU transforms, host-sampled λ, gather + lerp on the GPU) and
// AdaptiveAvgPool2d + view + Linear (resnet.py:279-281) — SURVEY.md K7, K15, K22.
#include "common.cuh"
using namespace fdt;

namespace {

struct AugParams { float mean[3], inv_std[3]; int H, W, pad, train, lam_mode; };

__device__ __forceinline__ void sample_aug(const Philox& rng, uint64_t off, int b, int pad, int& dx, int& dy, int& flip) {
    uint4 r = rng((uint64_t)b, off);
    dx = (int)(r.x % (2 * pad + 1)) - pad;
    dy = (int)(r.y % (2 * pad + 1)) - pad;
    flip = (int)(r.z & 1u);
}
__device__ __forceinline__ float fetch(const uint8_t* img, int h, int w, int c, int dx, int dy, int flip, const AugParams& a) {
    int ys = h + dy, xs = w + dx;
    if (flip) xs = (a.W - 1) - w + dx;
    float v = 0.f;
    if (ys >= 0 && ys < a.H && xs >= 0 && xs < a.W) v = img[(ys * a.W + xs) * 3 + c] * (1.f / 255.f);
    return (v - a.mean[c]) * a.inv_std[c];
}

__global__ void __launch_bounds__(256)
input_aug_mixup_kernel(const uint8_t* __restrict__ x, const int64_t* __restrict__ perm, const float* __restrict__ lam,
                       int lam_n, const int64_t* __restrict__ seed_off, __nv_bfloat16* __restrict__ out, int B,
                       const AugParams a) {
    const int b = blockIdx.y;
    const int npix = a.H * a.W;
    const int pb = perm ? (int)perm[b] : b;
    int dx0 = 0, dy0 = 0, f0 = 0, dx1 = 0, dy1 = 0, f1 = 0;
    if (a.train) {
        Philox rng((uint64_t)seed_off[0]);
        sample_aug(rng, (uint64_t)seed_off[1], b, a.pad, dx0, dy0, f0);
        sample_aug(rng, (uint64_t)seed_off[1], pb, a.pad, dx1, dy1, f1);
    }
    float l = lam ? (lam_n == 1 ? lam[0] : lam[b]) : 1.f;
    if (a.lam_mode == 1 && lam) l = 1.f / (1.f + __expf(-l));          // meta-mixup: the array holds theta, lam = sigmoid(theta)
    const uint8_t* i0 = x + (size_t)b * npix * 3;
    const uint8_t* i1 = x + (size_t)pb * npix * 3;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < npix * 3; i += gridDim.x * blockDim.x) {
        const int c = i % 3, p = i / 3, w = p % a.W, h = p / a.W;
        float v = fetch(i0, h, w, c, dx0, dy0, f0, a);
        if (perm) v = l * v + (1.f - l) * fetch(i1, h, w, c, dx1, dy1, f1, a);
        out[(size_t)b * npix * 3 + i] = __float2bfloat16(v);
    }
}

// Backward of the blend w.r.t. the per-sample mixing weight (meta-mixup, resnet50_test.py:388-401: lam_b = sigmoid(theta_b)):
//   dlam[b] += Σ_pixels g[b, ·] · (v0 − v1),   v0 / v1 = the augmented, normalised partners recomputed from the uint8 images
// with the forward's (seed, offset) — nothing of the forward is stored.  lam_mode 1: chain through the sigmoid
// (dtheta[b] += dlam · lam (1 − lam), lam = sigmoid(lamsrc[b])); lam_mode 0: plain dlam.
__global__ void __launch_bounds__(256)
input_aug_mixup_bwd_kernel(const uint8_t* __restrict__ x, const int64_t* __restrict__ perm, const float* __restrict__ lamsrc,
                           int lam_mode, const int64_t* __restrict__ seed_off, const __nv_bfloat16* __restrict__ g,
                           float* __restrict__ dlam, const AugParams a) {
    const int b = blockIdx.y;
    const int npix = a.H * a.W;
    const int pb = (int)perm[b];
    int dx0 = 0, dy0 = 0, f0 = 0, dx1 = 0, dy1 = 0, f1 = 0;
    if (a.train) {
        Philox rng((uint64_t)seed_off[0]);
        sample_aug(rng, (uint64_t)seed_off[1], b, a.pad, dx0, dy0, f0);
        sample_aug(rng, (uint64_t)seed_off[1], pb, a.pad, dx1, dy1, f1);
    }
    const uint8_t* i0 = x + (size_t)b * npix * 3;
    const uint8_t* i1 = x + (size_t)pb * npix * 3;
    float acc = 0.f;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < npix * 3; i += gridDim.x * blockDim.x) {
        const int c = i % 3, p = i / 3, w = p % a.W, h = p / a.W;
        const float d = fetch(i0, h, w, c, dx0, dy0, f0, a) - fetch(i1, h, w, c, dx1, dy1, f1, a);
        acc += __bfloat162float(g[(size_t)b * npix * 3 + i]) * d;
    }
    acc = warp_sum(acc);
    __shared__ float red[8];
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.f;
        for (int i = 0; i < 8; ++i) t += red[i];
        if (lam_mode == 1) { const float l = 1.f / (1.f + __expf(-lamsrc[b])); t *= l * (1.f - l); }
        atomicAdd(dlam + b, t);
    }
}

// x [N, HW, C] bf16 → pooled [N, C] fp32, logits [N, K] = pooled·Wᵀ + bias; one block per sample
__global__ void __launch_bounds__(256)
pool_fc_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ W, const float* __restrict__ bias,
                   float* __restrict__ pooled, float* __restrict__ logits, int HW, int C, int K) {
    extern __shared__ float sp[];           // [C]
    const int n = blockIdx.x;
    const __nv_bfloat16* xn = x + (size_t)n * HW * C;
    const float inv = 1.f / HW;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        float s = 0.f;
        for (int p = 0; p < HW; ++p) s += __bfloat162float(xn[(size_t)p * C + c]);
        s *= inv;
        sp[c] = s;
        pooled[(size_t)n * C + c] = s;
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int k = warp; k < K; k += blockDim.x >> 5) {
        float s = 0.f;
        for (int c = lane; c < C; c += 32) s += sp[c] * W[(size_t)k * C + c];
        s = warp_sum(s);
        if (lane == 0) logits[(size_t)n * K + k] = s + (bias ? bias[k] : 0.f);
    }
}
// dlogits [N,K] → dx [N,HW,C] bf16 (= dpooled / HW broadcast), dW [K,C] += dlogitsᵀ·pooled, db [K] += Σ dlogits
__global__ void __launch_bounds__(256)
pool_fc_bwd_kernel(const float* __restrict__ dlogits, const float* __restrict__ W, const float* __restrict__ pooled,
                   __nv_bfloat16* __restrict__ dx, float* __restrict__ dW, float* __restrict__ db, int HW, int C, int K) {
    __shared__ float dl[32];
    const int n = blockIdx.x;
    if (threadIdx.x < K) {
        dl[threadIdx.x] = dlogits[(size_t)n * K + threadIdx.x];
        if (db) atomicAdd(db + threadIdx.x, dl[threadIdx.x]);
    }
    __syncthreads();
    const float inv = 1.f / HW;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        float g = 0.f;
        const float pc = pooled[(size_t)n * C + c];
        for (int k = 0; k < K; ++k) {
            g += dl[k] * W[(size_t)k * C + c];
            atomicAdd(dW + (size_t)k * C + c, dl[k] * pc);
        }
        const __nv_bfloat16 gv = __float2bfloat16(g * inv);
        if (dx) for (int p = 0; p < HW; ++p) dx[((size_t)n * HW + p) * C + c] = gv;
    }
}

}  // namespace

// x: uint8 [B,H,W,3]; perm: int64 [B] or null (no mixup); lam: fp32 [1] or [B] or null; seed_off: int64 (seed, offset),
// required when train != 0; out: bf16 [B,H,W,3].
FDT_API int fdt_input_aug_mixup_ex(const void* x, const int64_t* perm, const float* lam, int lam_n, int lam_mode,
                                   const int64_t* seed_off, void* out, int B, int H, int W, int pad, int train,
                                   const float* mean3, const float* std3, cudaStream_t st) {
    AugParams a;
    for (int c = 0; c < 3; ++c) { a.mean[c] = mean3[c]; a.inv_std[c] = 1.f / std3[c]; }
    a.H = H; a.W = W; a.pad = pad; a.train = train; a.lam_mode = lam_mode;
    if (train && !seed_off) return -1;
    dim3 grid((H * W * 3 + 255) / 256, B);
    input_aug_mixup_kernel<<<grid, 256, 0, st>>>((const uint8_t*)x, perm, lam, lam_n, seed_off, (__nv_bfloat16*)out, B, a);
    FDT_CHECK_LAUNCH();
    return 0;
}

FDT_API int fdt_input_aug_mixup(const void* x, const int64_t* perm, const float* lam, int lam_n, const int64_t* seed_off,
                                void* out, int B, int H, int W, int pad, int train, const float* mean3, const float* std3,
                                cudaStream_t st) {
    return fdt_input_aug_mixup_ex(x, perm, lam, lam_n, 0, seed_off, out, B, H, W, pad, train, mean3, std3, st);
}

// dlam [B] (+=, per-sample; zero it first for a fresh gradient): gradient of the blended batch w.r.t. the mixing weights.
// g: bf16 [B, H, W, 3] = dL/d(out of fdt_input_aug_mixup_ex); seed_off: the SAME (seed, offset) pair the forward used.
FDT_API int fdt_input_aug_mixup_bwd(const void* x, const int64_t* perm, const float* lamsrc, int lam_mode,
                                    const int64_t* seed_off, const void* g, float* dlam, int B, int H, int W, int pad, int train,
                                    const float* mean3, const float* std3, cudaStream_t st) {
    if (!perm || !lamsrc || !dlam || (train && !seed_off)) return -1;
    AugParams a;
    for (int c = 0; c < 3; ++c) { a.mean[c] = mean3[c]; a.inv_std[c] = 1.f / std3[c]; }
    a.H = H; a.W = W; a.pad = pad; a.train = train; a.lam_mode = lam_mode;
    dim3 grid(2, B);
    input_aug_mixup_bwd_kernel<<<grid, 256, 0, st>>>((const uint8_t*)x, perm, lamsrc, lam_mode, seed_off,
                                                     (const __nv_bfloat16*)g, dlam, a);
    FDT_CHECK_LAUNCH();
    return 0;
}

FDT_API int fdt_pool_fc_fwd(const void* x, const float* W, const float* bias, float* pooled, float* logits, int N, int HW,
                            int C, int K, cudaStream_t st) {
    if (K > 32 || C * sizeof(float) > 48 * 1024) return -1;
    pool_fc_fwd_kernel<<<N, 256, C * sizeof(float), st>>>((const __nv_bfloat16*)x, W, bias, pooled, logits, HW, C, K);
    FDT_CHECK_LAUNCH();
    return 0;
}

// dW [K,C] and db [K] are accumulated into (zero them first for a fresh gradient); dx may be null.
FDT_API int fdt_pool_fc_bwd(const float* dlogits, const float* W, const float* pooled, void* dx, float* dW, float* db,
                            int N, int HW, int C, int K, cudaStream_t st) {
    if (K > 32) return -1;
    pool_fc_bwd_kernel<<<N, 256, 0, st>>>(dlogits, W, pooled, (__nv_bfloat16*)dx, dW, db, HW, C, K);
    FDT_CHECK_LAUNCH();
    return 0;
}

// ---- device-side mixup draws ---------------------------------------------------------------------------------------
// lam ~ Beta(alpha, alpha) and a uniform random permutation of 0..B-1 in ONE tiny launch (the reference samples Beta on the
// host with `.item()` and calls torch.randperm — resnet50_test.py:359-362; the torch equivalents on the device are ~18
// small kernels: two Gamma draws, a sort-based randperm, index arithmetic).  One thread does the work: it is O(B).
namespace {
__device__ float philox_uniform(const Philox& rng, uint64_t off, uint32_t& ctr) {       // (0, 1]
    const uint4 r = rng((uint64_t)ctr++, off);
    return (float)((r.x >> 8) + 1u) * (1.0f / 16777216.0f);
}
__device__ float philox_normal(const Philox& rng, uint64_t off, uint32_t& ctr) {
    const float u1 = philox_uniform(rng, off, ctr), u2 = philox_uniform(rng, off, ctr);
    return sqrtf(-2.f * logf(u1)) * cospif(2.f * u2);
}
// Marsaglia & Tsang (2000); shape < 1 through Gamma(a) = Gamma(a + 1) * U^(1/a)
__device__ float philox_gamma(const Philox& rng, uint64_t off, uint32_t& ctr, float a) {
    float boost = 1.f;
    if (a < 1.f) { boost = powf(philox_uniform(rng, off, ctr), 1.f / a); a += 1.f; }
    const float d = a - 1.f / 3.f, c = rsqrtf(9.f * d);
    for (int it = 0; it < 64; ++it) {
        const float x = philox_normal(rng, off, ctr);
        float v = 1.f + c * x;
        if (v <= 0.f) continue;
        v = v * v * v;
        const float u = philox_uniform(rng, off, ctr);
        if (logf(u) < 0.5f * x * x + d - d * v + d * logf(v)) return boost * d * v;
    }
    return boost * d;
}
// One block.  Thread 0 draws lam (rejection sampling: a data-dependent number of Philox calls), then the B - 1 swap indices
// of the Fisher-Yates shuffle — one Philox call each, the serial part of the old single-thread form — are generated by all
// threads at their fixed counters, and thread 0 only walks the swaps in shared memory.  Same draws, same permutation.
constexpr int kPermMax = 4096;
__global__ void __launch_bounds__(256)
mixup_sample_kernel(const int64_t* __restrict__ seed_off, float alpha, int B, float* __restrict__ lam,
                    int64_t* __restrict__ perm) {
    __shared__ int s_perm[kPermMax];
    __shared__ int s_j[kPermMax];
    __shared__ uint32_t s_ctr;
    Philox rng((uint64_t)seed_off[0]);
    const uint64_t off = (uint64_t)seed_off[1];
    if (threadIdx.x == 0) {
        uint32_t ctr = 0x40000000u;                      // a counter range of its own (the augmentation draws use b < 2^30)
        if (lam) {
            float l = alpha;
            if (alpha > 0.f) {
                const float g0 = philox_gamma(rng, off, ctr, alpha), g1 = philox_gamma(rng, off, ctr, alpha);
                l = g0 / fmaxf(g0 + g1, 1e-20f);
            }
            lam[0] = l;
        }
        s_ctr = ctr;
    }
    if (!perm) return;
    __syncthreads();
    const uint32_t ctr0 = s_ctr;
    if (B <= kPermMax) {
        for (int i = threadIdx.x; i < B; i += blockDim.x) {
            s_perm[i] = i;
            if (i > 0) {                                 // the swap for position i is the (B - 1 - i)-th draw
                const uint4 r = rng((uint64_t)(ctr0 + (uint32_t)(B - 1 - i)), off);
                s_j[i] = (int)(((uint64_t)r.x * (uint64_t)(i + 1)) >> 32);
            }
        }
        __syncthreads();
        if (threadIdx.x == 0)
            for (int i = B - 1; i > 0; --i) {            // Fisher-Yates
                const int j = s_j[i];
                const int t = s_perm[i]; s_perm[i] = s_perm[j]; s_perm[j] = t;
            }
        __syncthreads();
        for (int i = threadIdx.x; i < B; i += blockDim.x) perm[i] = s_perm[i];
    } else if (threadIdx.x == 0) {
        uint32_t ctr = ctr0;
        for (int i = 0; i < B; ++i) perm[i] = i;
        for (int i = B - 1; i > 0; --i) {
            const uint4 r = rng((uint64_t)ctr++, off);
            const int j = (int)(((uint64_t)r.x * (uint64_t)(i + 1)) >> 32);
            const int64_t t = perm[i]; perm[i] = perm[j]; perm[j] = t;
        }
    }
}
}  // namespace

// lam [1] ~ Beta(alpha, alpha) (alpha <= 0: lam = alpha, as the reference) and perm [B] = a uniform random permutation,
// both optional (NULL); seed_off: device (seed, offset) pair.
FDT_API int fdt_mixup_sample(const int64_t* seed_off, float alpha, int B, float* lam, int64_t* perm, cudaStream_t st) {
    if (!seed_off || B < 0) return -1;
    mixup_sample_kernel<<<1, 256, 0, st>>>(seed_off, alpha, B, lam, perm);
    FDT_CHECK_LAUNCH();
    return 0;
}
