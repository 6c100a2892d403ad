// Flat-buffer optimizer kernels: one launch over the whole (or the rank-local) parameter range.
//
// Parity: the reference walks 69 / 105 tensors in Python for every one of: zero_grad, GradScaler
// unscale + inf check, clip_grad_norm_, weight decay, momentum, parameter update
// (resnet50_test.py:531,544-548; ngd_optimizer.py:478-506) and MADGRAD's per-tensor loop
// (SURVEY.md K20, K21).  With all tensors in one flat buffer each of these is a single kernel.
#include "common.cuh"
using namespace fdt;

namespace {

__global__ void sqnorm_kernel(const float* __restrict__ g, int64_t n, float* __restrict__ out) {
    float s = 0.f;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x * 4;
    for (int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4; i < n; i += stride) {
        if (i + 3 < n) {
            float4 v = *reinterpret_cast<const float4*>(g + i);
            s += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
        } else {
            for (int64_t j = i; j < n; ++j) s += g[j] * g[j];
        }
    }
    s = warp_sum(s);
    __shared__ float red[8];
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.f;
        for (int i = 0; i < (int)(blockDim.x >> 5); ++i) t += red[i];
        atomicAdd(out, t);
    }
}

// g *= min(1, max_norm / (sqrt(sq) + 1e-6)); non-finite norm → gradients zeroed and *skipped += 1
__global__ void clip_scale_kernel(float* __restrict__ g, int64_t n, const float* __restrict__ sq, float max_norm,
                                  float* __restrict__ skipped) {
    const float s = *sq;
    const bool finite = isfinite(s);
    float coef = finite ? (max_norm > 0.f ? fminf(1.f, max_norm / (sqrtf(s) + 1e-6f)) : 1.f) : 0.f;
    if (skipped && !finite && blockIdx.x == 0 && threadIdx.x == 0) *skipped += 1.f;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x * 4;
    for (int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4; i < n; i += stride) {
        if (i + 3 < n) {
            float4 v = *reinterpret_cast<float4*>(g + i);
            v.x = finite ? v.x * coef : 0.f; v.y = finite ? v.y * coef : 0.f;
            v.z = finite ? v.z * coef : 0.f; v.w = finite ? v.w * coef : 0.f;
            *reinterpret_cast<float4*>(g + i) = v;
        } else {
            for (int64_t j = i; j < n; ++j) g[j] = finite ? g[j] * coef : 0.f;
        }
    }
}

// buf = first ? d : m·buf + (1−damp)·d ;  step = nesterov ? d + m·buf : buf ;  p −= lr·step
__global__ void sgd_momentum_kernel(float* __restrict__ p, const float* __restrict__ d, float* __restrict__ buf,
                                    int64_t n, const float* __restrict__ lr_ptr, float lr_val, float momentum,
                                    float dampening, int nesterov, int first) {
    const float lr = lr_ptr ? *lr_ptr : lr_val;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float di = d[i];
        float step = di;
        if (momentum != 0.f) {
            const float b = first ? di : momentum * buf[i] + (1.f - dampening) * di;
            buf[i] = b;
            step = nesterov ? di + momentum * b : b;
        }
        p[i] -= lr * step;
    }
}

// MADGRAD (Defazio & Jelassi 2021), momentum form; lamb = lr·sqrt(k+1) passed in.
__global__ void madgrad_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ v,
                               float* __restrict__ s, float* __restrict__ x0, int64_t n, float lamb, float ck,
                               float decay, float eps, int mirror) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        float gi = g[i] + decay * p[i];
        float vi = v[i] + lamb * gi * gi;
        v[i] = vi;
        const float rms = cbrtf(vi) + eps;
        float z;
        if (mirror) { z = x0[i] - lamb * gi / rms; x0[i] = z; }
        else { const float si = s[i] + lamb * gi; s[i] = si; z = x0[i] - si / rms; }
        p[i] = (1.f - ck) * p[i] + ck * z;
    }
}

// Device-hyper-parameter variants: lr / momentum / dampening (SGD) and lr / step count / momentum (MADGRAD) are read
// from device memory, so a captured CUDA graph follows LR and momentum schedules without re-capture.
// hyper = {lr, momentum, dampening}
__global__ void sgd_momentum_dev_kernel(float* __restrict__ p, const float* __restrict__ d, float* __restrict__ buf,
                                        int64_t n, const float* __restrict__ hyper, int nesterov, int first,
                                        __nv_bfloat16* __restrict__ shadow) {
    const float lr = hyper[0], momentum = hyper[1], dampening = hyper[2];
    const int64_t stride = (int64_t)gridDim.x * blockDim.x * 4;
    for (int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4; i < n; i += stride) {
        if (i + 3 < n) {
            const float4 dv = *reinterpret_cast<const float4*>(d + i);
            float4 pv = *reinterpret_cast<float4*>(p + i);
            float4 st = dv;
            if (buf) {
                float4 b = first ? dv : *reinterpret_cast<const float4*>(buf + i);
                if (!first) {
                    b.x = momentum * b.x + (1.f - dampening) * dv.x; b.y = momentum * b.y + (1.f - dampening) * dv.y;
                    b.z = momentum * b.z + (1.f - dampening) * dv.z; b.w = momentum * b.w + (1.f - dampening) * dv.w;
                }
                *reinterpret_cast<float4*>(buf + i) = b;
                if (nesterov) { st.x = dv.x + momentum * b.x; st.y = dv.y + momentum * b.y; st.z = dv.z + momentum * b.z; st.w = dv.w + momentum * b.w; }
                else st = b;
            }
            pv.x -= lr * st.x; pv.y -= lr * st.y; pv.z -= lr * st.z; pv.w -= lr * st.w;
            *reinterpret_cast<float4*>(p + i) = pv;
            if (shadow) {          // the bf16 operand copy of the updated parameters, in the same pass
                *reinterpret_cast<__nv_bfloat162*>(shadow + i) = __floats2bfloat162_rn(pv.x, pv.y);
                *reinterpret_cast<__nv_bfloat162*>(shadow + i + 2) = __floats2bfloat162_rn(pv.z, pv.w);
            }
        } else {
            for (int64_t j = i; j < n; ++j) {
                const float dj = d[j];
                float st = dj;
                if (buf) {
                    const float b = first ? dj : momentum * buf[j] + (1.f - dampening) * dj;
                    buf[j] = b;
                    st = nesterov ? dj + momentum * b : b;
                }
                p[j] -= lr * st;
                if (shadow) shadow[j] = __float2bfloat16(p[j]);
            }
        }
    }
}

// g = finite(sq) ? g * min(1, max_norm / (sqrt(sq) + 1e-6)) + wd * p : 0      (clip + weight decay in ONE pass;
// the reference does clip_grad_norm_ and then `d_p.add_(p, alpha=wd)` per tensor: resnet50_test.py:546, ngd_optimizer.py:478)
__global__ void clip_wd_kernel(float* __restrict__ g, const float* __restrict__ p, int64_t n, const float* __restrict__ sq,
                               float max_norm, float wd, float* __restrict__ skipped) {
    const float s = *sq;
    const bool finite = isfinite(s);
    const float coef = finite ? (max_norm > 0.f ? fminf(1.f, max_norm / (sqrtf(s) + 1e-6f)) : 1.f) : 0.f;
    const float w = finite ? wd : 0.f;
    if (skipped && !finite && blockIdx.x == 0 && threadIdx.x == 0) *skipped += 1.f;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x * 4;
    for (int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4; i < n; i += stride) {
        if (i + 3 < n) {
            float4 v = *reinterpret_cast<float4*>(g + i);
            float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
            if (w != 0.f) q = *reinterpret_cast<const float4*>(p + i);
            v.x = finite ? v.x * coef + w * q.x : 0.f; v.y = finite ? v.y * coef + w * q.y : 0.f;
            v.z = finite ? v.z * coef + w * q.z : 0.f; v.w = finite ? v.w * coef + w * q.w : 0.f;
            *reinterpret_cast<float4*>(g + i) = v;
        } else {
            for (int64_t j = i; j < n; ++j) g[j] = finite ? g[j] * coef + (w != 0.f ? w * p[j] : 0.f) : 0.f;
        }
    }
}

// MADGRAD / MirrorMADGRAD with device-resident hyper-parameters: hyper = {lr (+eps already added), k, momentum}.
// lamb = lr * sqrt(k + 1).  madgrad_tick_kernel advances k afterwards (stream-ordered), so graph replays keep counting.
__global__ void madgrad_dev_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ v,
                                   float* __restrict__ s, float* __restrict__ x0, int64_t n, const float* __restrict__ hyper,
                                   float decay, float eps, int mirror) {
    const float lamb = hyper[0] * sqrtf(hyper[1] + 1.f);
    const float ck = 1.f - hyper[2];
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float gi = g[i] + decay * p[i];
        const float vi = v[i] + lamb * gi * gi;
        v[i] = vi;
        const float rms = cbrtf(vi) + eps;
        float z;
        if (mirror) { z = x0[i] - lamb * gi / rms; x0[i] = z; }
        else { const float si = s[i] + lamb * gi; s[i] = si; z = x0[i] - si / rms; }
        p[i] = (1.f - ck) * p[i] + ck * z;
    }
}
__global__ void madgrad_tick_kernel(float* hyper) { hyper[1] += 1.f; }

__global__ void cast_bf16_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ y, int64_t n) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x * 4;
    for (int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4; i < n; i += stride) {
        if (i + 3 < n) {
            float4 v = *reinterpret_cast<const float4*>(x + i);
            *reinterpret_cast<__nv_bfloat162*>(y + i) = __floats2bfloat162_rn(v.x, v.y);
            *reinterpret_cast<__nv_bfloat162*>(y + i + 2) = __floats2bfloat162_rn(v.z, v.w);
        } else {
            for (int64_t j = i; j < n; ++j) y[j] = __float2bfloat16(x[j]);
        }
    }
}

inline int grid_for(int64_t n, int per_thread) {
    int64_t b = (n / per_thread + 255) / 256;
    if (b < 1) b = 1;
    if (b > 148 * 16) b = 148 * 16;
    return (int)b;
}

}  // namespace

// *out must be zero on entry.
FDT_API int fdt_sqnorm(const float* g, int64_t n, float* out, cudaStream_t st) {
    sqnorm_kernel<<<grid_for(n, 4), 256, 0, st>>>(g, n, out);
    FDT_CHECK_LAUNCH();
    return 0;
}

FDT_API int fdt_clip_scale(float* g, int64_t n, const float* sq, float max_norm, float* skipped, cudaStream_t st) {
    clip_scale_kernel<<<grid_for(n, 4), 256, 0, st>>>(g, n, sq, max_norm, skipped);
    FDT_CHECK_LAUNCH();
    return 0;
}

FDT_API int fdt_sgd_momentum(float* p, const float* d, float* buf, int64_t n, const float* lr_ptr, float lr_val,
                             float momentum, float dampening, int nesterov, int first, cudaStream_t st) {
    sgd_momentum_kernel<<<grid_for(n, 1), 256, 0, st>>>(p, d, buf, n, lr_ptr, lr_val, momentum, dampening, nesterov, first);
    FDT_CHECK_LAUNCH();
    return 0;
}

FDT_API int fdt_madgrad(float* p, const float* g, float* v, float* s, float* x0, int64_t n, float lamb, float ck,
                        float decay, float eps, int mirror, cudaStream_t st) {
    madgrad_kernel<<<grid_for(n, 1), 256, 0, st>>>(p, g, v, s, x0, n, lamb, ck, decay, eps, mirror);
    FDT_CHECK_LAUNCH();
    return 0;
}

FDT_API int fdt_sgd_momentum_dev(float* p, const float* d, float* buf, int64_t n, const float* hyper, int nesterov, int first,
                                 void* shadow_bf16, cudaStream_t st) {
    if (!hyper) return -1;
    sgd_momentum_dev_kernel<<<grid_for(n, 4), 256, 0, st>>>(p, d, buf, n, hyper, nesterov, first, (__nv_bfloat16*)shadow_bf16);
    FDT_CHECK_LAUNCH();
    return 0;
}

FDT_API int fdt_clip_wd(float* g, const float* p, int64_t n, const float* sq, float max_norm, float wd, float* skipped,
                        cudaStream_t st) {
    clip_wd_kernel<<<grid_for(n, 4), 256, 0, st>>>(g, p, n, sq, max_norm, wd, skipped);
    FDT_CHECK_LAUNCH();
    return 0;
}

FDT_API int fdt_madgrad_dev(float* p, const float* g, float* v, float* s, float* x0, int64_t n, float* hyper, float decay,
                            float eps, int mirror, int tick, cudaStream_t st) {
    if (!hyper) return -1;
    madgrad_dev_kernel<<<grid_for(n, 1), 256, 0, st>>>(p, g, v, s, x0, n, hyper, decay, eps, mirror);
    FDT_CHECK_LAUNCH();
    if (tick) { madgrad_tick_kernel<<<1, 1, 0, st>>>(hyper); FDT_CHECK_LAUNCH(); }
    return 0;
}

FDT_API int fdt_cast_bf16(const float* x, void* y, int64_t n, cudaStream_t st) {
    cast_bf16_kernel<<<grid_for(n, 4), 256, 0, st>>>(x, (__nv_bfloat16*)y, n);
    FDT_CHECK_LAUNCH();
    return 0;
}
