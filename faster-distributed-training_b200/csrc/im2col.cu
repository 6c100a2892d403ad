// Explicit im2col / col2im for the convolutions the implicit-GEMM kernel does not cover (stride 2, the
// 3-channel stem, odd feature-map widths): they become plain tcgen05 GEMMs on a [pixels, kh·kw·C] matrix.
//
//   A[p, (r·kw + s)·C + c] = x[n, ho·stride + r − pad, wo·stride + s − pad, c]      (0 outside, 0 in pad columns)
//
// col2im is written as a GATHER over input pixels (each dx element sums the ≤ kh·kw entries of dA that
// reference it), so it needs no atomics and is deterministic.
//
// Parity: the strided nn.Conv2d layers and the stem of the reference's ResNet (resnet.py:204-205, 221-224,
// 238-240), which call cuDNN.
#include "common.cuh"
using namespace fdt;

namespace {

struct ConvGeom { int N, H, W, C, k, stride, pad, Ho, Wo, Kp; };

// vector path: C % 8 == 0; one thread per (pixel, tap, 8 channels)
__global__ void __launch_bounds__(256)
im2col_vec_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ A, const ConvGeom g) {
    const int cv = g.C >> 3, taps = g.k * g.k;
    const int64_t total = (int64_t)g.N * g.Ho * g.Wo * taps * cv;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int c8 = (int)(i % cv);
        int64_t t = i / cv;
        const int tap = (int)(t % taps); t /= taps;
        const int wo = (int)(t % g.Wo); t /= g.Wo;
        const int ho = (int)(t % g.Ho);
        const int n = (int)(t / g.Ho);
        const int r = tap / g.k, s = tap - r * g.k;
        const int h = ho * g.stride + r - g.pad, w = wo * g.stride + s - g.pad;
        bf16x8 v; v.u = make_uint4(0u, 0u, 0u, 0u);
        if (h >= 0 && h < g.H && w >= 0 && w < g.W) v = ld8(x + (((int64_t)n * g.H + h) * g.W + w) * g.C + c8 * 8);
        const int64_t p = ((int64_t)n * g.Ho + ho) * g.Wo + wo;
        st8(A + p * g.Kp + (int64_t)tap * g.C + c8 * 8, v);
    }
}
// scalar path (stem, C = 3): one thread per (pixel, column of the padded row)
__global__ void __launch_bounds__(256)
im2col_scalar_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ A, const ConvGeom g) {
    const int K = g.k * g.k * g.C;
    const int64_t total = (int64_t)g.N * g.Ho * g.Wo * g.Kp;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int col = (int)(i % g.Kp);
        int64_t p = i / g.Kp;
        __nv_bfloat16 v = __float2bfloat16(0.f);
        if (col < K) {
            const int c = col % g.C, tap = col / g.C;
            const int r = tap / g.k, s = tap - r * g.k;
            const int wo = (int)(p % g.Wo); int64_t t = p / g.Wo;
            const int ho = (int)(t % g.Ho); const int n = (int)(t / g.Ho);
            const int h = ho * g.stride + r - g.pad, w = wo * g.stride + s - g.pad;
            if (h >= 0 && h < g.H && w >= 0 && w < g.W) v = x[(((int64_t)n * g.H + h) * g.W + w) * g.C + c];
        }
        A[i] = v;
    }
}

// small-C path (the 3-channel stem): one thread per (pixel, 8 columns of the padded row) → one 16-byte store instead of
// eight 2-byte ones; whole chunks beyond k·k·C are just zero stores.  Kp % 8 == 0.
__global__ void __launch_bounds__(256)
im2col_smallc_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ A, const ConvGeom g) {
    const int K = g.k * g.k * g.C, chunks = g.Kp >> 3;
    const int64_t total = (int64_t)g.N * g.Ho * g.Wo * chunks;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int ch = (int)(i % chunks);
        const int64_t p = i / chunks;
        float f[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] = 0.f;
        if (ch * 8 < K) {
            const int wo = (int)(p % g.Wo); const int64_t t = p / g.Wo;
            const int ho = (int)(t % g.Ho); const int n = (int)(t / g.Ho);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int col = ch * 8 + j;
                if (col < K) {
                    const int c = col % g.C, tap = col / g.C;
                    const int r = tap / g.k, s = tap - r * g.k;
                    const int h = ho * g.stride + r - g.pad, w = wo * g.stride + s - g.pad;
                    if (h >= 0 && h < g.H && w >= 0 && w < g.W) f[j] = __bfloat162float(x[(((int64_t)n * g.H + h) * g.W + w) * g.C + c]);
                }
            }
        }
        st8(A + p * g.Kp + ch * 8, pack8(f));
    }
}

// dx[n,h,w,c] = Σ_{(r,s): (h+pad−r) % stride == 0, ...} dA[p(ho,wo), (r·kw+s)·C + c]
template <bool kVec>
__global__ void __launch_bounds__(256)
col2im_kernel(const __nv_bfloat16* __restrict__ dA, __nv_bfloat16* __restrict__ dx, const ConvGeom g) {
    const int cv = kVec ? (g.C >> 3) : g.C;
    const int64_t total = (int64_t)g.N * g.H * g.W * cv;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int c = (int)(i % cv);
        int64_t t = i / cv;
        const int w = (int)(t % g.W); t /= g.W;
        const int h = (int)(t % g.H);
        const int n = (int)(t / g.H);
        float acc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = 0.f;
        for (int r = 0; r < g.k; ++r) {
            const int hn = h + g.pad - r;
            if (hn < 0 || hn % g.stride) continue;
            const int ho = hn / g.stride;
            if (ho >= g.Ho) continue;
            for (int s = 0; s < g.k; ++s) {
                const int wn = w + g.pad - s;
                if (wn < 0 || wn % g.stride) continue;
                const int wo = wn / g.stride;
                if (wo >= g.Wo) continue;
                const int64_t p = ((int64_t)n * g.Ho + ho) * g.Wo + wo;
                const __nv_bfloat16* src = dA + p * g.Kp + (int64_t)(r * g.k + s) * g.C;
                if (kVec) {
                    float f[8];
                    unpack8(ld8(src + c * 8), f);
#pragma unroll
                    for (int j = 0; j < 8; ++j) acc[j] += f[j];
                } else {
                    acc[0] += __bfloat162float(src[c]);
                }
            }
        }
        if (kVec) st8(dx + i * 8, pack8(acc));
        else dx[i] = __float2bfloat16(acc[0]);
    }
}

inline int blocks_for(int64_t n) {
    int64_t b = (n + 255) / 256;
    if (b > 148 * 16) b = 148 * 16;
    return (int)(b < 1 ? 1 : b);
}

}  // namespace

// A: [N·Ho·Wo, Kp] bf16, Kp >= k·k·C (columns beyond k·k·C are zero-filled; Kp == k·k·C when C % 8 == 0).
FDT_API int fdt_im2col(const void* x, void* A, int N, int H, int W, int C, int k, int stride, int pad, int Ho, int Wo, int Kp,
                       cudaStream_t st) {
    ConvGeom g{N, H, W, C, k, stride, pad, Ho, Wo, Kp};
    if (C % 8 == 0) {
        if (Kp != k * k * C) return -1;
        im2col_vec_kernel<<<blocks_for((int64_t)N * Ho * Wo * k * k * (C / 8)), 256, 0, st>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)A, g);
    } else {
        if (Kp < k * k * C) return -1;
        if (Kp % 8 == 0)
            im2col_smallc_kernel<<<blocks_for((int64_t)N * Ho * Wo * (Kp / 8)), 256, 0, st>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)A, g);
        else
            im2col_scalar_kernel<<<blocks_for((int64_t)N * Ho * Wo * Kp), 256, 0, st>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)A, g);
    }
    FDT_CHECK_LAUNCH();
    return 0;
}

FDT_API int fdt_col2im(const void* dA, void* dx, int N, int H, int W, int C, int k, int stride, int pad, int Ho, int Wo, int Kp,
                       cudaStream_t st) {
    ConvGeom g{N, H, W, C, k, stride, pad, Ho, Wo, Kp};
    if (C % 8 == 0)
        col2im_kernel<true><<<blocks_for((int64_t)N * H * W * (C / 8)), 256, 0, st>>>((const __nv_bfloat16*)dA, (__nv_bfloat16*)dx, g);
    else
        col2im_kernel<false><<<blocks_for((int64_t)N * H * W * C), 256, 0, st>>>((const __nv_bfloat16*)dA, (__nv_bfloat16*)dx, g);
    FDT_CHECK_LAUNCH();
    return 0;
}
