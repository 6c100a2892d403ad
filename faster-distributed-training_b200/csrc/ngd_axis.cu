// NGD preconditioning of a SHORT tensor axis (the 3-wide kernel axes of a conv weight, a 10- or 4-wide class
// axis): D <= 12, R = rank <= 6.  The gradient is used where it lies — as the [A, D, B] view of the physical
// tensor (A = extent before the axis, B = extent after it) — so no transposed copy is made, and one pass over
// it produces everything the step needs:
//     h = W x,  x̂ = x − Wᵀ h      per (a, b) fibre, in registers
//     Σx², Σx̂²                    → sums[2n]   (norm restoration, applied by fdt_ngd_rescale)
//     J = Σ h xᵀ  [R, D]          → J          (Fisher statistics, refresh steps only)
// The reference runs these axes as two cuBLAS SGEMMs with an inner dimension of 3 plus a transpose per factor
// (ngd_optimizer.py:195-218 via _precondition_directions3); here it is one bandwidth-bound kernel.
#include "common.cuh"

using namespace fdt;

namespace {

template <int D, int R, int V>
__global__ void __launch_bounds__(256)
ngd_small_axis_kernel(const float* __restrict__ X, const float* __restrict__ W, float* __restrict__ Xhat,
                      float* __restrict__ sums, float* __restrict__ J, int n, int64_t A, int64_t B) {
    const int f = blockIdx.y;
    const int64_t M = A * D * B;
    const float* x = X + f * M;
    float* xh = Xhat + f * M;
    float w[R][D];
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
        for (int d = 0; d < D; ++d) w[r][d] = W[((int64_t)f * R + r) * D + d];
    float jacc[R][D];
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
        for (int d = 0; d < D; ++d) jacc[r][d] = 0.f;
    float sq_in = 0.f, sq_out = 0.f;
    const int64_t Bv = B / V;                 // V consecutive b per thread (B % V == 0)
    const int64_t P = A * Bv;
    for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < P; p += (int64_t)gridDim.x * blockDim.x) {
        const int64_t a = p / Bv, b = (p - a * Bv) * V;
        const int64_t base = a * D * B + b;
        float xv[D][V];
#pragma unroll
        for (int d = 0; d < D; ++d) {
            if constexpr (V == 4) {
                const float4 t = *reinterpret_cast<const float4*>(x + base + d * B);
                xv[d][0] = t.x; xv[d][1] = t.y; xv[d][2] = t.z; xv[d][3] = t.w;
            } else {
                xv[d][0] = x[base + d * B];
            }
        }
#pragma unroll
        for (int v = 0; v < V; ++v) {
            float h[R];
#pragma unroll
            for (int r = 0; r < R; ++r) {
                float s = 0.f;
#pragma unroll
                for (int d = 0; d < D; ++d) s = fmaf(w[r][d], xv[d][v], s);
                h[r] = s;
            }
#pragma unroll
            for (int d = 0; d < D; ++d) {
                const float xin = xv[d][v];
                float s = 0.f;
#pragma unroll
                for (int r = 0; r < R; ++r) s = fmaf(w[r][d], h[r], s);
                const float o = xin - s;
                sq_in = fmaf(xin, xin, sq_in);
                sq_out = fmaf(o, o, sq_out);
                if (J) {
#pragma unroll
                    for (int r = 0; r < R; ++r) jacc[r][d] = fmaf(h[r], xin, jacc[r][d]);
                }
                xv[d][v] = o;
            }
        }
#pragma unroll
        for (int d = 0; d < D; ++d) {
            if constexpr (V == 4)
                *reinterpret_cast<float4*>(xh + base + d * B) = make_float4(xv[d][0], xv[d][1], xv[d][2], xv[d][3]);
            else
                xh[base + d * B] = xv[d][0];
        }
    }
    // block reduction: Σx², Σx̂² and (refresh steps) the R x D statistics
    __shared__ float red[8][R * D + 2];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    sq_in = warp_sum(sq_in); sq_out = warp_sum(sq_out);
    if (lane == 0) { red[warp][R * D] = sq_in; red[warp][R * D + 1] = sq_out; }
    if (J) {
#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
            for (int d = 0; d < D; ++d) {
                const float t = warp_sum(jacc[r][d]);
                if (lane == 0) red[warp][r * D + d] = t;
            }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < R * D + 2; i += blockDim.x) {
        if (i < R * D && !J) continue;
        float t = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) t += red[k][i];
        if (i < R * D) atomicAdd(J + (int64_t)f * R * D + i, t);
        else atomicAdd(sums + (i == R * D ? f : n + f), t);
    }
}

template <int D, int R>
int launch_small(const float* X, const float* W, float* Xhat, float* sums, float* J, int n, int64_t A, int64_t B, cudaStream_t st) {
    const bool vec = (B % 4 == 0) && ((reinterpret_cast<uintptr_t>(X) | reinterpret_cast<uintptr_t>(Xhat)) % 16 == 0) && D * R <= 18;
    const int64_t P = A * (vec ? B / 4 : B);
    int64_t chunks = (P + 256 * 2 - 1) / (256 * 2);
    if (chunks < 1) chunks = 1;
    const int64_t cap = (148 * 8 + n - 1) / n;
    if (chunks > cap) chunks = cap;
    if (vec) ngd_small_axis_kernel<D, R, 4><<<dim3((unsigned)chunks, n), 256, 0, st>>>(X, W, Xhat, sums, J, n, A, B);
    else ngd_small_axis_kernel<D, R, 1><<<dim3((unsigned)chunks, n), 256, 0, st>>>(X, W, Xhat, sums, J, n, A, B);
    FDT_CHECK_LAUNCH();
    return 0;
}

}  // namespace

// X, Xhat: [n, A, D, B] fp32 contiguous; W: [n, R, D]; sums: [2n]; J: [n, R, D] or NULL.  sums and J are zeroed here.
// Covered: 2 <= D <= 12 with R == (D + 1) / 2 (the default rank); returns -1 for anything else.
FDT_API int fdt_ngd_small_axis(const float* X, const float* W, float* Xhat, float* sums, float* J, int n, int64_t A, int D,
                               int64_t B, int R, cudaStream_t st) {
    if (D < 2 || D > 12 || R != (D + 1) / 2 || n <= 0) return -1;
    cudaMemsetAsync(sums, 0, sizeof(float) * 2 * n, st);
    if (J) cudaMemsetAsync(J, 0, sizeof(float) * (size_t)n * R * D, st);
    switch (D) {
#define FDT_CASE(d) case d: return launch_small<d, (d + 1) / 2>(X, W, Xhat, sums, J, n, A, B, st);
        FDT_CASE(2) FDT_CASE(3) FDT_CASE(4) FDT_CASE(5) FDT_CASE(6) FDT_CASE(7) FDT_CASE(8) FDT_CASE(9) FDT_CASE(10) FDT_CASE(11)
        FDT_CASE(12)
#undef FDT_CASE
    }
    return -1;
}

namespace {
// K[f] = J Jᵀ, L[f] = J Wᵀ for the short-axis factors (R <= 6, D <= 12): one thread per output element.
__global__ void small_kl_kernel(const float* __restrict__ J, const float* __restrict__ W, float* __restrict__ K,
                                float* __restrict__ L, int n, int R, int D) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n * R * R) return;
    const int f = i / (R * R), rc = i - f * R * R, r = rc / R, c = rc - r * R;
    const float* jr = J + ((size_t)f * R + r) * D;
    const float* jc = J + ((size_t)f * R + c) * D;
    const float* wc = W + ((size_t)f * R + c) * D;
    float k = 0.f, l = 0.f;
    for (int d = 0; d < D; ++d) { k = fmaf(jr[d], jc[d], k); l = fmaf(jr[d], wc[d], l); }
    K[i] = k;
    L[i] = l;
}
}  // namespace

// The R x R Fisher statistics of the short-axis factors: K [n,R,R] = J Jᵀ, L [n,R,R] = J Wᵀ (J, W: [n,R,D] contiguous).
// ref: ngd_optimizer.py:205-218 (two torch.mm per factor).
FDT_API int fdt_ngd_small_kl(const float* J, const float* W, float* K, float* L, int n, int R, int D, cudaStream_t st) {
    if (n <= 0 || R < 1 || R > 16 || D < 1 || D > 32) return -1;
    const int total = n * R * R;
    small_kl_kernel<<<(total + 127) / 128, 128, 0, st>>>(J, W, K, L, n, R, D);
    FDT_CHECK_LAUNCH();
    return 0;
}

// ---- whole small factors in shared memory -------------------------------------------------------------------------------
// Factors whose gradient matrix is small and whose rows are not 16-byte aligned (the 3-channel stem's [64, 27] view, a
// [B, 1, 1, 1] meta-mixup theta, odd-length vectors) cannot be fed to TMA; they are also far too small for tensor cores.
// One CTA per factor keeps X, W and H in shared memory and does the complete step of ngd_optimizer.py:138-218 for it:
//   H = X Wᵀ, X̂ = X − H W, p0 = ΣX², p1 = ΣX̂², out = X̂·sqrt(p0 / (p1 + 1e-30)) (X if p1 is NaN), and optionally J = Hᵀ X.
namespace {
constexpr int kSmallMaxElems = 24576;     // N*D + R*D + N*R floats in total: 96 KB of shared memory

__global__ void __launch_bounds__(256)
ngd_small_matrix_kernel(const float* __restrict__ X, const float* __restrict__ W, int ldw, float* __restrict__ out,
                        float* __restrict__ p0_out, float* __restrict__ J, int ldj, int N, int D, int R, int x_trans) {
    extern __shared__ float sm[];
    float* sX = sm;                    // [N][D]
    float* sW = sm + N * D;            // [R][D]
    float* sH = sW + R * D;            // [N][R]
    __shared__ float red0[8], red1[8];
    __shared__ float s_scale, s_nan;
    const int f = blockIdx.x, tid = threadIdx.x;
    const float* Xf = X + (size_t)f * N * D;
    const float* Wf = W + (size_t)f * R * ldw;
    for (int i = tid; i < N * D; i += 256) {
        const int r = i / D, d = i - r * D;
        sX[i] = x_trans ? Xf[(size_t)d * N + r] : Xf[i];
    }
    for (int i = tid; i < R * D; i += 256) { const int r = i / D, d = i - r * D; sW[i] = Wf[(size_t)r * ldw + d]; }
    __syncthreads();
    for (int i = tid; i < N * R; i += 256) {
        const int row = i / R, r = i - row * R;
        float a = 0.f;
        for (int d = 0; d < D; ++d) a = fmaf(sX[row * D + d], sW[r * D + d], a);
        sH[i] = a;
    }
    __syncthreads();
    float p0 = 0.f, p1 = 0.f;
    float* outf = out + (size_t)f * N * D;
    // X̂ overwrites nothing yet: keep it in registers per element, two passes (norms first, then the rescaled store)
    for (int i = tid; i < N * D; i += 256) {
        const int row = i / D, d = i - row * D;
        float a = sX[i];
        p0 = fmaf(a, a, p0);
        float h = 0.f;
        for (int r = 0; r < R; ++r) h = fmaf(sH[row * R + r], sW[r * D + d], h);
        a -= h;
        p1 = fmaf(a, a, p1);
    }
    p0 = warp_sum(p0); p1 = warp_sum(p1);
    if ((tid & 31) == 0) { red0[tid >> 5] = p0; red1[tid >> 5] = p1; }
    __syncthreads();
    if (tid == 0) {
        float a = 0.f, b = 0.f;
        for (int i = 0; i < 8; ++i) { a += red0[i]; b += red1[i]; }
        p0_out[f] = a;
        s_nan = (b != b) ? 1.f : 0.f;
        s_scale = sqrtf(a / (b + 1e-30f));
    }
    __syncthreads();
    const float scale = s_scale;
    const bool use_x = s_nan != 0.f;
    for (int i = tid; i < N * D; i += 256) {
        const int row = i / D, d = i - row * D;
        float h = 0.f;
        for (int r = 0; r < R; ++r) h = fmaf(sH[row * R + r], sW[r * D + d], h);
        const float v = use_x ? sX[i] : (sX[i] - h) * scale;
        if (x_trans) outf[(size_t)d * N + row] = v; else outf[i] = v;
    }
    if (J) {
        float* Jf = J + (size_t)f * R * ldj;
        for (int i = tid; i < R * D; i += 256) {
            const int r = i / D, d = i - r * D;
            float a = 0.f;
            for (int row = 0; row < N; ++row) a = fmaf(sH[row * R + r], sX[row * D + d], a);
            Jf[(size_t)r * ldj + d] = a;
        }
    }
}
}  // namespace

// X [n, N, D] (x_trans == 0) or [n, D, N] (x_trans != 0); W [n, R, ldw] (ldw >= D: padded pitch); out: same layout as X
// (may NOT alias X); p0 [n] = ΣX² per factor; J [n, R, ldj] or NULL.  N*D + R*D + N*R <= 24576.
FDT_API int fdt_ngd_small_matrix(const float* X, const float* W, int ldw, float* out, float* p0, float* J, int ldj, int n,
                                 int N, int D, int R, int x_trans, cudaStream_t st) {
    if (n <= 0 || N < 1 || D < 1 || R < 1 || ldw < D || (J && ldj < D)) return -1;
    if ((int64_t)N * D + (int64_t)R * D + (int64_t)N * R > kSmallMaxElems) return -1;
    const size_t smem = sizeof(float) * ((size_t)N * D + (size_t)R * D + (size_t)N * R);
    static bool attr = false;
    if (!attr) {
        cudaError_t e = cudaFuncSetAttribute(ngd_small_matrix_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmallMaxElems * 4);
        if (e != cudaSuccess) return (int)e;
        attr = true;
    }
    ngd_small_matrix_kernel<<<n, 256, smem, st>>>(X, W, ldw, out, p0, J, ldj, N, D, R, x_trans);
    FDT_CHECK_LAUNCH();
    return 0;
}
