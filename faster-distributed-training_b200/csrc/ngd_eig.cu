// NGD Fisher-factor refresh: Z assembly + batched symmetric eigensolver + all scalar algebra,
// one CTA per factor, everything in shared memory, no host round trips.
//
// Parity: ngd_optimizer.py:225-317 of the reference, where each factor costs a
// cuSOLVER syevd call, ~25 tiny ATen kernels and 7 `.item()` host syncs.  Inputs per factor are the
// statistics K = J Jᵀ and L (R×R), the current d (R), rho and tr(X Xᵀ); outputs are
//     A (R×R)  with  W_{t+1} = A · (J + w_coeff ∘ W_t),   w_coeff (R),   d_{t+1} (R),   rho_{t+1}.
// The eigensolver is a two-sided cyclic Jacobi with round-robin (tournament) pair ordering: all
// R/2 rotations of a step are independent, so the CTA applies them in parallel (columns of Z and V,
// then rows of Z).  R <= 80 so Z and V (2 x 80 x 81 floats) fit in 52 KB of shared memory.
#include "common.cuh"
using namespace fdt;

namespace {

constexpr int kMaxR = 80;
constexpr int kThreads = 1024;                               // one CTA per factor and ~100 long (R = 80) problems on 148 SMs:
                                                              // the rotations of a step are spread over the whole SM
constexpr int kMaxIt = (kMaxR / 2 * kMaxR + kThreads - 1) / kThreads;   // (pair, index) slots per thread and step
constexpr float kEpsilon = 1.0e-10f;   // ngd_optimizer.py:37
constexpr float kDelta = 5.0e-4f;      // ngd_optimizer.py:38

__device__ __forceinline__ float block_sum(float v, float* scratch) {
    v = warp_sum(v);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
    if (l == 0) scratch[w] = v;
    __syncthreads();
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < kThreads / 32; ++i) t += scratch[i];
    return t;
}
__device__ __forceinline__ float block_max(float v, float* scratch) {
    v = warp_max(v);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
    if (l == 0) scratch[w] = v;
    __syncthreads();
    float t = -INFINITY;
#pragma unroll
    for (int i = 0; i < kThreads / 32; ++i) t = fmaxf(t, scratch[i]);
    return t;
}

__device__ __forceinline__ void
ngd_eig_body(const float* __restrict__ Kg, const float* __restrict__ Lg, const float* __restrict__ dg,
             const float* __restrict__ rhog, const float* __restrict__ trg,
             float* __restrict__ Ag, float* __restrict__ wcg, float* __restrict__ d1g, float* __restrict__ rho1g,
             const int f, int R, int D, float N, float alpha, float eta, int max_sweeps) {
    extern __shared__ float sm[];
    const int ld = R + 1 + ((R & 1) ? 1 : 0);        // padded leading dim, odd => conflict-free column walks
    const int n2 = (R + 1) & ~1;                     // even problem size (dummy index R when R is odd)
    float* Z = sm;                                   // [n2][ld]
    float* V = Z + n2 * ld;                          // [n2][ld]
    float* ise = V + n2 * ld;                        // inv_sqrt_e [R]
    float* dpr = ise + kMaxR;                        // d + rho     [R]
    float* cs = dpr + kMaxR;                         // rotation (c, s) per pair [n2]
    float* lam = cs + 2 * kMaxR;                     // eigenvalues, then sqrt_c
    float* leftv = lam + kMaxR;
    float* scratch = leftv + kMaxR;                  // [kThreads / 32]
    __shared__ int order[kMaxR];
    __shared__ int pp[kMaxR / 2 + 1], qq[kMaxR / 2 + 1];

    const int tid = threadIdx.x;
    const float* K = Kg + (int64_t)f * R * R;
    const float* L = Lg + (int64_t)f * R * R;
    const float* d = dg + (int64_t)f * R;
    const float rho = rhog[f];
    const float trXXt = trg[f];

    // ---- scalars that precede the eigenproblem (ngd_optimizer.py:225-260)
    float v = 0.f, tk = 0.f;
    for (int i = tid; i < R; i += kThreads) { v += d[i]; tk += K[i * R + i]; }
    const float d_sum = block_sum(v, scratch);
    const float z_scale = fmaxf(1.f, block_sum(tk, scratch));
    const float beta = rho * (1.f + alpha) + alpha * d_sum / D;
    for (int i = tid; i < R; i += kThreads) {
        const float e = 1.f / (beta / d[i] + 1.f);
        ise[i] = rsqrtf(e);
        dpr[i] = d[i] + rho;
    }
    __syncthreads();
    const float eN = eta / N;
    const float c1 = eN * eN / z_scale, c2 = eN * (1.f - eta) / z_scale, c3 = (1.f - eta) * (1.f - eta) / z_scale;
    for (int idx = tid; idx < n2 * n2; idx += kThreads) {
        const int i = idx / n2, j = idx % n2;
        float z = 0.f;
        if (i < R && j < R) {
            const float kij = 0.5f * (K[i * R + j] + K[j * R + i]);
            const float lij = 0.5f * (L[i * R + j] + L[j * R + i]);
            const float w = ise[i] * ise[j];
            z = kij * w * c1 + lij * w * (dpr[i] + dpr[j]) * c2;
            if (i == j) z += c3 * dpr[i] * dpr[i];
        }
        Z[i * ld + j] = z;
        V[i * ld + j] = (i == j) ? 1.f : 0.f;
    }
    __syncthreads();

    // ---- parallel cyclic Jacobi
    const int half = n2 / 2;
    // the (pair k, index i) slots a thread updates are the same in every step: no divisions in the hot loop
    int kk[kMaxIt], ii[kMaxIt];
#pragma unroll
    for (int it = 0; it < kMaxIt; ++it) {
        const int idx = tid + it * kThreads;
        kk[it] = idx < half * n2 ? idx / n2 : -1;
        ii[it] = idx % n2;
    }
    for (int sweep = 0; sweep < max_sweeps; ++sweep) {
        float off = 0.f, dia = 0.f;
        for (int idx = tid; idx < R * R; idx += kThreads) {
            const int i = idx / R, j = idx % R;
            const float z = Z[i * ld + j];
            if (i == j) dia += z * z; else off += z * z;
        }
        off = block_sum(off, scratch);
        dia = block_sum(dia, scratch);
        if (off <= 1e-13f * dia || off == 0.f) break;
        for (int step = 0; step < n2 - 1; ++step) {
            if (tid < half) {
                int p, q;
                if (tid == 0) { p = n2 - 1; q = step; }
                else { p = (step + tid) % (n2 - 1); q = (step - tid + (n2 - 1)) % (n2 - 1); }
                if (p > q) { int t = p; p = q; q = t; }
                float c = 1.f, s = 0.f;
                if (q < R) {
                    const float zpq = Z[p * ld + q];
                    if (fabsf(zpq) > 1e-37f) {
                        const float tau = (Z[q * ld + q] - Z[p * ld + p]) / (2.f * zpq);
                        const float t = (tau >= 0.f ? 1.f : -1.f) / (fabsf(tau) + sqrtf(1.f + tau * tau));
                        c = rsqrtf(1.f + t * t);
                        s = t * c;
                    }
                }
                pp[tid] = p; qq[tid] = q; cs[2 * tid] = c; cs[2 * tid + 1] = s;
            }
            __syncthreads();
            // columns of Z and V:  [col_p, col_q] <- [c·col_p − s·col_q,  s·col_p + c·col_q]
#pragma unroll
            for (int it = 0; it < kMaxIt; ++it) {
                const int k = kk[it], i = ii[it];
                if (k < 0) continue;
                const int p = pp[k], q = qq[k];
                const float c = cs[2 * k], s = cs[2 * k + 1];
                const float zp = Z[i * ld + p], zq = Z[i * ld + q];
                Z[i * ld + p] = c * zp - s * zq;
                Z[i * ld + q] = s * zp + c * zq;
                const float vp = V[i * ld + p], vq = V[i * ld + q];
                V[i * ld + p] = c * vp - s * vq;
                V[i * ld + q] = s * vp + c * vq;
            }
            __syncthreads();
            // rows of Z
#pragma unroll
            for (int it = 0; it < kMaxIt; ++it) {
                const int k = kk[it], j = ii[it];
                if (k < 0) continue;
                const int p = pp[k], q = qq[k];
                const float c = cs[2 * k], s = cs[2 * k + 1];
                const float zp = Z[p * ld + j], zq = Z[q * ld + j];
                Z[p * ld + j] = c * zp - s * zq;
                Z[q * ld + j] = s * zp + c * zq;
            }
            __syncthreads();
        }
    }

    // ---- sort eigenvalues descending (rank by counting), floors, rho/d update (ngd_optimizer.py:265-295)
    for (int i = tid; i < R; i += kThreads) lam[i] = Z[i * ld + i];
    __syncthreads();
    for (int i = tid; i < R; i += kThreads) {
        const float li = lam[i];
        int r = 0;
        for (int j = 0; j < R; ++j) { const float lj = lam[j]; r += (lj > li) || (lj == li && j < i); }
        order[r] = i;
    }
    __syncthreads();
    const float c_floor = (rho * (1.f - eta)) * (rho * (1.f - eta)) / z_scale;
    float sc = 0.f;
    if (tid < R) {
        sc = sqrtf(fmaxf(lam[order[tid]], c_floor)) * sqrtf(z_scale);   // sqrt_c, sorted position = tid
    }
    __syncthreads();
    if (tid < R) leftv[tid] = sc;          // stash sqrt_c by sorted position
    const float sum_sqrt_c = block_sum(tid < R ? sc : 0.f, scratch);
    const float max_sqrt_c = block_max(tid < R ? sc : -INFINITY, scratch);
    float rho1 = (1.f / (D - R)) * (eN * trXXt + (1.f - eta) * (D * rho + d_sum) - sum_sqrt_c);
    const float floor_val = fmaxf(kEpsilon, kDelta * max_sqrt_c);
    float d1 = 0.f;
    if (tid < R) d1 = fmaxf(sc - rho1, floor_val);
    if (rho1 < floor_val) rho1 = floor_val;
    const float d1_sum = block_sum(tid < R ? d1 : 0.f, scratch);
    const float beta1 = rho1 * (1.f + alpha) + alpha * d1_sum / D;
    if (tid < R) {
        const float e1 = 1.f / (beta1 / d1 + 1.f);
        d1g[(int64_t)f * R + tid] = d1;
        wcg[(int64_t)f * R + tid] = ((1.f - eta) / eN) * dpr[tid];
        lam[tid] = eN * sqrtf(e1) / sc;     // left_i  (lam reused; order[] already consumed for values)
    }
    if (tid == 0) rho1g[f] = rho1;
    __syncthreads();
    // A[i][j] = U[j][i] · left_i · inv_sqrt_e_j ,  U[:, i] = V[:, order[i]]
    float* A = Ag + (int64_t)f * R * R;
    for (int idx = tid; idx < R * R; idx += kThreads) {
        const int i = idx / R, j = idx % R;
        A[idx] = V[j * ld + order[i]] * lam[i] * ise[j];
    }
}

__global__ void __launch_bounds__(kThreads)
ngd_eig_kernel(const float* __restrict__ Kg, const float* __restrict__ Lg, const float* __restrict__ dg,
               const float* __restrict__ rhog, const float* __restrict__ trg,
               float* __restrict__ Ag, float* __restrict__ wcg, float* __restrict__ d1g, float* __restrict__ rho1g,
               int R, int D, float N, float alpha, float eta, int max_sweeps) {
    ngd_eig_body(Kg, Lg, dg, rhog, trg, Ag, wcg, d1g, rho1g, blockIdx.x, R, D, N, alpha, eta, max_sweeps);
}

// One descriptor per factor (= per CTA): factors of every shape group of the model in ONE launch.
struct NgdEigDesc {
    const float *K, *L, *d, *rho, *tr;
    float *A, *wc, *d1, *rho1;
    int R, D;
    float N, alpha, eta;
    int pad;
};
static_assert(sizeof(NgdEigDesc) == 96, "descriptor layout is mirrored in fdt/ops/cuda_ops.py");

__global__ void __launch_bounds__(kThreads)
ngd_eig_multi_kernel(const NgdEigDesc* __restrict__ descs, int max_sweeps) {
    const NgdEigDesc d = descs[blockIdx.x];
    ngd_eig_body(d.K, d.L, d.d, d.rho, d.tr, d.A, d.wc, d.d1, d.rho1, 0, d.R, d.D, d.N, d.alpha, d.eta, max_sweeps);
}

}  // namespace

// One launch refreshes `n` same-shaped factors.  K, L: [n,R,R]; d: [n,R]; rho, trXXt: [n];
// outputs A: [n,R,R]; w_coeff, d1: [n,R]; rho1: [n].
FDT_API int fdt_ngd_eig_update(const float* K, const float* L, const float* d, const float* rho, const float* trXXt,
                               float* A, float* w_coeff, float* d1, float* rho1, int n, int R, int D, float N,
                               float alpha, float eta, int max_sweeps, cudaStream_t st) {
    if (R < 1 || R > kMaxR || R >= D) return -1;
    const int n2 = (R + 1) & ~1;
    const int ld = R + 1 + ((R & 1) ? 1 : 0);
    const size_t smem = sizeof(float) * (2 * (size_t)n2 * ld + 6 * kMaxR + 32);
    cudaError_t e = cudaFuncSetAttribute(ngd_eig_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    if (e != cudaSuccess) return (int)e;
    ngd_eig_kernel<<<n, kThreads, smem, st>>>(K, L, d, rho, trXXt, A, w_coeff, d1, rho1, R, D, N, alpha, eta, max_sweeps);
    FDT_CHECK_LAUNCH();
    return 0;
}

// Heterogeneous batch: `count` descriptors (device memory, 96 bytes each, see NgdEigDesc), one CTA each.
FDT_API int fdt_ngd_eig_update_multi(const void* descs, int count, int max_sweeps, cudaStream_t st) {
    if (count <= 0) return 0;
    const size_t smem = sizeof(float) * (2 * (size_t)kMaxR * (kMaxR + 1) + 6 * kMaxR + 32);
    cudaError_t e = cudaFuncSetAttribute(ngd_eig_multi_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    if (e != cudaSuccess) return (int)e;
    ngd_eig_multi_kernel<<<count, kThreads, smem, st>>>((const NgdEigDesc*)descs, max_sweeps);
    FDT_CHECK_LAUNCH();
    return 0;
}

// ---------------------------------------------------------------------------------------------
// Fused element-wise tails of the NGD step (ngd_optimizer.py:151-168 and the state hand-over at
// :322-324), which in tensor-op form cost ~10 + ~15 tiny launches per factor batch.
namespace {

// sums[f] = Σ X_f², sums[n + f] = Σ X̂_f²  (sums zeroed by the caller); grid (chunks, n)
__global__ void __launch_bounds__(256)
ngd_sumsq2_kernel(const float* __restrict__ X, const float* __restrict__ Xh, float* __restrict__ sums, int n, int64_t M) {
    const int f = blockIdx.y;
    const float* x = X + (int64_t)f * M;
    const float* xh = Xh + (int64_t)f * M;
    float a = 0.f, b = 0.f;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < M; i += (int64_t)gridDim.x * blockDim.x) {
        const float u = x[i], v = xh[i];
        a += u * u; b += v * v;
    }
    a = warp_sum(a); b = warp_sum(b);
    __shared__ float ra[8], rb[8];
    if ((threadIdx.x & 31) == 0) { ra[threadIdx.x >> 5] = a; rb[threadIdx.x >> 5] = b; }
    __syncthreads();
    if (threadIdx.x == 0) {
        float ta = 0.f, tb = 0.f;
        for (int i = 0; i < 8; ++i) { ta += ra[i]; tb += rb[i]; }
        atomicAdd(sums + f, ta);
        atomicAdd(sums + n + f, tb);
    }
}
// out_f = isnan(p1) ? X_f : X̂_f · sqrt(p0 / (p1 + 1e-30)),  p0 = p_in[f], p1 = p_out[f].  The choice is uniform per factor,
// so only one of the two inputs is read; 16 bytes per access when the factor length allows it.
template <bool kVec>
__global__ void __launch_bounds__(256)
ngd_rescale_kernel(const float* __restrict__ X, const float* Xh, const float* __restrict__ p_in, const float* __restrict__ p_out,
                   float* out, int64_t M) {                 // out may alias Xh (in-place rescale): no __restrict__ on the pair
    const int f = blockIdx.y;
    const float p0 = p_in[f], p1 = p_out[f];
    const bool bad = isnan(p1);
    const float sc = bad ? 1.f : sqrtf(p0 / (p1 + 1.0e-30f));
    const float* src = (bad ? X : Xh) + (int64_t)f * M;
    float* o = out + (int64_t)f * M;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    if (kVec) {
        const float4* s4 = reinterpret_cast<const float4*>(src);
        float4* o4 = reinterpret_cast<float4*>(o);
        for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < (M >> 2); i += stride) {
            float4 v = s4[i];
            v.x *= sc; v.y *= sc; v.z *= sc; v.w *= sc;
            o4[i] = v;
        }
    } else {
        for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < M; i += stride) o[i] = src[i] * sc;
    }
}

struct NgdCommitDesc {          // one per factor; mirrored in fdt/ops/cuda_ops.py
    const float *W1, *d1, *rho1;
    float *W, *d, *rho;
    int R, D;
};
static_assert(sizeof(NgdCommitDesc) == 56, "descriptor layout is mirrored in python");

// a factor whose refresh produced a non-finite number keeps its previous (W, d, rho).
// Two passes so that a large factor (the 80 x 30522 token-embedding factor) is spread over many CTAs:
// pass 1 raises bad[f] if anything is non-finite, pass 2 copies where bad[f] == 0.   grid = (chunks, factors)
__global__ void __launch_bounds__(256)
ngd_check_kernel(const NgdCommitDesc* __restrict__ descs, int* __restrict__ bad_flags) {
    const NgdCommitDesc c = descs[blockIdx.y];
    const int nW = c.R * c.D;
    int bad = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nW; i += gridDim.x * blockDim.x) bad |= !isfinite(c.W1[i]);
    if (blockIdx.x == 0) {
        for (int i = threadIdx.x; i < c.R; i += blockDim.x) bad |= !isfinite(c.d1[i]);
        if (threadIdx.x == 0) bad |= !isfinite(c.rho1[0]);
    }
    bad = __syncthreads_or(bad);
    if (bad && threadIdx.x == 0) atomicOr(bad_flags + blockIdx.y, 1);
}
__global__ void __launch_bounds__(256)
ngd_copy_kernel(const NgdCommitDesc* __restrict__ descs, const int* __restrict__ bad_flags) {
    if (bad_flags[blockIdx.y]) return;
    const NgdCommitDesc c = descs[blockIdx.y];
    const int nW = c.R * c.D;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nW; i += gridDim.x * blockDim.x) c.W[i] = c.W1[i];
    if (blockIdx.x == 0) {
        for (int i = threadIdx.x; i < c.R; i += blockDim.x) c.d[i] = c.d1[i];
        if (threadIdx.x == 0) c.rho[0] = c.rho1[0];
    }
}


// ---- W_{t+1} = A (J + w_coeff ∘ W_t) for every factor in one launch, fused with the finiteness check of the
// hand-over (ngd_optimizer.py:311-324: torch.mm + add per factor).  grid = (column chunks, factors).
struct NgdFinishDesc {          // NgdCommitDesc + the operands of W1; mirrored in fdt/ops/cuda_ops.py
    float *W1;
    const float *d1, *rho1;
    float *W, *d, *rho;
    const float *A, *J, *wc;
    int R, D;
};
static_assert(sizeof(NgdFinishDesc) == 80, "descriptor layout is mirrored in python");
constexpr int kW1Cols = 64;

__global__ void __launch_bounds__(256)
ngd_w1_check_kernel(const NgdFinishDesc* __restrict__ descs, int* __restrict__ bad_flags) {
    const NgdFinishDesc c = descs[blockIdx.y];
    const int R = c.R, D = c.D;
    extern __shared__ float sm[];
    float* As = sm;                         // [R][R + 1]
    float* Bs = sm + kMaxR * (kMaxR + 1);   // [R][kW1Cols]
    int bad = 0;
    bool loaded = false;
    for (int c0 = blockIdx.x * kW1Cols; c0 < D; c0 += gridDim.x * kW1Cols) {
        if (!loaded) {
            for (int i = threadIdx.x; i < R * R; i += blockDim.x) As[(i / R) * (R + 1) + i % R] = c.A[i];
            loaded = true;
        }
        const int nc = min(kW1Cols, D - c0);
        __syncthreads();
        for (int i = threadIdx.x; i < R * kW1Cols; i += blockDim.x) {
            const int r = i / kW1Cols, cc = i % kW1Cols;
            Bs[i] = cc < nc ? fmaf(c.wc[r], c.W[(int64_t)r * D + c0 + cc], c.J[(int64_t)r * D + c0 + cc]) : 0.f;
        }
        __syncthreads();
        for (int i = threadIdx.x; i < R * kW1Cols; i += blockDim.x) {
            const int r = i / kW1Cols, cc = i % kW1Cols;
            if (cc >= nc) continue;
            float acc = 0.f;
            for (int j = 0; j < R; ++j) acc = fmaf(As[r * (R + 1) + j], Bs[j * kW1Cols + cc], acc);
            c.W1[(int64_t)r * D + c0 + cc] = acc;
            bad |= !isfinite(acc);
        }
    }
    if (blockIdx.x == 0) {
        for (int i = threadIdx.x; i < R; i += blockDim.x) bad |= !isfinite(c.d1[i]);
        if (threadIdx.x == 0) bad |= !isfinite(c.rho1[0]);
    }
    bad = __syncthreads_or(bad);
    if (bad && threadIdx.x == 0) atomicOr(bad_flags + blockIdx.y, 1);
}
__global__ void __launch_bounds__(256)
ngd_finish_copy_kernel(const NgdFinishDesc* __restrict__ descs, const int* __restrict__ bad_flags) {
    if (bad_flags[blockIdx.y]) return;
    const NgdFinishDesc c = descs[blockIdx.y];
    const int nW = c.R * c.D;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nW; i += gridDim.x * blockDim.x) c.W[i] = c.W1[i];
    if (blockIdx.x == 0) {
        for (int i = threadIdx.x; i < c.R; i += blockDim.x) c.d[i] = c.d1[i];
        if (threadIdx.x == 0) c.rho[0] = c.rho1[0];
    }
}

}  // namespace

// W1 = A (J + wc ∘ W) for `count` factors (80-byte NgdFinishDesc each, device memory), then the guarded hand-over
// W←W1, d←d1, rho←rho1.  flags: `count` ints of caller scratch; max_D / max_elems size the grids.
FDT_API int fdt_ngd_finish_multi(const void* descs, int count, int* flags, int max_D, int max_elems, cudaStream_t st) {
    if (count <= 0) return 0;
    static bool attr = false;
    const size_t smem = sizeof(float) * ((size_t)kMaxR * (kMaxR + 1) + (size_t)kMaxR * kW1Cols);
    if (!attr) {
        cudaError_t e = cudaFuncSetAttribute(ngd_w1_check_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return (int)e;
        attr = true;
    }
    int cchunks = (max_D + kW1Cols - 1) / kW1Cols;
    if (cchunks > 32) cchunks = 32;
    if (cchunks < 1) cchunks = 1;
    int chunks = (max_elems + 256 * 8 - 1) / (256 * 8);
    if (chunks < 1) chunks = 1;
    if (chunks > 64) chunks = 64;
    cudaMemsetAsync(flags, 0, sizeof(int) * count, st);
    ngd_w1_check_kernel<<<dim3(cchunks, count), 256, smem, st>>>((const NgdFinishDesc*)descs, flags);
    ngd_finish_copy_kernel<<<dim3(chunks, count), 256, 0, st>>>((const NgdFinishDesc*)descs, flags);
    FDT_CHECK_LAUNCH();
    return 0;
}

// sums: [2n] zero on entry.  X, Xh: [n, M] contiguous fp32.
FDT_API int fdt_ngd_sumsq2(const float* X, const float* Xh, float* sums, int n, int64_t M, cudaStream_t st) {
    int chunks = (int)((M + 256 * 8 - 1) / (256 * 8));
    if (chunks < 1) chunks = 1;
    const int cap = (148 * 8 + n - 1) / n;
    if (chunks > cap) chunks = cap;
    ngd_sumsq2_kernel<<<dim3(chunks, n), 256, 0, st>>>(X, Xh, sums, n, M);
    FDT_CHECK_LAUNCH();
    return 0;
}

// p_in / p_out: [n] each — Σ of squares of the reference input and of Xh.  With a chain of factors applied to one tensor
// they come from different links: the norm the reference restores after every axis (ngd_optimizer.py:168) is the norm of the
// chain's first input, so one rescale at the end with (first Σin, last Σout) equals the per-axis rescales.
FDT_API int fdt_ngd_rescale2(const float* X, const float* Xh, const float* p_in, const float* p_out, float* out, int n, int64_t M,
                             cudaStream_t st) {
    if (n <= 0 || M <= 0) return n < 0 || M < 0 ? -1 : 0;
    const bool vec = M % 4 == 0 && ((reinterpret_cast<uintptr_t>(X) | reinterpret_cast<uintptr_t>(Xh) | reinterpret_cast<uintptr_t>(out)) & 15) == 0;
    const int64_t work = vec ? M / 4 : M;
    int chunks = (int)((work + 256 * 4 - 1) / (256 * 4));
    if (chunks < 1) chunks = 1;
    const int cap = (148 * 8 + n - 1) / n;
    if (chunks > cap) chunks = cap;
    if (vec) ngd_rescale_kernel<true><<<dim3(chunks, n), 256, 0, st>>>(X, Xh, p_in, p_out, out, M);
    else ngd_rescale_kernel<false><<<dim3(chunks, n), 256, 0, st>>>(X, Xh, p_in, p_out, out, M);
    FDT_CHECK_LAUNCH();
    return 0;
}
// sums: [2n] = Σ X_f² then Σ X̂_f²
FDT_API int fdt_ngd_rescale(const float* X, const float* Xh, const float* sums, float* out, int n, int64_t M, cudaStream_t st) {
    return fdt_ngd_rescale2(X, Xh, sums, sums + n, out, n, M, st);
}

// flags: `count` ints of caller-provided scratch.  max_elems = largest R*D among the factors (sizes the grid).
FDT_API int fdt_ngd_commit_multi(const void* descs, int count, int* flags, int max_elems, cudaStream_t st) {
    if (count <= 0) return 0;
    int chunks = (max_elems + 256 * 8 - 1) / (256 * 8);
    if (chunks < 1) chunks = 1;
    if (chunks > 64) chunks = 64;
    cudaMemsetAsync(flags, 0, sizeof(int) * count, st);
    ngd_check_kernel<<<dim3(chunks, count), 256, 0, st>>>((const NgdCommitDesc*)descs, flags);
    ngd_copy_kernel<<<dim3(chunks, count), 256, 0, st>>>((const NgdCommitDesc*)descs, flags);
    FDT_CHECK_LAUNCH();
    return 0;
}
