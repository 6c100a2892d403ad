// Row-wise kernels for the transformer: reference-semantics LayerNorm, fused embedding
// gather-sum-scale, dropout+residual, and the fused two-target (mixup) cross-entropy.
//
// Parity: custom LayerNorm a·(x−mean)/(std_unbiased+eps)+b (transformer.py:230-242; 5-6 ATen launches
// each way in the reference); Embeddings (pos + token + segment)·sqrt(d)
// (transformer.py:150-156, K12); λ·CE(y_a)+(1−λ)·CE(y_b) (resnet50_test.py:451-457, K16).
#include "common.cuh"
using namespace fdt;

namespace {

// ------------------------------------------------------------------ LayerNorm (one warp per row)
// x fp32 [M, D] → y bf16 [M, D]; stats[M,2] = (mean, 1/(std+eps)).  D % 128 == 0, D <= 1024.
template <int kD>
__global__ void __launch_bounds__(256)
ln_fwd_kernel(const float* __restrict__ x, const float* __restrict__ a, const float* __restrict__ b,
              __nv_bfloat16* __restrict__ y, float* __restrict__ stats, int M, float eps) {
    constexpr int kPer = kD / 32;
    const int row = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= M) return;
    const float* xr = x + (int64_t)row * kD;
    float v[kPer];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < kPer / 4; ++i) {
        float4 t = *reinterpret_cast<const float4*>(xr + (i * 32 + lane) * 4);
        v[4 * i] = t.x; v[4 * i + 1] = t.y; v[4 * i + 2] = t.z; v[4 * i + 3] = t.w;
        s += t.x + t.y + t.z + t.w;
    }
    const float mean = warp_sum(s) * (1.f / kD);
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < kPer; ++i) { float d = v[i] - mean; q += d * d; }
    const float sd = sqrtf(warp_sum(q) * (1.f / (kD - 1)));
    const float r = 1.f / (sd + eps);
    if (lane == 0) { stats[2 * row] = mean; stats[2 * row + 1] = r; }
    __nv_bfloat16* yr = y + (int64_t)row * kD;
#pragma unroll
    for (int i = 0; i < kPer / 4; ++i) {
        const int c = (i * 32 + lane) * 4;
        float4 av = *reinterpret_cast<const float4*>(a + c), bv = *reinterpret_cast<const float4*>(b + c);
        __nv_bfloat162 o0 = __floats2bfloat162_rn(av.x * (v[4 * i] - mean) * r + bv.x, av.y * (v[4 * i + 1] - mean) * r + bv.y);
        __nv_bfloat162 o1 = __floats2bfloat162_rn(av.z * (v[4 * i + 2] - mean) * r + bv.z, av.w * (v[4 * i + 3] - mean) * r + bv.w);
        *reinterpret_cast<__nv_bfloat162*>(yr + c) = o0;
        *reinterpret_cast<__nv_bfloat162*>(yr + c + 2) = o1;
    }
}

// dx (fp32, ACCUMULATED into dx_acc which already holds the residual-branch gradient) and da, db.
// g bf16 [M,D] is the gradient w.r.t. the LayerNorm output.
template <int kD>
__global__ void __launch_bounds__(256)
ln_bwd_kernel(const float* __restrict__ x, const float* __restrict__ a, const float* __restrict__ stats,
              const __nv_bfloat16* __restrict__ g, float* __restrict__ dx_acc, float* __restrict__ da,
              float* __restrict__ db, int M, float eps, int rows_per_block) {
    constexpr int kPer = kD / 32;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float pa[kPer], pb[kPer];
#pragma unroll
    for (int i = 0; i < kPer; ++i) pa[i] = pb[i] = 0.f;
    const int r0 = blockIdx.x * rows_per_block, r1 = min(M, r0 + rows_per_block);
    for (int row = r0 + warp; row < r1; row += 8) {
        const float mean = stats[2 * row], r = stats[2 * row + 1];
        const float sd = 1.f / r - eps;
        const float* xr = x + (int64_t)row * kD;
        const __nv_bfloat16* gr = g + (int64_t)row * kD;
        float xc[kPer], dh[kPer];
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int i = 0; i < kPer / 4; ++i) {
            const int c = (i * 32 + lane) * 4;
            float4 t = *reinterpret_cast<const float4*>(xr + c);
            float4 av = *reinterpret_cast<const float4*>(a + c);
            float2 g0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(gr + c));
            float2 g1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(gr + c + 2));
            float gg[4] = {g0.x, g0.y, g1.x, g1.y}, xx[4] = {t.x, t.y, t.z, t.w}, aa[4] = {av.x, av.y, av.z, av.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float d = xx[j] - mean;
                xc[4 * i + j] = d;
                dh[4 * i + j] = gg[j] * aa[j];
                pa[4 * i + j] += gg[j] * d * r;
                pb[4 * i + j] += gg[j];
                s1 += dh[4 * i + j];
                s2 += dh[4 * i + j] * d;
            }
        }
        s1 = warp_sum(s1) * (1.f / kD);
        s2 = warp_sum(s2);
        const float k2 = s2 * r * r / ((kD - 1) * fmaxf(sd, 1e-20f));
        float* dr = dx_acc + (int64_t)row * kD;
#pragma unroll
        for (int i = 0; i < kPer / 4; ++i) {
            const int c = (i * 32 + lane) * 4;
            float4 acc = *reinterpret_cast<float4*>(dr + c);
            acc.x += r * (dh[4 * i] - s1) - xc[4 * i] * k2;
            acc.y += r * (dh[4 * i + 1] - s1) - xc[4 * i + 1] * k2;
            acc.z += r * (dh[4 * i + 2] - s1) - xc[4 * i + 2] * k2;
            acc.w += r * (dh[4 * i + 3] - s1) - xc[4 * i + 3] * k2;
            *reinterpret_cast<float4*>(dr + c) = acc;
        }
    }
    __shared__ float red[8][kD + 1];
    // reduce da / db across the 8 warps, then one atomic per column per block
    for (int pass = 0; pass < 2; ++pass) {
#pragma unroll
        for (int i = 0; i < kPer / 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) red[warp][(i * 32 + lane) * 4 + j] = pass ? pb[4 * i + j] : pa[4 * i + j];
        __syncthreads();
        for (int c = threadIdx.x; c < kD; c += 256) {
            float t = 0.f;
#pragma unroll
            for (int w = 0; w < 8; ++w) t += red[w][c];
            atomicAdd((pass ? db : da) + c, t);
        }
        __syncthreads();
    }
}

// ------------------------------------------------------------------ embedding
// out fp32 [B*L, D] = (pos[l] + tok[tokens] + seg[types]) * scale
__global__ void embed_fwd_kernel(const int64_t* __restrict__ tokens, const int64_t* __restrict__ types,
                                 const float* __restrict__ tok_w, const float* __restrict__ pos_w,
                                 const float* __restrict__ seg_w, float* __restrict__ out, int L, int D, float scale) {
    const int row = blockIdx.x;
    const int l = row % L;
    const float* t = tok_w + tokens[row] * (int64_t)D;
    const float* p = pos_w + (int64_t)l * D;
    const float* s = seg_w + types[row] * (int64_t)D;
    float* o = out + (int64_t)row * D;
    for (int c = threadIdx.x * 4; c < D; c += blockDim.x * 4) {
        float4 a = *reinterpret_cast<const float4*>(t + c), b = *reinterpret_cast<const float4*>(p + c),
               d = *reinterpret_cast<const float4*>(s + c);
        float4 r = make_float4((a.x + b.x + d.x) * scale, (a.y + b.y + d.y) * scale, (a.z + b.z + d.z) * scale, (a.w + b.w + d.w) * scale);
        *reinterpret_cast<float4*>(o + c) = r;
    }
}
// scatter-add of g·scale into the three tables (fp32 atomics; token rows are sparse)
__global__ void embed_bwd_kernel(const int64_t* __restrict__ tokens, const int64_t* __restrict__ types,
                                 const float* __restrict__ g, float* __restrict__ dtok, float* __restrict__ dpos,
                                 float* __restrict__ dseg, int L, int D, float scale) {
    const int row = blockIdx.x;
    const int l = row % L;
    float* t = dtok + tokens[row] * (int64_t)D;
    float* p = dpos + (int64_t)l * D;
    float* s = dseg + types[row] * (int64_t)D;
    const float* gr = g + (int64_t)row * D;
    for (int c = threadIdx.x; c < D; c += blockDim.x) {
        const float v = gr[c] * scale;
        atomicAdd(t + c, v);
        atomicAdd(p + c, v);
        atomicAdd(s + c, v);
    }
}

// ------------------------------------------------------------------ dropout + residual
// out fp32 = res + dropout(x)   (x fp32 or bf16); mask regenerated from (seed, offset) in backward
template <typename T>
__global__ void dropout_add_kernel(const T* __restrict__ x, const float* __restrict__ res, float* __restrict__ out,
                                   int64_t n, float p, const int64_t* __restrict__ seed_off) {
    const int64_t i4 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i4 >= n) return;
    Philox rng((uint64_t)seed_off[0]);
    uint4 r = rng((uint64_t)(i4 >> 2), (uint64_t)seed_off[1]);
    const uint32_t rr[4] = {r.x, r.y, r.z, r.w};
    const float keep = 1.f / (1.f - p);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        if (i4 + j < n) {
            float v = (float)x[i4 + j];
            v = (p > 0.f) ? (u32_to_unit(rr[j]) >= p ? v * keep : 0.f) : v;
            out[i4 + j] = (res ? res[i4 + j] : 0.f) + v;
        }
    }
}
// dx = dropout_mask(g)  (g fp32 → T)
template <typename T>
__global__ void dropout_bwd_kernel(const float* __restrict__ g, T* __restrict__ dx, int64_t n, float p,
                                   const int64_t* __restrict__ seed_off) {
    const int64_t i4 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i4 >= n) return;
    Philox rng((uint64_t)seed_off[0]);
    uint4 r = rng((uint64_t)(i4 >> 2), (uint64_t)seed_off[1]);
    const uint32_t rr[4] = {r.x, r.y, r.z, r.w};
    const float keep = 1.f / (1.f - p);
#pragma unroll
    for (int j = 0; j < 4; ++j)
        if (i4 + j < n) {
            float v = g[i4 + j];
            dx[i4 + j] = (T)((p > 0.f) ? (u32_to_unit(rr[j]) >= p ? v * keep : 0.f) : v);
        }
}

// ------------------------------------------------------------------ two-target cross entropy
// logits fp32 [B, C<=32]; lam scalar (lam_n==1) or per-sample.  loss_sum += Σ_i loss_i / B ;
// dlogits = (softmax − lam·1[y_a] − (1−lam)·1[y_b]) / B ; correct += Σ_i lam·[argmax==y_a] + (1−lam)·[argmax==y_b]
__global__ void mixup_ce_kernel(const float* __restrict__ logits, const int64_t* __restrict__ ya,
                                const int64_t* __restrict__ yb, const float* __restrict__ lam, int lam_n,
                                float* __restrict__ out /*[2]: loss, correct*/, float* __restrict__ dlogits,
                                float* __restrict__ dlam, int B, int C) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    float loss = 0.f, corr = 0.f;
    if (i < B) {
        const float* z = logits + (int64_t)i * C;
        float m = -INFINITY; int am = 0;
        for (int c = 0; c < C; ++c) if (z[c] > m) { m = z[c]; am = c; }
        float se = 0.f;
        for (int c = 0; c < C; ++c) se += __expf(z[c] - m);
        const float lse = m + __logf(se);
        const float l = lam_n == 1 ? lam[0] : lam[i];
        const int a = (int)ya[i], b = (int)yb[i];
        loss = (l * (lse - z[a]) + (1.f - l) * (lse - z[b])) / B;
        corr = l * (am == a) + (1.f - l) * (am == b);
        if (dlam) dlam[i] = (z[b] - z[a]) / B;            // d/dlam of lam·(lse − z_a) + (1 − lam)·(lse − z_b), per-sample lam
        if (dlogits) {
            float* d = dlogits + (int64_t)i * C;
            const float invB = 1.f / B;
            for (int c = 0; c < C; ++c) {
                float p = __expf(z[c] - lse);
                d[c] = (p - (c == a ? l : 0.f) - (c == b ? 1.f - l : 0.f)) * invB;
            }
        }
    }
    loss = warp_sum(loss); corr = warp_sum(corr);
    if ((threadIdx.x & 31) == 0) { atomicAdd(out, loss); atomicAdd(out + 1, corr); }
}

// ---- out = dropout(act(z)) in one pass (bf16 in / out), 8 elements per thread; the mask is the Philox stream of
// (seed, offset) indexed by element / 4 — the same stream linear_bwd_prep regenerates in the backward pass.
// ref: the FFN's gelu → nn.Dropout pair (transformer.py:296-306), two elementwise kernels + a stored mask there.
__device__ __forceinline__ void keep8(const int64_t* __restrict__ seed_off, int64_t elem0, float p, float* m) {
    Philox rng((uint64_t)seed_off[0]);
    const uint4 a = rng((uint64_t)(elem0 >> 2), (uint64_t)seed_off[1]);
    const uint4 b = rng((uint64_t)(elem0 >> 2) + 1, (uint64_t)seed_off[1]);
    const uint32_t rr[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
    const float keep = 1.f / (1.f - p);
#pragma unroll
    for (int j = 0; j < 8; ++j) m[j] = u32_to_unit(rr[j]) >= p ? keep : 0.f;
}
__global__ void __launch_bounds__(256)
act_dropout_fwd_kernel(const __nv_bfloat16* __restrict__ z, __nv_bfloat16* __restrict__ out, int64_t n8, int act, float p,
                       const int64_t* __restrict__ seed_off) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (int64_t)gridDim.x * blockDim.x) {
        float f[8], m[8];
        unpack8(ld8(z + i * 8), f);
        if (p > 0.f) keep8(seed_off, i * 8, p, m);
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] = act_fwd(f[j], act) * (p > 0.f ? m[j] : 1.f);
        st8(out + i * 8, pack8(f));
    }
}

// ---- Linear backward preparation: gz = gout ∘ act'(saved) as the bf16 GEMM operand, and the bias gradient
// db += Σ_rows gz in the same pass (ref: the autograd of F.linear + activation — an elementwise chain, a cast and
// a reduction per layer).  `saved` is the pre-activation z for GELU, the output for ReLU / tanh.
// block = 32 column groups of 8 x 8 row lanes; grid = (column slabs of 256, row chunks).
template <bool kF32In>
__global__ void __launch_bounds__(256)
linear_bwd_prep_kernel(const void* __restrict__ gout, const __nv_bfloat16* __restrict__ saved, __nv_bfloat16* __restrict__ gz,
                       float* __restrict__ db, int64_t T, int N, int act, float p, const int64_t* __restrict__ seed_off) {
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int col = (blockIdx.x * 32 + tx) * 8;
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    if (col < N) {
        // four rows per trip: their loads are issued together (a thread otherwise has one 16-byte load in flight at a time,
        // far too little to cover the memory latency with ~900 resident threads per SM)
        constexpr int kU = 4;
        const int64_t step = (int64_t)gridDim.y * 8;
        for (int64_t r0 = (int64_t)blockIdx.y * 8 + ty; r0 < T; r0 += kU * step) {
            float g[kU][8];
            bf16x8 sv8[kU];
#pragma unroll
            for (int u = 0; u < kU; ++u) {
                const int64_t r = r0 + u * step;
                if (r >= T) break;
                const int64_t off = r * N + col;
                if (kF32In) {
                    const float4 a = *reinterpret_cast<const float4*>(static_cast<const float*>(gout) + off);
                    const float4 b = *reinterpret_cast<const float4*>(static_cast<const float*>(gout) + off + 4);
                    g[u][0] = a.x; g[u][1] = a.y; g[u][2] = a.z; g[u][3] = a.w; g[u][4] = b.x; g[u][5] = b.y; g[u][6] = b.z; g[u][7] = b.w;
                } else {
                    unpack8(ld8(static_cast<const __nv_bfloat16*>(gout) + off), g[u]);
                }
                if (act != ACT_NONE) sv8[u] = ld8(saved + off);
            }
#pragma unroll
            for (int u = 0; u < kU; ++u) {
                const int64_t r = r0 + u * step;
                if (r >= T) break;
                const int64_t off = r * N + col;
                if (p > 0.f) {              // the forward applied dropout after the activation: same Philox stream
                    float m[8];
                    keep8(seed_off, off, p, m);
#pragma unroll
                    for (int j = 0; j < 8; ++j) g[u][j] *= m[j];
                }
                if (act != ACT_NONE) {
                    float sv[8];
                    unpack8(sv8[u], sv);
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        float d;
                        if (act == ACT_GELU) {
                            d = gelu_bwd_from_in(sv[j]);
                        } else {
                            d = act_bwd_from_out(sv[j], act);
                        }
                        g[u][j] *= d;
                    }
                }
                if (gz) st8(gz + off, pack8(g[u]));
                if (db) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) acc[j] += gz ? __bfloat162float(__float2bfloat16(g[u][j])) : g[u][j];
                }
            }
        }
    }
    if (db == nullptr) return;
    __shared__ float red[8][32 * 8 + 1];
#pragma unroll
    for (int j = 0; j < 8; ++j) red[ty][tx * 8 + j] = acc[j];
    __syncthreads();
    const int c = threadIdx.x;                // one of the slab's 256 columns
    const int gc = blockIdx.x * 256 + c;
    if (gc < N) {
        float t = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) t += red[k][c];
        atomicAdd(db + gc, t);
    }
}

}  // namespace

// gout: [T, N] bf16 or fp32; saved: [T, N] bf16 or NULL (act == 0); gz: [T, N] bf16 out or NULL (only the bias
// gradient is wanted); db: [N] fp32, ACCUMULATED into (+=), or NULL.  N % 8 == 0.
static int linear_bwd_prep_impl(const void* gout, int gout_is_f32, const void* saved, void* gz, float* db, int64_t T, int N,
                                int act, float p, const int64_t* seed_off, cudaStream_t st) {
    if (N % 8 || (act != ACT_NONE && saved == nullptr) || (p > 0.f && seed_off == nullptr)) return -1;
    const int slabs = (N + 255) / 256;
    int chunks = (int)((T + 63) / 64);
    const int cap = (148 * 8 + slabs - 1) / slabs;
    if (chunks > cap) chunks = cap;
    if (chunks < 1) chunks = 1;
    dim3 grid(slabs, chunks);
    if (gout_is_f32)
        linear_bwd_prep_kernel<true><<<grid, 256, 0, st>>>(gout, (const __nv_bfloat16*)saved, (__nv_bfloat16*)gz, db, T, N, act, p, seed_off);
    else
        linear_bwd_prep_kernel<false><<<grid, 256, 0, st>>>(gout, (const __nv_bfloat16*)saved, (__nv_bfloat16*)gz, db, T, N, act, p, seed_off);
    FDT_CHECK_LAUNCH();
    return 0;
}

FDT_API int fdt_linear_bwd_prep(const void* gout, int gout_is_f32, const void* saved, void* gz, float* db, int64_t T, int N,
                                int act, cudaStream_t st) {
    return linear_bwd_prep_impl(gout, gout_is_f32, saved, gz, db, T, N, act, 0.f, nullptr, st);
}
// same for a layer whose forward was out = dropout(act(z), p): the mask of (seed, offset) is regenerated here
FDT_API int fdt_linear_bwd_prep_dropout(const void* gout, int gout_is_f32, const void* saved, void* gz, float* db, int64_t T,
                                        int N, int act, float p, const int64_t* seed_off, cudaStream_t st) {
    return linear_bwd_prep_impl(gout, gout_is_f32, saved, gz, db, T, N, act, p, seed_off, st);
}
// out = dropout(act(z), p), bf16, n % 8 == 0; seed_off = {seed, offset} (device) or NULL when p == 0
FDT_API int fdt_act_dropout_fwd(const void* z, void* out, int64_t n, int act, float p, const int64_t* seed_off, cudaStream_t st) {
    if (n % 8 || (p > 0.f && seed_off == nullptr) || p >= 1.f) return -1;
    const int64_t n8 = n / 8;
    int64_t blocks = (n8 + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    if (blocks < 1) blocks = 1;
    act_dropout_fwd_kernel<<<(int)blocks, 256, 0, st>>>((const __nv_bfloat16*)z, (__nv_bfloat16*)out, n8, act, p, seed_off);
    FDT_CHECK_LAUNCH();
    return 0;
}

FDT_API int fdt_ln_fwd(const float* x, const float* a, const float* b, void* y, float* stats, int M, int D, float eps,
                       cudaStream_t st) {
    const int grid = (M + 7) / 8;
    if (D == 512) ln_fwd_kernel<512><<<grid, 256, 0, st>>>(x, a, b, (__nv_bfloat16*)y, stats, M, eps);
    else if (D == 256) ln_fwd_kernel<256><<<grid, 256, 0, st>>>(x, a, b, (__nv_bfloat16*)y, stats, M, eps);
    else if (D == 128) ln_fwd_kernel<128><<<grid, 256, 0, st>>>(x, a, b, (__nv_bfloat16*)y, stats, M, eps);
    else if (D == 1024) ln_fwd_kernel<1024><<<grid, 256, 0, st>>>(x, a, b, (__nv_bfloat16*)y, stats, M, eps);
    else return -1;
    FDT_CHECK_LAUNCH();
    return 0;
}

// da / db must be zero on entry (or hold a value to accumulate into); dx_acc is accumulated into.
FDT_API int fdt_ln_bwd(const float* x, const float* a, const float* stats, const void* g, float* dx_acc, float* da,
                       float* db, int M, int D, float eps, cudaStream_t st) {
    int rpb = (M + 148 * 2 - 1) / (148 * 2);
    rpb = (rpb + 7) / 8 * 8;
    if (rpb < 8) rpb = 8;
    const int grid = (M + rpb - 1) / rpb;
#define LNB(DD) ln_bwd_kernel<DD><<<grid, 256, 0, st>>>(x, a, stats, (const __nv_bfloat16*)g, dx_acc, da, db, M, eps, rpb)
    if (D == 512) LNB(512); else if (D == 256) LNB(256); else if (D == 128) LNB(128); else if (D == 1024) LNB(1024); else return -1;
#undef LNB
    FDT_CHECK_LAUNCH();
    return 0;
}

FDT_API int fdt_embed_fwd(const int64_t* tokens, const int64_t* types, const float* tok_w, const float* pos_w,
                          const float* seg_w, float* out, int rows, int L, int D, float scale, cudaStream_t st) {
    if (D % 4) return -1;
    embed_fwd_kernel<<<rows, 128, 0, st>>>(tokens, types, tok_w, pos_w, seg_w, out, L, D, scale);
    FDT_CHECK_LAUNCH();
    return 0;
}

FDT_API int fdt_embed_bwd(const int64_t* tokens, const int64_t* types, const float* g, float* dtok, float* dpos,
                          float* dseg, int rows, int L, int D, float scale, cudaStream_t st) {
    embed_bwd_kernel<<<rows, 128, 0, st>>>(tokens, types, g, dtok, dpos, dseg, L, D, scale);
    FDT_CHECK_LAUNCH();
    return 0;
}

FDT_API int fdt_dropout_add(const void* x, int x_is_bf16, const float* res, float* out, int64_t n, float p,
                            const int64_t* seed_off, cudaStream_t st) {
    const int grid = (int)((n / 4 + 255) / 256) + 1;
    if (x_is_bf16) dropout_add_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)x, res, out, n, p, seed_off);
    else dropout_add_kernel<float><<<grid, 256, 0, st>>>((const float*)x, res, out, n, p, seed_off);
    FDT_CHECK_LAUNCH();
    return 0;
}

FDT_API int fdt_dropout_bwd(const float* g, void* dx, int dx_is_bf16, int64_t n, float p, const int64_t* seed_off,
                            cudaStream_t st) {
    const int grid = (int)((n / 4 + 255) / 256) + 1;
    if (dx_is_bf16) dropout_bwd_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(g, (__nv_bfloat16*)dx, n, p, seed_off);
    else dropout_bwd_kernel<float><<<grid, 256, 0, st>>>(g, (float*)dx, n, p, seed_off);
    FDT_CHECK_LAUNCH();
    return 0;
}

// out[2] must be zero on entry.
FDT_API int fdt_mixup_ce(const float* logits, const int64_t* ya, const int64_t* yb, const float* lam, int lam_n,
                         float* out, float* dlogits, int B, int C, cudaStream_t st) {
    mixup_ce_kernel<<<(B + 127) / 128, 128, 0, st>>>(logits, ya, yb, lam, lam_n, out, dlogits, nullptr, B, C);
    FDT_CHECK_LAUNCH();
    return 0;
}

// same, plus dlam [B] = d loss / d lam_i for a per-sample (learnable) mixing weight (lam_n == B)
FDT_API int fdt_mixup_ce_ex(const float* logits, const int64_t* ya, const int64_t* yb, const float* lam, int lam_n,
                            float* out, float* dlogits, float* dlam, int B, int C, cudaStream_t st) {
    if (dlam && lam_n != B) return -1;
    mixup_ce_kernel<<<(B + 127) / 128, 128, 0, st>>>(logits, ya, yb, lam, lam_n, out, dlogits, dlam, B, C);
    FDT_CHECK_LAUNCH();
    return 0;
}

// ------------------------------------------------------------------ pooled classification head (transformer.py:61-84, 94-101)
// CLS gather, pooled-vector dropout and manifold mixup as three small kernels around the two tcgen05 GEMMs of the head
// (the reference: index + Linear + tanh + dropout + randperm gather + lerp = ~10 ATen launches each way).
namespace {

// cls[b, :] = bf16(x[b, 0, :])            x fp32 [B, L, D]
__global__ void cls_gather_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ cls, int B, int L, int D) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;          // one thread per 4 elements
    if (i >= B * (D / 4)) return;
    const int b = i / (D / 4), c = (i - b * (D / 4)) * 4;
    const float4 v = *reinterpret_cast<const float4*>(x + ((int64_t)b * L) * D + c);
    *reinterpret_cast<uint2*>(cls + (int64_t)b * D + c) = make_uint2(f2_to_bf2(v.x, v.y), f2_to_bf2(v.z, v.w));
}
// dx[b, l, :] = (l == 0) ? g[b, :] : 0     dx fp32 [B, L, D] (the gradient buffer the residual stream then accumulates into)
__global__ void cls_scatter_kernel(const float* __restrict__ g, float* __restrict__ dx, int B, int L, int D) {
    const int64_t n4 = (int64_t)B * L * (D / 4);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t row = i / (D / 4);
        const int c = (int)(i - row * (D / 4)) * 4;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (row % L == 0) v = *reinterpret_cast<const float4*>(g + (row / L) * D + c);
        *reinterpret_cast<float4*>(dx + row * D + c) = v;
    }
}
__device__ __forceinline__ float head_keep(const Philox& rng, uint64_t off, int row, int c, int D, float p, float inv_keep) {
    if (p <= 0.f) return 1.f;
    const uint4 r = rng((uint64_t)row * (D / 4) + (c >> 2), off);
    const uint32_t rr[4] = {r.x, r.y, r.z, r.w};
    return u32_to_unit(rr[c & 3]) >= p ? inv_keep : 0.f;
}
// out[i, :] = lam * drop(z[i, :]) + (1 - lam) * drop(z[perm[i], :])     z, out bf16 [B, D]; perm == NULL: no mixing (eval)
__global__ void head_mix_fwd_kernel(const __nv_bfloat16* __restrict__ z, const int64_t* __restrict__ perm,
                                    const float* __restrict__ lam, __nv_bfloat16* __restrict__ out, int B, int D, float p,
                                    const int64_t* __restrict__ seed_off) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B * D) return;
    const int b = i / D, c = i - b * D;
    Philox rng(seed_off ? (uint64_t)seed_off[0] : 0ull);
    const uint64_t off = seed_off ? (uint64_t)seed_off[1] : 0ull;
    const float ik = 1.f / (1.f - p);
    float v = __bfloat162float(z[i]) * head_keep(rng, off, b, c, D, p, ik);
    if (perm) {
        const int pb = (int)perm[b];
        const float l = lam[0];
        v = l * v + (1.f - l) * __bfloat162float(z[(int64_t)pb * D + c]) * head_keep(rng, off, pb, c, D, p, ik);
    }
    out[i] = __float2bfloat16(v);
}
// dz[j, :] = keep(j, :) * (lam * g[j, :] + (1 - lam) * Σ_{i: perm[i] = j} g[i, :])     dz fp32 [B, D], ZERO on entry
__global__ void head_mix_bwd_kernel(const __nv_bfloat16* __restrict__ g, const int64_t* __restrict__ perm,
                                    const float* __restrict__ lam, float* __restrict__ dz, int B, int D, float p,
                                    const int64_t* __restrict__ seed_off) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B * D) return;
    const int b = i / D, c = i - b * D;
    Philox rng(seed_off ? (uint64_t)seed_off[0] : 0ull);
    const uint64_t off = seed_off ? (uint64_t)seed_off[1] : 0ull;
    const float ik = 1.f / (1.f - p);
    const float gv = __bfloat162float(g[i]);
    if (perm) {
        const int pb = (int)perm[b];
        const float l = lam[0];
        atomicAdd(dz + i, l * gv * head_keep(rng, off, b, c, D, p, ik));
        atomicAdd(dz + (int64_t)pb * D + c, (1.f - l) * gv * head_keep(rng, off, pb, c, D, p, ik));
    } else {
        dz[i] = gv * head_keep(rng, off, b, c, D, p, ik);
    }
}

}  // namespace

FDT_API int fdt_cls_gather(const float* x, void* cls_bf16, int B, int L, int D, cudaStream_t st) {
    if (D % 4 || B < 1) return -1;
    cls_gather_kernel<<<(B * (D / 4) + 255) / 256, 256, 0, st>>>(x, (__nv_bfloat16*)cls_bf16, B, L, D);
    FDT_CHECK_LAUNCH();
    return 0;
}
FDT_API int fdt_cls_scatter(const float* g, float* dx, int B, int L, int D, cudaStream_t st) {
    if (D % 4 || B < 1) return -1;
    int64_t blocks = ((int64_t)B * L * (D / 4) + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    cls_scatter_kernel<<<(int)blocks, 256, 0, st>>>(g, dx, B, L, D);
    FDT_CHECK_LAUNCH();
    return 0;
}
FDT_API int fdt_head_mix_fwd(const void* z, const int64_t* perm, const float* lam, void* out, int B, int D, float p,
                             const int64_t* seed_off, cudaStream_t st) {
    if (D % 4 || B < 1 || p < 0.f || p >= 1.f) return -1;
    head_mix_fwd_kernel<<<(B * D + 255) / 256, 256, 0, st>>>((const __nv_bfloat16*)z, perm, lam, (__nv_bfloat16*)out, B, D, p,
                                                              p > 0.f ? seed_off : nullptr);
    FDT_CHECK_LAUNCH();
    return 0;
}
FDT_API int fdt_head_mix_bwd(const void* g, const int64_t* perm, const float* lam, float* dz, int B, int D, float p,
                             const int64_t* seed_off, cudaStream_t st) {
    if (D % 4 || B < 1 || p < 0.f || p >= 1.f) return -1;
    cudaMemsetAsync(dz, 0, sizeof(float) * (size_t)B * D, st);
    head_mix_bwd_kernel<<<(B * D + 255) / 256, 256, 0, st>>>((const __nv_bfloat16*)g, perm, lam, dz, B, D, p,
                                                              p > 0.f ? seed_off : nullptr);
    FDT_CHECK_LAUNCH();
    return 0;
}
