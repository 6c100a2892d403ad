// Channel-wise (batch-norm family) kernels on NHWC bf16 activations viewed as [P, C].
//
// Parity: the normalisation maths of the reference's FusedConvBN2DFunction (resnet.py:75-100,
// batch_norm_backward :37-69) and of nn.BatchNorm2d for its 7 "real" BN layers — there ~5 (fwd)
// and ~15 (bwd) ATen launches per layer over NCHW tensors.  Here each
// direction is two bandwidth-bound passes plus a C-sized finalise:
//   fwd : chan_reduce2 (Σy, Σy²)  → bn_finalize_fwd → chan_affine_act  (normalise [+res] + act)
//   bwd : chan_reduce2 (g = gout·act'(out), Σg, Σg·y) → bn_finalize_bwd → chan_affine_act (A·g + B·y + C)
// (the forward reduction disappears once the conv epilogue emits the sums).
#include "common.cuh"
#include <cstdio>
using namespace fdt;

namespace {

constexpr int kThreads = 256;

// cp.async ring: 16-byte global → shared copies that hold no registers while in flight (see chan_affine_act_ring_kernel)
constexpr int kRing = 8;                 // copies per input in flight per thread, affine kernels
constexpr int kRing2 = 6;                // same, reduction kernels
__device__ __forceinline__ void cp_async16(uint32_t smem_addr, const void* g) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr), "l"(g) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int kN>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(kN) : "memory"); }

// the ring pays a fixed prologue per thread: only tensors of at least this many 16-byte vectors take it
inline int64_t ring_min_vecs() {
    static const int64_t v = [] { const char* e = getenv("FDT_CHAN_RING_MIN"); return e ? atoll(e) : (int64_t)1 << 20; }();
    return v;
}

// shift of the forward affine map ŷ = scale·y + shift.  One spelling (a single fused multiply-add) for the forward kernels and
// for the backward kernels that re-derive the ReLU mask from y: the sign of scale·y + shift must come out the same.
__device__ __forceinline__ float bn_shift(float beta, float mean, float scale) { return __fmaf_rn(-mean, scale, beta); }

// Σ_p a[p,c]  and  Σ_p a[p,c]·b[p,c]   where a = g (optionally g = gout·act'(out), optionally stored).
// 2-D grid: blockIdx.y = tile of kTPR·8 channels, blockIdx.x = slab of rows.  A thread owns one
// 8-channel vector column (16-byte loads) and walks its slab two rows at a time; partial sums are
// combined with warp shuffles (threads of a warp that share a column), then across the 8 warps in
// shared memory, and leave the block as one atomicAdd per channel.
// kMask: 0 = none; 1 = act'(out_act) read from the activated output; 2 = ReLU mask re-derived from b = y, the conv output,
// as (scale·y + shift > 0) with the forward's coefficients (mctx = the forward ctx, mbeta = beta or NULL): neither the
// activated output is read nor the masked gradient stored.
template <int kTPR, int kMask, bool kSame, bool kRingMode = false>
__global__ void __launch_bounds__(kThreads, kMask == 2 ? 4 : 0)
chan_reduce2_kernel(const __nv_bfloat16* __restrict__ a, const __nv_bfloat16* __restrict__ b,
                    const __nv_bfloat16* __restrict__ out_act, __nv_bfloat16* __restrict__ g_store,
                    float* __restrict__ sums /*[2,C]*/, int64_t P, int C, int act, int rows_per_block,
                    const float* __restrict__ mctx, const float* __restrict__ mbeta) {
    pdl_wait();
    pdl_launch_dependents();
    constexpr int kLanes = kThreads / kTPR;
    const int cv = threadIdx.x % kTPR, rl = threadIdx.x / kTPR;
    const int c0 = (blockIdx.y * kTPR + cv) * 8;
    const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
    const int64_t r1 = min(P, r0 + rows_per_block);
    float s1[8], s2[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) s1[i] = s2[i] = 0.f;
    float msc[8], msh[8];
    if (kMask == 2) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            msc[i] = mctx[2 * C + c0 + i];
            msh[i] = bn_shift(mbeta ? mbeta[c0 + i] : 0.f, mctx[c0 + i], msc[i]);
        }
    }

    auto body = [&](int64_t r) {
        const int64_t off = r * C + c0;
        float fa[8], fb[8];
        unpack8(ld8(a + off), fa);
        if (!kSame) unpack8(ld8(b + off), fb);
        if (kMask == 1) {
            float fo[8];
            unpack8(ld8(out_act + off), fo);
#pragma unroll
            for (int i = 0; i < 8; ++i) fa[i] *= act_bwd_from_out(fo[i], act);
            if (g_store) {
                bf16x8 pk = pack8(fa);
                st8(g_store + off, pk);
                unpack8(pk, fa);          // reduce exactly what later passes will read
            }
        } else if (kMask == 2) {
#pragma unroll
            for (int i = 0; i < 8; ++i) fa[i] = fmaf(msc[i], fb[i], msh[i]) > 0.f ? fa[i] : 0.f;
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) { s1[i] += fa[i]; s2[i] += fa[i] * (kSame ? fa[i] : fb[i]); }
    };
    // four rows in flight per thread (issue the loads of all four before the dependent math)
    auto load = [&](int64_t r, bf16x8& va, bf16x8& vb, bf16x8& vo) {
        const int64_t off = r * C + c0;
        va = ld8(a + off);
        if (!kSame) vb = ld8(b + off);
        if (kMask == 1) vo = ld8(out_act + off);
    };
    auto consume = [&](int64_t r, const bf16x8& va, const bf16x8& vb, const bf16x8& vo) {
        float fa[8], fb[8];
        unpack8(va, fa);
        if (!kSame) unpack8(vb, fb);
        if (kMask == 1) {
            float fo[8];
            unpack8(vo, fo);
#pragma unroll
            for (int i = 0; i < 8; ++i) fa[i] *= act_bwd_from_out(fo[i], act);
            if (g_store) {
                bf16x8 pk = pack8(fa);
                st8(g_store + r * C + c0, pk);
                unpack8(pk, fa);
            }
        } else if (kMask == 2) {
#pragma unroll
            for (int i = 0; i < 8; ++i) fa[i] = fmaf(msc[i], fb[i], msh[i]) > 0.f ? fa[i] : 0.f;
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) { s1[i] += fa[i]; s2[i] += fa[i] * (kSame ? fa[i] : fb[i]); }
    };
    if (kRingMode) {
        // inputs streamed through a per-thread shared-memory ring (kRing2 rows per input in flight, no registers held)
        extern __shared__ uint4 ring[];                                    // [kRing2][kS][kThreads]
        constexpr int kS = 1 + (kSame ? 0 : 1) + (kMask == 1 ? 1 : 0);
        const int64_t first = r0 + rl;
        const int64_t count = first < r1 ? (r1 - first + kLanes - 1) / kLanes : 0;
        const uint32_t slot0 = (uint32_t)__cvta_generic_to_shared(ring + threadIdx.x);
        auto issue = [&](int64_t n) {
            const int64_t off = (first + n * kLanes) * C + c0;
            const uint32_t sa = slot0 + (uint32_t)((n % kRing2) * kS) * kThreads * 16;
            cp_async16(sa, a + off);
            if (!kSame) cp_async16(sa + kThreads * 16, b + off);
            if (kMask == 1) cp_async16(sa + (kSame ? 1 : 2) * kThreads * 16, out_act + off);
        };
#pragma unroll
        for (int k = 0; k < kRing2; ++k) {
            if (k < count) issue(k);
            cp_async_commit();
        }
        for (int64_t n = 0; n < count; ++n) {
            cp_async_wait<kRing2 - 1>();
            const uint4* sl = ring + (size_t)((n % kRing2) * kS) * kThreads + threadIdx.x;
            bf16x8 va, vb, vo;
            va.u = sl[0];
            if (!kSame) vb.u = sl[kThreads];
            if (kMask == 1) vo.u = sl[(kSame ? 1 : 2) * kThreads];
            if (n + kRing2 < count) issue(n + kRing2);
            cp_async_commit();
            consume(first + n * kLanes, va, vb, vo);
        }
    } else {
    int64_t r = r0 + rl;
    for (; r + 3 * kLanes < r1; r += 4 * kLanes) {
        bf16x8 va[4], vb[4], vo[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) load(r + u * kLanes, va[u], vb[u], vo[u]);
#pragma unroll
        for (int u = 0; u < 4; ++u) consume(r + u * kLanes, va[u], vb[u], vo[u]);
    }
    for (; r < r1; r += kLanes) body(r);
    }

    // threads of one warp that share a channel column: lane = k·kTPR + cv
#pragma unroll
    for (int off = kTPR; off < 32; off <<= 1) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            s1[i] += __shfl_xor_sync(0xffffffffu, s1[i], off);
            s2[i] += __shfl_xor_sync(0xffffffffu, s2[i], off);
        }
    }
    constexpr int kCols = (kTPR < 32 ? kTPR : 32) * 8;       // channels covered by one warp
    __shared__ __align__(16) float red[2][kThreads / 32][kCols];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane < kTPR) {
#pragma unroll
        for (int i = 0; i < 8; ++i) { red[0][warp][(lane % kTPR) * 8 + i] = s1[i]; red[1][warp][(lane % kTPR) * 8 + i] = s2[i]; }
    }
    __syncthreads();
    // every block of the grid lands on the same 2·C addresses at about the same time: 16-byte reductions (4 channels per
    // operation) quarter the number of operations the L2 has to serialise per address range
    if ((reinterpret_cast<uintptr_t>(sums) & 15) == 0 && (C & 3) == 0) {
        if (threadIdx.x < kCols / 4) {
            float t1[4] = {0.f, 0.f, 0.f, 0.f}, t2[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int w = 0; w < kThreads / 32; ++w) {
                const float4 a = *reinterpret_cast<const float4*>(&red[0][w][threadIdx.x * 4]);
                const float4 b2 = *reinterpret_cast<const float4*>(&red[1][w][threadIdx.x * 4]);
                t1[0] += a.x; t1[1] += a.y; t1[2] += a.z; t1[3] += a.w;
                t2[0] += b2.x; t2[1] += b2.y; t2[2] += b2.z; t2[3] += b2.w;
            }
            const int c = blockIdx.y * kTPR * 8 + threadIdx.x * 4;
            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(sums + c), "f"(t1[0]), "f"(t1[1]), "f"(t1[2]), "f"(t1[3])
                         : "memory");
            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(sums + C + c), "f"(t2[0]), "f"(t2[1]), "f"(t2[2]), "f"(t2[3])
                         : "memory");
        }
        return;
    }
    if (threadIdx.x < kCols) {
        float t1 = 0.f, t2 = 0.f;
#pragma unroll
        for (int w = 0; w < kThreads / 32; ++w) { t1 += red[0][w][threadIdx.x]; t2 += red[1][w][threadIdx.x]; }
        const int c = blockIdx.y * kTPR * 8 + threadIdx.x;
        atomicAdd(&sums[c], t1);
        atomicAdd(&sums[C + c], t2);
    }
}

// Where the affine kernels get their per-channel coefficients from.  mode 0: a ready [3,C] table.  mode 1 / 2: derived
// on the fly from the batch sums — the work of bn_finalize_fwd / bn_finalize_bwd, redone by every CTA for the few
// channels it needs (a handful of flops) so that the separate finalize launch disappears; the CTA with
// blockIdx.x == 0 also writes the side outputs (forward: ctx + running statistics; backward: dgamma / dbeta).
struct BnDerive {
    int mode, form;
    const float* sums;                     // mode 1: Σy, Σy² ; mode 2: Σg, Σg·y
    float* ctx;                            // mode 1: written [5,C]; mode 2: read
    const float* gamma; const float* beta;
    float* running_mean; float* running_var;
    float* dgamma; float* dbeta;
    float N, eps, momentum;
};

// form 0: reference FusedConvBN  (unbiased var, eps added to the std, no affine)      resnet.py:85-100
// form 1: nn.BatchNorm2d training (biased var, eps inside sqrt, affine, running-stat update)
__device__ __forceinline__ void bn_coef_fwd(const BnDerive& d, int c, int C, bool writer, float& A, float& B, float& Cc) {
    const float g = d.gamma ? d.gamma[c] : 1.f, b = d.beta ? d.beta[c] : 0.f;
    const float s = d.sums[c], ss = d.sums[C + c];
    const float mean = s / d.N;
    const float m2 = fmaxf(ss - s * mean, 0.f);
    float inv, k;
    if (d.form == 0) {
        const float sd = sqrtf(m2 / (d.N - 1.f));
        inv = 1.f / (sd + d.eps);
        k = (sd + d.eps) / ((d.N - 1.f) * fmaxf(sd, 1e-20f));
    } else {
        inv = rsqrtf(m2 / d.N + d.eps);
        k = 1.f / d.N;
        if (writer && d.running_mean) {
            d.running_mean[c] = (1.f - d.momentum) * d.running_mean[c] + d.momentum * mean;
            d.running_var[c] = (1.f - d.momentum) * d.running_var[c] + d.momentum * (m2 / fmaxf(d.N - 1.f, 1.f));
        }
    }
    const float scale = g * inv;
    A = scale; B = 0.f; Cc = bn_shift(b, mean, scale);
    if (writer) { d.ctx[c] = mean; d.ctx[C + c] = inv; d.ctx[2 * C + c] = scale; d.ctx[3 * C + c] = k; d.ctx[4 * C + c] = g; }
}
// sums = (Σg, Σg·y) → dy = A·g + B·y + Cc ;  dgamma = Σg·ŷ, dbeta = Σg
__device__ __forceinline__ void bn_coef_bwd(const BnDerive& d, int c, int C, bool writer, float& A, float& B, float& Cc) {
    const float mean = d.ctx[c], inv = d.ctx[C + c], scale = d.ctx[2 * C + c], k = d.ctx[3 * C + c];
    const float sg = d.sums[c], sgy = d.sums[C + c];
    const float sgyh = inv * (sgy - mean * sg);          // Σ g·ŷ
    if (writer && d.dgamma) { d.dgamma[c] = sgyh; d.dbeta[c] = sg; }
    if (d.form == 2) { A = scale; B = 0.f; Cc = 0.f; return; }
    const float b = sg / d.N, cc = k * sgyh;
    A = scale; B = -scale * cc * inv; Cc = -scale * b + scale * cc * inv * mean;
}
__device__ __forceinline__ void bn_coef(const BnDerive& d, const float* __restrict__ coef, int c, int C, bool writer,
                                        float& A, float& B, float& Cc) {
    if (d.mode == 1) bn_coef_fwd(d, c, C, writer, A, B, Cc);
    else if (d.mode == 2) bn_coef_bwd(d, c, C, writer, A, B, Cc);
    else { A = coef[c]; B = coef[C + c]; Cc = coef[2 * C + c]; }
}

// out = act(A[c]·x1 + B[c]·x2 + Cc[c] + res);  x2 / res optional.  Grid-stride over 16-byte vectors,
// four independent vectors in flight per thread, coefficients staged in shared memory.
template <bool kX2, bool kRes>
__global__ void __launch_bounds__(kThreads)
chan_affine_act_kernel(const __nv_bfloat16* __restrict__ x1, const __nv_bfloat16* __restrict__ x2,
                       const __nv_bfloat16* __restrict__ res, __nv_bfloat16* __restrict__ out,
                       const float* __restrict__ coef /*[3,C]: A,B,Cc*/, int64_t total_vecs, int C, int act, const BnDerive d) {
    extern __shared__ float sc[];
    pdl_wait();
    pdl_launch_dependents();
    for (int i = threadIdx.x; i < C; i += kThreads) bn_coef(d, coef, i, C, blockIdx.x == 0, sc[i], sc[C + i], sc[2 * C + i]);
    __syncthreads();
    const int vecC = C >> 3;
    const int64_t stride = (int64_t)gridDim.x * kThreads * 4;
    for (int64_t base = (int64_t)blockIdx.x * kThreads * 4 + threadIdx.x; base < total_vecs; base += stride) {
        bf16x8 v1[4], v2[4], vr[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int64_t v = base + (int64_t)j * kThreads;
            if (v < total_vecs) {
                v1[j] = ld8(x1 + v * 8);
                if (kX2) v2[j] = ld8(x2 + v * 8);
                if (kRes) vr[j] = ld8(res + v * 8);
            }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int64_t v = base + (int64_t)j * kThreads;
            if (v < total_vecs) {
                const int c = (int)(v % vecC) * 8;
                float f1[8], o[8];
                unpack8(v1[j], f1);
#pragma unroll
                for (int i = 0; i < 8; ++i) o[i] = fmaf(sc[c + i], f1[i], sc[2 * C + c + i]);
                if (kX2) {
                    float f2[8];
                    unpack8(v2[j], f2);
#pragma unroll
                    for (int i = 0; i < 8; ++i) o[i] = fmaf(sc[C + c + i], f2[i], o[i]);
                }
                if (kRes) {
                    float fr[8];
                    unpack8(vr[j], fr);
#pragma unroll
                    for (int i = 0; i < 8; ++i) o[i] += fr[i];
                }
#pragma unroll
                for (int i = 0; i < 8; ++i) o[i] = act_fwd(o[i], act);
                st8(out + v * 8, pack8(o));
            }
        }
    }
}

// Fast path (C/8 divides 256, i.e. every power-of-two channel count up to 2048): with a grid stride that is a
// multiple of C/8 vectors a thread always meets the SAME 8 channels, so its 24 coefficients live in registers
// and the loop body is nothing but 128-bit loads, FMAs and 128-bit stores (no shared memory, no index division).
// kMaskY (backward of a ReLU unit without a residual input; needs kX2, x2 = y): x1 = the incoming gradient is first masked
// by (scale·y + shift > 0), the forward's own expression — the activated output is not read and no masked copy exists.
template <bool kX2, bool kRes, int kAct, bool kMaskY = false>
__global__ void __launch_bounds__(kThreads, kMaskY ? 3 : 0)
chan_affine_act_fast_kernel(const __nv_bfloat16* __restrict__ x1, const __nv_bfloat16* __restrict__ x2,
                            const __nv_bfloat16* __restrict__ res, __nv_bfloat16* __restrict__ out,
                            const float* __restrict__ coef, int64_t total_vecs, int C, const BnDerive d) {
    pdl_wait();
    pdl_launch_dependents();
    const int vecC = C >> 3;
    const int c = (threadIdx.x % vecC) * 8;          // blockIdx.x * 1024 and j * 256 are multiples of vecC
    float A[8], B[8], Cc[8];
    const bool writer = blockIdx.x == 0 && threadIdx.x < vecC;      // these threads cover every channel exactly once
#pragma unroll
    for (int i = 0; i < 8; ++i) bn_coef(d, coef, c + i, C, writer, A[i], B[i], Cc[i]);
    float Msh[8];                                      // forward shift; the forward scale is A (bn_coef_bwd: A = scale)
    if (kMaskY) {
#pragma unroll
        for (int i = 0; i < 8; ++i) Msh[i] = bn_shift(d.beta ? d.beta[c + i] : 0.f, d.ctx[c + i], d.ctx[2 * C + c + i]);
    }
    const int64_t stride = (int64_t)gridDim.x * kThreads * 4;
    for (int64_t base = (int64_t)blockIdx.x * kThreads * 4 + threadIdx.x; base < total_vecs; base += stride) {
        bf16x8 v1[4], v2[4], vr[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int64_t v = base + (int64_t)j * kThreads;
            if (v < total_vecs) {
                v1[j] = ld8(x1 + v * 8);
                if (kX2) v2[j] = ld8(x2 + v * 8);
                if (kRes) vr[j] = ld8(res + v * 8);
            }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int64_t v = base + (int64_t)j * kThreads;
            if (v < total_vecs) {
                float f1[8], f2[8], o[8];
                unpack8(v1[j], f1);
                if (kX2) unpack8(v2[j], f2);
                if (kMaskY) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) f1[i] = fmaf(A[i], f2[i], Msh[i]) > 0.f ? f1[i] : 0.f;
                }
#pragma unroll
                for (int i = 0; i < 8; ++i) o[i] = fmaf(A[i], f1[i], Cc[i]);
                if (kX2) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) o[i] = fmaf(B[i], f2[i], o[i]);
                }
                if (kRes) {
                    float fr[8];
                    unpack8(vr[j], fr);
#pragma unroll
                    for (int i = 0; i < 8; ++i) o[i] += fr[i];
                }
#pragma unroll
                for (int i = 0; i < 8; ++i) o[i] = act_fwd(o[i], kAct);
                st8(out + v * 8, pack8(o));
            }
        }
    }
}

// Same map with the inputs streamed through a per-thread ring in shared memory (cp.async, 16 bytes per copy, kRing copies
// per input in flight per thread).  A register-staged load keeps 4 vectors per input in flight per thread — with the ~800
// resident threads per SM the registers allow that is ~100 KB per SM, at the knee of what HBM3e's latency needs; the ring
// holds 2-3x as much in flight and holds no registers while it waits.  Every thread reads only its own slots: no barriers.

template <bool kX2, bool kRes, int kAct, bool kMaskY = false>
__global__ void __launch_bounds__(kThreads)
chan_affine_act_ring_kernel(const __nv_bfloat16* __restrict__ x1, const __nv_bfloat16* __restrict__ x2,
                            const __nv_bfloat16* __restrict__ res, __nv_bfloat16* __restrict__ out,
                            const float* __restrict__ coef, int64_t total_vecs, int C, const BnDerive d) {
    constexpr int kS = 1 + (kX2 ? 1 : 0) + (kRes ? 1 : 0);             // input streams
    extern __shared__ uint4 ring[];                                    // [kRing][kS][kThreads]
    pdl_wait();
    pdl_launch_dependents();
    const int vecC = C >> 3;
    const int c = (threadIdx.x % vecC) * 8;          // the vector stride below is a multiple of vecC: a thread keeps its channels
    float A[8], B[8], Cc[8];
    const bool writer = blockIdx.x == 0 && threadIdx.x < vecC;      // these threads cover every channel exactly once
#pragma unroll
    for (int i = 0; i < 8; ++i) bn_coef(d, coef, c + i, C, writer, A[i], B[i], Cc[i]);
    float Msh[8];
    if (kMaskY) {
#pragma unroll
        for (int i = 0; i < 8; ++i) Msh[i] = bn_shift(d.beta ? d.beta[c + i] : 0.f, d.ctx[c + i], d.ctx[2 * C + c + i]);
    }
    const int64_t step = (int64_t)gridDim.x * kThreads;
    const int64_t first = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    const int64_t count = first < total_vecs ? (total_vecs - first + step - 1) / step : 0;    // vectors of this thread
    const uint32_t slot0 = (uint32_t)__cvta_generic_to_shared(ring + threadIdx.x);
    auto issue = [&](int64_t n) {
        const int64_t v = first + n * step;
        const uint32_t sa = slot0 + (uint32_t)((n % kRing) * kS) * kThreads * 16;
        cp_async16(sa, x1 + v * 8);
        if (kX2) cp_async16(sa + kThreads * 16, x2 + v * 8);
        if (kRes) cp_async16(sa + (kX2 ? 2 : 1) * kThreads * 16, res + v * 8);
    };
#pragma unroll
    for (int k = 0; k < kRing; ++k) {
        if (k < count) issue(k);
        cp_async_commit();
    }
    for (int64_t n = 0; n < count; ++n) {
        cp_async_wait<kRing - 1>();
        const uint4* sl = ring + (size_t)((n % kRing) * kS) * kThreads + threadIdx.x;
        bf16x8 v1, v2, vr;
        v1.u = sl[0];
        if (kX2) v2.u = sl[kThreads];
        if (kRes) vr.u = sl[(kX2 ? 2 : 1) * kThreads];
        if (n + kRing < count) issue(n + kRing);       // the slot was just read by its only user: refill it
        cp_async_commit();
        float f1[8], f2[8], o[8];
        unpack8(v1, f1);
        if (kX2) unpack8(v2, f2);
        if (kMaskY) {
#pragma unroll
            for (int i = 0; i < 8; ++i) f1[i] = fmaf(A[i], f2[i], Msh[i]) > 0.f ? f1[i] : 0.f;
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) o[i] = fmaf(A[i], f1[i], Cc[i]);
        if (kX2) {
#pragma unroll
            for (int i = 0; i < 8; ++i) o[i] = fmaf(B[i], f2[i], o[i]);
        }
        if (kRes) {
            float fr[8];
            unpack8(vr, fr);
#pragma unroll
            for (int i = 0; i < 8; ++i) o[i] += fr[i];
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) o[i] = act_fwd(o[i], kAct);
        st8(out + (first + n * step) * 8, pack8(o));
    }
}

// form 0: reference FusedConvBN  (unbiased var, eps added to the std, no affine)      resnet.py:85-100
// form 1: nn.BatchNorm2d training (biased var, eps inside sqrt, affine, running-stat update)
// form 2: nn.BatchNorm2d eval     (running stats)
// ctx[5,C] = mean, inv (1/denominator, no gamma), scale (A of the fwd affine), k (bwd coefficient), gamma
__global__ void bn_finalize_fwd_kernel(const float* __restrict__ sums, float* __restrict__ coef, float* __restrict__ ctx,
                                       const float* __restrict__ gamma, const float* __restrict__ beta,
                                       float* __restrict__ running_mean, float* __restrict__ running_var,
                                       int C, float N, float eps, float momentum, int form) {
    pdl_wait();
    pdl_launch_dependents();
    int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    float mean, inv, k, g = gamma ? gamma[c] : 1.f, b = beta ? beta[c] : 0.f;
    if (form == 2) {
        mean = running_mean[c];
        inv = rsqrtf(running_var[c] + eps);
        k = 0.f;
    } else {
        float s = sums[c], ss = sums[C + c];
        mean = s / N;
        float m2 = fmaxf(ss - s * mean, 0.f);
        if (form == 0) {
            float sd = sqrtf(m2 / (N - 1.f));
            inv = 1.f / (sd + eps);
            k = (sd + eps) / ((N - 1.f) * fmaxf(sd, 1e-20f));
        } else {
            float var_b = m2 / N;
            inv = rsqrtf(var_b + eps);
            k = 1.f / N;
            if (running_mean) {
                running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
                running_var[c] = (1.f - momentum) * running_var[c] + momentum * (m2 / fmaxf(N - 1.f, 1.f));
            }
        }
    }
    float scale = g * inv;
    coef[c] = scale; coef[C + c] = 0.f; coef[2 * C + c] = bn_shift(b, mean, scale);
    ctx[c] = mean; ctx[C + c] = inv; ctx[2 * C + c] = scale; ctx[3 * C + c] = k; ctx[4 * C + c] = g;
}

// sums = (Σg, Σg·y) → dy = A·g + B·y + Cc ;  dgamma = Σg·ŷ, dbeta = Σg
__global__ void bn_finalize_bwd_kernel(const float* __restrict__ sums, const float* __restrict__ ctx,
                                       float* __restrict__ coef, float* __restrict__ dgamma, float* __restrict__ dbeta,
                                       int C, float N, int form) {
    pdl_wait();
    pdl_launch_dependents();
    int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    float mean = ctx[c], inv = ctx[C + c], scale = ctx[2 * C + c], k = ctx[3 * C + c];
    float sg = sums[c], sgy = sums[C + c];
    float sgyh = inv * (sgy - mean * sg);          // Σ g·ŷ
    if (dgamma) { dgamma[c] = sgyh; dbeta[c] = sg; }
    if (form == 2) {                                 // eval: statistics are constants
        coef[c] = scale; coef[C + c] = 0.f; coef[2 * C + c] = 0.f;
        return;
    }
    float a = scale, b = sg / N, cc = k * sgyh;
    coef[c] = a;
    coef[C + c] = -a * cc * inv;
    coef[2 * C + c] = -a * b + a * cc * inv * mean;
}

template <int kTPR>
int launch_reduce2(const void* a, const void* b, const void* out_act, void* g_store, float* sums,
                   int64_t P, int C, int act, cudaStream_t st, const float* mctx = nullptr, const float* mbeta = nullptr) {
    constexpr int kLanes = kThreads / kTPR;
    const int ctiles = (C / 8) / kTPR;
    static const bool ring_on = [] { const char* e = getenv("FDT_CHAN_RING"); return !(e && e[0] == '0'); }();
    const bool use_ring = ring_on && P * (int64_t)(C / 8) >= ring_min_vecs();
    const int mode = mctx ? 2 : ((out_act && act != ACT_NONE) ? 1 : 0);
    const bool same = mode == 0 && a == b;
    const int streams = 1 + (same ? 0 : 1) + (mode == 1 ? 1 : 0);
    const size_t rsm = use_ring ? (size_t)kRing2 * streams * kThreads * 16 : 0;
    int per_sm = 4;
    if (use_ring && (int)((220 * 1024) / rsm) < per_sm) per_sm = (int)((220 * 1024) / rsm);
    int64_t slabs = (148 * per_sm + ctiles - 1) / ctiles;
    int64_t rpb = (P + slabs - 1) / slabs;
    const int64_t unit = 4 * kLanes;
    if (rpb < unit) rpb = unit;
    rpb = (rpb + unit - 1) / unit * unit;
    dim3 grid((unsigned)((P + rpb - 1) / rpb), ctiles);
    auto A = (const __nv_bfloat16*)a; auto B = (const __nv_bfloat16*)b; auto O = (const __nv_bfloat16*)out_act;
    auto G = (__nv_bfloat16*)g_store;
    const __nv_bfloat16* no_in = nullptr; __nv_bfloat16* no_out = nullptr;
#define FDT_RED(MODE, SAME, IN, OUT) do { \
        if (use_ring) { auto kern = chan_reduce2_kernel<kTPR, MODE, SAME, true>; \
            cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rsm); \
            fdt::launch(kern, grid, dim3(kThreads), rsm, st, A, B, IN, OUT, sums, P, C, act, (int)rpb, mctx, mbeta); } \
        else fdt::launch(chan_reduce2_kernel<kTPR, MODE, SAME, false>, grid, dim3(kThreads), 0, st, A, B, IN, OUT, sums, P, C, act, (int)rpb, mctx, mbeta); \
    } while (0)
    if (mode == 2) FDT_RED(2, false, no_in, no_out);
    else if (mode == 1) FDT_RED(1, false, O, G);
    else if (same) FDT_RED(0, true, no_in, no_out);
    else FDT_RED(0, false, no_in, no_out);
#undef FDT_RED
    return 0;
}

}  // namespace

// sums[2,C] is cleared by this call (stream-ordered memset) before the reduction.
static int reduce2_impl(bool zero, const void* a, const void* b, const void* out_act, void* g_store, float* sums,
                             int64_t P, int C, int act, cudaStream_t st, const float* mctx = nullptr, const float* mbeta = nullptr) {
    if (C % 8 || C <= 0) return -1;
    if (zero) cudaMemsetAsync(sums, 0, sizeof(float) * 2 * C, st);
    const int vecC = C / 8;
    if (vecC % 32 == 0) launch_reduce2<32>(a, b, out_act, g_store, sums, P, C, act, st, mctx, mbeta);
    else if (vecC % 16 == 0) launch_reduce2<16>(a, b, out_act, g_store, sums, P, C, act, st, mctx, mbeta);
    else if (vecC % 8 == 0) launch_reduce2<8>(a, b, out_act, g_store, sums, P, C, act, st, mctx, mbeta);
    else if (vecC % 4 == 0) launch_reduce2<4>(a, b, out_act, g_store, sums, P, C, act, st, mctx, mbeta);
    else if (vecC % 2 == 0) launch_reduce2<2>(a, b, out_act, g_store, sums, P, C, act, st, mctx, mbeta);
    else launch_reduce2<1>(a, b, out_act, g_store, sums, P, C, act, st, mctx, mbeta);
    FDT_CHECK_LAUNCH();
    return 0;
}

// sums [2C] = Σ a·act'(out), Σ a·act'(out)·b per channel (act' dropped when out_act is NULL); sums is cleared first.
FDT_API int fdt_chan_reduce2(const void* a, const void* b, const void* out_act, void* g_store, float* sums,
                             int64_t P, int C, int act, cudaStream_t st) {
    return reduce2_impl(true, a, b, out_act, g_store, sums, P, C, act, st);
}
// same, accumulating into sums (the caller cleared it — e.g. one memset for a whole model's statistics arena)
FDT_API int fdt_chan_reduce2_acc(const void* a, const void* b, const void* out_act, void* g_store, float* sums,
                                 int64_t P, int C, int act, cudaStream_t st) {
    return reduce2_impl(false, a, b, out_act, g_store, sums, P, C, act, st);
}

static int affine_launch(const void* x1, const void* x2, const void* res, void* out, const float* coef,
                         int64_t P, int C, int act, const BnDerive& d, cudaStream_t st, bool masky = false) {
    if (C % 8 || C <= 0 || 3 * (size_t)C * sizeof(float) > 48 * 1024) return -1;
    const int64_t total = P * (C / 8);
    int64_t blocks = (total + kThreads * 4 - 1) / (kThreads * 4);
    if (blocks > 148 * 8) blocks = 148 * 8;
    const size_t smem = 3 * (size_t)C * sizeof(float);
    auto X1 = (const __nv_bfloat16*)x1; auto X2 = (const __nv_bfloat16*)x2; auto R = (const __nv_bfloat16*)res;
    auto O = (__nv_bfloat16*)out;
    if (masky) {
        if (kThreads % (C / 8) != 0 || !x2 || res || act != ACT_NONE || d.mode != 2) return -2;
        fdt::launch(chan_affine_act_fast_kernel<true, false, ACT_NONE, true>, dim3((int)blocks), dim3(kThreads), 0, st, X1, X2, R, O, coef,
                    total, C, d);
        FDT_CHECK_LAUNCH();
        return 0;
    }
    static const bool use_ring = [] { const char* e = getenv("FDT_CHAN_RING"); return !(e && e[0] == '0'); }();
    if (use_ring && total >= ring_min_vecs() && kThreads % (C / 8) == 0 && (act == ACT_NONE || act == ACT_RELU || act == ACT_CELU)) {
        const int streams = 1 + (x2 ? 1 : 0) + (res ? 1 : 0);
        const size_t rsm = (size_t)kRing * streams * kThreads * 16;
        int64_t rblocks = (total + kThreads - 1) / kThreads;
        const int per_sm = (int)((220 * 1024) / rsm) < 8 ? (int)((220 * 1024) / rsm) : 8;
        if (rblocks > 148 * per_sm) rblocks = 148 * per_sm;
#define FDT_RING(X2F, RF, ACTV) do { auto kern = chan_affine_act_ring_kernel<X2F, RF, ACTV>; \
            cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rsm); \
            fdt::launch(kern, dim3((int)rblocks), dim3(kThreads), rsm, st, X1, X2, R, O, coef, total, C, d); } while (0)
#define FDT_RING_ACT(X2F, RF) do { if (act == ACT_RELU) FDT_RING(X2F, RF, ACT_RELU); else if (act == ACT_CELU) FDT_RING(X2F, RF, ACT_CELU); else FDT_RING(X2F, RF, ACT_NONE); } while (0)
        if (x2 && res) FDT_RING_ACT(true, true);
        else if (x2) FDT_RING_ACT(true, false);
        else if (res) FDT_RING_ACT(false, true);
        else FDT_RING_ACT(false, false);
#undef FDT_RING_ACT
#undef FDT_RING
        FDT_CHECK_LAUNCH();
        return 0;
    }
    if (kThreads % (C / 8) == 0 && (act == ACT_NONE || act == ACT_RELU || act == ACT_CELU)) {
#define FDT_FAST(X2F, RF, ACTV) fdt::launch(chan_affine_act_fast_kernel<X2F, RF, ACTV>, dim3((int)blocks), dim3(kThreads), 0, st, X1, X2, R, O, coef, total, C, d)
#define FDT_FAST_ACT(X2F, RF) do { if (act == ACT_RELU) FDT_FAST(X2F, RF, ACT_RELU); else if (act == ACT_CELU) FDT_FAST(X2F, RF, ACT_CELU); else FDT_FAST(X2F, RF, ACT_NONE); } while (0)
        if (x2 && res) FDT_FAST_ACT(true, true);
        else if (x2) FDT_FAST_ACT(true, false);
        else if (res) FDT_FAST_ACT(false, true);
        else FDT_FAST_ACT(false, false);
#undef FDT_FAST_ACT
#undef FDT_FAST
        FDT_CHECK_LAUNCH();
        return 0;
    }
    if (x2 && res) fdt::launch(chan_affine_act_kernel<true, true>, dim3((int)blocks), dim3(kThreads), smem, st, X1, X2, R, O, coef, total, C, act, d);
    else if (x2) fdt::launch(chan_affine_act_kernel<true, false>, dim3((int)blocks), dim3(kThreads), smem, st, X1, X2, R, O, coef, total, C, act, d);
    else if (res) fdt::launch(chan_affine_act_kernel<false, true>, dim3((int)blocks), dim3(kThreads), smem, st, X1, X2, R, O, coef, total, C, act, d);
    else fdt::launch(chan_affine_act_kernel<false, false>, dim3((int)blocks), dim3(kThreads), smem, st, X1, X2, R, O, coef, total, C, act, d);
    FDT_CHECK_LAUNCH();
    return 0;
}

FDT_API int fdt_chan_affine_act(const void* x1, const void* x2, const void* res, void* out, const float* coef,
                                int64_t P, int C, int act, cudaStream_t st) {
    BnDerive d{};
    return affine_launch(x1, x2, res, out, coef, P, C, act, d, st);
}

// Normalise [+ residual] + activation with the coefficients derived in the same kernel from scratch[0 .. 2C) =
// Σy, Σy² (forms 0 and 1): one launch.  Writes ctx [5,C] (and the running statistics) for the backward pass.
FDT_API int fdt_bn_act_fwd_fused(const void* y, const void* res, void* out, float* ctx, const float* sums,
                                 const float* gamma, const float* beta, float* running_mean, float* running_var,
                                 int64_t P, int C, float eps, float momentum, int form, int act, cudaStream_t st) {
    if (form != 0 && form != 1) return -1;
    BnDerive d{};
    d.mode = 1; d.form = form; d.sums = sums; d.ctx = ctx; d.gamma = gamma; d.beta = beta;
    d.running_mean = running_mean; d.running_var = running_var; d.N = (float)P; d.eps = eps; d.momentum = momentum;
    return affine_launch(y, nullptr, res, out, nullptr, P, C, act, d, st);
}

// Backward counterpart: dy = A·g + B·y + Cc with the coefficients derived from sums = (Σg, Σg·y) and the forward ctx;
// dgamma / dbeta (may be NULL) are written by the same launch.
FDT_API int fdt_bn_bwd_apply(const void* g, const void* y, void* dy, const float* sums, const float* ctx, float* dgamma,
                             float* dbeta, int64_t P, int C, int form, cudaStream_t st) {
    BnDerive d{};
    d.mode = 2; d.form = form; d.sums = sums; d.ctx = const_cast<float*>(ctx); d.dgamma = dgamma; d.dbeta = dbeta; d.N = (float)P;
    return affine_launch(g, y, nullptr, dy, nullptr, P, C, ACT_NONE, d, st);
}

// ---- backward of a ReLU unit WITHOUT a residual input, in two passes that never touch the activated output: the mask is
// re-derived from y as (scale·y + shift > 0) with the forward's coefficients (ctx; beta = the BatchNorm's beta or NULL).
// Per unit this saves the read of `out` and the write of the masked gradient of the generic path (2 of its 7 tensor passes).
// C / 8 must divide 256 (every power-of-two channel count up to 2048); forms 0 and 1.
//   pass 1: sums[2C] (+)= Σ g·m, Σ g·m·y      pass 2: dy = A·(g·m) + B·y + Cc, dgamma / dbeta
FDT_API int fdt_chan_reduce2_masky(const void* g, const void* y, const float* ctx, const float* beta, float* sums, int accumulate,
                                   int64_t P, int C, cudaStream_t st) {
    if (!ctx || g == y) return -1;
    return reduce2_impl(!accumulate, g, y, nullptr, nullptr, sums, P, C, ACT_RELU, st, ctx, beta);
}
FDT_API int fdt_bn_bwd_apply_masky(const void* g, const void* y, void* dy, const float* sums, const float* ctx, const float* beta,
                                   float* dgamma, float* dbeta, int64_t P, int C, int form, cudaStream_t st) {
    if (form != 0 && form != 1) return -1;
    BnDerive d{};
    d.mode = 2; d.form = form; d.sums = sums; d.ctx = const_cast<float*>(ctx); d.beta = beta; d.dgamma = dgamma; d.dbeta = dbeta;
    d.N = (float)P;
    return affine_launch(g, y, nullptr, dy, nullptr, P, C, ACT_NONE, d, st, true);
}

FDT_API int fdt_bn_finalize_fwd(const float* sums, float* coef, float* ctx, const float* gamma, const float* beta,
                                float* running_mean, float* running_var, int C, float N, float eps, float momentum,
                                int form, cudaStream_t st) {
    fdt::launch(bn_finalize_fwd_kernel, dim3((C + 127) / 128), dim3(128), 0, st, sums, coef, ctx, gamma, beta, running_mean,
                running_var, C, N, eps, momentum, form);
    FDT_CHECK_LAUNCH();
    return 0;
}

FDT_API int fdt_bn_finalize_bwd(const float* sums, const float* ctx, float* coef, float* dgamma, float* dbeta,
                                int C, float N, int form, cudaStream_t st) {
    fdt::launch(bn_finalize_bwd_kernel, dim3((C + 127) / 128), dim3(128), 0, st, sums, ctx, coef, dgamma, dbeta, C, N, form);
    FDT_CHECK_LAUNCH();
    return 0;
}

// Fused forward of one normalisation unit: statistics → per-channel coefficients → normalise
// [+ residual] + activation, as three stream-ordered launches behind ONE call.  `scratch` is
// caller-provided working memory of 5*C floats (sums[2C] | coef[3C]); its contents on entry do not
// matter and are garbage afterwards.  `ctx` (5*C floats) is an OUTPUT the backward pass reads.
// form 0: reference FusedConvBN (gamma/beta/running_* unused) · 1: BatchNorm2d train · 2: BatchNorm2d eval.
FDT_API int fdt_bn_act_fwd(const void* y, const void* res, void* out, float* ctx, float* scratch,
                           const float* gamma, const float* beta, float* running_mean, float* running_var,
                           int64_t P, int C, float eps, float momentum, int form, int act, cudaStream_t st) {
    if (C % 8) return -1;
    float* sums = scratch;
    float* coef = scratch + 2 * (size_t)C;
    int rc;
    if (form != 2) {
        rc = fdt_chan_reduce2(y, y, nullptr, nullptr, sums, P, C, ACT_NONE, st);
        if (rc) return rc;
    }
    rc = fdt_bn_finalize_fwd(sums, coef, ctx, gamma, beta, running_mean, running_var, C, (float)P, eps, momentum, form, st);
    if (rc) return rc;
    return fdt_chan_affine_act(y, nullptr, res, out, coef, P, C, act, st);
}

// Same unit when the producer of `y` already accumulated the per-channel Σy, Σy² into scratch[0 .. 2C) (the conv
// GEMM's epilogue does, csrc/gemm_sm100.cu): coefficients → normalise [+ residual] + activation, two launches and
// one pass over y less than fdt_bn_act_fwd.  Not for form 2 (eval: no batch statistics).
FDT_API int fdt_bn_act_fwd_presummed(const void* y, const void* res, void* out, float* ctx, float* scratch,
                                     const float* gamma, const float* beta, float* running_mean, float* running_var,
                                     int64_t P, int C, float eps, float momentum, int form, int act, cudaStream_t st) {
    if (C % 8 || form == 2) return -1;
    return fdt_bn_act_fwd_fused(y, res, out, ctx, scratch, gamma, beta, running_mean, running_var, P, C, eps, momentum, form, act, st);
}

// ---------------------------------------------------------------------------------------------
// fdt_bn_act_fwd_noalloc — the same fused normalisation forward as fdt_bn_act_fwd, for callers
// that have no working memory to hand over and are not allowed to go and get some.
//
// Who that is: a host program that has already given the whole device to a memory pool of its own
// (so a cudaMalloc from inside a library would fail or fragment that pool), a plain C caller with no
// tensor allocator at hand, or a latency-sensitive loop that cannot afford cudaMalloc / cudaFree,
// which may synchronise the whole device.  For them the 5*C floats of working memory that
// fdt_bn_act_fwd takes from its caller live in the library instead: one file-scope array in device
// memory, sized for at most kNoAllocMaxC channels (5 * 4096 floats = 80 KiB).  Every GPU that loads
// this library carries that array until the process exits.
//
// What to know before using it:
//  * C must be a multiple of 8 and at most 4096; a larger C is refused with a message on stderr and
//    the return value -2.  Nothing is launched in that case.
//  * The first call on each device looks up the array's device address (cudaGetSymbolAddress, with
//    that device current) and remembers it.  Make that first call outside any stream capture.
//  * All calls on a device share the one array and it keeps whatever the previous call left in it.
//    The part that has to start from zero (the running sums) is cleared on the stream by every call,
//    so back-to-back calls on ONE stream are fine.  Two calls that can overlap on the GPU — different
//    streams, or different host threads without ordering between them — would race on the array;
//    the caller has to serialise them.
//  * No cudaMalloc, cudaFree or other allocating routine is called, on the first call or later.
namespace {
constexpr int kNoAllocMaxC = 4096;
constexpr int kNoAllocMaxDevices = 64;
static __device__ __align__(256) float g_bn_scratch[5 * kNoAllocMaxC];
float* g_bn_scratch_addr[kNoAllocMaxDevices] = {nullptr};   // per device, filled on first use
}  // namespace

FDT_API int fdt_bn_act_fwd_noalloc(const void* y, const void* res, void* out, float* ctx,
                                   const float* gamma, const float* beta, float* running_mean, float* running_var,
                                   int64_t P, int C, float eps, float momentum, int form, int act, cudaStream_t st) {
    if (C <= 0 || C % 8) return -1;
    if (C > kNoAllocMaxC) {
        fprintf(stderr, "fdt_bn_act_fwd_noalloc: C = %d channels exceeds the built-in working memory "
                        "(limit %d channels); use fdt_bn_act_fwd with caller-provided scratch\n", C, kNoAllocMaxC);
        return -2;
    }
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= kNoAllocMaxDevices) return -3;
    float* scratch = g_bn_scratch_addr[dev];
    if (scratch == nullptr) {
        // In host code `g_bn_scratch` is only a handle for the symbol, not a device address: ask the
        // runtime where the array lives on the current device and keep the answer.
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_bn_scratch);
        if (e != cudaSuccess) return (int)e;
        scratch = static_cast<float*>(p);
        g_bn_scratch_addr[dev] = scratch;
    }
    return fdt_bn_act_fwd(y, res, out, ctx, scratch, gamma, beta, running_mean, running_var,
                          P, C, eps, momentum, form, act, st);
}
