// Fused multi-head attention forward / backward on tcgen05 for sequences of up to 512 tokens, d_head = 64.
//
//   P = dropout(softmax(Q Kᵀ / sqrt(d) + key mask)),   O = P V
//
// Forward: one CTA per (128-query tile, batch, head) walks the key/value tiles with an online softmax (running max and
// sum per query row in registers; the partial P·V of each tile is read back from tensor memory and folded into a
// register accumulator with the usual rescale), the next K / V tile streaming in by TMA while the current one is used.
// Backward: one CTA per (128-key tile, batch, head) keeps K, V resident, accumulates dK, dV in tensor memory over the query
// tiles, and emits its dQ contribution per query tile (stored directly when there is a single key tile, fp32 red.add
// into an accumulation buffer otherwise).  Q, K, V (and dO) are read straight out of the projection outputs —
// [B·L, heads·64] slices of a packed [B·L, 3·heads·64] QKV matrix or separate matrices, any row stride — with 2-D TMA boxes
// at column offset head·64, and O / dQ / dK / dV are written back the same way, so there is no head transpose anywhere.  S = QKᵀ and all other products live in
// tensor memory; the softmax warps (one thread per query row) read S with tcgen05.ld, write the bf16
// probabilities into shared memory in the 128B-swizzled layout the next MMA expects, and the second MMA
// consumes them.  Every product of the backward pass is expressed with K-major or MN-major operand
// descriptors on the SAME shared-memory tiles (P and dS are read both ways), so nothing is transposed:
//
//   fwd : S = Q Kᵀ                 (A K-major, B K-major)         O  = P V      (A K-major, B MN-major)
//   bwd : S = Q Kᵀ, dP = dO Vᵀ     (K-major / K-major)            dV = Pᵀ dO    (A MN-major, B MN-major)
//         dQ = dS K                (A K-major, B MN-major)        dK = dSᵀ Q    (A MN-major, B MN-major)
//
// Parity: ScaledDotProduct / MultiheadAttention of the reference (transformer.py:180-227): two batched
// cuBLAS GEMMs + softmax + dropout kernels that materialise the B·h·L·L probabilities.
// The reference's mask fill value of -1e-9 (Q11) is available as `mask_compat`.
#include "common.cuh"
#include "sm100.cuh"
#include <cstdio>
#include <mutex>

using namespace fdt;
using namespace sm100;


namespace {

constexpr int kT = 128;        // tile: queries = keys = 128
constexpr int kD = 64;         // head dim
constexpr int kMaxL = 512;     // longest sequence (the reference's maxlen, transformer.py:35)
constexpr int kThreads = 192;
constexpr float kLog2e = 1.4426950408889634f;

struct AttnParams {
    int B, H, L;
    int ldq, ldo, ldg;             // row strides (elements): q/k/v inputs; o and dO; dq/dk/dv outputs
    const int64_t* key_mask;       // [B, L] (nonzero = keep) or null
    int mask_compat;
    float scale, p_drop;
    const int64_t* seed_off;       // device (seed, offset) pair for Philox, or null when p_drop == 0
    __nv_bfloat16* O;              // fwd out
    float* lse;                    // [B, H, L]
    const __nv_bfloat16* dO;       // bwd in (also via TMA)
    const __nv_bfloat16* Oin;      // bwd in
    __nv_bfloat16 *dQ, *dK, *dV;   // bwd out
    float* dQacc;                  // bwd: fp32 [B·L, H·64] accumulation buffer (zero on entry) when L > 128, else null
};

// byte offset of element (row, col) inside a [128 rows][64 bf16] 128B-swizzled atom
__device__ __forceinline__ uint32_t swz(int row, int col) {
    return (uint32_t)(row * 128 + ((((col >> 3) ^ (row & 7)) << 4) | ((col & 7) << 1)));
}

// dropout keep-scales of 8 consecutive keys (global key index 8*col8 ..) for global query row qrow of (batch, head) bh: one
// Philox call yields 128 random bits = eight 16-bit draws (threshold p * 65536: p is matched to 2^-16, the same in forward and
// backward) — half the generator work of one 32-bit draw per element, and the generator is a third of the softmax warps' work
__device__ __forceinline__ void drop_scale8(const Philox& rng, uint64_t off, int bh, int qrow, int col8, float p, float inv_keep,
                                            float* m8) {
    const uint4 r = rng(((uint64_t)bh * kMaxL + qrow) * (kMaxL / 8) + col8, off);
    const uint32_t thr = (uint32_t)(p * 65536.f);
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        m8[2 * i] = (w[i] & 0xFFFFu) >= thr ? inv_keep : 0.f;
        m8[2 * i + 1] = (w[i] >> 16) >= thr ? inv_keep : 0.f;
    }
}

// additive bias / replacement for every key of batch b (evaluated once per CTA into shared memory; keys >= L get -inf)
__device__ __forceinline__ void load_key_bias(const AttnParams& p, int b, float* kbias, uint8_t* kfill) {
    for (int j = threadIdx.x; j < kMaxL; j += kThreads) {
        float bias = 0.f; uint8_t fill = 0;
        if (j >= p.L) bias = -INFINITY;
        else if (p.key_mask && p.key_mask[(size_t)b * p.L + j] == 0) {
            if (p.mask_compat) fill = 1; else bias = -INFINITY;
        }
        kbias[j] = bias; kfill[j] = fill;
    }
}

constexpr int kFwdSmem = 4 * 16384 + 16384;                    // Q, K, V, P (2 atoms)
constexpr int kAuxBytes = 16 * 8 + 16 + kMaxL * 4 + kMaxL;     // barriers, tmem ptr, key bias, key fill

// ------------------------------------------------------------------ forward
__global__ void __launch_bounds__(kThreads, 2)
attn_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                const __grid_constant__ CUtensorMap tmV, const AttnParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sQ = smem;                 // 16 KB each
    uint8_t* sK = smem + 16384;
    uint8_t* sV = smem + 32768;
    uint8_t* sP = smem + 49152;         // 2 atoms (keys 0-63 | 64-127), 32 KB
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kFwdSmem);   // [0] Q, [1] K, [2] V, [3] S ready, [4] P ready, [5] PV done
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 16);
    float* kbias = reinterpret_cast<float*>(tmem_ptr + 4);
    uint8_t* kfill = reinterpret_cast<uint8_t*>(kbias + kMaxL);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int qt = blockIdx.x, bh = blockIdx.y, b = bh / p.H, h = bh % p.H;
    const int row0 = b * p.L, col0 = h * kD;
    const int nkv = (p.L + kT - 1) / kT;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmQ); prefetch_tmap(&tmK); prefetch_tmap(&tmV);
        mbar_init(&bars[0], 1); mbar_init(&bars[1], 1); mbar_init(&bars[2], 1);
        mbar_init(&bars[3], 1); mbar_init(&bars[4], 128); mbar_init(&bars[5], 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<256>(tmem_ptr);
    load_key_bias(p, b, kbias, kfill);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tm = *tmem_ptr;
    const uint32_t tmS = tm, tmO = tm + 128;

    if (warp == 0) {
        if (lane == 0) {
            mbar_arrive_expect_tx(&bars[0], 16384);
            tma_load_2d(sQ, &tmQ, &bars[0], col0, row0 + qt * kT);
            for (int j = 0; j < nkv; ++j) {
                if (j > 0) mbar_wait(&bars[3], (j - 1) & 1);            // S_{j-1} done: K buffer is free
                mbar_arrive_expect_tx(&bars[1], 16384);
                tma_load_2d(sK, &tmK, &bars[1], col0, row0 + j * kT);
                if (j > 0) mbar_wait(&bars[5], (j - 1) & 1);            // P·V_{j-1} done: V buffer is free
                mbar_arrive_expect_tx(&bars[2], 16384);
                tma_load_2d(sV, &tmV, &bars[2], col0, row0 + j * kT);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            const uint64_t dq = make_smem_desc(smem_u32(sQ), 16, 1024), dk = make_smem_desc(smem_u32(sK), 16, 1024);
            const uint64_t dv = make_smem_desc(smem_u32(sV), 8192, 1024);       // B = V, N(d)-major, K = keys
            mbar_wait(&bars[0], 0);
            for (int j = 0; j < nkv; ++j) {
                mbar_wait(&bars[1], j & 1);
                if (j > 0) mbar_wait(&bars[4], (j - 1) & 1);            // softmax has consumed S_{j-1}
                tc_fence_after();
#pragma unroll
                for (int k = 0; k < 4; ++k) umma_bf16(tmS, dq + 2 * k, dk + 2 * k, make_idesc_bf16(128, 128, 0, 0), k != 0);
                umma_commit(&bars[3]);
                mbar_wait(&bars[4], j & 1);                             // probabilities of tile j are in shared memory
                mbar_wait(&bars[2], j & 1);
                tc_fence_after();
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const uint64_t dp = make_smem_desc(smem_u32(sP) + (k >> 2) * 16384, 16, 1024) + 2 * (k & 3);
                    umma_bf16(tmO, dp, dv + 128 * k, make_idesc_bf16(128, 64, 0, 1), k != 0);
                }
                umma_commit(&bars[5]);
            }
        }
    } else {
        const int q = warp & 3, row = q * 32 + lane;
        const int qrow = qt * kT + row;                          // query index inside the sequence
        const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
        const float sl2 = p.scale * kLog2e;
        Philox rng(p.seed_off ? (uint64_t)p.seed_off[0] : 0ull);
        const uint64_t off = p.seed_off ? (uint64_t)p.seed_off[1] : 0ull;
        const float inv_keep = 1.f / (1.f - p.p_drop);
        float m_run = -INFINITY, l_run = 0.f;
        float o_acc[kD];
#pragma unroll
        for (int i = 0; i < kD; ++i) o_acc[i] = 0.f;
#pragma unroll 1
        for (int j = 0; j < nkv; ++j) {
            const float* kb = kbias + j * kT;
            const uint8_t* kf = kfill + j * kT;
            mbar_wait(&bars[3], j & 1);
            tc_fence_after();
            // pass 1: tile row maximum (scaled log2 units)
            float mx = m_run;
#pragma unroll 1
            for (int c = 0; c < 4; ++c) {
                uint32_t r[32];
                tmem_ld_32x32(tmS + lane_addr + c * 32, r);
                tmem_ld_wait();
#pragma unroll
                for (int jj = 0; jj < 32; ++jj) {
                    const int key = c * 32 + jj;
                    float s = kf[key] ? -1e-9f : __uint_as_float(r[jj]);
                    mx = fmaxf(mx, s * sl2 + kb[key]);
                }
            }
            const float m_new = (mx == -INFINITY) ? 0.f : mx;
            const float alpha = (m_run == -INFINITY) ? 0.f : exp2f(m_run - m_new);
            // pass 2: exponentials, row sum, dropout, bf16 P into swizzled shared memory
            // (the P buffer is free: P·V_{j-1} completed before O_{j-1} was read at the end of the previous iteration)
            float sum = 0.f;
#pragma unroll 1
            for (int c = 0; c < 4; ++c) {
                uint32_t r[32];
                tmem_ld_32x32(tmS + lane_addr + c * 32, r);
                tmem_ld_wait();
                float pv[32];
#pragma unroll
                for (int jj = 0; jj < 32; ++jj) {
                    const int key = c * 32 + jj;
                    float s = kf[key] ? -1e-9f : __uint_as_float(r[jj]);
                    const float e = exp2f(s * sl2 + kb[key] - m_new);
                    sum += e;
                    pv[jj] = e;
                }
                if (p.p_drop > 0.f) {
#pragma unroll
                    for (int jj = 0; jj < 32; jj += 8) {
                        float m8[8];
                        drop_scale8(rng, off, bh, qrow, j * 16 + c * 4 + (jj >> 3), p.p_drop, inv_keep, m8);
#pragma unroll
                        for (int u = 0; u < 8; ++u) pv[jj + u] *= m8[u];
                    }
                }
                uint8_t* atom = sP + (c >> 1) * 16384;
#pragma unroll
                for (int jj = 0; jj < 32; jj += 8) {
                    const int col = (c & 1) * 32 + jj;
                    *reinterpret_cast<uint4*>(atom + swz(row, col)) = pack8(pv + jj).u;
                }
            }
            tc_fence_before();
            fence_proxy_async();                                // generic-proxy smem writes → visible to the MMA
            mbar_arrive(&bars[4]);
            l_run = l_run * alpha + sum;
            m_run = m_new;
            mbar_wait(&bars[5], j & 1);
            tc_fence_after();
#pragma unroll
            for (int c = 0; c < 2; ++c) {               // unrolled: o_acc stays in registers
                uint32_t r[32];
                tmem_ld_32x32(tmO + lane_addr + c * 32, r);
                tmem_ld_wait();
#pragma unroll
                for (int jj = 0; jj < 32; ++jj) o_acc[c * 32 + jj] = o_acc[c * 32 + jj] * alpha + __uint_as_float(r[jj]);
            }
            tc_fence_before();
        }
        const float inv = l_run > 0.f ? 1.f / l_run : 0.f;
        if (qrow < p.L) {
            p.lse[(size_t)bh * p.L + qrow] = (m_run + log2f(l_run)) / kLog2e;   // natural-log units of the SCALED scores
            __nv_bfloat16* o = p.O + (size_t)(row0 + qrow) * p.ldo + col0;
#pragma unroll
            for (int i = 0; i < kD; ++i) o_acc[i] *= inv;
#pragma unroll
            for (int i = 0; i < kD; i += 8) st8(o + i, pack8(o_acc + i));
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<256>(tm);
}

// ------------------------------------------------------------------ backward
constexpr int kBwdSmem = 4 * 16384 + 2 * 32768;               // Q, K, V, dO, P (2 atoms), dS (2 atoms)

__global__ void __launch_bounds__(kThreads, 1)
attn_bwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmdO, const AttnParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sQ = smem;
    uint8_t* sK = smem + 16384;
    uint8_t* sV = smem + 32768;
    uint8_t* sdO = smem + 49152;
    uint8_t* sP = smem + 65536;          // 32 KB
    uint8_t* sdS = smem + 98304;         // 32 KB
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kBwdSmem);   // [0] K,V  [1] Q,dO  [2] S & dP ready  [3] P & dS ready  [4] grads done
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 16);
    float* kbias = reinterpret_cast<float*>(tmem_ptr + 4);
    uint8_t* kfill = reinterpret_cast<uint8_t*>(kbias + kMaxL);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int kt = blockIdx.x, bh = blockIdx.y, b = bh / p.H, h = bh % p.H;
    const int row0 = b * p.L, col0 = h * kD;
    const int nq = (p.L + kT - 1) / kT;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmQ); prefetch_tmap(&tmK); prefetch_tmap(&tmV); prefetch_tmap(&tmdO);
        mbar_init(&bars[0], 1); mbar_init(&bars[1], 1); mbar_init(&bars[2], 1); mbar_init(&bars[3], 128); mbar_init(&bars[4], 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<512>(tmem_ptr);
    load_key_bias(p, b, kbias, kfill);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tm = *tmem_ptr;
    const uint32_t tmS = tm, tmdP = tm + 128, tmdV = tm + 256, tmdK = tm + 320, tmdQ = tm + 384;

    if (warp == 0) {
        if (lane == 0) {
            mbar_arrive_expect_tx(&bars[0], 2 * 16384);
            tma_load_2d(sK, &tmK, &bars[0], col0, row0 + kt * kT);
            tma_load_2d(sV, &tmV, &bars[0], col0, row0 + kt * kT);
            for (int i = 0; i < nq; ++i) {
                if (i > 0) mbar_wait(&bars[4], (i - 1) & 1);            // the gradient MMAs of tile i-1 have read Q, dO
                mbar_arrive_expect_tx(&bars[1], 2 * 16384);
                tma_load_2d(sQ, &tmQ, &bars[1], col0, row0 + i * kT);
                tma_load_2d(sdO, &tmdO, &bars[1], col0, row0 + i * kT);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t aQ = smem_u32(sQ), aK = smem_u32(sK), aV = smem_u32(sV), adO = smem_u32(sdO);
            const uint32_t aP = smem_u32(sP), adS = smem_u32(sdS);
            const uint32_t iKK = make_idesc_bf16(128, 128, 0, 0);
            const uint32_t iMM = make_idesc_bf16(128, 64, 1, 1), iKM = make_idesc_bf16(128, 64, 0, 1);
            mbar_wait(&bars[0], 0);
            for (int i = 0; i < nq; ++i) {
                mbar_wait(&bars[1], i & 1);
                tc_fence_after();
                {   // S = Q Kᵀ ; dP = dO Vᵀ   (all K-major, K = d = 64)
                    const uint64_t dq = make_smem_desc(aQ, 16, 1024), dk = make_smem_desc(aK, 16, 1024);
                    const uint64_t ddo = make_smem_desc(adO, 16, 1024), dv = make_smem_desc(aV, 16, 1024);
#pragma unroll
                    for (int k = 0; k < 4; ++k) umma_bf16(tmS, dq + 2 * k, dk + 2 * k, iKK, k != 0);
#pragma unroll
                    for (int k = 0; k < 4; ++k) umma_bf16(tmdP, ddo + 2 * k, dv + 2 * k, iKK, k != 0);
                }
                umma_commit(&bars[2]);
                mbar_wait(&bars[3], i & 1);
                tc_fence_after();
                // MN-major views: tile [128 q rows (K)][128 keys (M)] = 2 atoms 16 KB apart; 8-row groups 1 KB apart
                const uint64_t dPt = make_smem_desc(aP, 16384, 1024), dSt = make_smem_desc(adS, 16384, 1024);
                const uint64_t ddoN = make_smem_desc(adO, 8192, 1024);     // B = dO, N(d)-major, K = q
                const uint64_t dqN = make_smem_desc(aQ, 8192, 1024);       // B = Q,  N(d)-major, K = q
                const uint64_t dkN = make_smem_desc(aK, 8192, 1024);       // B = K,  N(d)-major, K = keys
#pragma unroll
                for (int k = 0; k < 8; ++k) umma_bf16(tmdV, dPt + 128 * k, ddoN + 128 * k, iMM, (i | k) != 0);   // dV += Pᵀ dO
#pragma unroll
                for (int k = 0; k < 8; ++k) umma_bf16(tmdK, dSt + 128 * k, dqN + 128 * k, iMM, (i | k) != 0);    // dK += dSᵀ Q
#pragma unroll
                for (int k = 0; k < 8; ++k) {                                                                  // dQ_i = dS K
                    const uint64_t ds = make_smem_desc(adS + (k >> 2) * 16384, 16, 1024) + 2 * (k & 3);
                    umma_bf16(tmdQ, ds, dkN + 128 * k, iKM, k != 0);
                }
                umma_commit(&bars[4]);
            }
        }
    } else {
        const int q = warp & 3, row = q * 32 + lane;
        const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
        const float sl2 = p.scale * kLog2e;
        Philox rng(p.seed_off ? (uint64_t)p.seed_off[0] : 0ull);
        const uint64_t off = p.seed_off ? (uint64_t)p.seed_off[1] : 0ull;
        const float inv_keep = 1.f / (1.f - p.p_drop);
        const float* kb = kbias + kt * kT;
        const uint8_t* kf = kfill + kt * kT;
#pragma unroll 1
        for (int i = 0; i < nq; ++i) {
            const int qrow = i * kT + row;
            const bool row_ok = qrow < p.L;
            // D_r = Σ_d dO[r,d]·O[r,d]
            float Dr = 0.f, lse2 = 0.f;
            if (row_ok) {
                const __nv_bfloat16* o = p.Oin + (size_t)(row0 + qrow) * p.ldo + col0;
                const __nv_bfloat16* g = p.dO + (size_t)(row0 + qrow) * p.ldo + col0;
#pragma unroll
                for (int j = 0; j < kD; j += 8) {
                    float fo[8], fg[8];
                    unpack8(ld8(o + j), fo);
                    unpack8(ld8(g + j), fg);
#pragma unroll
                    for (int u = 0; u < 8; ++u) Dr += fo[u] * fg[u];
                }
                lse2 = p.lse[(size_t)bh * p.L + qrow] * kLog2e;
            }
            mbar_wait(&bars[2], i & 1);
            tc_fence_after();
#pragma unroll 1
            for (int c = 0; c < 4; ++c) {
                uint32_t rs[32], rp[32];
                tmem_ld_32x32(tmS + lane_addr + c * 32, rs);
                tmem_ld_32x32(tmdP + lane_addr + c * 32, rp);
                tmem_ld_wait();
                float pd[32], ds[32];
#pragma unroll
                for (int j = 0; j < 32; j += 8) {
                    float m4[8] = {1.f, 1.f, 1.f, 1.f, 1.f, 1.f, 1.f, 1.f};
                    if (p.p_drop > 0.f) drop_scale8(rng, off, bh, qrow, kt * 16 + c * 4 + (j >> 3), p.p_drop, inv_keep, m4);
#pragma unroll
                    for (int u = 0; u < 8; ++u) {
                        const int key = c * 32 + j + u;
                        float s = kf[key] ? -1e-9f : __uint_as_float(rs[j + u]);
                        s = s * sl2 + kb[key];
                        const float pr = row_ok ? exp2f(s - lse2) : 0.f;          // softmax probability
                        const float dpe = __uint_as_float(rp[j + u]) * m4[u];     // dropout applied to dP
                        pd[j + u] = pr * m4[u];
                        // masked-by-fill keys (compat mode) have a constant score: no gradient flows to Q, K
                        ds[j + u] = kf[key] ? 0.f : pr * (dpe - Dr) * p.scale;
                    }
                }
                uint8_t* aP = sP + (c >> 1) * 16384;
                uint8_t* aS = sdS + (c >> 1) * 16384;
#pragma unroll
                for (int j = 0; j < 32; j += 8) {
                    const int col = (c & 1) * 32 + j;
                    *reinterpret_cast<uint4*>(aP + swz(row, col)) = pack8(pd + j).u;
                    *reinterpret_cast<uint4*>(aS + swz(row, col)) = pack8(ds + j).u;
                }
            }
            tc_fence_before();
            fence_proxy_async();
            mbar_arrive(&bars[3]);
            mbar_wait(&bars[4], i & 1);
            tc_fence_after();
            // this key tile's contribution to dQ of query tile i
#pragma unroll 1
            for (int c = 0; c < 2; ++c) {
                uint32_t r[32];
                tmem_ld_32x32(tmdQ + lane_addr + c * 32, r);       // warp-collective: executed by every lane
                tmem_ld_wait();
                if (row_ok) {
                    if (p.dQacc) {
                        float* dst = p.dQacc + (size_t)(row0 + qrow) * (p.H * kD) + col0 + c * 32;
#pragma unroll
                        for (int j = 0; j < 32; j += 4)
                            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + j), "f"(__uint_as_float(r[j])),
                                         "f"(__uint_as_float(r[j + 1])), "f"(__uint_as_float(r[j + 2])), "f"(__uint_as_float(r[j + 3]))
                                         : "memory");
                    } else {
                        float v[32];
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
                        __nv_bfloat16* dst = p.dQ + (size_t)(row0 + qrow) * p.ldg + col0 + c * 32;
#pragma unroll
                        for (int j = 0; j < 32; j += 8) st8(dst + j, pack8(v + j));
                    }
                }
            }
            tc_fence_before();
        }
        // dK, dV of this key tile (complete after the last query tile)
        {
            const int krow = kt * kT + row;
            const bool k_ok = krow < p.L;
            const size_t base = (size_t)(row0 + krow) * p.ldg + col0;
            const uint32_t srcs[2] = {tmdK, tmdV};
            __nv_bfloat16* dsts[2] = {p.dK, p.dV};
#pragma unroll 1
            for (int t = 0; t < 2; ++t) {
#pragma unroll 1
                for (int c = 0; c < 2; ++c) {
                    uint32_t r[32];
                    tmem_ld_32x32(srcs[t] + lane_addr + c * 32, r);
                    tmem_ld_wait();
                    if (k_ok) {
                        float v[32];
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
#pragma unroll
                        for (int j = 0; j < 32; j += 8) st8(dsts[t] + base + c * 32 + j, pack8(v + j));
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<512>(tm);
}

// dq (bf16, row stride ldg) = the fp32 accumulation buffer [rows, cols]
__global__ void attn_dq_convert_kernel(const float* __restrict__ acc, __nv_bfloat16* __restrict__ dq, int64_t rows, int cols, int ldg) {
    const int64_t n8 = rows * (cols / 8);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / (cols / 8);
        const int c = (int)(i - r * (cols / 8)) * 8;
        const float4 a = *reinterpret_cast<const float4*>(acc + r * cols + c);
        const float4 b2 = *reinterpret_cast<const float4*>(acc + r * cols + c + 4);
        const float v[8] = {a.x, a.y, a.z, a.w, b2.x, b2.y, b2.z, b2.w};
        st8(dq + r * ldg + c, pack8(v));
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn get_encode() {
    // The encode call is a DRIVER-API function: it needs a current context on the calling thread.  PyTorch
    // runs backward on worker threads whose first CUDA call may be this one, so bind the primary context
    // through the runtime first (once per thread).
    static thread_local bool ctx_ready = false;
    if (!ctx_ready) { cudaFree(0); ctx_ready = true; }
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    });
    return fn;
}
// [rows, cols] bf16 with row stride ld elements; box = 64 columns x 128 rows
bool encode_rows(CUtensorMap* m, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld) {
    if (!get_encode()) return false;
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {ld * 2};
    cuuint32_t box[2] = {64, 128};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = get_encode()(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
        fprintf(stderr, "fdt attention: cuTensorMapEncodeTiled failed (%d): ptr=%p rows=%llu cols=%llu ld=%llu\n", (int)r, ptr,
                (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)ld);
    return r == CUDA_SUCCESS;
}


}  // namespace

// q, k, v: bf16 [B·L, H·64] matrices with row stride `ldq` elements (e.g. three column slices of one packed [B·L, 3·H·64] QKV
// projection output, ldq = 3·H·64); o: bf16 [B·L, H·64] with row stride `ldo`; lse: fp32 [B, H, L].
// key_mask: int64 [B, L] (0 = masked) or null.  1 <= L <= 512.
FDT_API int fdt_attn_fwd(const void* q, const void* k, const void* v, void* o, float* lse, const int64_t* key_mask,
                         int B, int H, int L, int ldq, int ldo, float scale, float p_drop, const int64_t* seed_off,
                         int mask_compat, cudaStream_t st) {
    if (L < 1 || L > kMaxL || ldq % 8 || ldo % 8 || B < 1 || H < 1) return -1;
    CUtensorMap tq, tk, tv;
    const uint64_t rows = (uint64_t)B * L, cols = (uint64_t)H * kD;
    if (!encode_rows(&tq, q, rows, cols, ldq) || !encode_rows(&tk, k, rows, cols, ldq) || !encode_rows(&tv, v, rows, cols, ldq))
        return -4;
    AttnParams p{};
    p.B = B; p.H = H; p.L = L; p.ldq = ldq; p.ldo = ldo; p.key_mask = key_mask; p.mask_compat = mask_compat; p.scale = scale;
    p.p_drop = p_drop; p.seed_off = p_drop > 0.f ? seed_off : nullptr; p.O = (__nv_bfloat16*)o; p.lse = lse;
    const int smem = kFwdSmem + kAuxBytes + 1024;
    cudaError_t e = cudaFuncSetAttribute(attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return (int)e;
    attn_fwd_kernel<<<dim3((L + kT - 1) / kT, B * H), kThreads, smem, st>>>(tq, tk, tv, p);
    FDT_CHECK_LAUNCH();
    return 0;
}

// dq, dk, dv: bf16 outputs with row stride `ldg` (e.g. the three column slices of one packed [B·L, 3·H·64] gradient, the
// operand of the fused QKV data-/weight-gradient GEMMs).  dq_acc: fp32 [B·L, H·64] work buffer, required when L > 128
// (the key tiles of a sequence add their dQ contributions into it; cleared and converted here), ignored otherwise.
FDT_API int fdt_attn_bwd(const void* q, const void* k, const void* v, const void* o, const void* dO, const float* lse,
                         void* dq, void* dk, void* dv, float* dq_acc, const int64_t* key_mask, int B, int H, int L, int ldq,
                         int ldo, int ldg, float scale, float p_drop, const int64_t* seed_off, int mask_compat, cudaStream_t st) {
    if (L < 1 || L > kMaxL || ldq % 8 || ldo % 8 || ldg % 8 || B < 1 || H < 1) return -1;
    const int nkv = (L + kT - 1) / kT;
    if (nkv > 1 && !dq_acc) return -1;
    CUtensorMap tq, tk, tv, tdo;
    const uint64_t rows = (uint64_t)B * L, cols = (uint64_t)H * kD;
    if (!encode_rows(&tq, q, rows, cols, ldq) || !encode_rows(&tk, k, rows, cols, ldq) || !encode_rows(&tv, v, rows, cols, ldq) ||
        !encode_rows(&tdo, dO, rows, cols, ldo)) return -4;
    AttnParams p{};
    p.B = B; p.H = H; p.L = L; p.ldq = ldq; p.ldo = ldo; p.ldg = ldg; p.key_mask = key_mask; p.mask_compat = mask_compat;
    p.scale = scale; p.p_drop = p_drop;
    p.seed_off = p_drop > 0.f ? seed_off : nullptr; p.lse = const_cast<float*>(lse);
    p.dO = (const __nv_bfloat16*)dO; p.Oin = (const __nv_bfloat16*)o;
    p.dQ = (__nv_bfloat16*)dq; p.dK = (__nv_bfloat16*)dk; p.dV = (__nv_bfloat16*)dv;
    p.dQacc = nkv > 1 ? dq_acc : nullptr;
    if (p.dQacc) {
        cudaError_t e = cudaMemsetAsync(dq_acc, 0, sizeof(float) * rows * cols, st);
        if (e != cudaSuccess) return (int)e;
    }
    const int smem = kBwdSmem + kAuxBytes + 1024;
    cudaError_t e = cudaFuncSetAttribute(attn_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return (int)e;
    attn_bwd_kernel<<<dim3(nkv, B * H), kThreads, smem, st>>>(tq, tk, tv, tdo, p);
    FDT_CHECK_LAUNCH();
    if (p.dQacc) {
        const int64_t n8 = (int64_t)rows * (cols / 8);
        int blocks = (int)((n8 + 255) / 256);
        if (blocks > 148 * 8) blocks = 148 * 8;
        attn_dq_convert_kernel<<<blocks, 256, 0, st>>>(dq_acc, (__nv_bfloat16*)dq, (int64_t)rows, (int)cols, ldg);
        FDT_CHECK_LAUNCH();
    }
    return 0;
}
