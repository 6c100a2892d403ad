// tcgen05 / TMEM / TMA GEMM and implicit-GEMM convolution for sm_100a.
//
//   D[M, N] = A[M, K] · B[N, K]ᵀ        bf16 operands (K-major), fp32 accumulation in tensor memory
//
// * plain mode : A is a row-major [M, K] matrix (linear layers, 1x1 convolutions on NHWC activations);
// * conv mode  : A is gathered on the fly from an NHWC activation tensor — for every filter tap (r, s)
//   and 64-channel slice the TMA engine loads the 4-D box {64 ch, W, bh rows, bn images} shifted by
//   (s - pad, r - pad); out-of-bounds pixels are zero-filled by the hardware, which implements the
//   padding.  B is the weight tensor physically laid out [Cout][kh][kw][Cin] (K-major, K = kh·kw·Cin).
//
// Parity: F.conv2d / F.linear call sites of the reference (resnet.py:83,108,204,221; transformer.py:
// 168-176,213-226), which run cuDNN / cuBLAS kernels.
//
// Structure (one 128 x BN output tile per CTA, 192 threads):
//   warp 0      TMA producer   — fills a kStages-deep ring of {A 128x64, B BNx64} bf16 tiles (128B swizzle)
//   warp 1      MMA issuer     — allocates BN TMEM columns, one elected lane issues tcgen05.mma
//                                (M=128, N=BN, K=16) x4 per stage, tcgen05.commit frees the stage
//   warps 2..5  epilogue       — tcgen05.ld the fp32 accumulator (warp w owns TMEM lanes 32·(w%4)..),
//                                bias + activation, bf16 pack, 16-byte global stores; optional per-column
//                                Σ and Σ² (batch-norm statistics) reduced by warp shuffles + atomics.
#include "common.cuh"
#include "sm100.cuh"
#include <cstdio>
#include <mutex>
#include <unordered_map>

using namespace fdt;
using namespace sm100;

namespace {

constexpr int BM = 128, BK = 64, kThreads = 320;   // TMA warp + MMA warp + 2 epilogue warpgroups (one per TMEM accumulator)

struct GemmParams {
    int M, N, num_kb;
    __nv_bfloat16* out;      // [M, N] bf16 (may be null)
    float* out_f32;          // [M, N] fp32 (may be null)
    const float* bias;       // [N] or null
    float* sums;             // [2N] Σ, Σ² per column (atomics) or null
    int act;
    const __nv_bfloat16* addend;      // optional bf16 tensor added to the result before it is stored (same indexing as `out`):
                                      // the other branch's gradient at a residual fork, so the sum needs no pass of its own
    // conv geometry (conv mode only).  H, W, bh, bn describe the grid of OUTPUT pixels a tile row maps to
    // (row -> (n, oh, ow)); the A tensor is addressed at (in_mul * oh + tap_dh, tap_dw + [0, W) * in_mul).
    int H, W, bh, bn, cpb;            // cpb = channels-per-tap / 64 (k-blocks per tap)
    int in_mul;                       // 1, or the convolution stride (the tensor map then carries element strides)
    int ntaps;
    int tap_dh[9], tap_dw[9], tap_k[9];   // per tap: A-coordinate offsets and first column of the tap in B's K dimension
    // output mapping: omode 0 = row-major [M, N] (TMA store); 1 = "parity class" of a stride-2 data gradient: tile row
    // (n, i, j) is pixel (n, 2i + oph, 2j + opw) of an [*, OH, OW, N] tensor (direct stores)
    int omode, OH, OW, oph, opw;
};

// Column statistics (Σ, Σ² per output channel — the BatchNorm batch statistics of a conv output): each warp reduces its
// 32 rows with a shuffle butterfly, the four warps of an epilogue warpgroup meet in a scratch array, and one thread per
// column adds the tile's total to the warpgroup's own accumulator in shared memory — plain loads and stores, no atomics
// on shared memory (they are compare-and-swap loops for fp32).  The accumulators are flushed with one global atomic per
// touched column when the CTA is done.
constexpr int kSumCols = 1024;

template <int BN, int kStages>
struct SmemLayout {
    static constexpr int kABytes = BM * BK * 2;
    static constexpr int kBBytes = BN * BK * 2;
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kCOff = kStages * kStageBytes;           // bf16 output staging: BN/64 atoms of 16 KB, per warpgroup
    static constexpr int kCBytes = 2 * (BN / 64) * 16384;
    static constexpr int kBarOff = kCOff + kCBytes;
    static constexpr int kSumOff = kBarOff + (((2 * kStages + 4) * 8 + 16 + 15) & ~15);   // [2 warpgroups][Σ, Σ²][kSumCols]
    static constexpr int kScrOff = kSumOff + 2 * 2 * kSumCols * 4;                       // [2 warpgroups][4 warps][Σ, Σ²][BN]
    static constexpr int kTotal = kScrOff + 2 * 4 * 2 * BN * 4 + 1024;                   // + alignment slack
};
static_assert(SmemLayout<128, 4>::kTotal <= 232448 && SmemLayout<64, 7>::kTotal <= 232448, "shared memory budget");

// Persistent: gridDim.x CTAs (<= #SMs) walk the tile list (m fastest, so neighbouring CTAs share the weight
// tile in L2).  The accumulator is double-buffered in tensor memory (2 x BN columns): while the epilogue warps
// drain tile i (tcgen05.ld → bias/act/stats → swizzled smem → TMA store), the MMA warp already runs tile i+1.
// Tile rasterisation: bands of 16 M-tiles, N-tiles inner → the ~148 tiles in flight cover a 16 x ~9 patch of the
// output, so both operands are re-used out of L2 instead of streaming one of them from HBM for every tile.
__device__ __forceinline__ void tile_coords(int tile, int tiles_m, int tiles_n, int& m_tile, int& n_tile) {
    constexpr int kBand = 16;
    const int per_band = kBand * tiles_n;
    const int b = tile / per_band, r = tile - b * per_band;
    const int rows = min(kBand, tiles_m - b * kBand);
    m_tile = b * kBand + r % rows;
    n_tile = r / rows;
}

// what the epilogue of an instantiation does besides the bf16 TMA store (EPI_ANY: decided at run time from GemmParams)
enum : int { EPI_STORE = 0, EPI_ADDEND = 1, EPI_SUMS = 2, EPI_ANY = 7 };

template <int BN, int kStages, bool kConv, int kEpi>
__global__ void __launch_bounds__(kThreads, 1)
gemm_nt_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const __grid_constant__ CUtensorMap tmC, const GemmParams p) {
    using L = SmemLayout<BN, kStages>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::kBarOff);
    uint64_t* empty_bar = full_bar + kStages;
    uint64_t* tmem_full_bar = empty_bar + kStages;       // [2]
    uint64_t* tmem_empty_bar = tmem_full_bar + 2;        // [2]
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tiles_m = (p.M + BM - 1) / BM, tiles_n = p.N / BN;
    const int num_tiles = tiles_m * tiles_n;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmA);
        prefetch_tmap(&tmB);
        prefetch_tmap(&tmC);
        for (int s = 0; s < kStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(&tmem_full_bar[a], 1); mbar_init(&tmem_empty_bar[a], 128); }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<2 * BN>(tmem_ptr);
    float* s_sums = reinterpret_cast<float*>(smem + L::kSumOff);
    float* s_scr = reinterpret_cast<float*>(smem + L::kScrOff);
    const bool smem_sums = p.sums != nullptr && p.N <= kSumCols;
    if (smem_sums)
        for (int i = threadIdx.x; i < 4 * kSumCols; i += kThreads) s_sums[i] = 0.f;
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;
    pdl_wait();                    // everything above overlapped the previous kernel's tail; its outputs are visible from here on
    pdl_launch_dependents();

    if (warp == 0) {
        if (lane == 0) {
            uint32_t it = 0;                                        // running k-block counter across tiles
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                int m_tile, n_tile;
                tile_coords(tile, tiles_m, tiles_n, m_tile, n_tile);
                int n0 = 0, h0 = 0;
                if (kConv) {
                    if (p.bn == 1) { const int tpi = p.H / p.bh; n0 = m_tile / tpi; h0 = (m_tile % tpi) * p.bh; }
                    else { n0 = m_tile * p.bn; }
                }
                for (int kb = 0; kb < p.num_kb; ++kb, ++it) {
                    const int s = it % kStages;
                    const uint32_t ph = (it / kStages) & 1;
                    mbar_wait(&empty_bar[s], ph ^ 1);
                    uint8_t* sA = smem + s * L::kStageBytes;
                    uint8_t* sB = sA + L::kABytes;
                    mbar_arrive_expect_tx(&full_bar[s], L::kStageBytes);
                    if (kConv) {
                        const int tap = kb / p.cpb, cb = kb - tap * p.cpb;
                        tma_load_4d(sA, &tmA, &full_bar[s], cb * BK, p.tap_dw[tap], p.in_mul * h0 + p.tap_dh[tap], n0);
                        tma_load_2d(sB, &tmB, &full_bar[s], p.tap_k[tap] + cb * BK, n_tile * BN);
                    } else {
                        tma_load_2d(sA, &tmA, &full_bar[s], kb * BK, m_tile * BM);
                        tma_load_2d(sB, &tmB, &full_bar[s], kb * BK, n_tile * BN);
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc_bf16(BM, BN, 0, 0);
            uint32_t it = 0, t = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++t) {
                const uint32_t acc = t & 1, use = t >> 1;           // accumulator buffer and how often it was used before
                mbar_wait(&tmem_empty_bar[acc], (use & 1) ^ 1);     // epilogue has drained it
                tc_fence_after();
                const uint32_t tm = tmem_base + acc * BN;
                for (int kb = 0; kb < p.num_kb; ++kb, ++it) {
                    const int s = it % kStages;
                    const uint32_t ph = (it / kStages) & 1;
                    mbar_wait(&full_bar[s], ph);
                    tc_fence_after();
                    const uint32_t a_addr = smem_u32(smem + s * L::kStageBytes);
                    const uint32_t b_addr = a_addr + L::kABytes;
                    const uint64_t da = make_smem_desc(a_addr, 16, 1024);
                    const uint64_t db = make_smem_desc(b_addr, 16, 1024);
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k)      // +32 bytes (= 2 in the >>4 address field) per K=16 slice
                        umma_bf16(tm, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
                    umma_commit(&empty_bar[s]);
                }
                umma_commit(&tmem_full_bar[acc]);
            }
        }
    } else {
        // ---------------- epilogue: warpgroup wg (warps 2-5 / 6-9) drains accumulator buffer wg, i.e. every other tile.
        // What the epilogue does is fixed per instantiation (kEpi != EPI_ANY) for the shapes a ResNet step is made of, so
        // the per-chunk code is straight-line: with two warps per scheduler there is nothing to hide a branch behind.
        constexpr bool kAny = kEpi == EPI_ANY;
        const bool has_bias = kAny && p.bias != nullptr;
        const bool has_act = kAny && p.act != ACT_NONE;
        const bool has_addend = kAny ? p.addend != nullptr : (kEpi & EPI_ADDEND) != 0;
        const bool has_sums = kAny ? p.sums != nullptr : (kEpi & EPI_SUMS) != 0;
        const bool has_f32 = kAny && p.out_f32 != nullptr;
        const bool direct = kAny && p.out != nullptr && p.omode == 1;
        const bool tma_out = kAny ? (p.out != nullptr && p.omode == 0) : true;
        const int q = warp & 3;
        const int wg = (warp - 2) >> 2;
        const int bar_id = 1 + wg;
        const int tw = threadIdx.x - (64 + wg * 128);              // thread index inside the warpgroup
        const bool is_leader = tw == 0;                            // first thread of the warpgroup issues the TMA stores
        uint8_t* sC = smem + L::kCOff + wg * (L::kCBytes / 2);
        constexpr int kAtoms = BN / 64;                            // 64-column (128-byte) slabs of the staging tile
        const int N = p.N;
        uint32_t t = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++t) {
            if ((int)(t & 1) != wg) continue;
            int m_tile, n_tile;
            tile_coords(tile, tiles_m, tiles_n, m_tile, n_tile);
            const uint32_t acc = t & 1, use = t >> 1;
            const uint32_t tm = tmem_base + acc * BN + ((uint32_t)(q * 32) << 16);
            const int r_in = q * 32 + lane;
            const int row = m_tile * BM + r_in;
            const bool row_ok = row < p.M;
            size_t orow = (size_t)row;                              // row of `out` / `addend` this thread's row maps to
            if (kConv && direct && row_ok) {
                const int hw = p.H * p.W;
                const int n = row / hw, rem = row - n * hw;
                const int i = rem / p.W, j = rem - i * p.W;
                orow = ((size_t)n * p.OH + 2 * i + p.oph) * p.OW + 2 * j + p.opw;
            }
            if ((tma_out || has_sums) && t > 1) {
                // slab 0 of the staging tile may still be read by the store this warpgroup issued for its previous tile
                // (the store of the last slab, issued after it, is allowed to stay in flight)
                if (is_leader && tma_out) {
                    if (kAtoms > 1) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
                    else asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
                }
                asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
            }
            // the residual addend does not depend on the accumulator: its 64 bytes per row and chunk are requested one chunk
            // ahead (the first ones before the wait for the MMA), so the global-load latency is not paid inside every chunk
            bf16x8 adn[2][4];
            const __nv_bfloat16* ad_row = p.addend + orow * N + n_tile * BN;
            if (has_addend && row_ok) {
#pragma unroll
                for (int j = 0; j < 4; ++j) adn[0][j] = ld8(ad_row + 8 * j);
            }
            mbar_wait(&tmem_full_bar[acc], use & 1);
            tc_fence_after();
            if (kEpi == EPI_SUMS) {
                // ---- store + BatchNorm statistics.  The accumulator is read in the MMA-fragment layout (16x256b): a thread
                // holds 4 rows x 16 columns of a 64-column slab, adds its rows in registers and meets the 7 other threads
                // that hold the same columns in 3 shuffle stages (14 exchanges per sum; the row-per-thread layout needs 31).
                const int g = lane >> 2, tq = lane & 3;                 // row group / column pair inside a block of 8
                bool ok[4];
#pragma unroll
                for (int rr = 0; rr < 4; ++rr) ok[rr] = m_tile * BM + q * 32 + g + 8 * rr < p.M;
#pragma unroll
                for (int a = 0; a < kAtoms; ++a) {
                    uint32_t r0[32], r1[32];                            // rows g, g + 8 (r0) and g + 16, g + 24 (r1)
                    tmem_ld_16x256b_x8(tm + a * 64, r0);
                    tmem_ld_16x256b_x8(tm + a * 64 + (16u << 16), r1);
                    tmem_ld_wait();
                    if (a == kAtoms - 1) {               // whole accumulator is in registers: hand it back to the MMA warp
                        tc_fence_before();
                        mbar_arrive(&tmem_empty_bar[acc]);
                    }
                    float s1[16], s2[16];                               // index 2 i + e: column 8 i + 2 tq + e of the slab
                    if (ok[3]) {                                        // all four rows exist (every tile but a ragged last one)
#pragma unroll
                        for (int i = 0; i < 8; ++i)
#pragma unroll
                            for (int e = 0; e < 2; ++e) {
                                const float v0 = __uint_as_float(r0[4 * i + e]), v1 = __uint_as_float(r0[4 * i + 2 + e]);
                                const float v2 = __uint_as_float(r1[4 * i + e]), v3 = __uint_as_float(r1[4 * i + 2 + e]);
                                s1[2 * i + e] = (v0 + v1) + (v2 + v3);
                                s2[2 * i + e] = fmaf(v0, v0, v1 * v1) + fmaf(v2, v2, v3 * v3);
                            }
                    } else {
#pragma unroll
                        for (int i = 0; i < 8; ++i)
#pragma unroll
                            for (int e = 0; e < 2; ++e) {
                                const float v0 = ok[0] ? __uint_as_float(r0[4 * i + e]) : 0.f;
                                const float v1 = ok[1] ? __uint_as_float(r0[4 * i + 2 + e]) : 0.f;
                                const float v2 = ok[2] ? __uint_as_float(r1[4 * i + e]) : 0.f;
                                const float v3 = ok[3] ? __uint_as_float(r1[4 * i + 2 + e]) : 0.f;
                                s1[2 * i + e] = (v0 + v1) + (v2 + v3);
                                s2[2 * i + e] = fmaf(v0, v0, v1 * v1) + fmaf(v2, v2, v3 * v3);
                            }
                    }
                    // transpose-reduce over lane bits 4, 3, 2: afterwards s[0], s[1] are the totals of column block i = g
#pragma unroll
                    for (int off = 16, n = 8; off >= 4; off >>= 1, n >>= 1) {
                        const bool upper = (lane & off) != 0;
#pragma unroll
                        for (int j = 0; j < n; ++j) {
                            const float send1 = upper ? s1[j] : s1[j + n];
                            const float send2 = upper ? s2[j] : s2[j + n];
                            const float recv1 = __shfl_xor_sync(0xffffffffu, send1, off);
                            const float recv2 = __shfl_xor_sync(0xffffffffu, send2, off);
                            s1[j] = (upper ? s1[j + n] : s1[j]) + recv1;
                            s2[j] = (upper ? s2[j + n] : s2[j]) + recv2;
                        }
                    }
                    {   // columns 8 g + 2 tq + {0, 1} = 2 lane + {0, 1} of the slab
                        float* sc = s_scr + ((wg * 4 + q) * 2) * BN + a * 64 + 2 * lane;
                        *reinterpret_cast<float2*>(sc) = make_float2(s1[0], s1[1]);
                        *reinterpret_cast<float2*>(sc + BN) = make_float2(s2[0], s2[1]);
                    }
                    // bf16 pairs → the 128B-swizzled staging slab (row & 7 == g for all four rows of this thread)
                    uint8_t* atom = sC + a * 16384;
#pragma unroll
                    for (int rr = 0; rr < 4; ++rr) {
                        const uint32_t* rs = (rr & 2) ? r1 : r0;
                        const uint32_t rowoff = (uint32_t)(q * 32 + g + 8 * rr) * 128 + tq * 4;
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const __nv_bfloat162 h = __floats2bfloat162_rn(__uint_as_float(rs[4 * i + 2 * (rr & 1)]),
                                                                           __uint_as_float(rs[4 * i + 2 * (rr & 1) + 1]));
                            *reinterpret_cast<__nv_bfloat162*>(atom + rowoff + ((i ^ g) << 4)) = h;
                        }
                    }
                    // the slab is complete: store it now, so its drain overlaps the conversion of the next slab
                    fence_proxy_async();
                    if (a + 1 < kAtoms && t > 1 && is_leader)          // the next slab must be free before anyone writes it
                        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
                    asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");          // the 4 warps of this warpgroup
                    if (is_leader) {
                        tma_store_2d(&tmC, sC + a * 16384, n_tile * BN + a * 64, m_tile * BM);
                        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                    }
                }
            } else {
            // two register buffers: the TMEM load of chunk c+1 is in flight while chunk c is converted and stored
            uint32_t rbuf[2][32];
            tmem_ld_32x32(tm, rbuf[0]);
#pragma unroll
            for (int c = 0; c < BN / 32; ++c) {
                tmem_ld_wait();
                if (c + 1 < BN / 32) tmem_ld_32x32(tm + (c + 1) * 32, rbuf[(c + 1) & 1]);
                else {                                   // whole accumulator is in registers: hand it back to the MMA warp
                    tc_fence_before();
                    mbar_arrive(&tmem_empty_bar[acc]);
                }
                if (has_addend && row_ok && c + 1 < BN / 32) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) adn[(c + 1) & 1][j] = ld8(ad_row + (c + 1) * 32 + 8 * j);
                }
                const uint32_t* r = rbuf[c & 1];
                const int col0 = n_tile * BN + c * 32;
                float v[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
                if (has_bias) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] += __ldg(p.bias + col0 + j);
                }
                if (has_act) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = act_fwd(v[j], p.act);
                }
                if (has_addend && row_ok) {
#pragma unroll
                    for (int j = 0; j < 32; j += 8) {
                        float f[8];
                        unpack8(adn[c & 1][j >> 3], f);
#pragma unroll
                        for (int u = 0; u < 8; ++u) v[j + u] += f[u];
                    }
                }
                if (has_sums) {
                    // per-column Σ and Σ² over the 32 rows of this warp: butterfly transpose-reduce —
                    // after 5 exchange steps lane j holds the total of column j.
                    float s1[32], s2[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) { const float tv = row_ok ? v[j] : 0.f; s1[j] = tv; s2[j] = tv * tv; }
#pragma unroll
                    for (int off = 16; off >= 1; off >>= 1) {
                        const bool upper = (lane & off) != 0;
#pragma unroll
                        for (int j = 0; j < off; ++j) {
                            const float send1 = upper ? s1[j] : s1[j + off];
                            const float send2 = upper ? s2[j] : s2[j + off];
                            const float recv1 = __shfl_xor_sync(0xffffffffu, send1, off);
                            const float recv2 = __shfl_xor_sync(0xffffffffu, send2, off);
                            s1[j] = (upper ? s1[j + off] : s1[j]) + recv1;
                            s2[j] = (upper ? s2[j + off] : s2[j]) + recv2;
                        }
                    }
                    float* sc = s_scr + ((wg * 4 + q) * 2) * BN + c * 32 + lane;
                    sc[0] = s1[0];
                    sc[BN] = s2[0];
                }
                if (direct) {
                    if (row_ok) {
                        __nv_bfloat16* o = p.out + orow * N + col0;
#pragma unroll
                        for (int j = 0; j < 32; j += 8) st8(o + j, pack8(v + j));
                    }
                } else if (tma_out) {
                    // bf16 tile → shared memory in the 128B-swizzled layout of the output tensor map
                    uint8_t* atom = sC + (c >> 1) * 16384;
#pragma unroll
                    for (int j = 0; j < 32; j += 8) {
                        const int col = (c & 1) * 32 + j;
                        const uint32_t off = (uint32_t)(r_in * 128 + ((((col >> 3) ^ (r_in & 7)) << 4)));
                        *reinterpret_cast<uint4*>(atom + off) = pack8(v + j).u;
                    }
                }
                if (has_f32 && row_ok) {
                    float* o = p.out_f32 + (size_t)row * N + col0;
#pragma unroll
                    for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(o + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
                }
                if (tma_out && (c & 1)) {
                    // a 64-column slab is complete: store it now, so its drain overlaps the conversion of the next slab
                    const int a = c >> 1;
                    fence_proxy_async();
                    if (a + 1 < kAtoms && t > 1 && is_leader)          // the next slab must be free before anyone writes it
                        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
                    asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");          // the 4 warps of this warpgroup
                    if (is_leader) {
                        tma_store_2d(&tmC, sC + a * 16384, n_tile * BN + a * 64, m_tile * BM);
                        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                    }
                }
            }
            }
            if (has_sums) {
                // the four warps' column partials of this tile are in the scratch array (ordered by the slab barrier above)
                if (!tma_out) asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
                if (tw < BN) {
                    const float* sc = s_scr + wg * 8 * BN + tw;
                    const float a = (sc[0] + sc[2 * BN]) + (sc[4 * BN] + sc[6 * BN]);
                    const float b = (sc[BN] + sc[3 * BN]) + (sc[5 * BN] + sc[7 * BN]);
                    const int col = n_tile * BN + tw;
                    if (smem_sums) {
                        float* ws = s_sums + wg * 2 * kSumCols;
                        ws[col] += a;
                        ws[kSumCols + col] += b;
                    } else {
                        atomicAdd(p.sums + col, a);
                        atomicAdd(p.sums + N + col, b);
                    }
                }
            }
        }
        if (has_sums) {
            asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
            if (smem_sums) {
                const float* ws = s_sums + wg * 2 * kSumCols;
                for (int i = tw; i < N; i += 128) {
                    const float a = ws[i], b = ws[kSumCols + i];
                    if (a != 0.f || b != 0.f) { atomicAdd(p.sums + i, a); atomicAdd(p.sums + N + i, b); }
                }
            }
        }
        if (tma_out && is_leader) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<2 * BN>(tmem_base);
}

// ------------------------------------------------------------------ host side: tensor maps
// cuTensorMapEncodeTiled is a driver-API function; it is resolved through the runtime
// (cudaGetDriverEntryPoint) so the library has no link-time dependency on libcuda and can be
// loaded (and built) on a machine without a GPU driver.
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
#define cuTensorMapEncodeTiled(...) (get_encode() ? get_encode()(__VA_ARGS__) : CUDA_ERROR_NOT_INITIALIZED)

bool encode_2d(CUtensorMap* m, const void* ptr, uint64_t inner, uint64_t outer, uint32_t box_inner, uint32_t box_outer) {
    cuuint64_t dims[2] = {inner, outer};
    cuuint64_t strides[1] = {inner * 2};
    cuuint32_t box[2] = {box_inner, box_outer};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = cuTensorMapEncodeTiled(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS;
}
// [Nb, H, W, C] bf16; the box covers bw x bh x bn PIXELS taken every `es`-th column / row (traversal stride: a strided
// convolution reads every es-th input pixel; the box extent in tensor elements is then es x as large)
bool encode_nhwc(CUtensorMap* m, const void* ptr, int Nb, int H, int W, int C, int bw, int bh, int bn, int es = 1) {
    cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)Nb};
    cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
    cuuint32_t box[4] = {64, (cuuint32_t)(bw * es), (cuuint32_t)(bh * es), (cuuint32_t)bn};
    cuuint32_t estr[4] = {1, (cuuint32_t)es, (cuuint32_t)es, 1};
    CUresult r = cuTensorMapEncodeTiled(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(ptr), dims, strides, box, estr,
                                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS;
}

template <int BN, bool kConv, int kEpi>
int launch_epi(const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmParams& p, cudaStream_t st) {
    CUtensorMap tmC;
    if (p.out) { if (!encode_2d(&tmC, p.out, p.N, p.M, 64, BM)) return -4; }
    else tmC = tmB;
    constexpr int kStages = (BN == 128) ? 4 : 7;      // one persistent CTA per SM: operand ring + 2 staging tiles
    using L = SmemLayout<BN, kStages>;
    auto kern = gemm_nt_kernel<BN, kStages, kConv, kEpi>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotal);
    if (e != cudaSuccess) return (int)e;
    const int num_tiles = ((p.M + BM - 1) / BM) * (p.N / BN);
    static int num_sms = 0;
    if (!num_sms) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev); }
    const int grid = num_tiles < num_sms ? num_tiles : num_sms;
    if (fdt::launch(kern, dim3(grid), dim3(kThreads), L::kTotal, st, tmA, tmB, tmC, p) != cudaSuccess) return (int)cudaGetLastError();
    FDT_CHECK_LAUNCH();
    return 0;
}

// The common epilogues (plain store, store + residual addend, store + column statistics) have their own instantiations.
template <int BN, bool kConv>
int launch(const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmParams& p, cudaStream_t st) {
    const bool simple = p.out && !p.out_f32 && !p.bias && p.act == ACT_NONE && p.omode == 0 && !(p.sums && p.addend);
    static const bool force_any = [] { const char* e = getenv("FDT_GEMM_EPI_ANY"); return e && e[0] == '1'; }();
    if (!simple || force_any) return launch_epi<BN, kConv, EPI_ANY>(tmA, tmB, p, st);
    if (p.sums) return launch_epi<BN, kConv, EPI_SUMS>(tmA, tmB, p, st);
    if (p.addend) return launch_epi<BN, kConv, EPI_ADDEND>(tmA, tmB, p, st);
    return launch_epi<BN, kConv, EPI_STORE>(tmA, tmB, p, st);
}

// 128-wide N tiles unless that leaves more than half of the SMs without a tile (deep, narrow problems such as the
// 4x4-feature-map convolutions): then 64-wide tiles double the tile count.
int pick_bn(int M, int N) {
    if (N % 128) return 64;
    const int tiles128 = ((M + BM - 1) / BM) * (N / 128);
    return tiles128 <= 74 ? 64 : 128;
}

}  // namespace

// out[M,N] (bf16 and/or fp32) = act(A[M,K] · B[N,K]ᵀ + bias) [+ addend].  K % 64 == 0, N % 64 == 0, 16-byte aligned rows.
// sums (optional, [2N], zeroed by the caller): per-column Σ and Σ² of the stored values.
// addend (optional, bf16 [M,N], may alias out_bf16): added before the store.
FDT_API int fdt_gemm_nt_ex(const void* A, const void* B, void* out_bf16, float* out_f32, const float* bias, float* sums,
                           const void* addend, int M, int N, int K, int act, cudaStream_t st) {
    if (M <= 0 || N % 64 || K % 64 || K <= 0) return -1;
    CUtensorMap tmA, tmB;
    const int BN = pick_bn(M, N);
    if (!encode_2d(&tmA, A, K, M, BK, BM) || !encode_2d(&tmB, B, K, N, BK, BN)) return -4;
    GemmParams p{};
    p.M = M; p.N = N; p.num_kb = K / BK; p.out = (__nv_bfloat16*)out_bf16; p.out_f32 = out_f32; p.bias = bias;
    p.sums = sums; p.act = act; p.addend = (const __nv_bfloat16*)addend;
    return BN == 128 ? launch<128, false>(tmA, tmB, p, st) : launch<64, false>(tmA, tmB, p, st);
}

FDT_API int fdt_gemm_nt(const void* A, const void* B, void* out_bf16, float* out_f32, const float* bias, float* sums,
                        int M, int N, int K, int act, cudaStream_t st) {
    return fdt_gemm_nt_ex(A, B, out_bf16, out_f32, bias, sums, nullptr, M, N, K, act, st);
}

namespace {
// tile geometry of a conv-mode launch over an OH x OW grid of output pixels: 128 pixels = bh rows x OW (x bn images)
bool tile_geom(int OH, int OW, int& bh, int& bn) {
    if (OW > 128 || OW < 1 || (OW & (OW - 1)) || (128 % OW)) return false;
    bh = 128 / OW; bn = 1;
    if (bh > OH) { if (bh % OH) return false; bn = bh / OH; bh = OH; }
    else if (OH % bh) return false;
    return true;
}
}  // namespace

// y[Nb,OH,OW,Cout] = conv(x[Nb,H,W,Cin], w[Cout,k,k,Cin]) [+ addend], stride 1 or 2, NHWC bf16, OH = (H + 2 pad - k)/stride + 1.
// Implicit GEMM: per tap one 4-D TMA box of the input shifted by (s - pad, r - pad) — for stride 2 the tensor map walks
// every second row / column (element strides), so neither an im2col matrix nor a strided copy is ever materialised.
// Cin % 64 == 0, Cout % 64 == 0, OW a power of two <= 128, OH*OW a multiple or a divisor of the 128-pixel tile.
FDT_API int fdt_conv_ex(const void* x, const void* w, void* y, float* sums, const void* addend, int Nb, int H, int W, int Cin,
                        int Cout, int ksize, int pad, int stride, cudaStream_t st) {
    if (Cin % 64 || Cout % 64 || ksize < 1 || ksize > 3 || (stride != 1 && stride != 2)) return -1;
    const int OH = (H + 2 * pad - ksize) / stride + 1, OW = (W + 2 * pad - ksize) / stride + 1;
    int bh, bn;
    if (!tile_geom(OH, OW, bh, bn)) return -1;
    if (stride == 2 && ((H & 1) || (W & 1))) return -1;
    const int M = Nb * OH * OW;
    CUtensorMap tmA, tmB;
    const int BN = pick_bn(M, Cout);
    const int K = ksize * ksize * Cin;
    if (!encode_nhwc(&tmA, x, Nb, H, W, Cin, OW, bh, bn, stride) || !encode_2d(&tmB, w, K, Cout, BK, BN)) return -4;
    GemmParams p{};
    p.M = M; p.N = Cout; p.out = (__nv_bfloat16*)y; p.sums = sums; p.act = ACT_NONE; p.addend = (const __nv_bfloat16*)addend;
    p.H = OH; p.W = OW; p.bh = bh; p.bn = bn; p.cpb = Cin / BK; p.in_mul = stride;
    p.ntaps = ksize * ksize;
    for (int r = 0; r < ksize; ++r)
        for (int sx = 0; sx < ksize; ++sx) {
            const int t = r * ksize + sx;
            p.tap_dh[t] = r - pad; p.tap_dw[t] = sx - pad; p.tap_k[t] = t * Cin;
        }
    p.num_kb = p.ntaps * p.cpb;
    return BN == 128 ? launch<128, true>(tmA, tmB, p, st) : launch<64, true>(tmA, tmB, p, st);
}

FDT_API int fdt_conv_fwd(const void* x, const void* w, void* y, float* sums, int Nb, int H, int W, int Cin, int Cout,
                         int ksize, int pad, cudaStream_t st) {
    return fdt_conv_ex(x, w, y, sums, nullptr, Nb, H, W, Cin, Cout, ksize, pad, 1, st);
}

// Data gradient of a stride-2 convolution, without col2im: dx[Nb,H,W,Cin] (+)= Σ_taps dy[Nb,H/2,W/2,Cout] · w.
// The input pixels fall into four parity classes (h % 2, w % 2); each class receives contributions from a fixed subset of
// the taps (3x3, pad 1: 1, 2, 2 and 4 taps; 1x1: only the even/even class), and within a class the gather from dy is a
// dense stride-1 pattern — so every class is one implicit-GEMM launch over the dy grid whose epilogue stores straight to
// the class's pixels.  wt: the tap-rotated transposed filter [Cin][k][k][Cout] (the layout the stride-1 dgrad uses).
// accumulate != 0: dx already holds the other branch's gradient and the result is ADDED to it in the epilogue (classes
// without taps are then left alone); accumulate == 0: dx is overwritten (classes without taps are zero-filled).
FDT_API int fdt_conv_dgrad_s2(const void* dy, const void* wt, void* dx, int accumulate, int Nb, int H, int W, int Cin, int Cout,
                              int ksize, int pad, cudaStream_t st) {
    if (Cin % 64 || Cout % 64 || (H & 1) || (W & 1) || ksize < 1 || ksize > 3) return -1;
    const int OH = (H + 2 * pad - ksize) / 2 + 1, OW = (W + 2 * pad - ksize) / 2 + 1;
    if (OH != H / 2 || OW != W / 2) return -1;
    int bh, bn;
    if (!tile_geom(OH, OW, bh, bn)) return -1;
    const int M = Nb * OH * OW;
    const int BN = pick_bn(M, Cin);
    const int K = ksize * ksize * Cout;
    CUtensorMap tmA, tmB;
    if (!encode_nhwc(&tmA, dy, Nb, OH, OW, Cout, OW, bh, bn, 1) || !encode_2d(&tmB, wt, K, Cin, BK, BN)) return -4;
    bool need_clear = false;
    for (int cls = 0; cls < 4 && !accumulate; ++cls) {
        const int ph = cls >> 1, pw = cls & 1;
        int nh = 0, nw = 0;
        for (int r = 0; r < ksize; ++r) nh += ((ph + pad - r) % 2 == 0);
        for (int s = 0; s < ksize; ++s) nw += ((pw + pad - s) % 2 == 0);
        if (nh * nw == 0) need_clear = true;
    }
    if (need_clear) {
        cudaError_t e = cudaMemsetAsync(dx, 0, (size_t)Nb * H * W * Cin * 2, st);
        if (e != cudaSuccess) return (int)e;
    }
    for (int cls = 0; cls < 4; ++cls) {
        const int ph = cls >> 1, pw = cls & 1;
        GemmParams p{};
        p.M = M; p.N = Cin; p.out = (__nv_bfloat16*)dx; p.act = ACT_NONE;
        p.addend = accumulate ? (const __nv_bfloat16*)dx : nullptr;
        p.H = OH; p.W = OW; p.bh = bh; p.bn = bn; p.cpb = Cout / BK; p.in_mul = 1;
        p.omode = 1; p.OH = H; p.OW = W; p.oph = ph; p.opw = pw;
        int nt = 0;
        for (int r = 0; r < ksize; ++r) {
            if ((ph + pad - r) % 2) continue;
            for (int s = 0; s < ksize; ++s) {
                if ((pw + pad - s) % 2) continue;
                // x row h = 2i + ph receives dy row (h + pad - r) / 2 = i + (ph + pad - r) / 2 through tap r
                p.tap_dh[nt] = (ph + pad - r) / 2; p.tap_dw[nt] = (pw + pad - s) / 2;
                p.tap_k[nt] = ((ksize - 1 - r) * ksize + (ksize - 1 - s)) * Cout;      // position in the ROTATED filter
                ++nt;
            }
        }
        if (nt == 0) continue;
        p.ntaps = nt; p.num_kb = nt * p.cpb;
        const int rc = BN == 128 ? launch<128, true>(tmA, tmB, p, st) : launch<64, true>(tmA, tmB, p, st);
        if (rc) return rc;
    }
    return 0;
}
