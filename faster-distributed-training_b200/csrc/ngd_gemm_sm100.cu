// Batched fp32-accurate tcgen05 GEMMs for the NGD preconditioner ("3xTF32": every fp32 operand is split into a
// tf32-exact high part and a low remainder in shared memory, and three kind::tf32 MMAs — hi·hi, lo·hi, hi·lo —
// accumulate in fp32 tensor memory; relative error ~2^-22 per product, the level of an fp32 FMA chain).
// One launch per same-shaped factor batch (blockIdx.z = factor).
//
// Per factor (X: the gradient matrix, N rows x D; W: the R x D inverse-Fisher factor, R <= 128):
//     H  = X Wᵀ                       "project"  (store)
//     X̂  = X − H W                    "deflate"  (subtract epilogue, + Σx², Σx̂² for the norm restoration)
//     J  = Hᵀ X                       "stats"    (long reduction over the N rows: split-K + red.add)
//
// Pipeline per CTA (one 128 x BN output tile):  TMA warp → [full] → 4 converter/epilogue warps → [conv] → MMA
// thread → [empty].  The converter warps do two jobs on each landed stage:
//   * split:      hi = x rounded to tf32 (exactly representable), lo = x − hi, written to a second tile;
//   * transpose:  kind::tf32 only takes K-major tiles with the plain 128B swizzle, but half of the operands of the
//                 three products are stored with the M/N index contiguous (Xᵀ for a tensor's first axis, W in the
//                 deflation, H and X in the statistics).  Those land un-swizzled as [32 k][128 m] boxes and the
//                 converter writes them out in the K-major 128B-swizzled layout — no transposed copy in global memory.
//
// Parity: the torch.mm chain of OnlineNaturalGradient._precondition_directions3 (ngd_optimizer.py:195-218;
// one cuBLAS SGEMM launch per product per factor — SURVEY.md K17).
#include "common.cuh"
#include "sm100.cuh"
#include <cstdio>
#include <cstdlib>
#include <mutex>

using namespace fdt;
using namespace sm100;

namespace {

constexpr int BM = 128, BKE = 32 /* fp32 elements = 128 B */, kThreads = 192;
enum Epi : int { EPI_STORE = 0, EPI_SUB = 1, EPI_RED = 2 };

struct TfParams {
    int M, N, K;                // GEMM extents per batch item: D[M,N] = Σ_k A[m,k] B[n,k]
    int BN;                     // N tile (multiple of 16, <= 128)
    int a_mn, b_mn;             // operand storage: 0 = K contiguous ([rows][K]), 1 = M/N contiguous ([K][rows])
    int epi;
    int terms;                  // 3: fp32-accurate split products; 1: plain tf32
    float* out; const float* cin;
    int64_t ldc, batch_stride_c;
    float* sums;                // EPI_SUB: [2 * batches] Σcin², Σout²
    int batches;
    int num_kb, kb_per_split;
    int stages, stage_bytes, a_raw_off, b_hi_off, b_lo_off, b_raw_off;   // stage layout (bytes from the stage base)
};

__device__ __forceinline__ void tma_load_3d(void* smem, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                 ::"r"(smem_u32(smem)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
                 : "memory");
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
}
__host__ __device__ constexpr uint32_t make_idesc_tf32(uint32_t M, uint32_t N) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((N >> 3) << 17) | ((M >> 4) << 24);      // K-major A and B
}

constexpr int kATile = BM * 128;            // 128 rows x 128 B

// byte offset of the 16-byte chunk (row r, k-quad q) in a K-major tile with the 128B swizzle
__device__ __forceinline__ uint32_t sw128(int r, int q) { return (uint32_t)((r >> 3) * 1024 + (r & 7) * 128 + ((q ^ (r & 7)) << 4)); }

// hi = x rounded to nearest tf32 (10 explicit mantissa bits), lo = x − hi (exact in fp32, |lo| <= 2^-11 |x|, either
// sign — so the rounding errors of the low-order products do not accumulate in one direction over a long K)
__device__ __forceinline__ float tf32_rn(float x) { return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u); }
__device__ __forceinline__ void split4(const float4 v, float4& hi, float4& lo) {
    hi.x = tf32_rn(v.x); lo.x = v.x - hi.x;
    hi.y = tf32_rn(v.y); lo.y = v.y - hi.y;
    hi.z = tf32_rn(v.z); lo.z = v.z - hi.z;
    hi.w = tf32_rn(v.w); lo.w = v.w - hi.w;
}

// converter: `rows` x 32 fp32 tile → hi / lo tiles in the K-major swizzled layout.  t = 0..127.
//   mn == 0: the tile already is K-major swizzled in `hi` (TMA wrote it there): split in place.
//   mn != 0: the tile sits un-swizzled in `raw` as [32 k][raw_pitch floats]: transpose while splitting.
__device__ __forceinline__ void convert_tile(uint8_t* hi, uint8_t* lo, const uint8_t* raw, int rows, int raw_pitch, int mn,
                                             bool want_lo, int t) {
    if (!mn) {
        for (int i = t; i < rows * 8; i += 128) {
            float4* ph = reinterpret_cast<float4*>(hi + i * 16);
            float4 h, l;
            split4(*ph, h, l);
            if (want_lo) { *ph = h; *reinterpret_cast<float4*>(lo + i * 16) = l; }
        }
    } else {
        const float* src = reinterpret_cast<const float*>(raw);
        for (int r = t; r < rows; r += 128) {
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                float4 v, h, l;
                v.x = src[(4 * q + 0) * raw_pitch + r]; v.y = src[(4 * q + 1) * raw_pitch + r];
                v.z = src[(4 * q + 2) * raw_pitch + r]; v.w = src[(4 * q + 3) * raw_pitch + r];
                split4(v, h, l);
                const uint32_t off = sw128(r, q);
                *reinterpret_cast<float4*>(hi + off) = want_lo ? h : v;
                if (want_lo) *reinterpret_cast<float4*>(lo + off) = l;
            }
        }
    }
}

__global__ void __launch_bounds__(kThreads, 2)
tf32_gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const TfParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + p.stages * p.stage_bytes);
    uint64_t* conv_bar = full_bar + 4;
    uint64_t* empty_bar = conv_bar + 4;
    uint64_t* tmem_full_bar = empty_bar + 4;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tiles_n = (p.N + p.BN - 1) / p.BN;
    const int m_tile = blockIdx.x / tiles_n, n_tile = blockIdx.x % tiles_n;
    const int batch = blockIdx.z;
    const int kb0 = blockIdx.y * p.kb_per_split;
    const int nkb = min(p.num_kb, kb0 + p.kb_per_split) - kb0;
    const int S = p.stages;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmA); prefetch_tmap(&tmB);
        for (int s = 0; s < S; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&conv_bar[s], 128); mbar_init(&empty_bar[s], 1); }
        mbar_init(tmem_full_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<128>(tmem_ptr);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;
    const uint32_t b_bytes = (uint32_t)p.BN * 128;

    if (nkb > 0) {
        if (warp == 0) {
            if (lane == 0) {
                for (int i = 0; i < nkb; ++i) {
                    const int kb = kb0 + i, s = i % S;
                    mbar_wait(&empty_bar[s], ((i / S) & 1) ^ 1);
                    uint8_t* st = smem + s * p.stage_bytes;
                    mbar_arrive_expect_tx(&full_bar[s], kATile + b_bytes);
                    if (p.a_mn) tma_load_3d(st + p.a_raw_off, &tmA, &full_bar[s], m_tile * BM, kb * BKE, batch);   // [32 k][128 m]
                    else tma_load_3d(st, &tmA, &full_bar[s], kb * BKE, m_tile * BM, batch);                        // [128 m][32 k] swizzled
                    if (p.b_mn) tma_load_3d(st + p.b_raw_off, &tmB, &full_bar[s], n_tile * p.BN, kb * BKE, batch);
                    else tma_load_3d(st + p.b_hi_off, &tmB, &full_bar[s], kb * BKE, n_tile * p.BN, batch);
                }
            }
        } else if (warp == 1) {
            if (lane == 0) {
                const uint32_t idesc = make_idesc_tf32(BM, (uint32_t)p.BN);
                for (int i = 0; i < nkb; ++i) {
                    const int s = i % S;
                    mbar_wait(&conv_bar[s], (i / S) & 1);
                    tc_fence_after();
                    const uint32_t base = smem_u32(smem + s * p.stage_bytes);
                    // K-major, 128B swizzle: 8-row groups 1 KB apart, +32 B per K = 8 step
                    const uint64_t ah = make_smem_desc(base, 16, 1024), al = make_smem_desc(base + kATile, 16, 1024);
                    const uint64_t bh = make_smem_desc(base + p.b_hi_off, 16, 1024), bl = make_smem_desc(base + p.b_lo_off, 16, 1024);
#pragma unroll
                    for (int k = 0; k < BKE / 8; ++k) {
                        if (p.terms == 3) {
                            umma_tf32(tmem_base, al + 2 * k, bh + 2 * k, idesc, (i | k) != 0);      // small terms first
                            umma_tf32(tmem_base, ah + 2 * k, bl + 2 * k, idesc, 1);
                            umma_tf32(tmem_base, ah + 2 * k, bh + 2 * k, idesc, 1);
                        } else {
                            umma_tf32(tmem_base, ah + 2 * k, bh + 2 * k, idesc, (i | k) != 0);
                        }
                    }
                    umma_commit(&empty_bar[s]);
                }
                umma_commit(tmem_full_bar);
            }
        } else {
            // ---------------- converter: split (and transpose) every landed stage
            const int t = threadIdx.x - 64;
            const bool want_lo = p.terms == 3;
            for (int i = 0; i < nkb; ++i) {
                const int s = i % S;
                mbar_wait(&full_bar[s], (i / S) & 1);
                uint8_t* st = smem + s * p.stage_bytes;
                if (want_lo || p.a_mn) convert_tile(st, st + kATile, st + p.a_raw_off, BM, BM, p.a_mn, want_lo, t);
                if (want_lo || p.b_mn) convert_tile(st + p.b_hi_off, st + p.b_lo_off, st + p.b_raw_off, p.BN, p.BN, p.b_mn, want_lo, t);
                fence_proxy_async();                 // generic-proxy writes → visible to the tensor core's async-proxy reads
                mbar_arrive(&conv_bar[s]);
            }
            // ---------------- epilogue
            mbar_wait(tmem_full_bar, 0);
            tc_fence_after();
            const int q = warp & 3;
            const int m = m_tile * BM + q * 32 + lane;
            const bool m_ok = m < p.M;
            float* __restrict__ out = p.out + (size_t)batch * p.batch_stride_c;
            const float* __restrict__ cin = p.cin ? p.cin + (size_t)batch * p.batch_stride_c : nullptr;
            float sq_in = 0.f, sq_out = 0.f;
#pragma unroll 1
            for (int c = 0; c < (p.BN + 31) / 32; ++c) {
                uint32_t r[32];
                tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + c * 32, r);
                tmem_ld_wait();
                const int n0 = n_tile * p.BN + c * 32;
                if (!m_ok) continue;
                const size_t row = (size_t)m * p.ldc;
                float4 x[8];
                if (p.epi == EPI_SUB) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const int n = n0 + 4 * j;
                        x[j] = (n + 3 < p.N && c * 32 + 4 * j + 3 < p.BN) ? __ldg(reinterpret_cast<const float4*>(cin + row + n))
                                                                          : make_float4(0.f, 0.f, 0.f, 0.f);
                    }
                }
#pragma unroll
                for (int j = 0; j < 32; j += 4) {
                    const int n = n0 + j;
                    if (n + 3 < p.N && c * 32 + j + 3 < p.BN) {
                        float4 v = make_float4(__uint_as_float(r[j]), __uint_as_float(r[j + 1]), __uint_as_float(r[j + 2]),
                                               __uint_as_float(r[j + 3]));
                        if (p.epi == EPI_SUB) {
                            const float4 xx = x[j >> 2];
                            sq_in += xx.x * xx.x + xx.y * xx.y + xx.z * xx.z + xx.w * xx.w;
                            v = make_float4(xx.x - v.x, xx.y - v.y, xx.z - v.z, xx.w - v.w);
                            sq_out += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
                        }
                        if (p.epi == EPI_RED)
                            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(out + row + n), "f"(v.x), "f"(v.y),
                                         "f"(v.z), "f"(v.w) : "memory");
                        else *reinterpret_cast<float4*>(out + row + n) = v;
                    }
                }
            }
            if (p.epi == EPI_SUB && p.sums) {
                sq_in = warp_sum(sq_in); sq_out = warp_sum(sq_out);
                if (lane == 0) { atomicAdd(p.sums + batch, sq_in); atomicAdd(p.sums + p.batches + batch, sq_out); }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<128>(tmem_base);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn get_encode() {
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
// fp32 matrix batch [batches][outer][inner] (inner contiguous, row stride ld elements).
//   swizzled: box = 32 inner x box_outer, 128B swizzle (a K-major operand tile);
//   raw:      box = box_inner x 32 outer, no swizzle   (an M/N-major tile the converter warps transpose)
bool encode3(CUtensorMap* m, const float* ptr, uint64_t inner, uint64_t outer, uint64_t batches, uint64_t ld, uint64_t bstride,
             uint32_t box_inner, uint32_t box_outer, bool swizzled) {
    if (!get_encode()) return false;
    cuuint64_t dims[3] = {inner, outer, batches};
    cuuint64_t strides[2] = {ld * 4, bstride * 4};
    cuuint32_t box[3] = {box_inner, box_outer, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = get_encode()(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(ptr), dims, strides, box, estr,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, swizzled ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
        fprintf(stderr, "fdt ngd gemm: tensor-map encode failed (%d): inner=%llu outer=%llu batches=%llu ld=%llu box=%ux%u\n", (int)r,
                (unsigned long long)inner, (unsigned long long)outer, (unsigned long long)batches, (unsigned long long)ld,
                box_inner, box_outer);
    return r == CUDA_SUCCESS;
}
// operand with `rows` rows (M or N index) and `K` reduction length, tile of box_rows rows
bool encode_operand(CUtensorMap* m, const float* ptr, int mn, uint64_t rows, uint64_t K, uint64_t batches, uint32_t box_rows) {
    if (mn) return encode3(m, ptr, rows, K, batches, rows, rows * K, box_rows, 32, false);     // stored [K][rows]
    return encode3(m, ptr, K, rows, batches, K, rows * K, 32, box_rows, true);                 // stored [rows][K]
}

int g_terms = 0;
int terms() {
    if (g_terms == 0) {
        const char* e = getenv("FDT_NGD_TF32_TERMS");          // 3 (default): fp32-accurate split products; 1: plain tf32
        g_terms = (e && atoi(e) == 1) ? 1 : 3;
    }
    return g_terms;
}

int launch(const CUtensorMap& a, const CUtensorMap& b, TfParams p, int splits, cudaStream_t st) {
    p.terms = terms();
    const int bt = p.BN * 128;
    // stage: [A hi][A lo][B hi][B lo][A raw?][B raw?]; the lo tiles only exist for the 3-term product
    int off = kATile;
    if (p.terms == 3) off += kATile;
    p.b_hi_off = off; off += bt;
    p.b_lo_off = off; if (p.terms == 3) off += bt;
    p.a_raw_off = off; if (p.a_mn) off += kATile;
    p.b_raw_off = off; if (p.b_mn) off += bt;
    p.stage_bytes = (off + 1023) & ~1023;
    p.num_kb = (p.K + BKE - 1) / BKE;
    // The four converter / epilogue warps of a CTA are latency-bound (one warp per scheduler), so two CTAs share an SM:
    // the operand ring is sized to half of the shared memory (a ring that cannot fit twice gets a single stage at most).
    int stages = (110 * 1024) / p.stage_bytes;
    if (stages < 1) stages = (220 * 1024) / p.stage_bytes >= 1 ? 1 : 0;
    if (stages > 4) stages = 4;
    if (stages > p.num_kb) stages = p.num_kb < 1 ? 1 : p.num_kb;
    if (stages < 1) return -7;
    p.stages = stages;
    const int smem = stages * p.stage_bytes + 13 * 8 + 16 + 1024;
    static int attr = 0;
    if (attr < smem) {
        cudaError_t e = cudaFuncSetAttribute(tf32_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (e != cudaSuccess) return (int)e;
        attr = 227 * 1024;
    }
    if (splits < 1) splits = 1;
    if (splits > p.num_kb) splits = p.num_kb;
    p.kb_per_split = (p.num_kb + splits - 1) / splits;
    splits = (p.num_kb + p.kb_per_split - 1) / p.kb_per_split;
    const int tiles = ((p.M + BM - 1) / BM) * ((p.N + p.BN - 1) / p.BN);
    tf32_gemm_kernel<<<dim3(tiles, splits, p.batches), kThreads, smem, st>>>(a, b, p);
    FDT_CHECK_LAUNCH();
    return 0;
}

int pick_bn(int N) { return N >= 128 ? 128 : ((N + 15) / 16) * 16; }

}  // namespace

// Projection + deflation for `n` same-shaped factors, straight from the gradient where it lies:
//   X [n, N, D] (x_trans == 0) or the transposed gradient [n, D, N] (x_trans != 0);  W [n, R, D]
//   H out [n, N, R] = X Wᵀ;  Xhat out (same layout as X) = X − H W;  sums [2n], zero on entry → Σ X², Σ Xhat².
// Requirements: R % 4 == 0, R <= 128, D % 4 == 0, (x_trans ? N % 4 == 0 : true).
// Dw: row pitch (in elements) of W — D rounded up to a multiple of 4 with zero padding columns, so that a factor whose axis
// length is not a multiple of 4 (the 30522-row token embedding) still has 16-byte-aligned rows for TMA.  The gradient X is
// used where it lies with its true extent D: the rows a padded K loop reaches beyond D are zero-filled by the TMA unit.
// Only the transposed layout (x_trans, D = number of rows of X) can have D % 4 != 0.
FDT_API int fdt_ngd_project_ex(const float* X, const float* W, float* H, float* Xhat, float* sums, int n, int N, int D, int Dw,
                               int R, int x_trans, cudaStream_t st) {
    if (Dw == 0) Dw = D;
    if (R % 4 || R > 128 || Dw % 4 || Dw < D || (!x_trans && D % 4) || (x_trans && N % 4) || n <= 0) return -1;
    CUtensorMap tA, tB;
    TfParams p{};
    p.batches = n;
    // ---- H[N, R] = X Wᵀ :  M = N, N = R, K = D;  A = X (K-major, or [D][N] when transposed), B = W [R][D] K-major
    p.M = N; p.N = R; p.K = Dw; p.BN = pick_bn(R); p.epi = EPI_STORE; p.a_mn = x_trans ? 1 : 0; p.b_mn = 0;
    p.out = H; p.cin = nullptr; p.ldc = R; p.batch_stride_c = (int64_t)N * R; p.sums = nullptr;
    if (!encode_operand(&tA, X, p.a_mn, N, D, n, BM)) return -4;
    if (!encode_operand(&tB, W, 0, R, Dw, n, p.BN)) return -4;
    // A long reduction with few output tiles (the token embedding's first axis: N = 512 rows, K = 30522) would run on a
    // handful of CTAs: split K over the SMs and red.add the partial tiles into a zeroed H instead.
    int splits = 1;
    {
        const int tiles = ((N + BM - 1) / BM) * ((R + p.BN - 1) / p.BN) * n;
        const int num_kb = (Dw + BKE - 1) / BKE;
        if (tiles < 74 && num_kb >= 64) {
            splits = 148 / tiles;
            if (splits > num_kb / 8) splits = num_kb / 8;
            if (splits < 1) splits = 1;
        }
    }
    if (splits > 1) {
        cudaMemsetAsync(H, 0, sizeof(float) * (size_t)n * N * R, st);
        p.epi = EPI_RED;
    }
    int rc = launch(tA, tB, p, splits, st);
    if (rc) return rc;
    p.epi = EPI_SUB; p.out = Xhat; p.cin = X; p.sums = sums; p.K = R;
    p.batch_stride_c = (int64_t)N * D;
    if (!x_trans) {
        // ---- Xhat = X − H W :  M = N, N = D, K = R;  A = H [N][R] K-major, B[n = d, k = r] = W[r][d]: stored [K][rows]
        p.M = N; p.N = D; p.BN = pick_bn(D); p.a_mn = 0; p.b_mn = 1; p.ldc = D;
        if (!encode_operand(&tA, H, 0, N, R, n, BM)) return -4;
        if (!encode_operand(&tB, W, 1, Dw, R, n, p.BN)) return -4;
    } else {
        // ---- the gradient lies transposed ([D][N]): compute Xhatᵀ = Xᵀ − Wᵀ Hᵀ instead, so that the output tile is row-major
        // where it is stored and the epilogue moves 16 bytes per access:  M = D, N = N, K = R;  A[m = d, k = r] = W[r][d]
        // (stored [K][rows]), B = H [N][R] K-major.  Rows D .. Dw-1 of A are the zero padding and are never stored.
        p.M = D; p.N = N; p.BN = pick_bn(N); p.a_mn = 1; p.b_mn = 0; p.ldc = N;
        if (!encode_operand(&tA, W, 1, Dw, R, n, BM)) return -4;
        if (!encode_operand(&tB, H, 0, N, R, n, p.BN)) return -4;
    }
    return launch(tA, tB, p, 1, st);
}

FDT_API int fdt_ngd_project(const float* X, const float* W, float* H, float* Xhat, float* sums, int n, int N, int D, int R,
                            int x_trans, cudaStream_t st) {
    return fdt_ngd_project_ex(X, W, H, Xhat, sums, n, N, D, D, R, x_trans, st);
}

// J[n, R, D] = Hᵀ X with H [n, N, R] and X [n, N, D] (x_trans == 0) or [n, D, N] (x_trans != 0).  J is zeroed here;
// split-K partial tiles are red.add-ed.  R % 4 == 0, D % 4 == 0, (x_trans ? N % 4 == 0 : true).
// Dw: row pitch of J (see fdt_ngd_project_ex); columns D .. Dw-1 of J come out zero.
FDT_API int fdt_ngd_stats_j_ex(const float* H, const float* X, float* J, int n, int N, int D, int Dw, int R, int x_trans,
                               cudaStream_t st) {
    if (Dw == 0) Dw = D;
    if (R % 4 || R > 128 || Dw % 4 || Dw < D || (!x_trans && D % 4) || (x_trans && N % 4) || n <= 0) return -1;
    cudaMemsetAsync(J, 0, sizeof(float) * (size_t)n * R * Dw, st);
    CUtensorMap tA, tB;
    TfParams p{};
    p.batches = n;
    // M = R, N = D, K = N;  A[m = r, k = i] = H[i][r]: stored [K][rows];  B[n = d, k = i] = X[i][d] ([K][rows]) or X'[d][i] (K-major)
    // N extent = the padded Dw (a multiple of 4: the epilogue adds 4 columns at a time); X's tensor map keeps the true D, so the
    // padding columns receive zeros
    p.M = R; p.N = Dw; p.K = N; p.BN = pick_bn(Dw); p.epi = EPI_RED; p.a_mn = 1; p.b_mn = x_trans ? 0 : 1;
    p.out = J; p.cin = nullptr; p.ldc = Dw; p.batch_stride_c = (int64_t)R * Dw; p.sums = nullptr;
    if (!encode_operand(&tA, H, 1, R, N, n, BM)) return -4;
    if (!encode_operand(&tB, X, p.b_mn, D, N, n, p.BN)) return -4;
    const int tiles = ((Dw + p.BN - 1) / p.BN) * n;
    int splits = (148 + tiles - 1) / tiles;
    return launch(tA, tB, p, splits, st);
}

FDT_API int fdt_ngd_stats_j(const float* H, const float* X, float* J, int n, int N, int D, int R, int x_trans, cudaStream_t st) {
    return fdt_ngd_stats_j_ex(H, X, J, n, N, D, D, R, x_trans, st);
}

// L[n, R, R] = Hᵀ H with H [n, N, R] — the rank-deficient case N <= D of the Fisher statistics (ngd_optimizer.py:215-217),
// both operands read M/N-major where they lie; the reduction over N is split over CTAs and red.add-ed.  L is zeroed here.
FDT_API int fdt_ngd_stats_hh(const float* H, float* L, int n, int N, int R, cudaStream_t st) {
    if (R % 4 || R < 16 || R > 128 || n <= 0 || N <= 0) return -1;
    cudaMemsetAsync(L, 0, sizeof(float) * (size_t)n * R * R, st);
    CUtensorMap tA, tB;
    TfParams p{};
    p.batches = n;
    p.M = R; p.N = R; p.K = N; p.BN = pick_bn(R); p.epi = EPI_RED; p.a_mn = 1; p.b_mn = 1;
    p.out = L; p.cin = nullptr; p.ldc = R; p.batch_stride_c = (int64_t)R * R; p.sums = nullptr;
    if (!encode_operand(&tA, H, 1, R, N, n, BM)) return -4;
    if (!encode_operand(&tB, H, 1, R, N, n, p.BN)) return -4;
    const int splits = (148 + n - 1) / n;
    return launch(tA, tB, p, splits, st);
}

// The R x R Fisher statistics of a refresh:  K[n, R, R] = J Jᵀ and (optionally) L[n, R, R] = J Wᵀ, with J, W [n, R, D]
// (ngd_optimizer.py:205-214: two torch.mm per factor).  Both operands are K-major as stored; the long reduction over D is
// split over CTAs and red.add-ed.  K and L are zeroed here.  R % 4 == 0, 16 <= R <= 128, D % 4 == 0.
FDT_API int fdt_ngd_stats_kl(const float* J, const float* W, float* K, float* L, int n, int D, int R, cudaStream_t st) {
    if (R % 4 || R < 16 || R > 128 || D % 4 || n <= 0) return -1;
    cudaMemsetAsync(K, 0, sizeof(float) * (size_t)n * R * R, st);
    if (L) cudaMemsetAsync(L, 0, sizeof(float) * (size_t)n * R * R, st);
    CUtensorMap tA, tB;
    TfParams p{};
    p.batches = n;
    p.M = R; p.N = R; p.K = D; p.BN = pick_bn(R); p.epi = EPI_RED; p.a_mn = 0; p.b_mn = 0;
    p.cin = nullptr; p.ldc = R; p.batch_stride_c = (int64_t)R * R; p.sums = nullptr;
    if (!encode_operand(&tA, J, 0, R, D, n, BM)) return -4;
    if (!encode_operand(&tB, J, 0, R, D, n, p.BN)) return -4;
    const int splits = (148 + n - 1) / n;
    p.out = K;
    int rc = launch(tA, tB, p, splits, st);
    if (rc || !L) return rc;
    if (!encode_operand(&tB, W, 0, R, D, n, p.BN)) return -4;
    p.out = L;
    return launch(tA, tB, p, splits, st);
}

