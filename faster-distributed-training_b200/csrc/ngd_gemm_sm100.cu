// Batched TF32 tcgen05 GEMMs for the NGD preconditioner (fp32 operands in shared memory, kind::tf32,
// fp32 accumulation in tensor memory), one launch per same-shaped factor batch (blockIdx.z = factor).
//
// Per factor (X: the gradient matrix, N rows x D; W: the R x D inverse-Fisher factor, R <= 80):
//     H  = X Wᵀ                       "project"  (store)
//     X̂  = X − H W                    "deflate"  (subtract epilogue, + Σx², Σx̂² for the norm restoration)
//     J  = Hᵀ X                       "stats"    (long reduction over the N rows: split-K + red.add)
// For a tensor's FIRST axis the gradient is only available transposed (X' = Xᵀ, D x N, N contiguous); the
// same three products are then issued with MN-major operand descriptors / a transposed epilogue instead of
// materialising X — all four K-major / MN-major operand combinations are used.
//
// Parity: the torch.mm chain of OnlineNaturalGradient._precondition_directions3 (ngd_optimizer.py:195-218;
// one cuBLAS SGEMM launch per product per factor — SURVEY.md K17).
#include "common.cuh"
#include "sm100.cuh"
#include <cstdio>
#include <mutex>

using namespace fdt;
using namespace sm100;

namespace {

constexpr int BM = 128, BKE = 32 /* fp32 elements = 128 B */, kThreads = 192, kStages = 4;
enum Epi : int { EPI_STORE = 0, EPI_SUB = 1, EPI_RED = 2 };

struct TfParams {
    int M, N, K;                // GEMM extents per batch item (D = A·Bᵀ-style: D[M,N] = Σ_k A[m,k] B[n,k])
    int BN;                     // N tile (multiple of 16, <= 128)
    int a_mn, b_mn;             // operand majorness: 0 = K contiguous, 1 = M/N contiguous
    int epi;
    int c_trans;                // EPI_SUB/STORE: C/out element (m, n) lives at [n * ldc + m] instead of [m * ldc + n]
    float* out; const float* cin;
    int64_t ldc, batch_stride_c;
    float* sums;                // EPI_SUB: [2 * batches] Σcin², Σout²
    int batches;
    int num_kb, kb_per_split;
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
__host__ __device__ constexpr uint32_t make_idesc_tf32(uint32_t M, uint32_t N, uint32_t a_mn, uint32_t b_mn) {
    return (1u << 4) | (2u << 7) | (2u << 10) | (a_mn << 15) | (b_mn << 16) | ((N >> 3) << 17) | ((M >> 4) << 24);
}

constexpr int kABytes = BM * 128;           // 128 rows x 128 B (K-major) or 4 atoms of 32 k-rows x 128 B (MN-major)
constexpr int kBBytes = 128 * 128;
constexpr int kStageBytes = kABytes + kBBytes;
constexpr int kSmemTotal = kStages * kStageBytes + (2 * kStages + 1) * 8 + 16 + 1024;

__global__ void __launch_bounds__(kThreads, 1)
tf32_gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const TfParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kStages * kStageBytes);
    uint64_t* empty_bar = full_bar + kStages;
    uint64_t* tmem_full_bar = empty_bar + kStages;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tiles_n = (p.N + p.BN - 1) / p.BN;
    const int m_tile = blockIdx.x / tiles_n, n_tile = blockIdx.x % tiles_n;
    const int batch = blockIdx.z;
    const int kb0 = blockIdx.y * p.kb_per_split;
    const int nkb = min(p.num_kb, kb0 + p.kb_per_split) - kb0;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmA); prefetch_tmap(&tmB);
        for (int s = 0; s < kStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
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
                    const int kb = kb0 + i, s = i % kStages;
                    mbar_wait(&empty_bar[s], ((i / kStages) & 1) ^ 1);
                    uint8_t* sA = smem + s * kStageBytes;
                    uint8_t* sB = sA + kABytes;
                    mbar_arrive_expect_tx(&full_bar[s], kABytes + b_bytes);
                    if (p.a_mn) {       // [K rows][M contiguous]: 4 atoms of {32 m, 32 k}
#pragma unroll
                        for (int a = 0; a < 4; ++a) tma_load_3d(sA + a * 4096, &tmA, &full_bar[s], m_tile * BM + a * 32, kb * BKE, batch);
                    } else {
                        tma_load_3d(sA, &tmA, &full_bar[s], kb * BKE, m_tile * BM, batch);
                    }
                    if (p.b_mn) {
                        for (int a = 0; a < p.BN / 32; ++a) tma_load_3d(sB + a * 4096, &tmB, &full_bar[s], n_tile * p.BN + a * 32, kb * BKE, batch);
                    } else {
                        tma_load_3d(sB, &tmB, &full_bar[s], kb * BKE, n_tile * p.BN, batch);
                    }
                }
            }
        } else if (warp == 1) {
            if (lane == 0) {
                const uint32_t idesc = make_idesc_tf32(BM, (uint32_t)p.BN, (uint32_t)p.a_mn, (uint32_t)p.b_mn);
                for (int i = 0; i < nkb; ++i) {
                    const int s = i % kStages;
                    mbar_wait(&full_bar[s], (i / kStages) & 1);
                    tc_fence_after();
                    const uint32_t a_addr = smem_u32(smem + s * kStageBytes), b_addr = a_addr + kABytes;
                    // K-major: 8-row groups 1 KB apart, +32 B per K=8 step.  MN-major: 32-wide atoms 4 KB apart (LBO),
                    // 8-k-row groups 1 KB apart (SBO), +1 KB per K=8 step.
                    const uint64_t da = p.a_mn ? make_smem_desc(a_addr, 4096, 1024) : make_smem_desc(a_addr, 16, 1024);
                    const uint64_t db = p.b_mn ? make_smem_desc(b_addr, 4096, 1024) : make_smem_desc(b_addr, 16, 1024);
                    const uint32_t sa = p.a_mn ? 64 : 2, sb = p.b_mn ? 64 : 2;
#pragma unroll
                    for (int k = 0; k < BKE / 8; ++k) umma_tf32(tmem_base, da + sa * k, db + sb * k, idesc, (i | k) != 0);
                    umma_commit(&empty_bar[s]);
                }
                umma_commit(tmem_full_bar);
            }
        } else {
            mbar_wait(tmem_full_bar, 0);
            tc_fence_after();
            const int q = warp & 3;
            const int m = m_tile * BM + q * 32 + lane;
            const bool m_ok = m < p.M;
            float* out = p.out + (size_t)batch * p.batch_stride_c;
            const float* cin = p.cin ? p.cin + (size_t)batch * p.batch_stride_c : nullptr;
            float sq_in = 0.f, sq_out = 0.f;
#pragma unroll 1
            for (int c = 0; c < (p.BN + 31) / 32; ++c) {
                uint32_t r[32];
                tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + c * 32, r);
                tmem_ld_wait();
                const int n0 = n_tile * p.BN + c * 32;
                if (!m_ok) continue;
                if (p.c_trans) {
                    // element (m, n) at [n * ldc + m]: lanes (consecutive m) hit consecutive addresses
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const int n = n0 + j;
                        if (n < p.N && c * 32 + j < p.BN) {
                            const size_t idx = (size_t)n * p.ldc + m;
                            float v = __uint_as_float(r[j]);
                            if (p.epi == EPI_SUB) { const float x = cin[idx]; sq_in += x * x; v = x - v; sq_out += v * v; }
                            if (p.epi == EPI_RED) atomicAdd(out + idx, v); else out[idx] = v;
                        }
                    }
                } else {
                    const size_t row = (size_t)m * p.ldc;
#pragma unroll
                    for (int j = 0; j < 32; j += 4) {
                        const int n = n0 + j;
                        if (n + 3 < p.N && c * 32 + j + 3 < p.BN) {
                            float4 v = make_float4(__uint_as_float(r[j]), __uint_as_float(r[j + 1]), __uint_as_float(r[j + 2]),
                                                   __uint_as_float(r[j + 3]));
                            if (p.epi == EPI_SUB) {
                                const float4 x = *reinterpret_cast<const float4*>(cin + row + n);
                                sq_in += x.x * x.x + x.y * x.y + x.z * x.z + x.w * x.w;
                                v = make_float4(x.x - v.x, x.y - v.y, x.z - v.z, x.w - v.w);
                                sq_out += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
                            }
                            if (p.epi == EPI_RED)
                                asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(out + row + n), "f"(v.x), "f"(v.y),
                                             "f"(v.z), "f"(v.w) : "memory");
                            else *reinterpret_cast<float4*>(out + row + n) = v;
                        }
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
// fp32 matrix batch [batches][outer][inner] (inner contiguous, row stride ld elements); box = 32 inner x box_outer
bool encode3(CUtensorMap* m, const float* ptr, uint64_t inner, uint64_t outer, uint64_t batches, uint64_t ld, uint64_t bstride,
             uint32_t box_outer) {
    if (!get_encode()) return false;
    cuuint64_t dims[3] = {inner, outer, batches};
    cuuint64_t strides[2] = {ld * 4, bstride * 4};
    cuuint32_t box[3] = {32, box_outer, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = get_encode()(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(ptr), dims, strides, box, estr,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
        fprintf(stderr, "fdt ngd gemm: tensor-map encode failed (%d): inner=%llu outer=%llu batches=%llu ld=%llu\n", (int)r,
                (unsigned long long)inner, (unsigned long long)outer, (unsigned long long)batches, (unsigned long long)ld);
    return r == CUDA_SUCCESS;
}

int launch(const CUtensorMap& a, const CUtensorMap& b, TfParams p, int splits, cudaStream_t st) {
    static bool attr = false;
    if (!attr) {
        cudaError_t e = cudaFuncSetAttribute(tf32_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemTotal);
        if (e != cudaSuccess) return (int)e;
        attr = true;
    }
    p.num_kb = (p.K + BKE - 1) / BKE;
    if (splits < 1) splits = 1;
    if (splits > p.num_kb) splits = p.num_kb;
    p.kb_per_split = (p.num_kb + splits - 1) / splits;
    splits = (p.num_kb + p.kb_per_split - 1) / p.kb_per_split;
    const int tiles = ((p.M + BM - 1) / BM) * ((p.N + p.BN - 1) / p.BN);
    tf32_gemm_kernel<<<dim3(tiles, splits, p.batches), kThreads, kSmemTotal, st>>>(a, b, p);
    FDT_CHECK_LAUNCH();
    return 0;
}

int pick_bn(int N, int granule) { return N >= 128 ? 128 : ((N + granule - 1) / granule) * granule; }   // 16 (K-major B) or 32 (MN-major B: whole atoms)

}  // namespace

// All operands are K-major (the reduction index contiguous): with kind::tf32 the plain 128B swizzle is only
// usable that way, so the few transposed views that are needed (Wᵀ, Hᵀ, and Xᵀ where the gradient is stored the
// other way round) are small explicit copies made by the caller — W and H are R-wide, X is weight-sized.
//
// Projection + deflation for `n` same-shaped factors:
//   Xrows [n, N, D]   the gradient matrix with rows = N (for a first-axis factor: the transposed copy of X')
//   W [n, R, D], Wt [n, D, R] (= Wᵀ)
//   H out [n, N, R] = Xrows Wᵀ
//   Xhat out = Cin − H W, where Cin / Xhat are [n, N, D] (c_trans == 0) or stored transposed [n, D, N] (c_trans != 0)
//   sums [2n] zero on entry → Σ Cin², Σ Xhat² per factor.
// Requirements: R % 16 == 0, R <= 128, D % 4 == 0, (c_trans ? N % 4 == 0 : true).
FDT_API int fdt_ngd_project(const float* Xrows, const float* W, const float* Wt, float* H, const float* Cin, float* Xhat,
                            float* sums, int n, int N, int D, int R, int c_trans, cudaStream_t st) {
    if (R % 16 || R > 128 || D % 4 || (c_trans && N % 4)) return -1;
    CUtensorMap tA, tB;
    TfParams p{};
    p.batches = n; p.a_mn = 0; p.b_mn = 0;
    // ---- H[N, R] = Xrows Wᵀ :  M = N, N = R, K = D
    p.M = N; p.N = R; p.K = D; p.BN = pick_bn(R, 16); p.epi = EPI_STORE; p.c_trans = 0;
    p.out = H; p.cin = nullptr; p.ldc = R; p.batch_stride_c = (int64_t)N * R; p.sums = nullptr;
    if (!encode3(&tA, Xrows, D, N, n, D, (uint64_t)N * D, BM)) return -4;
    if (!encode3(&tB, W, D, R, n, D, (uint64_t)R * D, p.BN)) return -4;
    int rc = launch(tA, tB, p, 1, st);
    if (rc) return rc;
    // ---- Xhat = Cin − H W :  M = N, N = D, K = R ;  A = H [N, R], B = Wt [D, R]
    p.M = N; p.N = D; p.K = R; p.BN = pick_bn(D, 16); p.epi = EPI_SUB;
    p.out = Xhat; p.cin = Cin; p.sums = sums;
    if (c_trans) { p.c_trans = 1; p.ldc = N; } else { p.c_trans = 0; p.ldc = D; }
    p.batch_stride_c = (int64_t)N * D;
    if (!encode3(&tA, H, R, N, n, R, (uint64_t)N * R, BM)) return -4;
    if (!encode3(&tB, Wt, R, D, n, R, (uint64_t)D * R, p.BN)) return -4;
    return launch(tA, tB, p, 1, st);
}

// J[n, R, D] = Hᵀ X = Ht · Xtᵀ with Ht [n, R, N] and Xt [n, D, N] (both with the N rows contiguous).
// J must be zero on entry: split-K partial tiles are red.add-ed.  R % 4 == 0, D % 4 == 0, N % 4 == 0.
FDT_API int fdt_ngd_stats_j(const float* Ht, const float* Xt, float* J, int n, int N, int D, int R, cudaStream_t st) {
    if (R % 4 || R > 128 || D % 4 || N % 4) return -1;
    CUtensorMap tA, tB;
    TfParams p{};
    p.batches = n; p.a_mn = 0; p.b_mn = 0;
    p.M = R; p.N = D; p.K = N; p.BN = pick_bn(D, 16); p.epi = EPI_RED; p.c_trans = 0;
    p.out = J; p.cin = nullptr; p.ldc = D; p.batch_stride_c = (int64_t)R * D; p.sums = nullptr;
    if (!encode3(&tA, Ht, N, R, n, N, (uint64_t)R * N, BM)) return -4;
    if (!encode3(&tB, Xt, N, D, n, N, (uint64_t)D * N, p.BN)) return -4;
    const int tiles = ((D + p.BN - 1) / p.BN) * n;
    int splits = (148 * 2 + tiles - 1) / tiles;
    return launch(tA, tB, p, splits, st);
}
