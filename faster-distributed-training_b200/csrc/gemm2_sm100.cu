// 2-CTA tcgen05 GEMM for sm_100a:  D[M, N] = act(A[M, K] · B[N, K]ᵀ + bias) [+ addend],  bf16 operands, fp32 accumulation.
//
// Thread-block clusters of TWO CTAs (one TPC) cooperate on a 256 x 256 output tile with `tcgen05.mma.cta_group::2`:
//   * each CTA stages ITS 128 rows of A and ITS 128 rows (output columns) of B per 64-wide k-block — half the shared-memory
//     traffic per FLOP of the 1-CTA kernel (csrc/gemm_sm100.cu), and the 256-wide MMA the tensor cores need for full rate;
//   * both CTAs' TMA loads complete on the LEADER's (cluster rank 0) mbarrier (`cp.async.bulk.tensor...cta_group::2`);
//   * one elected thread of the leader issues the MMAs for the pair (M = 256: rows 0-127 accumulate in the leader's tensor
//     memory, rows 128-255 in the peer's); `tcgen05.commit...multicast::cluster` releases the operand stage / publishes the
//     accumulator in BOTH CTAs;
//   * every CTA drains its own 128 x 256 accumulator half (two epilogue warpgroups alternate over a double-buffered TMEM
//     accumulator: 2 x 256 columns = the whole tensor memory), bias / activation / addend in registers, bf16 tile through
//     swizzled shared memory and a TMA store;
//   * persistent: clusters walk the tile list, the epilogue of tile i overlaps the main loop of tile i+1.
//
// Parity: the large F.linear / 1x1 F.conv2d call sites of the reference (transformer.py:168-176, 213-226; resnet.py:83),
// cuBLAS / cuDNN there.  Used when M >= 512 and N % 256 == 0; everything else stays on gemm_sm100.cu.
#include "common.cuh"
#include "sm100.cuh"
#include <cstdio>
#include <mutex>

using namespace fdt;
using namespace sm100;

namespace {

constexpr int BM = 128;            // rows per CTA (256 per pair)
constexpr int BN = 256;            // tile width; each CTA stages BN / 2 rows of B
constexpr int BK = 64;
constexpr int kStages = 4;
constexpr int kThreads = 320;      // TMA warp + MMA warp + 2 epilogue warpgroups

struct G2Params {
    int M, N, num_kb;
    __nv_bfloat16* out;
    const float* bias;
    const __nv_bfloat16* addend;
    int act;
};

constexpr int kABytes = BM * BK * 2;                 // 16 KB
constexpr int kBBytes = (BN / 2) * BK * 2;           // 16 KB
constexpr int kStageBytes = kABytes + kBBytes;
constexpr int kCOff = kStages * kStageBytes;         // 128 KB
constexpr int kCBytes = 2 * 2 * 16384;               // per warpgroup: 2 atoms (128 rows x 128 cols), 64 KB in total
constexpr int kBarOff = kCOff + kCBytes;
constexpr int kSmemTotal = kBarOff + (2 * kStages + 4) * 8 + 16 + 1024;

__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `smem_addr` (a shared::cta address of this CTA) in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa(uint32_t smem_addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(rank));
    return r;
}
// TMA load whose completion is signalled on an mbarrier that may live in the PEER CTA of the pair
__device__ __forceinline__ void tma_load_2d_2cta(void* smem, const CUtensorMap* m, uint32_t bar_cluster_addr, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(smem_u32(smem)), "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_cluster_addr), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void mbar_arrive_remote(uint32_t bar_cluster_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(bar_cluster_addr) : "memory");
}
template <uint32_t kCols>
__device__ __forceinline__ void tmem_alloc_2cta(uint32_t* smem_dst) {   // whole warp, in BOTH CTAs of the pair
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "n"(kCols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
template <uint32_t kCols>
__device__ __forceinline__ void tmem_dealloc_2cta(uint32_t taddr) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(kCols) : "memory");
}
__device__ __forceinline__ void umma_bf16_2cta(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
// arrive (once all MMAs issued so far have completed) on the barrier at this offset in BOTH CTAs of the pair
__device__ __forceinline__ void umma_commit_2cta(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"((uint16_t)0b11) : "memory");
}

// m fastest within bands of 8 pair-tiles (2048 rows), so concurrently running clusters share B tiles in L2
__device__ __forceinline__ void tile_coords(int tile, int tiles_m, int tiles_n, int& m_tile, int& n_tile) {
    constexpr int kBand = 8;
    const int per_band = kBand * tiles_n;
    const int b = tile / per_band, r = tile - b * per_band;
    const int rows = min(kBand, tiles_m - b * kBand);
    m_tile = b * kBand + r % rows;
    n_tile = r / rows;
}

// what the epilogue of an instantiation does besides the bf16 TMA store (EPI_ANY: decided at run time from G2Params)
enum : int { EPI_STORE = 0, EPI_ADDEND = 1, EPI_BIAS = 2, EPI_ANY = 7 };

template <int kEpi>
__global__ void __launch_bounds__(kThreads, 1)
gemm2_nt_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                const __grid_constant__ CUtensorMap tmC, const G2Params p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kBarOff);      // used in the leader only
    uint64_t* empty_bar = full_bar + kStages;                               // per CTA
    uint64_t* tmem_full_bar = empty_bar + kStages;                          // [2], per CTA
    uint64_t* tmem_empty_bar = tmem_full_bar + 2;                           // [2], used in the leader only
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const bool leader = rank == 0;
    const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;
    const int tiles_m = (p.M + 2 * BM - 1) / (2 * BM), tiles_n = p.N / BN;
    const int num_tiles = tiles_m * tiles_n;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmA);
        prefetch_tmap(&tmB);
        prefetch_tmap(&tmC);
        for (int s = 0; s < kStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(&tmem_full_bar[a], 1); mbar_init(&tmem_empty_bar[a], 8); }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc_2cta<512>(tmem_ptr);
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();                 // both CTAs' barriers are initialised before anyone signals across the pair
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp == 0) {
        if (lane == 0) {
            uint32_t it = 0;
            for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
                int m_tile, n_tile;
                tile_coords(tile, tiles_m, tiles_n, m_tile, n_tile);
                const int row_a = m_tile * 2 * BM + (int)rank * BM;
                const int row_b = n_tile * BN + (int)rank * (BN / 2);
                for (int kb = 0; kb < p.num_kb; ++kb, ++it) {
                    const int s = it % kStages;
                    const uint32_t ph = (it / kStages) & 1;
                    mbar_wait(&empty_bar[s], ph ^ 1);
                    uint8_t* sA = smem + s * kStageBytes;
                    uint8_t* sB = sA + kABytes;
                    const uint32_t bar = mapa(smem_u32(&full_bar[s]), 0);      // the LEADER's "stage full" barrier
                    if (leader) mbar_arrive_expect_tx(&full_bar[s], 2 * kStageBytes);   // bytes of both CTAs' loads
                    tma_load_2d_2cta(sA, &tmA, bar, kb * BK, row_a);
                    tma_load_2d_2cta(sB, &tmB, bar, kb * BK, row_b);
                }
            }
        }
    } else if (warp == 1) {
        if (leader && lane == 0) {
            constexpr uint32_t idesc = make_idesc_bf16(2 * BM, BN, 0, 0);        // M = 256 across the pair
            uint32_t it = 0, t = 0;
            for (int tile = cluster_id; tile < num_tiles; tile += num_clusters, ++t) {
                const uint32_t acc = t & 1, use = t >> 1;
                mbar_wait(&tmem_empty_bar[acc], (use & 1) ^ 1);                 // both CTAs' epilogues have drained it
                tc_fence_after();
                const uint32_t tm = tmem_base + acc * BN;
                for (int kb = 0; kb < p.num_kb; ++kb, ++it) {
                    const int s = it % kStages;
                    const uint32_t ph = (it / kStages) & 1;
                    mbar_wait(&full_bar[s], ph);
                    tc_fence_after();
                    const uint32_t a_addr = smem_u32(smem + s * kStageBytes);
                    const uint64_t da = make_smem_desc(a_addr, 16, 1024);
                    const uint64_t db = make_smem_desc(a_addr + kABytes, 16, 1024);
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k)
                        umma_bf16_2cta(tm, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
                    umma_commit_2cta(&empty_bar[s]);                            // frees the stage in both CTAs
                }
                umma_commit_2cta(&tmem_full_bar[acc]);                          // accumulator ready in both CTAs
            }
        }
    } else {
        // What the epilogue does besides the bf16 store is fixed per instantiation (kEpi != EPI_ANY), so the per-chunk
        // code is straight-line: with two epilogue warps per scheduler there is nothing to hide a branch behind.
        constexpr bool kAny = kEpi == EPI_ANY;
        const bool has_bias = kAny ? p.bias != nullptr : (kEpi & EPI_BIAS) != 0;
        const bool has_act = kAny && p.act != ACT_NONE;
        const bool has_addend = kAny ? p.addend != nullptr : (kEpi & EPI_ADDEND) != 0;
        const int q = warp & 3;
        const int wg = (warp - 2) >> 2;
        const int bar_id = 1 + wg;
        const bool is_store_thread = threadIdx.x == 64 + wg * 128;
        uint8_t* sC = smem + kCOff + wg * (kCBytes / 2);            // two 64-column slabs, used alternately
        const uint32_t empty_remote0 = mapa(smem_u32(&tmem_empty_bar[0]), 0);
        const uint32_t empty_remote1 = mapa(smem_u32(&tmem_empty_bar[1]), 0);
        const int N = p.N;
        uint32_t t = 0, stores = 0;
        for (int tile = cluster_id; tile < num_tiles; tile += num_clusters, ++t) {
            if ((int)(t & 1) != wg) continue;
            int m_tile, n_tile;
            tile_coords(tile, tiles_m, tiles_n, m_tile, n_tile);
            const uint32_t acc = t & 1, use = t >> 1;
            mbar_wait(&tmem_full_bar[acc], use & 1);
            tc_fence_after();
            const uint32_t tm = tmem_base + acc * BN + ((uint32_t)(q * 32) << 16);
            const int row0 = m_tile * 2 * BM + (int)rank * BM;
            const int r_in = q * 32 + lane;
            const int row = row0 + r_in;
            const bool row_ok = row < p.M;
            // two register buffers: the TMEM load of chunk c+1 is in flight while chunk c is converted and stored
            uint32_t rbuf[2][32];
            tmem_ld_32x32(tm, rbuf[0]);
#pragma unroll
            for (int c = 0; c < BN / 32; ++c) {
                tmem_ld_wait();
                if (c + 1 < BN / 32) tmem_ld_32x32(tm + (c + 1) * 32, rbuf[(c + 1) & 1]);
                else {                                    // whole accumulator half is in registers: hand it back
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) {
                        if (leader) mbar_arrive(&tmem_empty_bar[acc]);
                        else mbar_arrive_remote(acc ? empty_remote1 : empty_remote0);
                    }
                }
                const uint32_t* r = rbuf[c & 1];
                const int col0 = n_tile * BN + c * 32;
                float v[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
                if (has_bias) {
                    if (kAny) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] += __ldg(p.bias + col0 + j);
                    } else {                                 // 16-byte aligned bias (checked by the launcher)
                        const float4* b4 = reinterpret_cast<const float4*>(p.bias + col0);
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const float4 b = __ldg(b4 + j);
                            v[4 * j] += b.x; v[4 * j + 1] += b.y; v[4 * j + 2] += b.z; v[4 * j + 3] += b.w;
                        }
                    }
                }
                if (has_act) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = act_fwd(v[j], p.act);
                }
                if (has_addend && row_ok) {
                    const __nv_bfloat16* ad = p.addend + (size_t)row * N + col0;
#pragma unroll
                    for (int j = 0; j < 32; j += 8) {
                        float f[8];
                        unpack8(ld8(ad + j), f);
#pragma unroll
                        for (int u = 0; u < 8; ++u) v[j + u] += f[u];
                    }
                }
                const int slab = c >> 1;                               // 64-column slab of the tile (0..3)
                uint8_t* atom = sC + ((stores + slab) & 1) * 16384;
#pragma unroll
                for (int j = 0; j < 32; j += 8) {
                    const int col = (c & 1) * 32 + j;
                    const uint32_t off = (uint32_t)(r_in * 128 + ((((col >> 3) ^ (r_in & 7)) << 4)));
                    *reinterpret_cast<uint4*>(atom + off) = pack8(v + j).u;
                }
                if (c & 1) {
                    // the slab is complete: store it now so that its drain overlaps the conversion of the next one.  Before
                    // the barrier the store thread makes sure the PREVIOUS slab's store has been read out of shared
                    // memory — that is the buffer the next slab is written to.
                    fence_proxy_async();
                    if (is_store_thread && stores + slab > 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
                    asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
                    if (is_store_thread) {
                        tma_store_2d(&tmC, atom, n_tile * BN + slab * 64, row0);
                        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                    }
                }
            }
            stores += BN / 64;
        }
        if (is_store_thread) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();                 // the peer's shared memory / tensor memory stay alive until the pair is done
    if (warp == 1) tmem_dealloc_2cta<512>(tmem_base);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn get_encode() {
    static thread_local bool ctx_ready = false;     // the encode call is a driver-API function: bind the primary context first
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
bool encode_2d(CUtensorMap* m, const void* ptr, uint64_t inner, uint64_t outer, uint32_t box_inner, uint32_t box_outer) {
    if (!get_encode()) return false;
    cuuint64_t dims[2] = {inner, outer};
    cuuint64_t strides[1] = {inner * 2};
    cuuint32_t box[2] = {box_inner, box_outer};
    cuuint32_t estr[2] = {1, 1};
    return get_encode()(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

}  // namespace

// out[M,N] bf16 = act(A[M,K] · B[N,K]ᵀ + bias) [+ addend]; K % 64 == 0, N % 256 == 0.  Launched as clusters of 2 CTAs
// (cudaLaunchAttributeClusterDimension), one cluster per 256 x 256 tile in flight, at most one CTA per SM.
FDT_API int fdt_gemm2_nt(const void* A, const void* B, void* out_bf16, const float* bias, const void* addend, int M, int N, int K,
                         int act, cudaStream_t st) {
    if (M <= 0 || N % BN || K % BK || K <= 0 || !out_bf16) return -1;
    CUtensorMap tmA, tmB, tmC;
    if (!encode_2d(&tmA, A, K, M, BK, BM) || !encode_2d(&tmB, B, K, N, BK, BN / 2) || !encode_2d(&tmC, out_bf16, N, M, 64, BM))
        return -4;
    G2Params p{};
    p.M = M; p.N = N; p.num_kb = K / BK; p.out = (__nv_bfloat16*)out_bf16; p.bias = bias;
    p.addend = (const __nv_bfloat16*)addend; p.act = act;
    static int num_sms = 0;
    if (!num_sms) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev); }
    const int num_tiles = ((M + 2 * BM - 1) / (2 * BM)) * (N / BN);
    int clusters = num_sms / 2;
    if (clusters > num_tiles) clusters = num_tiles;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * clusters); cfg.blockDim = dim3(kThreads); cfg.dynamicSmemBytes = kSmemTotal; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    // the common epilogues (plain store, + bias, + residual addend) have their own straight-line instantiations
    void (*kern)(CUtensorMap, CUtensorMap, CUtensorMap, G2Params) = gemm2_nt_kernel<EPI_ANY>;
    if (act == ACT_NONE && !(bias && addend)) {
        if (bias) { if ((reinterpret_cast<uintptr_t>(bias) & 15) == 0) kern = gemm2_nt_kernel<EPI_BIAS>; }
        else if (addend) kern = gemm2_nt_kernel<EPI_ADDEND>;
        else kern = gemm2_nt_kernel<EPI_STORE>;
    }
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemTotal);
    if (e != cudaSuccess) return (int)e;
    e = cudaLaunchKernelEx(&cfg, kern, tmA, tmB, tmC, p);
    if (e != cudaSuccess) return (int)e;
    FDT_CHECK_LAUNCH();
    return 0;
}
