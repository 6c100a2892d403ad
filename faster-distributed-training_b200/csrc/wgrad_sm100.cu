// Weight-gradient GEMM on tcgen05 with MN-major operands and split-K.
//
//   dW[co, tap, ci] = Σ_p dY[p, co] · X[p ⊕ tap, ci]          (p = pixel / token index, fp32 result)
//
// The reduction runs over pixels, and both operands are stored pixel-major with the channel index
// contiguous (NHWC) — i.e. the *M/N* index is the contiguous one.  tcgen05.mma reads such tiles directly
// (instruction-descriptor bits a_major = b_major = MN), so no transposed copy of the activations is ever
// made: TMA drops {64 pixels x 64 channels} boxes (128-byte rows, 128B swizzle) into shared memory and
// the matrix descriptors walk them 16 pixel-rows (one MMA K-step) at a time.  For a 3x3 filter the X box
// of tap (r, s) is the 4-D NHWC box shifted by (s - pad, r - pad); the hardware zero-fills the halo.
//
// The pixel range is split over blockIdx.y (split-K) so that also a 64x64x9 weight fills the GPU; every
// CTA adds its 128 x BN fp32 partial tile into dW (physical layout [Cout][kh][kw][Cin], the layout the
// flat gradient buffer uses) with vectorised red.global.add.
//
// Parity: the dW half of convolution_backward (resnet.py:21-34, a cuDNN call on transposed tensors) and
// the `grad_out_transpose @ activated` products of MLPScratch / nn.Linear (transformer.py:318-334).
#include "common.cuh"
#include "sm100.cuh"
#include <cstdlib>
#include <mutex>

using namespace fdt;
using namespace sm100;

namespace {

constexpr int BM = 128, BKP = 64 /* pixels per stage */, kThreads = 192;

struct WgradParams {
    int Cout, Cin, taps, kw, pad;
    int num_kb;            // total pixel blocks (P / 64)
    int kb_per_split;
    float* dw;             // [Cout][taps][Cin] fp32, accumulated into (local, peer or multicast address)
    float alpha;           // scale applied to the tile before it is added (1/world for a fused all-reduce)
    int multicast;         // 1: dw is a multicast (NVLS) address → multimem.red adds into EVERY rank's buffer
    float* acc;            // multicast + split-K: local fp32 staging ([Cout][taps][Cin], all zero between launches)
    int* counters;         // multicast + split-K: per-tile arrival counters (all zero between launches)
    int splits;
    int H, W, bh, bn;      // conv geometry of the 64-pixel box, on the grid of OUTPUT pixels (= dY's pixels)
    int in_mul;            // convolution stride: output row oh reads input row in_mul * oh + r - pad (element strides in tmX)
    int tiles_n;           // Cin tiles
};

template <int BN, int kStages>
struct WSmem {
    static constexpr int kABytes = BM * BKP * 2;       // two 8 KB atoms (64 px x 64 co)
    static constexpr int kBBytes = BN * BKP * 2;
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kBarOff = kStages * kStageBytes;
    static constexpr int kTotal = kBarOff + (2 * kStages + 1) * 8 + 16 + 1024;
};

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// In-switch all-reduce of a finished tile: one instruction adds 4 floats into the same offset of every
// GPU's gradient buffer (the address is a multicast mapping of the symmetric heap).
__device__ __forceinline__ void mc_red_add_v4(float* addr, float a, float b, float c, float d) {
    asm volatile("multimem.red.relaxed.sys.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d)
                 : "memory");
}

template <int BN, int kStages, bool kConv>
__global__ void __launch_bounds__(kThreads, 2)
wgrad_kernel(const __grid_constant__ CUtensorMap tmDY, const __grid_constant__ CUtensorMap tmX, const WgradParams p) {
    using L = WSmem<BN, kStages>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::kBarOff);
    uint64_t* empty_bar = full_bar + kStages;
    uint64_t* tmem_full_bar = empty_bar + kStages;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // blockIdx.x = (m_tile * tiles_n + n_tile) * taps + tap ; blockIdx.y = split
    const int tap = blockIdx.x % p.taps;
    const int mn = blockIdx.x / p.taps;
    const int n_tile = mn % p.tiles_n, m_tile = mn / p.tiles_n;
    const int kb0 = blockIdx.y * p.kb_per_split;
    const int kb1 = min(p.num_kb, kb0 + p.kb_per_split);
    const int nkb = kb1 - kb0;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmDY);
        prefetch_tmap(&tmX);
        for (int s = 0; s < kStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        mbar_init(tmem_full_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<BN>(tmem_ptr);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;
    pdl_wait();
    pdl_launch_dependents();

    if (nkb > 0) {
        if (warp == 0) {
            if (lane == 0) {
                const int r = tap / p.kw, sx = tap - r * p.kw;
                for (int i = 0; i < nkb; ++i) {
                    const int kb = kb0 + i;
                    const int s = i % kStages;
                    const uint32_t ph = (i / kStages) & 1;
                    mbar_wait(&empty_bar[s], ph ^ 1);
                    uint8_t* sA = smem + s * L::kStageBytes;
                    uint8_t* sB = sA + L::kABytes;
                    mbar_arrive_expect_tx(&full_bar[s], L::kStageBytes);
                    // A: dY[p0 .. p0+64, co0 .. co0+128] as two 64-channel atoms
                    tma_load_2d(sA, &tmDY, &full_bar[s], m_tile * BM, kb * BKP);
                    tma_load_2d(sA + 8192, &tmDY, &full_bar[s], m_tile * BM + 64, kb * BKP);
                    // B: X (shifted by the tap) [64 px, ci0 .. ci0+BN]
                    if (kConv) {
                        int n0, h0;
                        if (p.bn == 1) { const int tpi = p.H / p.bh; n0 = kb / tpi; h0 = (kb % tpi) * p.bh; }
                        else { n0 = kb * p.bn; h0 = 0; }
#pragma unroll
                        for (int j = 0; j < BN / 64; ++j)
                            tma_load_4d(sB + j * 8192, &tmX, &full_bar[s], n_tile * BN + j * 64, sx - p.pad, p.in_mul * h0 + r - p.pad, n0);
                    } else {
#pragma unroll
                        for (int j = 0; j < BN / 64; ++j)
                            tma_load_2d(sB + j * 8192, &tmX, &full_bar[s], n_tile * BN + j * 64, kb * BKP);
                    }
                }
            }
        } else if (warp == 1) {
            if (lane == 0) {
                constexpr uint32_t idesc = make_idesc_bf16(BM, BN, 1, 1);      // both operands MN-major
                for (int i = 0; i < nkb; ++i) {
                    const int s = i % kStages;
                    const uint32_t ph = (i / kStages) & 1;
                    mbar_wait(&full_bar[s], ph);
                    tc_fence_after();
                    const uint32_t a_addr = smem_u32(smem + s * L::kStageBytes);
                    const uint32_t b_addr = a_addr + L::kABytes;
                    // MN-major, 128B swizzle: 8-pixel-row groups 1024 B apart (SBO), 64-channel atoms 8192 B apart (LBO)
                    const uint64_t da = make_smem_desc(a_addr, 8192, 1024);
                    const uint64_t db = make_smem_desc(b_addr, 8192, 1024);
#pragma unroll
                    for (int k = 0; k < BKP / 16; ++k)     // 16 pixel rows = 2048 B = 128 in the >>4 address field
                        umma_bf16(tmem_base, da + 128 * k, db + 128 * k, idesc, (i | k) != 0);
                    umma_commit(&empty_bar[s]);
                }
                umma_commit(tmem_full_bar);
            }
        } else {
            mbar_wait(tmem_full_bar, 0);
            tc_fence_after();
            const int q = warp & 3;
            // The partial tile goes TMEM → registers → shared memory (the pipeline stages are free: every MMA, hence
            // every TMA load, has completed) so that each warp can add whole 512-byte row segments with one
            // red.global.add.v4 per lane: full-line atomics instead of 32 scattered 16-byte ones per instruction.
            constexpr int kPitch = BN + 4;                      // floats; +16 B keeps the 128-bit row writes conflict-free
            float* stg = reinterpret_cast<float*>(smem);
            fence_proxy_async();
            const int rloc = q * 32 + lane;
            const float al = p.multicast ? 1.f : p.alpha;       // multicast: alpha is applied when the finished tile is pushed
#pragma unroll 1
            for (int c = 0; c < BN / 32; ++c) {
                uint32_t r[32];
                tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + c * 32, r);
                tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 32; j += 4)
                    *reinterpret_cast<float4*>(stg + rloc * kPitch + c * 32 + j) =
                        make_float4(al * __uint_as_float(r[j]), al * __uint_as_float(r[j + 1]), al * __uint_as_float(r[j + 2]),
                                    al * __uint_as_float(r[j + 3]));
            }
            __syncwarp();
            {
                constexpr int kF4 = BN / 4;                     // float4 per tile row: 32 (BN = 128) or 16 (BN = 64)
                constexpr int kRows = 32 / kF4;                 // rows covered by one warp instruction
                float* base = p.multicast ? p.acc : p.dw;       // multicast: split-K partials meet in the LOCAL staging buffer
                const int c4 = lane % kF4;
#pragma unroll 4
                for (int rr = 0; rr < 32; rr += kRows) {
                    const int row = q * 32 + rr + lane / kF4;
                    const int co = m_tile * BM + row;
                    if (co < p.Cout) {
                        const float4 v = *reinterpret_cast<const float4*>(stg + row * kPitch + c4 * 4);
                        red_add_v4(base + ((size_t)co * p.taps + tap) * p.Cin + n_tile * BN + c4 * 4, v.x, v.y, v.z, v.w);
                    }
                }
            }
            if (p.multicast) {
                // "last split in" detection: the CTA that completes the tile pushes it to every rank with
                // one multimem.red per 16 bytes (in-switch all-reduce), then re-zeroes the staging tile.
                __shared__ int s_last;
                __threadfence();
                asm volatile("bar.sync 1, 128;" ::: "memory");          // the 4 epilogue warps
                if (threadIdx.x == 64) {
                    const int old = atomicAdd(p.counters + blockIdx.x, 1);
                    s_last = (old == p.splits - 1);
                    if (s_last) p.counters[blockIdx.x] = 0;
                }
                asm volatile("bar.sync 1, 128;" ::: "memory");
                if (s_last) {
                    __threadfence();
                    // coalesced push: one warp per output row, 32 lanes x 16 B = 512 contiguous bytes per instruction
                    const float al = p.alpha;
                    for (int rr = 0; rr < 32; ++rr) {
                        const int co2 = m_tile * BM + q * 32 + rr;
                        if (co2 >= p.Cout) break;
                        const size_t off = ((size_t)co2 * p.taps + tap) * p.Cin + n_tile * BN;
                        for (int j = lane * 4; j < BN; j += 128) {
                            float4 v = __ldcg(reinterpret_cast<const float4*>(p.acc + off + j));
                            __stcg(reinterpret_cast<float4*>(p.acc + off + j), make_float4(0.f, 0.f, 0.f, 0.f));
                            mc_red_add_v4(p.dw + off + j, al * v.x, al * v.y, al * v.z, al * v.w);
                        }
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<BN>(tmem_base);
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

bool encode_pc(CUtensorMap* m, const void* ptr, uint64_t C, uint64_t P) {   // [P, C] bf16, box 64 ch x 64 px
    if (!get_encode()) return false;
    cuuint64_t dims[2] = {C, P};
    cuuint64_t strides[1] = {C * 2};
    cuuint32_t box[2] = {64, 64};
    cuuint32_t estr[2] = {1, 1};
    return get_encode()(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}
bool encode_nhwc64(CUtensorMap* m, const void* ptr, int Nb, int H, int W, int C, int bw, int bh, int bn, int es = 1) {
    if (!get_encode()) return false;
    cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)Nb};
    cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
    cuuint32_t box[4] = {64, (cuuint32_t)(bw * es), (cuuint32_t)(bh * es), (cuuint32_t)bn};   // every es-th pixel: bw x bh pixels
    cuuint32_t estr[4] = {1, (cuuint32_t)es, (cuuint32_t)es, 1};
    return get_encode()(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(ptr), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// Staging for the fused all-reduce (multicast destination + split-K): one fp32 buffer as large as the
// largest weight seen so far and a tile-counter array, per device.  Both are all-zero between launches
// (the kernels clean up after themselves), so launches on one stream can share them.  Allocated on
// first use / growth — outside CUDA-graph capture, because the warm-up steps run eagerly first.
struct Staging { float* acc = nullptr; size_t elems = 0; int* counters = nullptr; int ncounters = 0; };
Staging g_staging[64];

int get_staging(WgradParams& p, int tiles) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return -5;
    Staging& s = g_staging[dev];
    const size_t need = (size_t)p.Cout * p.taps * p.Cin;
    if (s.elems < need) {
        if (s.acc) cudaFree(s.acc);
        if (cudaMalloc(&s.acc, need * sizeof(float)) != cudaSuccess) return -6;
        cudaMemset(s.acc, 0, need * sizeof(float));
        s.elems = need;
    }
    if (s.ncounters < tiles) {
        if (s.counters) cudaFree(s.counters);
        const int n = tiles < 4096 ? 4096 : tiles;
        if (cudaMalloc(&s.counters, n * sizeof(int)) != cudaSuccess) return -6;
        cudaMemset(s.counters, 0, n * sizeof(int));
        s.ncounters = n;
    }
    p.acc = s.acc;
    p.counters = s.counters;
    return 0;
}

template <int BN, bool kConv>
int launch(const CUtensorMap& a, const CUtensorMap& b, WgradParams p, cudaStream_t st) {
    constexpr int kStages = BN == 128 ? 3 : 4;             // <= 97 KB of shared memory: two CTAs per SM, so one CTA's prologue /
                                                           // epilogue overlaps the other's main loop
    using L = WSmem<BN, kStages>;
    auto kern = wgrad_kernel<BN, kStages, kConv>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotal);
    if (e != cudaSuccess) return (int)e;
    const int tiles_m = (p.Cout + BM - 1) / BM;
    p.tiles_n = p.Cin / BN;
    const int tiles = tiles_m * p.tiles_n * p.taps;
    static int target = 0;                                 // CTAs per launch: one resident wave (2 CTAs per SM)
    if (target == 0) {
        const char* e = getenv("FDT_WGRAD_CTAS");
        target = e ? atoi(e) : 296;
        if (target < 1) target = 296;
    }
    int splits = target / tiles;                           // round down: a few CTAs over one wave would cost a whole extra round
    if (splits > p.num_kb) splits = p.num_kb;
    if (splits < 1) splits = 1;
    p.kb_per_split = (p.num_kb + splits - 1) / splits;
    splits = (p.num_kb + p.kb_per_split - 1) / p.kb_per_split;
    p.splits = splits;
    if (p.multicast) {
        int rc = get_staging(p, tiles);
        if (rc) return rc;
    }
    if (fdt::launch(kern, dim3(tiles, splits), dim3(kThreads), L::kTotal, st, a, b, p) != cudaSuccess) return (int)cudaGetLastError();
    FDT_CHECK_LAUNCH();
    return 0;
}

}  // namespace

// dw[Cout][k][k][Cin] (fp32) += dY[Nb,OH,OW,Cout]ᵀ ⊛ X[Nb,H,W,Cin]; stride 1 or 2, NHWC bf16, OH = (H + 2 pad - k)/stride + 1.
// Cin % 64 == 0; OW a power of two <= 64; (Nb·OH·OW) % 64 == 0.  For stride 2 the X tensor map walks every second pixel.
// alpha scales the result; multicast != 0 means `dw` is an NVLS multicast address and the accumulation is a
// fused all-reduce (every rank's buffer receives alpha·tile).
FDT_API int fdt_conv_wgrad_ex(const void* dy, const void* x, float* dw, int Nb, int H, int W, int Cin, int Cout, int ksize,
                              int pad, int stride, float alpha, int multicast, cudaStream_t st) {
    if (stride != 1 && stride != 2) return -1;
    const int OH = (H + 2 * pad - ksize) / stride + 1, OW = (W + 2 * pad - ksize) / stride + 1;
    if (Cin % 64 || Cout % 8 || OW > 64 || OW < 1 || (OW & (OW - 1)) || (64 % OW)) return -1;
    if (stride == 2 && ((H & 1) || (W & 1))) return -1;
    int bh = 64 / OW, bn = 1;
    if (bh > OH) { if (bh % OH) return -1; bn = bh / OH; bh = OH; }
    else if (OH % bh) return -1;
    const int64_t P = (int64_t)Nb * OH * OW;
    // bn > 1 (maps smaller than 64 pixels): any batch size — a k-block is bn whole images, and the images / rows its two boxes
    // reach beyond the batch are zero-filled by the TMA unit
    if (bn == 1 && P % 64) return -1;
    CUtensorMap tmDY, tmX;
    if (!encode_pc(&tmDY, dy, Cout, P)) return -4;
    WgradParams p{};
    p.Cout = Cout; p.Cin = Cin; p.taps = ksize * ksize; p.kw = ksize; p.pad = pad; p.num_kb = (int)((P + 63) / 64); p.dw = dw;
    p.H = OH; p.W = OW; p.bh = bh; p.bn = bn; p.in_mul = stride; p.alpha = alpha; p.multicast = multicast;
    const bool wide = (Cin % 128 == 0);
    if (ksize == 1 && stride == 1) {
        if (!encode_pc(&tmX, x, Cin, P)) return -4;
        return wide ? launch<128, false>(tmDY, tmX, p, st) : launch<64, false>(tmDY, tmX, p, st);
    }
    if (!encode_nhwc64(&tmX, x, Nb, H, W, Cin, OW, bh, bn, stride)) return -4;
    return wide ? launch<128, true>(tmDY, tmX, p, st) : launch<64, true>(tmDY, tmX, p, st);
}

FDT_API int fdt_conv_wgrad(const void* dy, const void* x, float* dw, int Nb, int H, int W, int Cin, int Cout, int ksize,
                           int pad, float alpha, int multicast, cudaStream_t st) {
    return fdt_conv_wgrad_ex(dy, x, dw, Nb, H, W, Cin, Cout, ksize, pad, 1, alpha, multicast, st);
}

// dw[N][K] (fp32) += dY[M,N]ᵀ · X[M,K]   (linear layers).  K % 64 == 0, N % 8 == 0; any M (the rows a 64-row box reaches
// beyond M are zero-filled by the TMA unit and contribute nothing).
FDT_API int fdt_linear_wgrad(const void* dy, const void* x, float* dw, int M, int N, int K, float alpha, int multicast,
                             cudaStream_t st) {
    if (K % 64 || N % 8 || M < 1) return -1;
    CUtensorMap tmDY, tmX;
    if (!encode_pc(&tmDY, dy, N, M) || !encode_pc(&tmX, x, K, M)) return -4;
    WgradParams p{};
    p.Cout = N; p.Cin = K; p.taps = 1; p.kw = 1; p.pad = 0; p.num_kb = (M + 63) / 64; p.dw = dw;
    p.alpha = alpha; p.multicast = multicast;
    return (K % 128 == 0) ? launch<128, false>(tmDY, tmX, p, st) : launch<64, false>(tmDY, tmX, p, st);
}
