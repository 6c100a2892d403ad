// Data-gradient operand layouts of every weight in ONE launch per step.
//
// dX of a convolution / Linear is a GEMM whose B operand is the transposed weight.  The reference leaves that to
// cuDNN / cuBLAS per layer; a tcgen05 GEMM wants it K-major in memory, which used to cost a
// flip + transpose + copy per layer inside the backward pass.  Here all of them are produced right after the bf16
// shadow cast, from a descriptor table built once:
//   mode 0   src [O][T][I]  →  dst [I][T-1-t][O]     (implicit-GEMM dgrad: taps rotated by 180°; T = 1 for Linear / 1x1)
//   mode 1   src [O][T][I]  →  dst [T][I][O]         (explicit col2im dgrad: the [K, O] matrix)
#include "common.cuh"

using namespace fdt;
typedef __nv_bfloat16 bf16;

namespace {

struct WPrepDesc {              // mirrored in fdt/parallel/flat.py
    const bf16* src;
    bf16* dst;
    int O, T, I, mode;
};
static_assert(sizeof(WPrepDesc) == 32, "descriptor layout is mirrored in python");

constexpr int kTile = 64;

// tiles first, first + step, ... of weight d (a tile = 64 x 64 elements of one tap, transposed through shared memory)
__device__ __forceinline__ void dgrad_weights_body(const WPrepDesc& d, int first, int step) {
    __shared__ bf16 tile[kTile][kTile + 2];
    const int to = (d.O + kTile - 1) / kTile, ti = (d.I + kTile - 1) / kTile;
    const int total = d.T * to * ti;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;     // 32 x 8
    for (int id = first; id < total; id += step) {
        const int t = id / (to * ti), rem = id - t * to * ti;
        const int o0 = (rem / ti) * kTile, i0 = (rem % ti) * kTile;
        // load: rows = o, 2 consecutive i per thread
#pragma unroll
        for (int r = ty; r < kTile; r += 8) {
            const int o = o0 + r, i = i0 + 2 * tx;
            bf16 a = __float2bfloat16(0.f), b = a;
            if (o < d.O) {
                const bf16* s = d.src + ((size_t)o * d.T + t) * d.I + i;
                if (i + 1 < d.I && (d.I & 1) == 0) { const __nv_bfloat162 v = *reinterpret_cast<const __nv_bfloat162*>(s); a = v.x; b = v.y; }
                else { if (i < d.I) a = s[0]; if (i + 1 < d.I) b = s[1]; }
            }
            tile[r][2 * tx] = a; tile[r][2 * tx + 1] = b;
        }
        __syncthreads();
        // store: rows = i, 2 consecutive o per thread
        const int tt = d.mode == 0 ? d.T - 1 - t : t;
#pragma unroll
        for (int r = ty; r < kTile; r += 8) {
            const int i = i0 + r, o = o0 + 2 * tx;
            if (i < d.I && o < d.O) {
                bf16* q = d.dst + (d.mode == 0 ? ((size_t)i * d.T + tt) * d.O : ((size_t)tt * d.I + i) * d.O) + o;
                if (o + 1 < d.O && (d.O & 1) == 0) {
                    __nv_bfloat162 v; v.x = tile[2 * tx][r]; v.y = tile[2 * tx + 1][r];
                    *reinterpret_cast<__nv_bfloat162*>(q) = v;
                } else {
                    q[0] = tile[2 * tx][r];
                    if (o + 1 < d.O) q[1] = tile[2 * tx + 1][r];
                }
            }
        }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(256)
dgrad_weights_kernel(const WPrepDesc* __restrict__ descs) {
    dgrad_weights_body(descs[blockIdx.y], blockIdx.x, gridDim.x);
}

// Balanced form: one tile per block over the concatenated tile lists of all weights (a fixed number of blocks per weight
// leaves the 512x512x3x3 filters with 12 tiles per block and the 64x64 ones with idle blocks).  The (weight, tile) pair of
// every block comes from a table the caller built with the descriptors — a search through the descriptor list would be a
// chain of dependent loads as long as the list.
__global__ void __launch_bounds__(256)
dgrad_weights_map_kernel(const WPrepDesc* __restrict__ descs, const int2* __restrict__ tile_map) {
    const int2 m = tile_map[blockIdx.x];
    dgrad_weights_body(descs[m.x], m.y, 1 << 30);
}

}  // namespace

// descs: `count` 32-byte descriptors {src, dst, O, T, I, mode} in device memory.
FDT_API int fdt_dgrad_weights(const void* descs, int count, cudaStream_t st) {
    if (count <= 0) return 0;
    dgrad_weights_kernel<<<dim3(48, count), 256, 0, st>>>((const WPrepDesc*)descs);
    FDT_CHECK_LAUNCH();
    return 0;
}

// tile_map: total_tiles x {descriptor index, tile index inside that weight} (int32 pairs, device memory)
FDT_API int fdt_dgrad_weights_map(const void* descs, const void* tile_map, int total_tiles, cudaStream_t st) {
    if (total_tiles <= 0) return 0;
    dgrad_weights_map_kernel<<<total_tiles, 256, 0, st>>>((const WPrepDesc*)descs, (const int2*)tile_map);
    FDT_CHECK_LAUNCH();
    return 0;
}
