// In-kernel collectives over NVLink peer memory (symmetric heap), fused with the adjacent
// element-wise work of the training step.
//
// Every rank maps every peer's copy of the flat buffer (torch symmetric memory does the CUDA-VMM
// handle exchange; see fdt/parallel/symm.py).  The kernels below read and write those peer
// addresses directly — P2P ld/st through NVSwitch, or `multimem` (in-switch reduction / multicast,
// NVLS) when a multicast mapping exists — so a collective is ONE launch with the scaling, the
// gradient-norm partial and the bf16 operand cast folded into the same pass:
//
//   fdt_comm_all_reduce_mean   DDP gradients: two-shot all-reduce; rank r reduces chunk r across all
//                              peers, scales by 1/world, accumulates Σg² (for clipping) and stores the
//                              result into every peer's buffer.
//   fdt_comm_reduce_scatter_mean  FSDP gradients: rank r reduces its own region only (+ scale + Σg²,
//                              the Σ is made global through peer atomics — the norm the reference gets
//                              wrong under FSDP, transformer_test.py:269).
//   fdt_comm_all_gather_cast   FSDP parameters: pull every peer's region into the local fp32 buffer and
//                              emit the bf16 operand shadow in the same pass.
//
// Parity: DDP's bucketed NCCL all-reduce (resnet50_test.py:716), FSDP's all-gather / reduce-scatter
// (transformer_test.py:387-392) and the separate clip_grad_norm_ pass.
//
// Synchronisation: per-block flag barriers in a dedicated symmetric signal buffer
// (system-scope CAS: set my slot on the peer, consume the peer's slot on me).  Block b only talks
// to block b of the peers, so grids must be identical on all ranks and small enough to be
// co-resident (they are: <= 64 blocks).
#include "common.cuh"
#include <cstdlib>
using namespace fdt;

namespace {

constexpr int kMaxWorld = 8;
constexpr int kBlock = 512;

struct CommCtx {
    float* buf[kMaxWorld];        // peer views of the data buffer
    uint32_t* sig[kMaxWorld];     // peer views of the signal buffer
    float* scratch[kMaxWorld];    // peer views of a small fp32 scratch (Σg² slots)
    float* mc;                    // multicast view of the data buffer (NVLS) or null
    int rank, world;
    int sig_off, pad1;            // sig_off: first slot of this context's signal channel (collectives that may be in flight
                                  // at the same time — e.g. on two streams — use disjoint channels)
};

__device__ __forceinline__ void block_barrier(const CommCtx& c, int base) {
    __syncthreads();
    if (threadIdx.x < c.world) {
        __threadfence_system();
        uint32_t* remote = c.sig[threadIdx.x] + c.sig_off + base + c.rank;
        while (atomicCAS_system(remote, 0u, 1u) != 0u) {}
        uint32_t* local = c.sig[c.rank] + c.sig_off + base + threadIdx.x;
        while (atomicCAS_system(local, 1u, 0u) != 1u) {}
        __threadfence_system();
    }
    __syncthreads();
}

__device__ __forceinline__ float4 ld_cg(const float* p) { return __ldcg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ void st_cg(float* p, float4 v) { __stcg(reinterpret_cast<float4*>(p), v); }

__device__ __forceinline__ float4 mc_ld_reduce(const float* p) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void mc_st(float* p, float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}

__device__ __forceinline__ float block_sum512(float v) {
    __shared__ float red[kBlock / 32];
    v = warp_sum(v);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    float t = 0.f;
    if (threadIdx.x < kBlock / 32) t = red[threadIdx.x];
    if (threadIdx.x < 32) t = warp_sum(t);
    return t;      // valid in thread 0
}

// kMode 0: all-reduce (result to every peer); 1: reduce-scatter (result to the local region only).
// [lo, hi) is the element range this rank reduces (multiple of 4 elements, 16-byte aligned).
template <int kMode, bool kMulticast>
__global__ void __launch_bounds__(kBlock)
reduce_kernel(const CommCtx c, int64_t lo, int64_t hi, float scale) {
    const int nb = gridDim.x;
    block_barrier(c, blockIdx.x * c.world);                                        // peers' data is complete
    float sq = 0.f;
    const int64_t stride = (int64_t)nb * kBlock * 4;
    for (int64_t i = lo + ((int64_t)blockIdx.x * kBlock + threadIdx.x) * 4; i < hi; i += stride * 2) {
        const int64_t j = i + stride;
        const bool two = j < hi;
        float4 a0, a1 = make_float4(0.f, 0.f, 0.f, 0.f);
        if (kMulticast) {
            a0 = mc_ld_reduce(c.mc + i);
            if (two) a1 = mc_ld_reduce(c.mc + j);
        } else {
            float4 v0[kMaxWorld], v1[kMaxWorld];
#pragma unroll
            for (int p = 0; p < kMaxWorld; ++p) {
                if (p < c.world) {
                    v0[p] = ld_cg(c.buf[p] + i);
                    if (two) v1[p] = ld_cg(c.buf[p] + j);
                }
            }
            a0 = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int p = 0; p < kMaxWorld; ++p) {
                if (p < c.world) {
                    a0.x += v0[p].x; a0.y += v0[p].y; a0.z += v0[p].z; a0.w += v0[p].w;
                    if (two) { a1.x += v1[p].x; a1.y += v1[p].y; a1.z += v1[p].z; a1.w += v1[p].w; }
                }
            }
        }
        a0.x *= scale; a0.y *= scale; a0.z *= scale; a0.w *= scale;
        a1.x *= scale; a1.y *= scale; a1.z *= scale; a1.w *= scale;
        sq += a0.x * a0.x + a0.y * a0.y + a0.z * a0.z + a0.w * a0.w;
        sq += a1.x * a1.x + a1.y * a1.y + a1.z * a1.z + a1.w * a1.w;
        if (kMode == 1) {
            st_cg(c.buf[c.rank] + i, a0);
            if (two) st_cg(c.buf[c.rank] + j, a1);
        } else if (kMulticast) {
            mc_st(c.mc + i, a0);
            if (two) mc_st(c.mc + j, a1);
        } else {
#pragma unroll
            for (int p = 0; p < kMaxWorld; ++p) {
                if (p < c.world) {
                    st_cg(c.buf[p] + i, a0);
                    if (two) st_cg(c.buf[p] + j, a1);
                }
            }
        }
    }
    // Σg² of my range → every rank's scratch slot (so each rank ends with the global norm)
    sq = block_sum512(sq);
    if (threadIdx.x == 0) {
        for (int p = 0; p < c.world; ++p) atomicAdd_system(c.scratch[p], sq);
    }
    block_barrier(c, (nb + blockIdx.x) * c.world);                                 // all writes have landed
}

// All-gather by PUSH, fused with the bf16 operand cast.  Block b of rank r owns sub-slice b of region r:
// it pushes it to every peer (one multimem.st per 16 bytes when a multicast mapping exists, else one
// P2P store per peer), meets block b of every peer at a flag barrier — at which point sub-slice b of
// EVERY region has landed locally — and then converts exactly those sub-slices to bf16.
template <bool kMulticast>
__global__ void __launch_bounds__(kBlock)
all_gather_cast_kernel(const CommCtx c, int64_t region, __nv_bfloat16* __restrict__ shadow) {
    const int nb = gridDim.x;
    block_barrier(c, blockIdx.x * c.world);               // peers no longer read the stale regions
    const int64_t per = ((region / 4 + nb - 1) / nb) * 4;  // elements of one sub-slice (multiple of 4)
    const int64_t s0 = min(region, per * blockIdx.x), s1 = min(region, s0 + per);
    float* mine = c.buf[c.rank];
    const int64_t base = region * c.rank;
    for (int64_t i = s0 + threadIdx.x * 4; i < s1; i += kBlock * 4 * 4) {
        float4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int64_t k = i + (int64_t)u * kBlock * 4;
            if (k < s1) v[u] = ld_cg(mine + base + k);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int64_t k = i + (int64_t)u * kBlock * 4;
            if (k < s1) {
                if (kMulticast) mc_st(c.mc + base + k, v[u]);
                else {
#pragma unroll
                    for (int p = 0; p < kMaxWorld; ++p)
                        if (p < c.world && p != c.rank) st_cg(c.buf[p] + base + k, v[u]);
                }
            }
        }
    }
    block_barrier(c, (nb + blockIdx.x) * c.world);        // sub-slice b of every region is here
    if (shadow) {
        for (int r = 0; r < c.world; ++r) {
            const int64_t rb = region * r;
            for (int64_t i = s0 + threadIdx.x * 4; i < s1; i += kBlock * 4) {
                float4 v = ld_cg(mine + rb + i);
                __nv_bfloat162 h0 = __floats2bfloat162_rn(v.x, v.y), h1 = __floats2bfloat162_rn(v.z, v.w);
                uint2 pk = make_uint2(*reinterpret_cast<uint32_t*>(&h0), *reinterpret_cast<uint32_t*>(&h1));
                *reinterpret_cast<uint2*>(shadow + rb + i) = pk;
            }
        }
    }
}

// Pull variant (the default): block b reads sub-slice b of every PEER's region straight out of the peer's memory
// (P2P loads, 8 x 16 B in flight per thread), writes it into the local fp32 buffer and emits the bf16 shadow in the
// same pass — one remote read, no second sweep over the gathered buffer.  Measured against the push kernel above
// (multimem.st sustains ~200 GB/s per GPU): see profiles/r1_collective_sweep_*.
__global__ void __launch_bounds__(kBlock)
all_gather_pull_cast_kernel(const CommCtx c, int64_t region, __nv_bfloat16* __restrict__ shadow) {
    const int nb = gridDim.x;
    block_barrier(c, blockIdx.x * c.world);               // every rank's region is final (its optimizer step is done)
    const int64_t per = ((region / 4 + nb - 1) / nb) * 4;
    const int64_t s0 = min(region, per * blockIdx.x), s1 = min(region, s0 + per);
    float* mine = c.buf[c.rank];
    constexpr int U = 8;
    for (int k = 0; k < c.world; ++k) {
        const int r = (c.rank + k) % c.world;             // own region first, then the peers in a rotated order
        const float* src = c.buf[r];
        const int64_t rb = region * r;
        for (int64_t i = s0 + threadIdx.x * 4; i < s1; i += (int64_t)kBlock * 4 * U) {
            float4 v[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int64_t e = i + (int64_t)u * kBlock * 4;
                if (e < s1) v[u] = ld_cg(src + rb + e);
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int64_t e = i + (int64_t)u * kBlock * 4;
                if (e < s1) {
                    if (r != c.rank) st_cg(mine + rb + e, v[u]);
                    if (shadow) {
                        __nv_bfloat162 h0 = __floats2bfloat162_rn(v[u].x, v[u].y), h1 = __floats2bfloat162_rn(v[u].z, v[u].w);
                        *reinterpret_cast<uint2*>(shadow + rb + e) =
                            make_uint2(*reinterpret_cast<uint32_t*>(&h0), *reinterpret_cast<uint32_t*>(&h1));
                    }
                }
            }
        }
    }
    block_barrier(c, (nb + blockIdx.x) * c.world);        // nobody still reads my region when I go on to overwrite it
}

// The Σg² slot is complete only when the whole collective kernel has finished on this rank (every
// local block has passed its end barrier).  This 1-thread kernel, stream-ordered right after it, moves
// the value to ordinary memory and re-arms the slot for the next call.
__global__ void take_scalar_kernel(float* slot, float* out) {
    *out = *slot;
    *slot = 0.f;
}

__global__ void barrier_kernel(const CommCtx c) { block_barrier(c, blockIdx.x * c.world); }

// dst_mc[i] += alpha * src[i] on EVERY rank (multimem.red); for gradients produced outside the fused wgrad kernel
__global__ void __launch_bounds__(256)
mc_red_add_kernel(float* __restrict__ dst_mc, const float* __restrict__ src, int64_t n, float alpha) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x * 4;
    for (int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4; i < n; i += stride) {
        float4 v = *reinterpret_cast<const float4*>(src + i);
        asm volatile("multimem.red.relaxed.sys.global.add.v4.f32 [%0], {%1, %2, %3, %4};"
                     ::"l"(dst_mc + i), "f"(alpha * v.x), "f"(alpha * v.y), "f"(alpha * v.z), "f"(alpha * v.w) : "memory");
    }
}

int pick_blocks(int64_t elems) {
    int64_t b = (elems / 4 + kBlock * 2 - 1) / (kBlock * 2);
    if (b < 1) b = 1;
    if (b > 64) b = 64;
    return (int)b;
}

}  // namespace

// ctx: host pointer to a CommCtx (see fdt/parallel/symm.py for the mirrored ctypes layout).
// `blocks` must be the same on every rank (0 = derive from n, which is the same everywhere).
FDT_API int fdt_comm_all_reduce_mean(const void* ctx, int64_t n, int use_multicast, cudaStream_t st) {
    CommCtx c = *static_cast<const CommCtx*>(ctx);
    if (c.world < 2 || c.world > kMaxWorld || n % 4) return -1;
    int64_t chunk = ((n / 4 + c.world - 1) / c.world) * 4;
    const int64_t lo = min(n, chunk * c.rank), hi = min(n, lo + chunk);
    const int blocks = pick_blocks(chunk);
    const float scale = 1.f / c.world;
    if (use_multicast && c.mc) reduce_kernel<0, true><<<blocks, kBlock, 0, st>>>(c, lo, hi, scale);
    else reduce_kernel<0, false><<<blocks, kBlock, 0, st>>>(c, lo, hi, scale);
    FDT_CHECK_LAUNCH();
    return 0;
}

// [lo, hi): this rank's region; every rank must pass a region of the same length.
FDT_API int fdt_comm_reduce_scatter_mean(const void* ctx, int64_t lo, int64_t hi, int use_multicast, cudaStream_t st) {
    CommCtx c = *static_cast<const CommCtx*>(ctx);
    if (c.world < 2 || c.world > kMaxWorld || (lo % 4) || (hi % 4)) return -1;
    const int blocks = pick_blocks(hi - lo);
    const float scale = 1.f / c.world;
    if (use_multicast && c.mc) reduce_kernel<1, true><<<blocks, kBlock, 0, st>>>(c, lo, hi, scale);
    else reduce_kernel<1, false><<<blocks, kBlock, 0, st>>>(c, lo, hi, scale);
    FDT_CHECK_LAUNCH();
    return 0;
}

// Bucketed form of the reduce-scatter for overlap with the backward pass: this rank reduces [lo, hi) — its own slice of
// the bucket, which may be EMPTY and differs in length between ranks — with a grid of `blocks` CTAs that must be the same
// on every rank (block b pairs with block b of the peers).  Σg² accumulates in the scratch slot across calls; read it once
// with fdt_comm_take_scalar after the last bucket.
FDT_API int fdt_comm_reduce_scatter_range(const void* ctx, int64_t lo, int64_t hi, int blocks, int use_multicast, cudaStream_t st) {
    CommCtx c = *static_cast<const CommCtx*>(ctx);
    if (c.world < 2 || c.world > kMaxWorld || (lo % 4) || (hi % 4) || hi < lo || blocks < 1 || blocks > 64) return -1;
    const float scale = 1.f / c.world;
    if (use_multicast && c.mc) reduce_kernel<1, true><<<blocks, kBlock, 0, st>>>(c, lo, hi, scale);
    else reduce_kernel<1, false><<<blocks, kBlock, 0, st>>>(c, lo, hi, scale);
    FDT_CHECK_LAUNCH();
    return 0;
}

// Range form of the all-gather (+ bf16 shadow): every rank pulls [lo_r, hi_r) — absolute element offsets in the flat
// buffer, owned by rank r, the same table on every rank — from each owner.  Regions need not have equal length (ranks own
// whole tensors), and a table can describe the parameters of ONE model stage, so that the next stage's gather overlaps this
// stage's forward pass.
__global__ void __launch_bounds__(kBlock)
all_gather_range_kernel(const CommCtx c, const int64_t* __restrict__ sub, __nv_bfloat16* __restrict__ shadow) {
    const int nb = gridDim.x;
    block_barrier(c, blockIdx.x * c.world);
    float* mine = c.buf[c.rank];
    constexpr int U = 8;
    for (int k = 0; k < c.world; ++k) {
        const int r = (c.rank + k) % c.world;
        const int64_t lo = sub[2 * r], hi = sub[2 * r + 1];
        const int64_t len = hi - lo;
        const int64_t per = ((len / 4 + nb - 1) / nb) * 4;
        const int64_t s0 = lo + min(len, per * blockIdx.x), s1 = lo + min(len, per * blockIdx.x + per);
        const float* src = c.buf[r];
        for (int64_t i = s0 + threadIdx.x * 4; i < s1; i += (int64_t)kBlock * 4 * U) {
            float4 v[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int64_t e = i + (int64_t)u * kBlock * 4;
                if (e < s1) v[u] = ld_cg(src + e);
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int64_t e = i + (int64_t)u * kBlock * 4;
                if (e < s1) {
                    if (r != c.rank) st_cg(mine + e, v[u]);
                    if (shadow) {
                        __nv_bfloat162 h0 = __floats2bfloat162_rn(v[u].x, v[u].y), h1 = __floats2bfloat162_rn(v[u].z, v[u].w);
                        *reinterpret_cast<uint2*>(shadow + e) =
                            make_uint2(*reinterpret_cast<uint32_t*>(&h0), *reinterpret_cast<uint32_t*>(&h1));
                    }
                }
            }
        }
    }
    block_barrier(c, (nb + blockIdx.x) * c.world);
}

// Broadcast of ONE owner's large range [lo, hi) to every rank in two phases, so that the owner's NVLink egress is one copy of
// the range instead of world-1 copies (a 62 MB token embedding owned by one rank would otherwise take 7 x 62 MB out of a single
// GPU): the range is cut into `world` slices; phase 1: rank r pulls slice r from the owner; flag barrier; phase 2: rank r pulls
// slice s from rank s, for every s.  Block b handles portion b of every slice, so the per-block barrier (block b of every rank)
// covers exactly the data block b reads in phase 2.  The bf16 shadow of everything that lands (and of the owner's own copy)
// is written in the same pass.
__device__ __forceinline__ void copy_cast_span(const float* __restrict__ src, float* __restrict__ dst, __nv_bfloat16* __restrict__ shadow,
                                               int64_t s0, int64_t s1, bool store_f32) {
    constexpr int U = 4;
    for (int64_t i = s0 + threadIdx.x * 4; i < s1; i += (int64_t)kBlock * 4 * U) {
        float4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t e = i + (int64_t)u * kBlock * 4;
            if (e < s1) v[u] = ld_cg(src + e);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t e = i + (int64_t)u * kBlock * 4;
            if (e < s1) {
                if (store_f32) st_cg(dst + e, v[u]);
                if (shadow) {
                    __nv_bfloat162 h0 = __floats2bfloat162_rn(v[u].x, v[u].y), h1 = __floats2bfloat162_rn(v[u].z, v[u].w);
                    *reinterpret_cast<uint2*>(shadow + e) = make_uint2(*reinterpret_cast<uint32_t*>(&h0), *reinterpret_cast<uint32_t*>(&h1));
                }
            }
        }
    }
}

__global__ void __launch_bounds__(kBlock)
bcast_two_phase_kernel(const CommCtx c, int owner, int64_t lo, int64_t hi, __nv_bfloat16* __restrict__ shadow) {
    const int nb = gridDim.x;
    block_barrier(c, blockIdx.x * c.world);                           // the owner's range is final
    float* mine = c.buf[c.rank];
    const int64_t len = hi - lo;
    const int64_t slice = ((len / 4 + c.world - 1) / c.world) * 4;    // elements per slice (multiple of 4)
    const int64_t per = ((slice / 4 + nb - 1) / nb) * 4;              // elements of a slice handled by one block
    auto span = [&](int s, int64_t& a, int64_t& b) {
        const int64_t sl0 = lo + min(len, slice * s), sl1 = lo + min(len, slice * s + slice);
        a = min(sl1, sl0 + per * blockIdx.x);
        b = min(sl1, a + per);
    };
    int64_t a, b;
    if (c.rank != owner) {
        span(c.rank, a, b);
        copy_cast_span(c.buf[owner], mine, shadow, a, b, true);       // phase 1: my slice, from the owner
    }
    block_barrier(c, (nb + blockIdx.x) * c.world);                    // slice s (portion b) is valid on rank s
    for (int k = 1; k <= c.world; ++k) {
        const int s = (c.rank + k) % c.world;                         // rotated: spread the load over the peers
        if (s == c.rank && c.rank != owner) continue;                 // my own slice: copied (and cast) in phase 1
        span(s, a, b);
        const bool local = (c.rank == owner);                         // the owner already holds everything: cast only
        const int src_rank = local ? owner : s;
        copy_cast_span(c.buf[src_rank], mine, shadow, a, b, !local);
    }
    block_barrier(c, (2 * nb + blockIdx.x) * c.world);                // nobody still reads my copy when I go on to overwrite it
}

// [lo, hi): absolute element offsets (multiples of 4) of a range owned by rank `owner`.  blocks <= 85 (3 barriers per block).
FDT_API int fdt_comm_bcast_range(const void* ctx, int owner, int64_t lo, int64_t hi, void* shadow_bf16, int blocks, cudaStream_t st) {
    CommCtx c = *static_cast<const CommCtx*>(ctx);
    if (c.world < 2 || c.world > kMaxWorld || owner < 0 || owner >= c.world || (lo % 4) || (hi % 4) || hi < lo) return -1;
    if (blocks < 1 || blocks > 85) return -1;
    bcast_two_phase_kernel<<<blocks, kBlock, 0, st>>>(c, owner, lo, hi, (__nv_bfloat16*)shadow_bf16);
    FDT_CHECK_LAUNCH();
    return 0;
}

// sub: DEVICE pointer to 2*world int64 {lo_r, hi_r} (multiples of 4, absolute element offsets; hi_r == lo_r allowed).
FDT_API int fdt_comm_all_gather_range(const void* ctx, const int64_t* sub, void* shadow_bf16, int blocks, cudaStream_t st) {
    CommCtx c = *static_cast<const CommCtx*>(ctx);
    if (c.world < 2 || c.world > kMaxWorld || !sub || blocks < 1 || blocks > 128) return -1;
    all_gather_range_kernel<<<blocks, kBlock, 0, st>>>(c, sub, (__nv_bfloat16*)shadow_bf16);
    FDT_CHECK_LAUNCH();
    return 0;
}

FDT_API int fdt_comm_all_gather_cast(const void* ctx, int64_t region, void* shadow_bf16, int use_multicast, cudaStream_t st) {
    CommCtx c = *static_cast<const CommCtx*>(ctx);
    if (c.world < 2 || c.world > kMaxWorld || region % 4) return -1;
    static int mode = -1;                                  // FDT_AG_PUSH=1 restores the push (multimem.st / P2P store) kernel
    if (mode < 0) { const char* e = getenv("FDT_AG_PUSH"); mode = (e && e[0] == '1') ? 1 : 0; }
    if (mode == 0) {
        int64_t b = (region / 4 + kBlock * 2 - 1) / (kBlock * 2);
        const int blocks = (int)(b < 1 ? 1 : (b > 128 ? 128 : b));      // 2 * blocks * world signal slots <= 2048
        all_gather_pull_cast_kernel<<<blocks, kBlock, 0, st>>>(c, region, (__nv_bfloat16*)shadow_bf16);
        FDT_CHECK_LAUNCH();
        return 0;
    }
    const int blocks = pick_blocks(region * 2);
    if (use_multicast && c.mc) all_gather_cast_kernel<true><<<blocks, kBlock, 0, st>>>(c, region, (__nv_bfloat16*)shadow_bf16);
    else all_gather_cast_kernel<false><<<blocks, kBlock, 0, st>>>(c, region, (__nv_bfloat16*)shadow_bf16);
    FDT_CHECK_LAUNCH();
    return 0;
}

// *out = *slot; *slot = 0   (slot: this rank's Σg² scratch word; call right after a reduce collective)
FDT_API int fdt_comm_take_scalar(float* slot, float* out, cudaStream_t st) {
    take_scalar_kernel<<<1, 1, 0, st>>>(slot, out);
    FDT_CHECK_LAUNCH();
    return 0;
}

// Cross-rank barrier on the stream (one block): everything the peers enqueued before their call has completed.
FDT_API int fdt_comm_barrier(const void* ctx, cudaStream_t st) {
    CommCtx c = *static_cast<const CommCtx*>(ctx);
    if (c.world < 2 || c.world > kMaxWorld) return -1;
    barrier_kernel<<<1, 32, 0, st>>>(c);
    FDT_CHECK_LAUNCH();
    return 0;
}

// dst_mc (multicast address, 16-byte aligned, n % 4 == 0) += alpha * src on every rank.
FDT_API int fdt_comm_mc_red_add(float* dst_mc, const float* src, int64_t n, float alpha, cudaStream_t st) {
    if (n % 4) return -1;
    int64_t b = (n / 4 + 255) / 256;
    if (b > 148 * 4) b = 148 * 4;
    if (b < 1) b = 1;
    mc_red_add_kernel<<<(int)b, 256, 0, st>>>(dst_mc, src, n, alpha);
    FDT_CHECK_LAUNCH();
    return 0;
}
