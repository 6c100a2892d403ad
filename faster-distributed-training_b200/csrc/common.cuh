// Shared device helpers for the sm_100a kernels of faster-distributed-training_b200.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>

#define FDT_API extern "C" __attribute__((visibility("default")))

#define FDT_CHECK_LAUNCH() do { cudaError_t e__ = cudaGetLastError(); if (e__ != cudaSuccess) return (int)e__; } while (0)

#include <cstdlib>
#include <utility>

namespace fdt {

// ---- programmatic dependent launch (PDL).  A kernel launched through fdt::launch may be scheduled while its
// predecessor in the stream is still draining: its CTAs run their prologue (barrier init, TMEM allocation,
// tensor-map prefetch, shared-memory clearing) and then block in pdl_wait() until the predecessor has completed
// and its writes are visible.  Every kernel launched this way MUST call pdl_wait() before its first global-memory
// access.  The launch-to-launch bubble (~2 us x several hundred kernels per training step) disappears; the edges
// survive CUDA-graph capture as programmatic dependencies.  Opt-in (FDT_PDL=1): inside the captured training step it
// measured no gain (6.60 vs 6.57 ms/step on ResNet-50) — graph replay already hides the launch latency.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

inline bool pdl_enabled() {
    static int v = -1;
    if (v < 0) { const char* e = getenv("FDT_PDL"); v = (e && e[0] == '1') ? 1 : 0; }
    return v != 0;
}

template <typename... KArgs, typename... Args>
inline cudaError_t launch(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...);
}

enum Act : int { ACT_NONE = 0, ACT_RELU = 1, ACT_CELU = 2, ACT_GELU = 3, ACT_TANH = 4 };

// Φ(z) = ½(1 + erf(z/√2)) and e = exp(−z²/2) for the exact (erf) GELU of the reference (transformer.py:175, F.gelu) from ONE
// exponential: erf by Abramowitz & Stegun 7.1.26 (|error| <= 1.5e-7, far below the bf16 resolution of every tensor these
// values land in), erf(u) = 1 − t(a1 + t(a2 + t(a3 + t(a4 + t a5)))) exp(−u²), t = 1/(1 + p u), u = |z|/√2 — whose exponential
// is the exp(−z²/2) the derivative needs anyway.  ~14 instructions instead of erff + expf (~30); the tail is formed without
// cancellation (Φ = ½·poly·e for z < 0).
__device__ __forceinline__ float gelu_phi(float z, float& e) {
    const float u = fabsf(z) * 0.70710678118654752f;
    e = __expf(-0.5f * z * z);
    const float t = __fdividef(1.f, fmaf(0.3275911f, u, 1.f));
    const float poly = t * fmaf(t, fmaf(t, fmaf(t, fmaf(t, 1.061405429f, -1.453152027f), 1.421413741f), -0.284496736f), 0.254829592f);
    const float half_tail = 0.5f * poly * e;                 // ½(1 − erf(u))
    return z < 0.f ? half_tail : 1.f - half_tail;
}
constexpr float kCeluAlpha = 0.075f;   // resnet.py:240

__device__ __forceinline__ float act_fwd(float x, int act) {
    switch (act) {
        case ACT_RELU: return fmaxf(x, 0.f);
        case ACT_CELU: return x > 0.f ? x : kCeluAlpha * (__expf(x * (1.f / kCeluAlpha)) - 1.f);
        case ACT_GELU: { float e; return x * gelu_phi(x, e); }
        case ACT_TANH: return tanhf(x);
        default: return x;
    }
}
// derivative expressed through the OUTPUT o = act(x) (relu / celu / tanh) — lets the backward
// pass read the saved activation instead of the pre-activation
__device__ __forceinline__ float act_bwd_from_out(float o, int act) {
    switch (act) {
        case ACT_RELU: return o > 0.f ? 1.f : 0.f;
        case ACT_CELU: return o > 0.f ? 1.f : (o * (1.f / kCeluAlpha) + 1.f);
        case ACT_TANH: return 1.f - o * o;
        default: return 1.f;
    }
}
__device__ __forceinline__ float gelu_bwd_from_in(float x) {
    float e;
    const float phi = gelu_phi(x, e);
    return fmaf(x * 0.3989422804014327f, e, phi);        // Φ(x) + x·φ(x),  φ = exp(−x²/2)/√(2π)
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// 8 bf16 values moved as ONE 128-bit transaction (LDG.E.128 / STG.E.128).  The payload is a uint4 on
// purpose: a struct of four __nv_bfloat162 is copied member-wise (4 x 32-bit accesses).
struct alignas(16) bf16x8 { uint4 u; };

__device__ __forceinline__ float2 bf2_to_f2(uint32_t w) {
    __nv_bfloat162 h = *reinterpret_cast<__nv_bfloat162*>(&w);
    return __bfloat1622float2(h);
}
__device__ __forceinline__ uint32_t f2_to_bf2(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ void unpack8(const bf16x8& p, float* f) {
    float2 t;
    t = bf2_to_f2(p.u.x); f[0] = t.x; f[1] = t.y;
    t = bf2_to_f2(p.u.y); f[2] = t.x; f[3] = t.y;
    t = bf2_to_f2(p.u.z); f[4] = t.x; f[5] = t.y;
    t = bf2_to_f2(p.u.w); f[6] = t.x; f[7] = t.y;
}
__device__ __forceinline__ bf16x8 pack8(const float* f) {
    bf16x8 p;
    p.u = make_uint4(f2_to_bf2(f[0], f[1]), f2_to_bf2(f[2], f[3]), f2_to_bf2(f[4], f[5]), f2_to_bf2(f[6], f[7]));
    return p;
}
__device__ __forceinline__ bf16x8 ld8(const __nv_bfloat16* p) {
    bf16x8 v;
    v.u = __ldg(reinterpret_cast<const uint4*>(p));
    return v;
}
__device__ __forceinline__ void st8(__nv_bfloat16* p, const bf16x8& v) { *reinterpret_cast<uint4*>(p) = v.u; }

// Philox4x32-10 counter-based RNG (dropout masks, device-side Beta / permutation sampling)
// Philox4x32 with 7 rounds: the smallest round count that passes BigCrush (Salmon et al., "Parallel random numbers: as easy as
// 1, 2, 3", SC'11, table 2; 10 is the conservative library default).  The streams only feed dropout masks, augmentation and
// mixup draws, and forward / backward regenerate them from the same function, so the round count is a pure cost knob: the
// generator is a third of the work of the dropout-carrying kernels.
constexpr int kPhiloxRounds = 7;
struct Philox {
    uint32_t key[2];
    __device__ Philox(uint64_t seed) { key[0] = (uint32_t)seed; key[1] = (uint32_t)(seed >> 32); }
    __device__ __forceinline__ uint4 operator()(uint64_t ctr_lo, uint64_t ctr_hi) const {
        uint32_t c0 = (uint32_t)ctr_lo, c1 = (uint32_t)(ctr_lo >> 32), c2 = (uint32_t)ctr_hi, c3 = (uint32_t)(ctr_hi >> 32);
        uint32_t k0 = key[0], k1 = key[1];
#pragma unroll
        for (int r = 0; r < kPhiloxRounds; ++r) {
            uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
            uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
            uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
            c0 = n0; c1 = n1; c2 = n2; c3 = n3;
            k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
        }
        return make_uint4(c0, c1, c2, c3);
    }
};
__device__ __forceinline__ float u32_to_unit(uint32_t x) { return (x >> 8) * (1.0f / 16777216.0f); }   // [0,1)

}  // namespace fdt
