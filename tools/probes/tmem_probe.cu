// Probe: which (row, column) does each register of tcgen05.ld.16x256b.x8 hold?  TMEM is filled with tcgen05.st.32x32b
// (thread = row, register = column; value = row * 256 + column), read back with the 16x256b shape and dumped.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -o gpurun_out/tmem_probe tools/probes/tmem_probe.cu ; run on a B200
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__global__ void probe(uint32_t* out) {
    __shared__ uint32_t tmem_ptr;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 64;" ::"r"(smem_u32(&tmem_ptr)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t base = tmem_ptr;
    const uint32_t row = warp * 32 + lane;
    for (int half = 0; half < 2; ++half) {
        uint32_t v[32];
        for (int j = 0; j < 32; ++j) v[j] = row * 256 + half * 32 + j;
        const uint32_t taddr = base + ((uint32_t)(warp * 32) << 16) + half * 32;
        asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
                     "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
                     ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]),
                       "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]),
                       "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]),
                       "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31]) : "memory");
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    for (int h = 0; h < 2; ++h) {
        uint32_t r[32];
        const uint32_t taddr = base + ((uint32_t)(warp * 32 + h * 16) << 16);
        asm volatile("tcgen05.ld.sync.aligned.16x256b.x8.b32 "
                     "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                     "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                       "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
                       "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
                       "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                     : "r"(taddr) : "memory");
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        for (int j = 0; j < 32; ++j) out[((warp * 2 + h) * 32 + lane) * 32 + j] = r[j];
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 64;" ::"r"(base) : "memory");
}
int main() {
    uint32_t* d; cudaMalloc(&d, 4 * 2 * 32 * 32 * 4);
    probe<<<1, 128>>>(d);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
    static uint32_t h[4 * 2 * 32 * 32];
    cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    // check the expected mapping: thread t, register 4 i + 2 k + e  ->  row = 16 h + t / 4 + 8 k, column = 8 i + 2 (t % 4) + e
    int bad = 0;
    for (int w = 0; w < 4; ++w) for (int hh = 0; hh < 2; ++hh) for (int t = 0; t < 32; ++t) for (int j = 0; j < 32; ++j) {
        const int i = j / 4, k = (j / 2) & 1, e2 = j & 1;
        const uint32_t want = (uint32_t)(w * 32 + 16 * hh + t / 4 + 8 * k) * 256 + 8 * i + 2 * (t % 4) + e2;
        if (h[((w * 2 + hh) * 32 + t) * 32 + j] != want) ++bad;
    }
    printf("16x256b.x8 mapping mismatches vs expectation: %d\n", bad);
    for (int t = 0; t < 6; ++t) {
        printf("warp0 h0 thread %d:", t);
        for (int j = 0; j < 12; ++j) { uint32_t v = h[t * 32 + j]; printf(" (r%d,c%d)", v / 256, v % 256); }
        printf("\n");
    }
    return 0;
}
