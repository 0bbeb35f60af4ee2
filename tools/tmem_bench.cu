// Isolated timing of tcgen05.st / tcgen05.ld patterns (development aid).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void st16(uint32_t taddr, const uint32_t (&r)[16]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
        "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]) : "memory");
}
__device__ __forceinline__ void st32(uint32_t taddr, const uint32_t (&r)[16]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
        "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]) : "memory");
}
__device__ __forceinline__ void ld16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]) : "r"(taddr) : "memory");
}
template <int V>
__global__ void __launch_bounds__(128) bench(int iters, long long* out, uint32_t* sink) {
    __shared__ uint32_t tslot;
    if (threadIdx.x < 32) { asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tslot)), "r"(512)); asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;"); }
    asm volatile("tcgen05.fence::before_thread_sync;"); __syncthreads(); asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tb = tslot + ((uint32_t)((threadIdx.x >> 5) * 32) << 16);
    uint32_t r[16]; for (int j = 0; j < 16; ++j) r[j] = threadIdx.x * 16 + j;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        if (V == 0) { for (int a = 0; a < 4; ++a) { st16(tb + a * 64, r); st16(tb + a * 64 + 16, r); st16(tb + a * 64 + 32, r); st16(tb + a * 64 + 48, r); asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); } }
        if (V == 1) { for (int a = 0; a < 4; ++a) { st16(tb + a * 64, r); st16(tb + a * 64 + 16, r); st16(tb + a * 64 + 32, r); st16(tb + a * 64 + 48, r); } asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
        if (V == 2) { for (int a = 0; a < 4; ++a) { st32(tb + a * 64, r); st32(tb + a * 64 + 32, r); } asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
        if (V == 3) { for (int a = 0; a < 4; ++a) { ld16(tb + a * 64, r); ld16(tb + a * 64 + 16, r); ld16(tb + a * 64 + 32, r); ld16(tb + a * 64 + 48, r); } asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
        if (V == 4) { for (int a = 0; a < 4; ++a) { ld16(tb + a * 64, r); asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); r[0] += r[5]; } }
        if (V == 5) { st16(tb, r); asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
    }
    long long t1 = clock64();
    if ((threadIdx.x & 31) == 0) out[threadIdx.x >> 5] = t1 - t0;
    if (r[3] == 0x7fffffff) sink[0] = r[1];
    asm volatile("tcgen05.fence::before_thread_sync;"); __syncthreads();
    if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tslot), "r"(512));
}
int main() {
    long long* dout; uint32_t* ds; cudaMalloc(&dout, 64); cudaMalloc(&ds, 64); long long h[4]; const int iters = 2000;
#define RUN(V, what) bench<V><<<1, 128>>>(iters, dout, ds); printf("%s: %s ", what, cudaGetErrorString(cudaDeviceSynchronize())); cudaMemcpy(h, dout, 32, cudaMemcpyDeviceToHost); printf("cycles/iter per warp: %.0f %.0f %.0f %.0f\n", (double)h[0] / iters, (double)h[1] / iters, (double)h[2] / iters, (double)h[3] / iters);
    RUN(0, "16 x st16 (256 cols), wait::st per 4") RUN(1, "16 x st16, one wait") RUN(2, "8 x st32, one wait") RUN(3, "16 x ld16, one wait") RUN(4, "4 x (ld16 + wait)") RUN(5, "1 x (st16 + wait)")
    return 0;
}
