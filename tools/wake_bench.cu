// Wake-up latency of a warp blocked on an mbarrier as a function of how long it has been waiting
// (development aid).  Warp 0 delays `D` cycles, stamps clock64 and arrives; warp 1 (blocked in the wait)
// stamps clock64 when it resumes.  modes: 0 = test_wait spin, 1 = try_wait, 2 = try_wait + 10 ms hint.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
template <int MODE>
__device__ __noinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    if (MODE == 0) {
        uint32_t done = 0;
        while (!done) {
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(done) : "r"(bar), "r"(parity) : "memory");
        }
    } else if (MODE == 1) {
        asm volatile("{\n\t.reg .pred p;\n\tW1:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra D1;\n\tbra W1;\n\tD1:\n\t}" ::"r"(bar), "r"(parity) : "memory");
    } else {
        asm volatile("{\n\t.reg .pred p;\n\tW2:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n\t@p bra D2;\n\tbra W2;\n\tD2:\n\t}" ::"r"(bar), "r"(parity), "r"(0x989680) : "memory");
    }
}

template <int MODE>
__global__ void wake_kernel(int iters, int delay, int hogs, long long* out, float* sink) {
    __shared__ __align__(8) unsigned long long bars[2];
    __shared__ long long t_arrive;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t b0 = smem_u32(&bars[0]), b1 = smem_u32(&bars[1]);
    if (threadIdx.x == 0) { mbar_init(b0, 1); mbar_init(b1, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    __syncthreads();
    if (warp == 0) {
        for (int i = 0; i < iters; ++i) {
            const long long t0 = clock64();
            while (clock64() - t0 < delay) {}
            if (lane == 0) { t_arrive = clock64(); __threadfence_block(); mbar_arrive(b0); }
            mbar_wait<MODE>(b1, i & 1);
        }
    } else if (warp == 1) {
        long long total = 0;
        for (int i = 0; i < iters; ++i) {
            mbar_wait<MODE>(b0, i & 1);
            const long long t1 = clock64();
            total += t1 - *(volatile long long*)&t_arrive;
            __syncwarp();
            if (lane == 0) mbar_arrive(b1);
        }
        if (lane == 0) out[blockIdx.x] = total;
    } else if (warp < 2 + hogs) {
        float a = threadIdx.x, b = 1.0001f;
        const long long t0 = clock64();
        while (clock64() - t0 < (long long)iters * (delay + 400)) { a = fmaf(a, b, 0.5f); b = fmaf(b, 0.999f, 0.001f); }
        if (a == 123.f) sink[0] = a + b;
    }
}

template <int MODE>
void run(int delay, int hogs) {
    long long* d; float* s;
    cudaMalloc(&d, 148 * 8); cudaMalloc(&s, 4);
    const int iters = 2000;
    wake_kernel<MODE><<<148, 32 * (2 + hogs)>>>(iters, delay, hogs, d, s);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[148];
    cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    printf("mode %d delay %6d hogs %2d: %s  arrive->resume %.0f cycles\n", MODE, delay, hogs, cudaGetErrorString(e),
           (double)h[0] / iters);
    cudaFree(d); cudaFree(s);
}

int main() {
    for (int hogs : {0, 20})
        for (int delay : {100, 3000, 30000}) { run<0>(delay, hogs); run<1>(delay, hogs); run<2>(delay, hogs); }
    return 0;
}
