// Round-trip latency of an mbarrier hand-off between two warps of one CTA (development aid).
// modes: 0 = test_wait spin, 1 = try_wait (default time limit), 2 = try_wait with a 10 ms suspend hint.
// `hogs` extra warps run dependent FFMA chains to load the schedulers like the fused kernels do.
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
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
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
__global__ void hop_kernel(int iters, int hogs, long long* out, float* sink) {
    __shared__ __align__(8) unsigned long long bars[2];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t b0 = smem_u32(&bars[0]), b1 = smem_u32(&bars[1]);
    if (threadIdx.x == 0) { mbar_init(b0, 1); mbar_init(b1, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    __syncthreads();
    if (warp == 0) {
        const long long t0 = clock64();
        for (int i = 0; i < iters; ++i) {
            if (lane == 0) mbar_arrive(b0);
            mbar_wait<MODE>(b1, i & 1);
        }
        if (lane == 0) out[blockIdx.x] = clock64() - t0;
    } else if (warp == 1) {
        for (int i = 0; i < iters; ++i) {
            mbar_wait<MODE>(b0, i & 1);
            __syncwarp();
            if (lane == 0) mbar_arrive(b1);
        }
    } else if (warp < 2 + hogs) {
        float a = threadIdx.x, b = 1.0001f;
        for (int i = 0; i < iters * 40; ++i) { a = fmaf(a, b, 0.5f); b = fmaf(b, 0.999f, 0.001f); }
        if (a == 123.f) sink[0] = a + b;
    }
}

template <int MODE>
void run(int hogs) {
    long long* d; float* s;
    cudaMalloc(&d, 148 * 8); cudaMalloc(&s, 4);
    const int iters = 20000;
    hop_kernel<MODE><<<148, 32 * (2 + hogs)>>>(iters, hogs, d, s);
    cudaDeviceSynchronize();
    hop_kernel<MODE><<<148, 32 * (2 + hogs)>>>(iters, hogs, d, s);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[148];
    cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    printf("mode %d hogs %2d: %s  round trip %.0f cycles (one hop %.0f)\n", MODE, hogs, cudaGetErrorString(e),
           (double)h[0] / iters, (double)h[0] / iters / 2);
    cudaFree(d); cudaFree(s);
}

int main() {
    for (int hogs : {0, 8, 20}) { run<0>(hogs); run<1>(hogs); run<2>(hogs); }
    return 0;
}
