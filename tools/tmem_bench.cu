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
__device__ __forceinline__ void st64(uint32_t taddr, const uint32_t (&r)[64]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x64.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32, %33, %34, %35, %36, %37, %38, %39, %40, %41, %42, %43, %44, %45, %46, %47, %48, %49, %50, %51, %52, %53, %54, %55, %56, %57, %58, %59, %60, %61, %62, %63, %64};" ::"r"(taddr),
        "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]),
        "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31]),
        "r"(r[32]), "r"(r[33]), "r"(r[34]), "r"(r[35]), "r"(r[36]), "r"(r[37]), "r"(r[38]), "r"(r[39]), "r"(r[40]), "r"(r[41]), "r"(r[42]), "r"(r[43]), "r"(r[44]), "r"(r[45]), "r"(r[46]), "r"(r[47]),
        "r"(r[48]), "r"(r[49]), "r"(r[50]), "r"(r[51]), "r"(r[52]), "r"(r[53]), "r"(r[54]), "r"(r[55]), "r"(r[56]), "r"(r[57]), "r"(r[58]), "r"(r[59]), "r"(r[60]), "r"(r[61]), "r"(r[62]), "r"(r[63]) : "memory");
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
__device__ __forceinline__ uint64_t make_desc(uint32_t a) { uint64_t d = 0; d |= (uint64_t)((a >> 4) & 0x3FFF); d |= (uint64_t)1 << 16; d |= (uint64_t)64 << 32; d |= (uint64_t)1 << 46; d |= (uint64_t)2 << 61; return d; }
// warps 0-3 time tcgen05.st batches while warp 4 keeps the tensor pipe busy with tf32 MMAs (N columns)
__global__ void __launch_bounds__(160) st_under_mma(int N, int mma_on, int iters, long long* out, int wide, int ss_mode, int gap) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint32_t tslot; __shared__ volatile int stop;
    uint8_t* sm = smem + ((1024u - (smem_u32(smem) & 1023u)) & 1023u);
    for (int i = threadIdx.x; i < 32 * 1024 / 4; i += blockDim.x) ((float*)sm)[i] = 1.0f;
    if (threadIdx.x == 0) stop = 0;
    if (threadIdx.x < 32) { asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tslot)), "r"(512)); asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;"); }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;"); __syncthreads(); asm volatile("tcgen05.fence::after_thread_sync;");
    const int warp = threadIdx.x >> 5;
    if (warp < 4) {
        const uint32_t tb = tslot + ((uint32_t)(warp * 32) << 16) + 256;
        uint32_t r[16]; for (int j = 0; j < 16; ++j) r[j] = threadIdx.x * 16 + j;
        uint32_t w[64]; for (int j = 0; j < 64; ++j) w[j] = threadIdx.x * 64 + j;
        long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
            if (wide == 2) { } else if (wide) { for (int a = 0; a < 4; ++a) { st64(tb + a * 64, w); asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); } }
            else { for (int a = 0; a < 4; ++a) { st16(tb + a * 64, r); st16(tb + a * 64 + 16, r); st16(tb + a * 64 + 32, r); st16(tb + a * 64 + 48, r); asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); } }
            if (gap) __nanosleep(gap);
        }
        long long t1 = clock64();
        if ((threadIdx.x & 31) == 0) out[warp] = t1 - t0;
        __syncwarp();
        if (threadIdx.x == 0) stop = 1;
    } else if (mma_on) {
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | (8u << 24);
        const uint64_t bdesc = make_desc(smem_u32(sm));
        long long n = 0;
        while (!stop) {
            for (int i = 0; i < 8; ++i) {
                uint32_t pred;
                asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
                if (pred && !ss_mode) asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tslot), "r"(tslot + 128 + i * 8), "l"(bdesc + (uint64_t)((i & 3) * 2)), "r"(idesc), "r"(1) : "memory");
                if (pred && ss_mode) asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tslot), "l"(bdesc + 1024 + (uint64_t)((i & 3) * 2)), "l"(bdesc + (uint64_t)((i & 3) * 2)), "r"(idesc), "r"(1) : "memory");
            }
            n += 8;
        }
        if ((threadIdx.x & 31) == 0) out[4] = n;
    }
    asm volatile("tcgen05.fence::before_thread_sync;"); __syncthreads();
    if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tslot), "r"(512));
}
int main() {
    {
        long long* dout; cudaMalloc(&dout, 64); long long h[5]; const int iters = 2000;
        cudaFuncSetAttribute(st_under_mma, cudaFuncAttributeMaxDynamicSharedMemorySize, 40 * 1024);
        for (int cfg = 0; cfg < 10; ++cfg) {
            const int wide = cfg >= 8 ? 2 : (cfg & 1), ss = cfg >= 8 ? (cfg & 1) : ((cfg >> 1) & 1), gap = (cfg >> 2) ? 1000 : 0;
            cudaMemset(dout, 0, 64);
            st_under_mma<<<1, 160, 40 * 1024>>>(64, 1, iters, dout, wide, ss, gap);
            cudaError_t e = cudaDeviceSynchronize(); cudaMemcpy(h, dout, 40, cudaMemcpyDeviceToHost);
            printf("%s per iter%s, MMA %s N=64: st loop %.0f cycles/iter; cycles per MMA meanwhile: %.0f (%s)\n", wide == 2 ? "no stores" : wide ? "4 x st64" : "16 x st16", gap ? " + 1us gap" : "", ss ? "SS" : "TS", (double)h[0] / iters, (double)h[0] / (double)h[4], cudaGetErrorString(e));
        }
    }

    long long* dout; uint32_t* ds; cudaMalloc(&dout, 64); cudaMalloc(&ds, 64); long long h[4]; const int iters = 2000;
#define RUN(V, what) bench<V><<<1, 128>>>(iters, dout, ds); printf("%s: %s ", what, cudaGetErrorString(cudaDeviceSynchronize())); cudaMemcpy(h, dout, 32, cudaMemcpyDeviceToHost); printf("cycles/iter per warp: %.0f %.0f %.0f %.0f\n", (double)h[0] / iters, (double)h[1] / iters, (double)h[2] / iters, (double)h[3] / iters);
    RUN(0, "16 x st16 (256 cols), wait::st per 4") RUN(1, "16 x st16, one wait") RUN(2, "8 x st32, one wait") RUN(3, "16 x ld16, one wait") RUN(4, "4 x (ld16 + wait)") RUN(5, "1 x (st16 + wait)")
    return 0;
}
