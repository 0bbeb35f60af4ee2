// Microbenchmark: issue rate of tcgen05.mma kind::tf32 (M=128, K=8) for several N, A from TMEM (TS) or smem (SS),
// same accumulator vs alternating accumulators.  Development aid; build: nvcc -gencode arch=compute_100a,code=sm_100a
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t make_desc(uint32_t a) {
    uint64_t d = 0; d |= (uint64_t)((a >> 4) & 0x3FFF); d |= (uint64_t)1 << 16; d |= (uint64_t)64 << 32; d |= (uint64_t)1 << 46; d |= (uint64_t)2 << 61; return d; }
__device__ __forceinline__ void mma_ts(uint32_t d, uint32_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d), "r"(a), "l"(b), "r"(idesc), "r"(acc) : "memory"); }
__device__ __forceinline__ void mma_ss(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory"); }
__global__ void __launch_bounds__(128, 1) bench(int N, int mode, int iters, long long* out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint32_t tslot; __shared__ __align__(8) uint64_t bar;
    uint8_t* sm = smem + ((1024u - (smem_u32(smem) & 1023u)) & 1023u);
    for (int i = threadIdx.x; i < 64 * 1024 / 4; i += 128) ((float*)sm)[i] = 1.0f;
    if (threadIdx.x == 0) { asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar))); asm volatile("fence.mbarrier_init.release.cluster;"); }
    if (threadIdx.x < 32) { asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tslot)), "r"(512)); asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;"); }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;"); __syncthreads(); asm volatile("tcgen05.fence::after_thread_sync;");
    uint32_t tb = tslot;
    uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | (8u << 24);
    if (mode & 4) {
      if (threadIdx.x < 32) {
        uint64_t bdesc = make_desc(smem_u32(sm));
        long long t0 = clock64();
        for (int i = 0; i < iters; ++i) {
            uint32_t d = tb + ((mode & 1) ? (uint32_t)((i & 1) * N) : 0u);
            uint32_t pred;
            asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
            if (pred) mma_ts(d, tb + 256 + (i & 7) * 8, bdesc + (uint64_t)((i & 3) * 2), idesc, i > 1);
        }
        long long t1 = clock64();
        if (threadIdx.x == 0) {
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
        asm volatile("{\n\t.reg .pred p;\n\tW2:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n\t@p bra D2;\n\tbra W2;\n\tD2:\n\t}" ::"r"(smem_u32(&bar)) : "memory");
        long long t2 = clock64();
        out[0] = t1 - t0; out[1] = t2 - t0; }
      }
    } else if (threadIdx.x == 0) {
        uint64_t bdesc = make_desc(smem_u32(sm)); uint64_t adesc = make_desc(smem_u32(sm) + 32768);
        long long t0 = clock64();
        for (int i = 0; i < iters; ++i) {
            uint32_t d = tb + ((mode & 1) ? (uint32_t)((i & 1) * N) : 0u);
            if (mode & 2) mma_ss(d, adesc + (uint64_t)((i & 3) * 2), bdesc + (uint64_t)((i & 3) * 2), idesc, i > 1);
            else mma_ts(d, tb + 256 + (i & 7) * 8, bdesc + (uint64_t)((i & 3) * 2), idesc, i > 1);
        }
        long long t1 = clock64();
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
        asm volatile("{\n\t.reg .pred p;\n\tW:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n\t@p bra D;\n\tbra W;\n\tD:\n\t}" ::"r"(smem_u32(&bar)) : "memory");
        long long t2 = clock64();
        out[0] = t1 - t0; out[1] = t2 - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;"); __syncthreads();
    if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tb), "r"(512));
}
__global__ void __launch_bounds__(288, 1) bench2(int N, int noise, int iters, long long* out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint32_t tslot; __shared__ __align__(8) uint64_t bar; __shared__ volatile int stop;
    uint8_t* sm = smem + ((1024u - (smem_u32(smem) & 1023u)) & 1023u);
    for (int i = threadIdx.x; i < 64 * 1024 / 4; i += blockDim.x) ((float*)sm)[i] = 1.0f;
    if (threadIdx.x == 0) { stop = 0; asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar))); asm volatile("fence.mbarrier_init.release.cluster;"); }
    if (threadIdx.x < 32) { asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tslot)), "r"(512)); asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;"); }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;"); __syncthreads(); asm volatile("tcgen05.fence::after_thread_sync;");
    uint32_t tb = tslot;
    uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | (8u << 24);
    const int warp = threadIdx.x >> 5;
    if (warp == 0) {
        uint64_t bdesc = make_desc(smem_u32(sm));
        long long t0 = clock64();
        for (int i = 0; i < iters; ++i) {
            uint32_t pred;
            asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
            if (pred) mma_ts(tb, tb + 256 + (i & 7) * 8, bdesc + (uint64_t)((i & 3) * 2), idesc, i > 1);
            if ((noise & 8) && (i & 7) == 7) { if (pred) asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar) + 0) : "memory"); }
        }
        long long t1 = clock64();
        if (threadIdx.x == 0) {
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
            // note: with noise&8 the barrier saw earlier arrivals; just spin on the clock instead
            long long t2 = clock64();
            out[0] = t1 - t0; out[1] = t2 - t0; stop = 1;
        }
    } else if (warp >= 1 && warp <= 4) {
        const uint32_t lane_base = tb + ((uint32_t)((warp & 3) * 32) << 16);
        uint32_t r[16];
        for (int j = 0; j < 16; ++j) r[j] = threadIdx.x + j;
        while (!stop) {
            if (noise & 1) {
                asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(lane_base + 384),
                    "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]) : "memory");
                asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
            }
            if (noise & 2) {
                asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                    : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]) : "r"(lane_base + 448) : "memory");
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            }
        }
        if (r[3] == 0x7fffffff) out[2] = r[5];
    } else if (warp >= 5) {
        float acc = 0.f;
        while (!stop) {
            if (noise & 4) {
                for (int j = 0; j < 32; ++j) { float4 v = *reinterpret_cast<const float4*>(sm + 32768 + ((threadIdx.x * 16 + j * 512) & 16383)); acc += v.x + v.w; }
            }
        }
        if (acc == 12345.f) out[3] = 1;
    }
    asm volatile("tcgen05.fence::before_thread_sync;"); __syncthreads();
    if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tb), "r"(512));
}
int main() {
    {
        long long* d_out; cudaMalloc(&d_out, 64); long long h[2];
        cudaFuncSetAttribute(bench2, cudaFuncAttributeMaxDynamicSharedMemorySize, 80 * 1024);
        for (int noise : {0, 1, 2, 4, 8, 15})
            for (int N : {32, 64}) {
                bench2<<<1, 288, 80 * 1024>>>(N, noise, 512, d_out);
                cudaError_t e = cudaDeviceSynchronize();
                cudaMemcpy(h, d_out, 16, cudaMemcpyDeviceToHost);
                printf("contention noise=%2d (1=tmem st,2=tmem ld,4=lds,8=commit/8) N=%3d: issue %.1f cyc/mma (%s)\n", noise, N, (double)h[0] / 512, cudaGetErrorString(e));
            }
    }

    long long* d_out; cudaMalloc(&d_out, 16); long long h[2];
    cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, 80 * 1024);
    const int iters = 512;
    for (int mode = 0; mode < 6; ++mode)
        for (int N : {16, 32, 64, 128, 256}) {
            if ((mode & 1) && 2 * N > 256) continue;
            bench<<<1, 128, 80 * 1024>>>(N, mode, iters, d_out);
            cudaError_t e = cudaDeviceSynchronize();
            cudaMemcpy(h, d_out, 16, cudaMemcpyDeviceToHost);
            printf("%s %s N=%3d: issue %.1f cyc/mma, complete %.1f cyc/mma (%s)\n", (mode & 4) ? "TS-elect" : (mode & 2) ? "SS" : "TS",
                   (mode & 1) ? "alt-D " : "same-D", N, (double)h[0] / iters, (double)h[1] / iters, cudaGetErrorString(e));
        }
    return 0;
}
