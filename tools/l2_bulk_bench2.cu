// Variant of l2_bulk_bench: does a cluster launch (2 CTAs), a large shared-memory footprint or a long box change
// the latency of bulk copies from an L2-resident buffer?  usage: l2_bulk_bench2 slot_bytes slots
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile("{\n\t.reg .pred p;\n\tW:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n\t@p bra D;\n\tbra W;\n\tD:\n\t}" ::"r"(bar), "r"(parity), "r"(0x989680) : "memory");
}
template <bool CLUSTER>
__device__ __forceinline__ void body(const uint8_t* buf, size_t region_bytes, int slot_bytes, int S, int iters, int hold,
                                     unsigned long long* cycles, unsigned long long* lat) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t full[16], empty[16];
    __shared__ long long t_issue[16];
    if (threadIdx.x == 0) {
        for (int s = 0; s < S; ++s) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&full[s])));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&empty[s])));
        }
        asm volatile("fence.mbarrier_init.release.cluster;");
    }
    __syncthreads();
    const size_t nblk = region_bytes / slot_bytes;
    const size_t start = ((size_t)blockIdx.x * 7919u) % nblk;
    const long long t0 = clock64();
    if (threadIdx.x == 0) {
        for (int it = 0; it < iters; ++it) {
            const int s = it % S;
            mbar_wait(smem_u32(&empty[s]), ((it / S) & 1) ^ 1);
            t_issue[s] = clock64();
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&full[s])), "r"(slot_bytes) : "memory");
            const uint8_t* src = buf + ((start + it) % nblk) * (size_t)slot_bytes;
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem) + s * slot_bytes), "l"(src), "r"(slot_bytes), "r"(smem_u32(&full[s])) : "memory");
        }
    } else if (threadIdx.x == 32) {
        long long lsum = 0;
        for (int it = 0; it < iters; ++it) {
            const int s = it % S;
            mbar_wait(smem_u32(&full[s]), (it / S) & 1);
            const long long now = clock64();
            lsum += now - *(volatile long long*)&t_issue[s];
            while (clock64() - now < hold) {}
            asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&empty[s])) : "memory");
        }
        cycles[blockIdx.x] = (unsigned long long)(clock64() - t0);
        lat[blockIdx.x] = (unsigned long long)(lsum / iters);
    }
}
__global__ void __launch_bounds__(64, 1) pull_plain(const uint8_t* buf, size_t rb, int sb, int S, int iters, int hold, unsigned long long* c, unsigned long long* l) { body<false>(buf, rb, sb, S, iters, hold, c, l); }
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(64, 1) pull_cluster(const uint8_t* buf, size_t rb, int sb, int S, int iters, int hold, unsigned long long* c, unsigned long long* l) { body<true>(buf, rb, sb, S, iters, hold, c, l); }
int main(int argc, char** argv) {
    const int slot = argc > 1 ? atoi(argv[1]) : 16384;
    const int S = argc > 2 ? atoi(argv[2]) : 5;
    const int iters = 4000;
    uint8_t* buf;
    cudaMalloc(&buf, (size_t)4 << 20);
    cudaMemset(buf, 1, (size_t)4 << 20);
    unsigned long long *cyc, *lat;
    cudaMalloc(&cyc, 148 * 8); cudaMalloc(&lat, 148 * 8);
    cudaFuncSetAttribute(pull_plain, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    cudaFuncSetAttribute(pull_cluster, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    for (int cl = 0; cl < 2; ++cl)
        for (int bigsmem = 0; bigsmem < 2; ++bigsmem)
            for (int hold : {0, 500, 1000, 2000}) {
                const int smem = bigsmem ? 220 * 1024 : S * slot;
                for (int rep = 0; rep < 2; ++rep) {
                    if (cl) pull_cluster<<<148, 64, smem>>>(buf, (size_t)2 << 20, slot, S, iters, hold, cyc, lat);
                    else pull_plain<<<148, 64, smem>>>(buf, (size_t)2 << 20, slot, S, iters, hold, cyc, lat);
                    cudaError_t err = cudaDeviceSynchronize();
                    unsigned long long h[148], l[148];
                    cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost); cudaMemcpy(l, lat, sizeof(l), cudaMemcpyDeviceToHost);
                    double avg = 0, la = 0; for (int i = 0; i < 148; ++i) { avg += h[i]; la += l[i]; } avg /= 148; la /= 148;
                    if (rep) printf("cluster=%d smem=%3d KB hold=%4d: %.1f cycles/op  %.1f B/clk/SM  issue->landed %.0f cycles (%s)\n", cl, smem / 1024, hold, avg / iters, (double)iters * slot / avg, la, cudaGetErrorString(err));
                }
            }
    return 0;
}
