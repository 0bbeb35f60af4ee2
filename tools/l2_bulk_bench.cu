// How fast can every SM pull the SAME (L2-resident) operand through bulk copies into a shared-memory ring?
// modes: 0 = all SMs walk one shared buffer in the same order, 1 = same buffer, per-SM rotated start,
//        2 = every SM its own region (L2-resident in total), 3 = every SM its own region of a buffer larger than L2.
// usage: l2_bulk_bench [slot_bytes] [slots] [buffer_bytes_shared]
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile("{\n\t.reg .pred p;\n\tW:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n\t@p bra D;\n\tbra W;\n\tD:\n\t}" ::"r"(bar), "r"(parity), "r"(0x989680) : "memory");
}
__global__ void __launch_bounds__(64, 1) pull(const uint8_t* buf, size_t region_bytes, size_t region_stride, int rotate, int slot_bytes,
                                              int S, int iters, unsigned long long* cycles) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t full[16], empty[16];
    if (threadIdx.x == 0) {
        for (int s = 0; s < S; ++s) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&full[s])));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&empty[s])));
        }
        asm volatile("fence.mbarrier_init.release.cluster;");
    }
    __syncthreads();
    const uint8_t* base = buf + (size_t)blockIdx.x * region_stride;
    const size_t nblk = region_bytes / slot_bytes;
    const size_t start = rotate ? ((size_t)blockIdx.x * 7919u) % nblk : 0;
    const long long t0 = clock64();
    if (threadIdx.x == 0) {
        for (int it = 0; it < iters; ++it) {
            const int s = it % S;
            mbar_wait(smem_u32(&empty[s]), ((it / S) & 1) ^ 1);
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&full[s])), "r"(slot_bytes) : "memory");
            const uint8_t* src = base + ((start + it) % nblk) * (size_t)slot_bytes;
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem) + s * slot_bytes), "l"(src), "r"(slot_bytes), "r"(smem_u32(&full[s])) : "memory");
        }
    } else if (threadIdx.x == 32) {
        for (int it = 0; it < iters; ++it) {
            const int s = it % S;
            mbar_wait(smem_u32(&full[s]), (it / S) & 1);
            asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&empty[s])) : "memory");
        }
        cycles[blockIdx.x] = (unsigned long long)(clock64() - t0);
    }
}
__global__ void fill(float* p, size_t n, int wt) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        if (wt) asm volatile("st.global.wt.f32 [%0], %1;" ::"l"(p + i), "f"(1.f) : "memory");
        else p[i] = 1.f;
    }
}
int main(int argc, char** argv) {
    const int slot = argc > 1 ? atoi(argv[1]) : 16384;
    const int S = argc > 2 ? atoi(argv[2]) : 5;
    const size_t shared_bytes = argc > 3 ? (size_t)atoll(argv[3]) : (size_t)2 << 20;
    const int iters = 4000;
    uint8_t* buf;
    const size_t big = (size_t)148 * ((size_t)64 << 20);
    cudaMalloc(&buf, big);
    cudaMemset(buf, 1, big);
    unsigned long long* cyc;
    cudaMalloc(&cyc, 148 * 8);
    cudaFuncSetAttribute(pull, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    struct { const char* name; size_t region, stride; int rot; } modes[] = {
        {"shared buffer, same order", shared_bytes, 0, 0},
        {"shared buffer, rotated start", shared_bytes, 0, 1},
        {"private 512 KB regions (74 MB total, L2)", (size_t)512 << 10, (size_t)512 << 10, 0},
        {"private 64 MB regions (DRAM)", (size_t)64 << 20, (size_t)64 << 20, 0},
    };
    for (int dirty = 0; dirty < 3; ++dirty) {
      printf("---- shared region %s\n", dirty == 0 ? "clean (written long ago)" : dirty == 1 ? "freshly written by a kernel (plain stores)" : "freshly written with st.wt");
    for (auto& m : modes) {
        for (int rep = 0; rep < 2; ++rep) {
            if (dirty) { fill<<<148, 256>>>((float*)buf, (m.stride ? (size_t)148 * m.region : m.region) / 4 > ((size_t)80 << 20) ? ((size_t)80 << 20) : (m.stride ? (size_t)148 * m.region : m.region) / 4, dirty == 2); }
            cudaEvent_t e0, e1;
            cudaEventCreate(&e0); cudaEventCreate(&e1);
            cudaEventRecord(e0);
            pull<<<148, 64, S * slot>>>(buf, m.region, m.stride, m.rot, slot, S, iters, cyc);
            cudaEventRecord(e1);
            cudaError_t err = cudaDeviceSynchronize();
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            unsigned long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
            double avg = 0; for (auto c : h) avg += c; avg /= 148;
            if (rep) printf("%-44s slot %d x %d: %.3f ms  %.2f TB/s  %.1f B/clk/SM  (%s)\n", m.name, slot, S, ms,
                            148.0 * iters * slot / ms / 1e9, (double)iters * slot / avg, cudaGetErrorString(err));
        }
    }
    }
    return 0;
}
