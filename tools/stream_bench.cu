// How fast can one CTA per SM stream 128-row tiles of a [n,100] fp32 matrix into shared memory?
// (a) four 2-D TMA boxes per tile (3 x 32 columns SWIZZLE_128B + 8 columns SWIZZLE_32B), (b) one 1-D bulk copy per tile.
#include <cstdio>
#include <cstdint>
#include <cuda.h>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile("{\n\t.reg .pred p;\n\tW:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n\t@p bra D;\n\tbra W;\n\tD:\n\t}" ::"r"(bar), "r"(parity), "r"(0x989680) : "memory");
}
template <int MODE, int STAGES>
__global__ void __launch_bounds__(64, 1) stream(const __grid_constant__ CUtensorMap tmF, const __grid_constant__ CUtensorMap tmT,
                                                const float* X, long long num_tiles, int d, float* sink) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    __shared__ __align__(8) uint64_t full[STAGES], empty[STAGES];
    const int stage_bytes = MODE == 0 ? 3 * 16384 + 4096 : 128 * d * 4;
    if (threadIdx.x == 0) { for (int s = 0; s < STAGES; ++s) { asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&full[s]))); asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&empty[s]))); } asm volatile("fence.mbarrier_init.release.cluster;"); }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0) {
        if (lane == 0) {
            long long it = 0;
            for (long long tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
                const int s = it % STAGES;
                mbar_wait(smem_u32(&empty[s]), ((it / STAGES) & 1) ^ 1);
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&full[s])), "r"(stage_bytes) : "memory");
                const uint32_t dst = smem_u32(smem) + s * ((stage_bytes + 1023) & ~1023);
                if (MODE == 0) {
                    for (int a = 0; a < 3; ++a)
                        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst + a * 16384), "l"(&tmF), "r"(smem_u32(&full[s])), "r"(a * 32), "r"((int)(tile * 128)) : "memory");
                    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst + 3 * 16384), "l"(&tmT), "r"(smem_u32(&full[s])), "r"(96), "r"((int)(tile * 128)) : "memory");
                } else {
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(X + tile * 128 * d), "r"(stage_bytes), "r"(smem_u32(&full[s])) : "memory");
                }
            }
        }
    } else {
        float acc = 0.f; long long it = 0;
        for (long long tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
            const int s = it % STAGES;
            mbar_wait(smem_u32(&full[s]), (it / STAGES) & 1);
            acc += *reinterpret_cast<float*>(smem + s * ((stage_bytes + 1023) & ~1023) + lane * 4);
            __syncwarp();
            if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&empty[s])) : "memory");
        }
        if (acc == 0.12345f) sink[0] = acc;
    }
}
// Same stream, but every stage is held `hold` cycles after it has landed (one consumer warp per stage, so holds
// overlap like the per-tile processing chain of the fused kernels): throughput vs. chain length at 3 stages.
template <int STAGES>
__global__ void __launch_bounds__(32 * (STAGES + 1), 1) stream_hold(const __grid_constant__ CUtensorMap tmF, const __grid_constant__ CUtensorMap tmT,
                                                                   long long num_tiles, int hold, float* sink) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    __shared__ __align__(8) uint64_t full[STAGES], empty[STAGES];
    const int stage_bytes = 3 * 16384 + 4096;
    if (threadIdx.x == 0) { for (int s = 0; s < STAGES; ++s) { asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&full[s]))); asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&empty[s]))); } asm volatile("fence.mbarrier_init.release.cluster;"); }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0) {
        if (lane == 0) {
            long long it = 0;
            for (long long tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
                const int s = it % STAGES;
                mbar_wait(smem_u32(&empty[s]), ((it / STAGES) & 1) ^ 1);
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&full[s])), "r"(stage_bytes) : "memory");
                const uint32_t dst = smem_u32(smem) + s * stage_bytes;
                for (int a = 0; a < 3; ++a)
                    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst + a * 16384), "l"(&tmF), "r"(smem_u32(&full[s])), "r"(a * 32), "r"((int)(tile * 128)) : "memory");
                asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst + 3 * 16384), "l"(&tmT), "r"(smem_u32(&full[s])), "r"(96), "r"((int)(tile * 128)) : "memory");
            }
        }
    } else {
        const int s = warp - 1;
        float acc = 0.f; long long it = s;
        for (long long tile = blockIdx.x + (long long)s * gridDim.x; tile < num_tiles; tile += (long long)STAGES * gridDim.x, it += STAGES) {
            mbar_wait(smem_u32(&full[s]), (it / STAGES) & 1);
            const long long t0 = clock64();
            acc += *reinterpret_cast<float*>(smem + s * stage_bytes + lane * 4);
            while (clock64() - t0 < hold) {}
            __syncwarp();
            if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&empty[s])) : "memory");
        }
        if (acc == 0.12345f) sink[0] = acc;
    }
}
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main() {
    const long long n = 20000000; const int d = 100;
    float* X; cudaMalloc(&X, n * d * 4); cudaMemset(X, 0, n * d * 4); float* sink; cudaMalloc(&sink, 4);
    void* fp; cudaDriverEntryPointQueryResult q; cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q); EncodeTiledFn enc = (EncodeTiledFn)fp;
    CUtensorMap tmF, tmT; cuuint64_t dims[2] = {(cuuint64_t)d, (cuuint64_t)n}; cuuint64_t str[1] = {(cuuint64_t)d * 4}; cuuint32_t es[2] = {1, 1};
    cuuint32_t boxF[2] = {32, 128}, boxT[2] = {8, 128};
    enc(&tmF, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, X, dims, str, boxF, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    enc(&tmT, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, X, dims, str, boxT, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    const long long tiles = n / 128; cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
#define RUN(MODE, ST) { auto k = stream<MODE, ST>; cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024); \
    k<<<148, 64, ST * 53248 + 1024>>>(tmF, tmT, X, tiles, d, sink); cudaEventRecord(e0); k<<<148, 64, ST * 53248 + 1024>>>(tmF, tmT, X, tiles, d, sink); cudaEventRecord(e1); \
    cudaError_t err = cudaDeviceSynchronize(); float ms; cudaEventElapsedTime(&ms, e0, e1); \
    printf("%s, %d stages: %.3f ms, %.0f GB/s (%s)\n", MODE == 0 ? "4 TMA boxes/tile" : "1 bulk copy/tile", ST, ms, n * d * 4.0 / ms / 1e6, cudaGetErrorString(err)); }
    RUN(0, 2) RUN(0, 3) RUN(0, 4) RUN(1, 2) RUN(1, 3) RUN(1, 4)
#define RUNH(ST, HOLD) { auto k = stream_hold<ST>; cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024); \
    k<<<148, 32 * (ST + 1), ST * 53248 + 1024>>>(tmF, tmT, tiles, HOLD, sink); cudaEventRecord(e0); k<<<148, 32 * (ST + 1), ST * 53248 + 1024>>>(tmF, tmT, tiles, HOLD, sink); cudaEventRecord(e1); \
    cudaError_t err = cudaDeviceSynchronize(); float ms; cudaEventElapsedTime(&ms, e0, e1); \
    printf("%d stages held %5d cycles after landing: %.3f ms, %.0f GB/s, %.0f cycles/tile at 1.9 GHz (%s)\n", ST, HOLD, ms, n * d * 4.0 / ms / 1e6, ms * 1e-3 * 1.9e9 / (tiles / 148.0), cudaGetErrorString(err)); }
    RUNH(3, 0) RUNH(3, 1000) RUNH(3, 2000) RUNH(3, 3000) RUNH(3, 4000) RUNH(3, 6000) RUNH(3, 8000) RUNH(4, 4000) RUNH(4, 8000) RUNH(2, 4000)
    return 0;
}
