// b200km_predict_static: nearest-centroid labels without a single allocation.
//
// Some callers of this library are not in a position to allocate device memory: a transform task that
// shares its GPU with a memory pool which already owns all of HBM (cudaMalloc would fail), a serving
// thread with a latency budget (cudaMalloc takes a driver lock and may synchronise the whole device), or
// code that must not fail half-way for lack of memory.  The regular entry points grow a per-handle
// workspace with cudaMalloc on first use; this one keeps its scratch in a fixed array that is part of the
// loaded module instead.  Every GPU that loads libb200km.so holds that array until the process exits.
#include "../../include/b200km.h"
#include "common.cuh"

#include <cstdarg>
#include <cstdio>

using namespace b200km;

namespace {

constexpr int64_t kMaxRows = B200KM_STATIC_MAX_ROWS;
constexpr int kMaxCentroids = B200KM_STATIC_MAX_K;

// Layout of the scratch: iteration state | ||c||^2 [kMaxCentroids] | ||x||^2 [kMaxRows] | distances
// [kMaxRows] | near-tie work list [kMaxRows].
constexpr size_t kStateBytes = 256;
constexpr size_t kOffCnorm = kStateBytes;
constexpr size_t kOffXnorm = kOffCnorm + sizeof(float) * kMaxCentroids;
constexpr size_t kOffBest = kOffXnorm + sizeof(float) * kMaxRows;
constexpr size_t kOffWork = kOffBest + sizeof(float) * kMaxRows;
constexpr size_t kScratchBytes = kOffWork + sizeof(unsigned int) * kMaxRows;
static_assert(sizeof(IterState) <= kStateBytes, "state block too small");
static_assert(kScratchBytes <= (1u << 20), "the static scratch is meant to stay modest (<= 1 MiB)");

// In host code this name is only a handle for the runtime, not a device address: the address the
// kernels need comes from cudaGetSymbolAddress, once per device (below).
static __device__ __align__(256) unsigned char g_static_scratch[kScratchBytes];

constexpr int kMaxDevices = 64;
unsigned char* g_scratch_addr[kMaxDevices] = {};

thread_local char g_static_err[512] = "";

int fail(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_static_err, sizeof(g_static_err), fmt, ap);
    va_end(ap);
    return 1;
}

}  // namespace

extern "C" {

const char* b200km_static_last_error(void) { return g_static_err; }

int b200km_predict_static(const float* X, int64_t n, int d, const float* centroids, int k, int32_t* labels_out,
                          void* stream) {
    if (n < 0 || n > kMaxRows)
        return fail("n=%lld exceeds the static scratch (at most %lld rows per call; split the batch)", (long long)n,
                    (long long)kMaxRows);
    if (k <= 0 || k > kMaxCentroids)
        return fail("k=%d exceeds the static scratch (at most %d centroids)", k, kMaxCentroids);
    if (d <= 0 || d > (1 << 20)) return fail("d=%d out of range", d);
    if (n == 0) return 0;
    if (!X || !centroids || !labels_out) return fail("null pointer argument");

    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail("cudaGetDevice failed: %s", cudaGetErrorString(e));
    if (dev < 0 || dev >= kMaxDevices) return fail("device ordinal %d not supported (max %d)", dev, kMaxDevices);
    if (!g_scratch_addr[dev]) {
        // one-time set-up for this device (it is the current one): resolve the array's device address
        int major = 0;
        e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
        if (e != cudaSuccess) return fail("cudaDeviceGetAttribute failed: %s", cudaGetErrorString(e));
        if (major != 10) return fail("b200km is built for sm_100a only; device %d has compute capability %d.x", dev, major);
        unsigned char* ptr = nullptr;
        e = cudaGetSymbolAddress((void**)&ptr, g_static_scratch);
        if (e != cudaSuccess || !ptr)
            return fail("cudaGetSymbolAddress(g_static_scratch) failed on device %d: %s", dev, cudaGetErrorString(e));
        g_scratch_addr[dev] = ptr;
    }
    unsigned char* base = g_scratch_addr[dev];
    IterState* st = reinterpret_cast<IterState*>(base);
    float* cnorm = reinterpret_cast<float*>(base + kOffCnorm);
    float* xnorm = reinterpret_cast<float*>(base + kOffXnorm);
    float* best = reinterpret_cast<float*>(base + kOffBest);
    unsigned int* work = reinterpret_cast<unsigned int*>(base + kOffWork);

    cudaStream_t s = (cudaStream_t)stream;
    // the array keeps whatever the previous call left: the election ticket, the work-list counter and the
    // done flag have to start at zero
    e = cudaMemsetAsync(st, 0, kStateBytes, s);
    if (e != cudaSuccess) return fail("cudaMemsetAsync failed: %s", cudaGetErrorString(e));
    launch_rownorm(X, n, d, xnorm, s);
    launch_prep(centroids, k, d, cnorm, nullptr, 0, st, false, s);
    launch_assign_ffma(X, n, d, centroids, k, cnorm, xnorm, labels_out, best, st, work, false, s);
    launch_refine(X, n, d, centroids, k, labels_out, best, st, work, false, s);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail("kernel launch failed: %s", cudaGetErrorString(e));
    return 0;
}

}  // extern "C"
