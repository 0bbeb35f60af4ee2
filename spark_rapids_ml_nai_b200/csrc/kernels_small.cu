// Whole-fit kernel for the latency regime (BASELINE.json C1: 100k x 20, k = 8, 20 iterations): when one GPU's
// shard fits the shared memory of all SMs together, ONE cooperative launch runs every Lloyd iteration.
//
// Each CTA loads its rows into shared memory once and keeps them there for the whole fit.  Per iteration:
//   assign      one thread per row, squared distances (x - c)^2 in float64 straight from shared memory -
//               exact labels, lowest index on ties, no screening needed at this size;
//   accumulate  thread (cluster, column, group) adds its rows in row order (no atomics), the groups are added in
//               group order: this CTA's float64 partial sums and counts;
//   exchange    the partial goes to the CTA's slot in global memory, one grid-wide barrier;
//   update      EVERY CTA adds the slots in slot order and computes the new centroids itself (mean rounded to
//               float32, or kept when the count is 0), the shift^2 in index order and the stop test - the same
//               bits everywhere, so no second barrier and no broadcast.
// The reference's loop runs one kernel sequence plus a host read of the shift per iteration; here an
// iteration costs one grid barrier.  Results are bitwise reproducible.  Replaces KMeansMG.fit behind reference
// python/src/spark_rapids_ml/clustering.py:367-370 for small single-GPU fits.
#include "common.cuh"

#include <cooperative_groups.h>

#include <cfloat>
#include <cstdio>

namespace cg = cooperative_groups;

namespace b200km {

namespace {

thread_local char g_sm_err[512] = "";

constexpr int kSmThreads = 704;
constexpr int kSmMaxSmem = 200 * 1024;
constexpr int kSmMaxAcc = 2048;  // k*d + k + 1 float64 per slot at most

struct SmallParams {
    const float* X;
    int64_t n;
    int d, k;
    int rows_per_cta;  // ceil(n / grid)
    int groups;        // row groups of the accumulation
    int xs_stride;     // floats per row in shared memory (odd: conflict-free column walks)
    const float* init;
    int max_iter;
    double tol;
    double* slots;  // [2 parities][grid][acc_stride], then [grid] inertia partials, then [2 parities][acc_stride] totals
    unsigned int* bar;  // grid barrier counter (zero at launch)
    int acc_stride;
    float* C_out;         // [k, d] working + final centroids (device)
    int32_t* labels_out;  // may be null
    double* inertia_out;  // device scalar, may be null
    IterState* st;
};

// Nearest centroid of one row in float64: score_j = ||c_j||^2 - 2 x.c_j (four centroids at a time: independent FMA
// chains, one conversion of x per four centroids); ties keep the lowest index.  Returns the squared distance too.
__device__ __forceinline__ int nearest_f64(const float* __restrict__ x, const double* __restrict__ cs,
                                           const double* __restrict__ cn, int k, int d, double* best_out) {
    double best = DBL_MAX, xn = 0.0;
    int bi = 0;
    int j = 0;
    for (; j + 4 <= k; j += 4) {
        const double* c0 = cs + j * d;
        const double* c1 = c0 + d;
        const double* c2 = c1 + d;
        const double* c3 = c2 + d;
        double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0, q = 0.0;
        int i = 0;
        if ((d & 1) == 0) {
            // two dimensions per step: the centroid values come as 16-byte shared-memory loads (the loop is bound by
            // the number of shared-memory instructions, not by the FMAs)
            for (; i < d; i += 2) {
                const double xa = (double)x[i], xb = (double)x[i + 1];
                const double2 v0 = *reinterpret_cast<const double2*>(c0 + i);
                const double2 v1 = *reinterpret_cast<const double2*>(c1 + i);
                const double2 v2 = *reinterpret_cast<const double2*>(c2 + i);
                const double2 v3 = *reinterpret_cast<const double2*>(c3 + i);
                q = fma(xb, xb, fma(xa, xa, q));
                s0 = fma(xb, v0.y, fma(xa, v0.x, s0));
                s1 = fma(xb, v1.y, fma(xa, v1.x, s1));
                s2 = fma(xb, v2.y, fma(xa, v2.x, s2));
                s3 = fma(xb, v3.y, fma(xa, v3.x, s3));
            }
        }
        for (; i < d; ++i) {
            const double xv = (double)x[i];
            q = fma(xv, xv, q);
            s0 = fma(xv, c0[i], s0);
            s1 = fma(xv, c1[i], s1);
            s2 = fma(xv, c2[i], s2);
            s3 = fma(xv, c3[i], s3);
        }
        xn = q;
        s0 = fma(-2.0, s0, cn[j]);
        s1 = fma(-2.0, s1, cn[j + 1]);
        s2 = fma(-2.0, s2, cn[j + 2]);
        s3 = fma(-2.0, s3, cn[j + 3]);
        if (s0 < best) { best = s0; bi = j; }
        if (s1 < best) { best = s1; bi = j + 1; }
        if (s2 < best) { best = s2; bi = j + 2; }
        if (s3 < best) { best = s3; bi = j + 3; }
    }
    for (; j < k; ++j) {
        const double* c = cs + j * d;
        double sacc = 0.0, q = 0.0;
        for (int i = 0; i < d; ++i) {
            const double xv = (double)x[i];
            q = fma(xv, xv, q);
            sacc = fma(xv, c[i], sacc);
        }
        xn = q;
        sacc = fma(-2.0, sacc, cn[j]);
        if (sacc < best) { best = sacc; bi = j; }
    }
    const double dist = best + xn;
    *best_out = dist > 0.0 ? dist : 0.0;
    return bi;
}

// Grid-wide barrier for a cooperative launch (all CTAs resident): arrive on a monotonic counter, one thread polls.
// Lighter than a cooperative-groups grid sync (measured ~5500 cycles here); data written before it is read with
// __ldcg (L2) after it.
__device__ __forceinline__ void grid_barrier(unsigned int* bar, unsigned int target) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(bar, 1u);
        unsigned int v;
        do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(bar) : "memory");
        } while (v < target);
    }
    __syncthreads();
}

// Sum of one float64 per thread over the CTA in a fixed order (warp trees, then the warp sums in warp order);
// every thread gets the result.
__device__ __forceinline__ double block_sum_fixed(double v, double* red) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    double t = 0.0;
    for (int w = 0; w < kSmThreads / 32; ++w) t += red[w];
    __syncthreads();
    return t;
}

__global__ void __launch_bounds__(kSmThreads, 1) kmeans_small_fit_kernel(const SmallParams p) {
    unsigned int bar_target = 0;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int k = p.k, d = p.d, kd = k * d, acc_n = kd + k;
    // layout: centroids f64 [kd] | their norms f64 [k] | partial f64 [groups][acc_n] | reduce scratch f64 [512] |
    // rows f32 | labels u8
    double* cs = reinterpret_cast<double*>(smem_raw);
    double* cnrm = cs + kd;
    const int groups = p.groups;  // row groups of the accumulation = slot parts of the update
    double* part = cnrm + k;
    double* red = part + (size_t)groups * acc_n;
    float* xs = reinterpret_cast<float*>(red + kSmThreads);
    unsigned char* lab = reinterpret_cast<unsigned char*>(xs + (size_t)p.rows_per_cta * p.xs_stride);

    const int tid = threadIdx.x;
    const int64_t row0 = (int64_t)blockIdx.x * p.rows_per_cta;
    const int rows = (int)(p.n - row0 < p.rows_per_cta ? (p.n - row0 > 0 ? p.n - row0 : 0) : p.rows_per_cta);
    for (int i = tid; i < rows * d; i += kSmThreads) {
        const int r = i / d, c = i - r * d;
        xs[r * p.xs_stride + c] = p.X[(row0 + r) * d + c];
    }
    for (int i = tid; i < kd; i += kSmThreads) cs[i] = (double)p.init[i];
    __syncthreads();
    auto norms = [&]() {  // ||c_j||^2 in index order: identical in every CTA
        for (int j = tid; j < k; j += kSmThreads) {
            double q = 0.0;
            for (int i = 0; i < d; ++i) q = fma(cs[j * d + i], cs[j * d + i], q);
            cnrm[j] = q;
        }
        __syncthreads();
    };
    norms();

    int n_iter = 0;
    bool done = false;
    double shift2 = 0.0;
#ifdef B200KM_ENABLE_TRACE
    long long tph[6] = {0, 0, 0, 0, 0, 0}, tmark = clock64();
#define SM_MARK(i) do { const long long now_ = clock64(); tph[i] += now_ - tmark; tmark = now_; } while (0)
#else
#define SM_MARK(i) do { } while (0)
#endif
    for (int it = 0; it < p.max_iter && !done; ++it) {
        // ---- assign ----
        for (int r = tid; r < rows; r += kSmThreads) {
            double best;
            lab[r] = (unsigned char)nearest_f64(xs + r * p.xs_stride, cs, cnrm, k, d, &best);
        }
        __syncthreads();
        SM_MARK(0);
        // ---- accumulate: thread (group g, column c) owns column c of every cluster in partial g and adds the rows
        // g, g + groups, ... in order (each row element is touched once; no atomics, fixed order) ----
        for (int u = tid; u < groups * acc_n; u += kSmThreads) part[u] = 0.0;
        __syncthreads();
        for (int u = tid; u < groups * (d + 1); u += kSmThreads) {
            const int g = u / (d + 1), c = u - g * (d + 1);
            double* pg = part + g * acc_n;
            if (c < d) {
                for (int r = g; r < rows; r += groups) pg[lab[r] * d + c] += (double)xs[r * p.xs_stride + c];
            } else {
                for (int r = g; r < rows; r += groups) pg[kd + lab[r]] += 1.0;
            }
        }
        __syncthreads();
        SM_MARK(1);
        double* slot = p.slots + ((size_t)(it & 1) * gridDim.x + blockIdx.x) * p.acc_stride;
        for (int e = tid; e < acc_n; e += kSmThreads) {
            double s = 0.0;
            for (int g = 0; g < groups; ++g) s += part[g * acc_n + e];
            slot[e] = s;
        }
        SM_MARK(2);
        bar_target += gridDim.x;
        grid_barrier(p.bar, bar_target);
        SM_MARK(3);
        // ---- every CTA adds a slice of the elements over all slots, in slot order (one warp per element: lane l
        // takes slots l, l + 32, ..., then a fixed tree), a second barrier, and everyone reads the totals ----
        const double* base = p.slots + (size_t)(it & 1) * gridDim.x * p.acc_stride;
        double* totals = p.slots + (size_t)2 * gridDim.x * p.acc_stride + gridDim.x + (size_t)(it & 1) * p.acc_stride;
        {
            const int lane = tid & 31, warp = tid >> 5;
            for (int e = (int)blockIdx.x + warp * (int)gridDim.x; e < acc_n; e += (int)gridDim.x * (kSmThreads / 32)) {
                double t = 0.0;
                for (int b = lane; b < (int)gridDim.x; b += 32) t += __ldcg(base + (size_t)b * p.acc_stride + e);
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) t += __shfl_down_sync(0xffffffffu, t, o);
                if (lane == 0) totals[e] = t;
            }
        }
        bar_target += gridDim.x;
        grid_barrier(p.bar, bar_target);
        for (int e = tid; e < acc_n; e += kSmThreads) part[e] = __ldcg(totals + e);
        __syncthreads();
        SM_MARK(4);
        double sh = 0.0;
        for (int e = tid; e < kd; e += kSmThreads) {
            const int j = e / d;
            const double ssum = part[e], cnt = part[kd + j];
            const double old = cs[e];
            const double nw = cnt > 0.0 ? (double)(float)(ssum / cnt) : old;
            const double df = nw - old;
            sh += df * df;
            cs[e] = nw;  // each element is read (old) and written by the same thread only
        }
        shift2 = block_sum_fixed(sh, red);  // fixed order, identical in every CTA (also orders the writes of cs)
        norms();
        SM_MARK(5);
        ++n_iter;
        done = shift2 < p.tol;
    }
#ifdef B200KM_ENABLE_TRACE
    if (blockIdx.x == 0 && tid == 0 && n_iter > 0)
        printf("resident fit, cycles per iteration: assign %lld accumulate %lld slot %lld barrier %lld slots->part %lld update %lld\n",
               tph[0] / n_iter, tph[1] / n_iter, tph[2] / n_iter, tph[3] / n_iter, tph[4] / n_iter, tph[5] / n_iter);
#endif

    // ---- final assignment against the final centroids: labels and inertia ----
    if (p.labels_out || p.inertia_out) {
        double inert = 0.0;
        for (int r = tid; r < rows; r += kSmThreads) {
            double best;
            const int bi = nearest_f64(xs + r * p.xs_stride, cs, cnrm, k, d, &best);
            if (p.labels_out) p.labels_out[row0 + r] = bi;
            inert += best;
        }
        if (p.inertia_out) {
            const double t = block_sum_fixed(inert, red);
            if (tid == 0) p.slots[((size_t)2 * gridDim.x) * p.acc_stride + blockIdx.x] = t;
            bar_target += gridDim.x;
            grid_barrier(p.bar, bar_target);
            if (blockIdx.x == 0 && tid == 0) {
                double tot = 0.0;
                for (unsigned int b = 0; b < gridDim.x; ++b) tot += __ldcg(p.slots + ((size_t)2 * gridDim.x) * p.acc_stride + b);
                *p.inertia_out = tot;
            }
        }
    }
    if (blockIdx.x == 0) {
        for (int i = tid; i < kd; i += kSmThreads) p.C_out[i] = (float)cs[i];
        if (tid == 0) {
            p.st->n_iter = n_iter;
            p.st->done = done ? 1 : 0;
            p.st->shift2 = shift2;
        }
    }
}

struct SmallPlan {
    int grid, rows_per_cta, xs_stride, acc_stride, groups;
    size_t smem;
    bool ok;
};

SmallPlan small_plan(int64_t n, int d, int k) {
    SmallPlan pl = {};
    pl.ok = false;
    const int64_t acc_n = (int64_t)k * d + k;
    if (n < 1 || k > 255 || acc_n + 1 > kSmMaxAcc || acc_n > kSmThreads * 4) return pl;
    pl.grid = kNumSMs;
    if (n < pl.grid * 32) pl.grid = (int)((n + 31) / 32);
    pl.rows_per_cta = (int)((n + pl.grid - 1) / pl.grid);
    pl.grid = (int)((n + pl.rows_per_cta - 1) / pl.rows_per_cta);
    pl.xs_stride = d | 1;
    pl.acc_stride = (int)((acc_n + 1 + 7) & ~7);
    pl.groups = kSmThreads / (d + 1) > 0 ? kSmThreads / (d + 1) : 1;
    while (pl.groups > 1 && (size_t)pl.groups * acc_n * sizeof(double) > 40 * 1024) pl.groups--;
    pl.smem = sizeof(double) * ((size_t)k * d + k + (size_t)pl.groups * acc_n + kSmThreads) +
              sizeof(float) * (size_t)pl.rows_per_cta * pl.xs_stride + (size_t)pl.rows_per_cta + 64;
    pl.ok = pl.smem <= (size_t)kSmMaxSmem;
    return pl;
}

}  // namespace

const char* small_last_error() { return g_sm_err; }

bool small_fit_supported(int64_t n, int d, int k) { return small_plan(n, d, k).ok; }

size_t small_slots_bytes(int64_t n, int d, int k) {
    const SmallPlan pl = small_plan(n, d, k);
    return pl.ok ? sizeof(double) * ((size_t)2 * pl.grid * pl.acc_stride + pl.grid + 2 * (size_t)pl.acc_stride + 8) + 64 : 0;
}

// One cooperative launch: the whole Lloyd loop (and, when asked, the final labels / inertia).  slots: scratch of
// small_slots_bytes().  C_out receives the final centroids, st the iteration count / stop flag / last shift^2.
int small_fit(const float* X, int64_t n, int d, int k, const float* init, int max_iter, double tol, double* slots,
              float* C_out, int32_t* labels_out, double* inertia_out, IterState* st, cudaStream_t s) {
    const SmallPlan pl = small_plan(n, d, k);
    if (!pl.ok) {
        snprintf(g_sm_err, sizeof(g_sm_err), "shape n=%lld d=%d k=%d does not fit the shared-memory resident fit", (long long)n, d, k);
        return 1;
    }
    static PerDeviceOnce attr_set;  // the opt-in is a per-device attribute
    const int dev = PerDeviceOnce::current();
    if (dev < 0 || !attr_set.done[dev]) {
        if (cudaFuncSetAttribute(kmeans_small_fit_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmMaxSmem) != cudaSuccess) {
            snprintf(g_sm_err, sizeof(g_sm_err), "cannot raise dynamic shared memory to %d bytes", kSmMaxSmem);
            return 1;
        }
        if (dev >= 0) attr_set.done[dev] = true;
    }
    SmallParams p;
    p.X = X;
    p.n = n;
    p.d = d;
    p.k = k;
    p.rows_per_cta = pl.rows_per_cta;
    p.groups = pl.groups;
    p.xs_stride = pl.xs_stride;
    p.init = init;
    p.max_iter = max_iter;
    p.tol = tol;
    p.slots = slots;
    // the barrier counter lives behind the slots; it starts at zero
    p.bar = reinterpret_cast<unsigned int*>(slots + (size_t)2 * pl.grid * pl.acc_stride + pl.grid + 2 * (size_t)pl.acc_stride + 4);
    cudaMemsetAsync(p.bar, 0, sizeof(unsigned int), s);
    p.acc_stride = pl.acc_stride;
    p.C_out = C_out;
    p.labels_out = labels_out;
    p.inertia_out = inertia_out;
    p.st = st;
    void* args[] = {(void*)&p};
    cudaError_t e = cudaLaunchCooperativeKernel((void*)kmeans_small_fit_kernel, dim3(pl.grid), dim3(kSmThreads), args, pl.smem, s);
    count_launch();
    if (e != cudaSuccess) {
        snprintf(g_sm_err, sizeof(g_sm_err), "cooperative launch of the resident fit failed: %s", cudaGetErrorString(e));
        return 1;
    }
    return 0;
}

}  // namespace b200km
