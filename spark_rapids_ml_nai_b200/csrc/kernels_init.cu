// Device kernels of the k-means|| ("scalable k-means++") initialisation (orchestrated by b200km_init_centroids).
//
// Random numbers come from a counter-based hash keyed by (seed, stream, counter) - the counter of a row is its
// GLOBAL index, so the result does not depend on how the rows are sharded across ranks (the CPU checker under
// tests/ restates the same hash).  Everything that feeds a decision is reduced in a fixed order.
//
// Replaces cuML's initScalableKMeansPlusPlus behind init="scalable-k-means++" (reference
// python/src/spark_rapids_ml/clustering.py:72,117-119,290).
#include "common.cuh"

namespace b200km {

namespace {

__host__ __device__ __forceinline__ unsigned long long mix64(unsigned long long z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
__host__ __device__ __forceinline__ unsigned long long hash64(unsigned long long seed, unsigned long long stream,
                                                              unsigned long long counter) {
    return mix64(mix64(mix64(seed) ^ stream) ^ counter);
}
__host__ __device__ __forceinline__ double hash_uniform(unsigned long long seed, unsigned long long stream,
                                                        unsigned long long counter) {
    return (double)(hash64(seed, stream, counter) >> 11) * (1.0 / 9007199254740992.0);
}

__device__ __forceinline__ double block_sum_256(double v, double* s_red) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = v;
    __syncthreads();
    double t = 0.0;
    if (threadIdx.x == 0)
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += s_red[w];
    __syncthreads();
    return t;  // valid in thread 0
}

// Bernoulli sampling of one round: row i is picked when u(seed, round, global row) < scale * dist[i].
__global__ void __launch_bounds__(256) kpar_sample_kernel(const float* __restrict__ dist, int64_t n, long long row0,
                                                          unsigned long long seed, int stream, double scale,
                                                          long long* __restrict__ list, unsigned int* __restrict__ count,
                                                          unsigned int cap) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const double pr = scale * (double)dist[i];
        if (hash_uniform(seed, (unsigned long long)stream, (unsigned long long)(row0 + i)) < pr) {
            const unsigned int pos = atomicAdd(count, 1u);
            if (pos < cap) list[pos] = row0 + i;
        }
    }
}

__global__ void __launch_bounds__(256) min_inplace_kernel(float* __restrict__ dist, const float* __restrict__ other, int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        dist[i] = fminf(dist[i], other[i]);
}

// Fixed-order float64 sum of a float array: block b adds the slab [b*per, (b+1)*per) (thread-strided, then a
// fixed tree); the last block to finish adds the block sums in index order.
__global__ void __launch_bounds__(256) det_sum_kernel(const float* __restrict__ x, int64_t n, int64_t per,
                                                      double* __restrict__ partial, double* __restrict__ out,
                                                      unsigned int* __restrict__ ticket) {
    __shared__ double s_red[8];
    __shared__ bool last;
    const int64_t lo = (int64_t)blockIdx.x * per;
    const int64_t hi = lo + per < n ? lo + per : n;
    double t = 0.0;
    for (int64_t i = lo + threadIdx.x; i < hi; i += 256) t += (double)x[i];
    t = block_sum_256(t, s_red);
    if (threadIdx.x == 0) {
        partial[blockIdx.x] = t;
        __threadfence();
        last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (last && threadIdx.x == 0) {
        __threadfence();
        double s = 0.0;
        for (unsigned int b = 0; b < gridDim.x; ++b) s += ((volatile double*)partial)[b];
        *out = s;
        *ticket = 0;
    }
}

// out[j] = row X[ids[j] - row0] when this rank owns it, zeros otherwise (a sum all-reduce then assembles the rows)
__global__ void __launch_bounds__(256) gather_rows_kernel(const float* __restrict__ X, int64_t n, int d, long long row0,
                                                          const long long* __restrict__ ids, int m, float* __restrict__ out) {
    const int j = blockIdx.x;
    if (j >= m) return;
    const long long g = ids[j] - row0;
    const bool mine = g >= 0 && g < n;
    for (int c = threadIdx.x; c < d; c += blockDim.x) out[(size_t)j * d + c] = mine ? X[(size_t)g * d + c] : 0.f;
}

__global__ void __launch_bounds__(256) label_count_kernel(const int32_t* __restrict__ labels, int64_t n,
                                                          unsigned long long* __restrict__ counts) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        atomicAdd(counts + labels[i], 1ull);  // integers: the order does not matter
}

// ---- greedy weighted k-means++ over the m candidates (all state on the device, one block per kernel unless noted) ----
struct KppState {
    double total;      // sum of w * d2 (of w for the first pick)
    int trial[16];     // candidate index of every trial of this step
    double pot[16];    // potential after adding that trial
    int best;          // index into trial[]
    int n_chosen;
};

// P[j] = inclusive prefix of w[j] * d2[j] (d2 == nullptr: of w[j]); st->total = P[m-1].  One block, sequential tiles.
__global__ void __launch_bounds__(1024) kpp_prefix_kernel(const double* __restrict__ w, const double* __restrict__ d2, int m,
                                                          double* __restrict__ P, KppState* st) {
    __shared__ double s_warp[32];
    __shared__ double s_carry;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_carry = 0.0;
    __syncthreads();
    for (int base = 0; base < m; base += 1024) {
        const int j = base + tid;
        double v = j < m ? (d2 ? w[j] * d2[j] : w[j]) : 0.0;
        double incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const double t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            double ws = s_warp[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const double t = __shfl_up_sync(0xffffffffu, ws, o);
                if (lane >= o) ws += t;
            }
            s_warp[lane] = ws;
        }
        __syncthreads();
        const double mine = s_carry + (warp ? s_warp[warp - 1] : 0.0) + incl;
        if (j < m) P[j] = mine;
        __syncthreads();
        if (tid == 1023) s_carry = mine;
        __syncthreads();
    }
    if (tid == 0) st->total = s_carry;
}

// trial t = first j with P[j] > u_t * total (u_t from the hash stream of this step); total == 0 -> the lowest
// candidate not chosen yet, one trial.
__global__ void kpp_pick_kernel(const double* __restrict__ P, int m, unsigned long long seed, int stream, int trials,
                                const unsigned char* __restrict__ taken, KppState* st) {
    const int t = threadIdx.x;
    if (t >= trials) return;
    const double total = st->total;
    int idx;
    if (total <= 0.0) {
        idx = 0;
        while (idx < m - 1 && taken[idx]) ++idx;
    } else {
        const double target = hash_uniform(seed, (unsigned long long)stream, (unsigned long long)t) * total;
        int lo = 0, hi = m;  // first j with P[j] > target
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (P[mid] > target) hi = mid;
            else lo = mid + 1;
        }
        idx = lo < m - 1 ? lo : m - 1;
    }
    st->trial[t] = idx;
}

// Dt[t][j] = ||c_j - c_trial[t]||^2 in float64: one warp per (t, j), lanes over d in a fixed order
__global__ void __launch_bounds__(256) kpp_dist_kernel(const float* __restrict__ C, int m, int d, int trials,
                                                       const KppState* __restrict__ st, double* __restrict__ Dt) {
    const int lane = threadIdx.x & 31;
    const int64_t unit = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (unit >= (int64_t)trials * m) return;
    const int t = (int)(unit / m), j = (int)(unit - (int64_t)t * m);
    const float* a = C + (size_t)j * d;
    const float* b = C + (size_t)st->trial[t] * d;
    double s = 0.0;
    for (int i = lane; i < d; i += 32) {
        const double df = (double)a[i] - (double)b[i];
        s = fma(df, df, s);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
    if (lane == 0) Dt[(size_t)t * m + j] = s;
}

// pot[t] = sum_j w[j] * min(d2[j], Dt[t][j]) (d2 == nullptr: not needed); one block per trial, fixed order
__global__ void __launch_bounds__(256) kpp_pot_kernel(const double* __restrict__ w, const double* __restrict__ d2,
                                                      const double* __restrict__ Dt, int m, KppState* st) {
    __shared__ double s_red[8];
    const int t = blockIdx.x;
    double s = 0.0;
    for (int j = threadIdx.x; j < m; j += 256) s += w[j] * fmin(d2[j], Dt[(size_t)t * m + j]);
    s = block_sum_256(s, s_red);
    if (threadIdx.x == 0) st->pot[t] = s;
}

// the trial with the smallest potential (first on ties) becomes centre number n_chosen; d2 <- min(d2, its distances)
__global__ void __launch_bounds__(256) kpp_apply_kernel(int m, int trials, int use_pot, const double* __restrict__ Dt,
                                                        double* __restrict__ d2, int first, unsigned char* __restrict__ taken,
                                                        int* __restrict__ chosen, KppState* st) {
    __shared__ int s_best;
    if (threadIdx.x == 0) {
        int best = 0;
        if (use_pot)
            for (int t = 1; t < trials; ++t)
                if (st->pot[t] < st->pot[best]) best = t;
        s_best = best;
    }
    __syncthreads();
    const int best = s_best;
    for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < m; j += gridDim.x * blockDim.x) {
        const double v = Dt[(size_t)best * m + j];
        d2[j] = first ? v : fmin(d2[j], v);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        const int idx = st->trial[best];
        chosen[st->n_chosen] = idx;
        taken[idx] = 1;
        st->best = best;
        st->n_chosen += 1;
    }
}

__global__ void __launch_bounds__(256) gather_chosen_kernel(const float* __restrict__ C, int d, const int* __restrict__ chosen,
                                                            float* __restrict__ out) {
    const int j = blockIdx.x;
    for (int c = threadIdx.x; c < d; c += blockDim.x) out[(size_t)j * d + c] = C[(size_t)chosen[j] * d + c];
}

}  // namespace

unsigned long long init_hash64(unsigned long long seed, unsigned long long stream, unsigned long long counter) {
    return hash64(seed, stream, counter);
}
double init_hash_uniform(unsigned long long seed, unsigned long long stream, unsigned long long counter) {
    return hash_uniform(seed, stream, counter);
}

void launch_kpar_sample(const float* dist, int64_t n, long long row0, unsigned long long seed, int stream, double scale,
                        long long* list, unsigned int* count, unsigned int cap, cudaStream_t s) {
    int64_t blocks = (n + 255) / 256;
    if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
    if (blocks < 1) blocks = 1;
    kpar_sample_kernel<<<(unsigned)blocks, 256, 0, s>>>(dist, n, row0, seed, stream, scale, list, count, cap);
    count_launch();
}

void launch_min_inplace(float* dist, const float* other, int64_t n, cudaStream_t s) {
    int64_t blocks = (n + 255) / 256;
    if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
    if (blocks < 1) blocks = 1;
    min_inplace_kernel<<<(unsigned)blocks, 256, 0, s>>>(dist, other, n);
    count_launch();
}

// partial: >= 1024 doubles; ticket: a zeroed unsigned int
void launch_det_sum(const float* x, int64_t n, double* partial, double* out, unsigned int* ticket, cudaStream_t s) {
    int blocks = (int)((n + 65535) / 65536);
    if (blocks > 1024) blocks = 1024;
    if (blocks < 1) blocks = 1;
    const int64_t per = (n + blocks - 1) / blocks;
    det_sum_kernel<<<blocks, 256, 0, s>>>(x, n, per, partial, out, ticket);
    count_launch();
}

void launch_gather_rows(const float* X, int64_t n, int d, long long row0, const long long* ids, int m, float* out,
                        cudaStream_t s) {
    if (m <= 0) return;
    gather_rows_kernel<<<m, 256, 0, s>>>(X, n, d, row0, ids, m, out);
    count_launch();
}

void launch_label_count(const int32_t* labels, int64_t n, unsigned long long* counts, cudaStream_t s) {
    int64_t blocks = (n + 255) / 256;
    if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
    if (blocks < 1) blocks = 1;
    label_count_kernel<<<(unsigned)blocks, 256, 0, s>>>(labels, n, counts);
    count_launch();
}

size_t kpp_state_bytes() { return sizeof(KppState); }

// Greedy weighted k-means++ over the m candidate rows C [m, d] with weights w: writes the k chosen rows to out.
// Scratch (device): P [m] f64, d2 [m] f64, Dt [trials * m] f64, taken [m] u8 (zeroed), chosen [k] i32, state.
void run_weighted_kmeanspp(const float* C, int m, int d, int k, const double* w, unsigned long long seed, int trials,
                           double* P, double* d2, double* Dt, unsigned char* taken, int* chosen, void* state, float* out,
                           cudaStream_t s) {
    KppState* st = (KppState*)state;
    cudaMemsetAsync(st, 0, sizeof(KppState), s);
    cudaMemsetAsync(taken, 0, (size_t)m, s);
    const int apply_blocks = (m + 255) / 256 < kNumSMs ? (m + 255) / 256 : kNumSMs;
    for (int step = 0; step < k; ++step) {
        const bool first = step == 0;
        const int T = first ? 1 : trials;
        kpp_prefix_kernel<<<1, 1024, 0, s>>>(w, first ? nullptr : d2, m, P, st);
        kpp_pick_kernel<<<1, 32, 0, s>>>(P, m, seed, first ? 200 : 1000 + step, T, taken, st);
        kpp_dist_kernel<<<(unsigned)(((int64_t)T * m + 7) / 8), 256, 0, s>>>(C, m, d, T, st, Dt);
        if (!first) kpp_pot_kernel<<<T, 256, 0, s>>>(w, d2, Dt, m, st);
        kpp_apply_kernel<<<apply_blocks, 256, 0, s>>>(m, T, first ? 0 : 1, Dt, d2, first ? 1 : 0, taken, chosen, st);
        count_launch(first ? 4 : 5);
    }
    gather_chosen_kernel<<<k, 256, 0, s>>>(C, d, chosen, out);
    count_launch();
}

}  // namespace b200km
