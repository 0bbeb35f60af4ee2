// Generic SIMT path of the KMeans Lloyd iteration: works for any n, k, d.
//
//   rownorm      ||x||^2 per row, once per fit / predict call
//   cnorm        ||c||^2 per centroid (+ max), once per call for the initial centroids
//   assign_ffma  register-tiled FP32 FFMA  x.c^T  with a fused running (best, second-best, argmin)
//                epilogue; near-ties go to a work list
//   refine       float64 re-decision of the near-tie rows (lowest index wins exact ties)
//   update       per-cluster float64 sums / counts / inertia from (X, labels): label-sorted 128-row
//                tiles, FP32 run sums flushed into float64 accumulators (shared memory when k*d
//                fits, else global atomics)
//   finalize     mean update, keep-empty rule, shift^2, convergence flag, next ||c||^2
//
// This path replaces what cuML's KMeansMG.fit / .predict run behind
// reference python/src/spark_rapids_ml/clustering.py:367-370 and :493-497.
#include "common.cuh"

#include <cfloat>
#include <climits>

namespace b200km {

thread_local unsigned long long g_launches = 0;

namespace {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// True for exactly one block of the grid: the one that arrives last.  Resets the ticket.
__device__ __forceinline__ bool last_block_done(IterState* st) {
    __shared__ bool is_last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned int t = atomicAdd(&st->ticket, 1u);
        is_last = (t == gridDim.x - 1);
        if (is_last) st->ticket = 0;
    }
    __syncthreads();
    if (is_last) __threadfence();
    return is_last;
}

// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) rownorm_kernel(const float* __restrict__ X, int64_t n, int d,
                                                      float* __restrict__ xnorm) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
    for (int64_t r = warp; r < n; r += nwarps) {
        const float* row = X + r * d;
        float s = 0.f;
        for (int i = lane; i < d; i += 32) {
            float v = __ldg(row + i);
            s = fmaf(v, v, s);
        }
        s = warp_sum(s);
        if (lane == 0) xnorm[r] = s;
    }
}

// One block per centroid.  The last block folds max_j ||c_j||^2 and (optionally) resets the
// iteration state.
__global__ void __launch_bounds__(128) cnorm_kernel(const float* __restrict__ C, int k, int d,
                                                    float* __restrict__ cnorm, IterState* st,
                                                    int reset_iter) {
    __shared__ float red[4];
    const int j = blockIdx.x;
    float s = 0.f;
    for (int i = threadIdx.x; i < d; i += blockDim.x) {
        float v = C[(size_t)j * d + i];
        s = fmaf(v, v, s);
    }
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) cnorm[j] = red[0] + red[1] + red[2] + red[3];
    if (last_block_done(st)) {
        float m = 0.f;
        for (int q = threadIdx.x; q < k; q += blockDim.x) m = fmaxf(m, ((volatile float*)cnorm)[q]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
        __syncthreads();
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
        __syncthreads();
        if (threadIdx.x == 0) {
            st->cnorm_max = fmaxf(fmaxf(red[0], red[1]), fmaxf(red[2], red[3]));
            st->amb_count = 0;
            st->screen_refined = 0;
            for (int u = 0; u < 12; ++u) {
                const int lo = (u * k) >> 3, hi = ((u + 1) * k) >> 3;
                for (int li = 0; li < 4; ++li)
                    st->own[u * 4 + li] = (u < 8 && k <= 32 && lo + li < hi) ? (signed char)(lo + li) : (signed char)-1;
            }
            st->own_bins = kOwnBinsFirst;
            if (reset_iter) {
                st->done = 0;
                st->n_iter = 0;
                st->shift2 = 0.0;
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// assign_ffma: CTA tile 128 rows x (16*NJ) centroids, K chunks of 32, 256 threads, each thread an
// 8 x NJ register tile.  Rows of a thread are ty + 16*i, centroids tx + 16*j, so that the two
// ty values of a warp read adjacent As rows and 8 consecutive tx lanes cover all 32 banks.
constexpr int TM = 128;
constexpr int TK = 32;
constexpr int LDS = TK + 4;

struct Best {
    float b1, b2;
    int i1;
};

__device__ __forceinline__ void merge_best(Best& a, float ob1, int oi1, float ob2) {
    if (ob1 < a.b1 || (ob1 == a.b1 && oi1 < a.i1)) {
        a.b2 = fminf(a.b1, ob2);
        a.b1 = ob1;
        a.i1 = oi1;
    } else {
        a.b2 = fminf(a.b2, ob1);
    }
}

template <int NJ>
__global__ void __launch_bounds__(256, 2)
    assign_ffma_kernel(const float* __restrict__ X, int64_t n, int d, const float* __restrict__ C, int k,
                       const float* __restrict__ cnorm, const float* __restrict__ xnorm,
                       int32_t* __restrict__ labels, float* __restrict__ best, IterState* st,
                       unsigned int* __restrict__ worklist, int honor_done, int vec_ok) {
    constexpr int TN = 16 * NJ;
    __shared__ __align__(16) float As[TM][LDS];
    __shared__ __align__(16) float Bs[TN][LDS];

    if (honor_done && st->done) return;

    const int tid = threadIdx.x;
    const int tx = tid & 15;
    const int ty = tid >> 4;
    const int64_t row0 = (int64_t)blockIdx.x * TM;

    Best run[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        run[i].b1 = FLT_MAX;
        run[i].b2 = FLT_MAX;
        run[i].i1 = 0;
    }

    for (int n0 = 0; n0 < k; n0 += TN) {
        float acc[8][NJ];
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < NJ; ++j) acc[i][j] = 0.f;

        for (int k0 = 0; k0 < d; k0 += TK) {
            // ---- stage the X chunk [128 x 32] and the centroid chunk [TN x 32] ----
#pragma unroll
            for (int it = 0; it < 4; ++it) {
                const int r = (tid >> 3) + 32 * it;
                const int c4 = (tid & 7) * 4;
                const int64_t grow = row0 + r;
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (grow < n) {
                    const float* p = X + grow * d + k0 + c4;
                    if (vec_ok) {
                        if (k0 + c4 < d) v = __ldg(reinterpret_cast<const float4*>(p));
                    } else {
                        if (k0 + c4 + 0 < d) v.x = __ldg(p + 0);
                        if (k0 + c4 + 1 < d) v.y = __ldg(p + 1);
                        if (k0 + c4 + 2 < d) v.z = __ldg(p + 2);
                        if (k0 + c4 + 3 < d) v.w = __ldg(p + 3);
                    }
                }
                *reinterpret_cast<float4*>(&As[r][c4]) = v;
            }
            for (int idx = tid; idx < TN * 8; idx += 256) {
                const int r = idx >> 3;
                const int c4 = (idx & 7) * 4;
                const int gc = n0 + r;
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (gc < k) {
                    const float* p = C + (size_t)gc * d + k0 + c4;
                    if (vec_ok) {
                        if (k0 + c4 < d) v = *reinterpret_cast<const float4*>(p);
                    } else {
                        if (k0 + c4 + 0 < d) v.x = p[0];
                        if (k0 + c4 + 1 < d) v.y = p[1];
                        if (k0 + c4 + 2 < d) v.z = p[2];
                        if (k0 + c4 + 3 < d) v.w = p[3];
                    }
                }
                *reinterpret_cast<float4*>(&Bs[r][c4]) = v;
            }
            __syncthreads();
#pragma unroll
            for (int kk = 0; kk < TK; kk += 4) {
                float4 a[8], b[NJ];
#pragma unroll
                for (int i = 0; i < 8; ++i) a[i] = *reinterpret_cast<const float4*>(&As[ty + 16 * i][kk]);
#pragma unroll
                for (int j = 0; j < NJ; ++j) b[j] = *reinterpret_cast<const float4*>(&Bs[tx + 16 * j][kk]);
#pragma unroll
                for (int i = 0; i < 8; ++i)
#pragma unroll
                    for (int j = 0; j < NJ; ++j) {
                        acc[i][j] = fmaf(a[i].x, b[j].x, acc[i][j]);
                        acc[i][j] = fmaf(a[i].y, b[j].y, acc[i][j]);
                        acc[i][j] = fmaf(a[i].z, b[j].z, acc[i][j]);
                        acc[i][j] = fmaf(a[i].w, b[j].w, acc[i][j]);
                    }
            }
            __syncthreads();
        }

        // ---- epilogue: score = ||c||^2 - 2 x.c ; running (best, argmin, second) per row ----
        float cn[NJ];
#pragma unroll
        for (int j = 0; j < NJ; ++j) {
            const int c = n0 + tx + 16 * j;
            cn[j] = (c < k) ? __ldg(cnorm + c) : FLT_MAX;
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            Best t;
            t.b1 = FLT_MAX;
            t.b2 = FLT_MAX;
            t.i1 = INT_MAX;
#pragma unroll
            for (int j = 0; j < NJ; ++j) {
                const int c = n0 + tx + 16 * j;
                if (c < k) {
                    const float s = fmaf(-2.f, acc[i][j], cn[j]);
                    merge_best(t, s, c, FLT_MAX);
                }
            }
#pragma unroll
            for (int o = 8; o > 0; o >>= 1) {
                const float ob1 = __shfl_xor_sync(0xffffffffu, t.b1, o);
                const float ob2 = __shfl_xor_sync(0xffffffffu, t.b2, o);
                const int oi1 = __shfl_xor_sync(0xffffffffu, t.i1, o);
                merge_best(t, ob1, oi1, ob2);
            }
            if (t.i1 != INT_MAX) merge_best(run[i], t.b1, t.i1, t.b2);
        }
    }

    if (tx == 0) {
        const float cmax = st->cnorm_max;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int64_t r = row0 + ty + 16 * i;
            if (r < n) {
                const float xn = __ldg(xnorm + r);
                labels[r] = run[i].i1;
                best[r] = fmaxf(xn + run[i].b1, 0.f);
                if (k > 1 && (run[i].b2 - run[i].b1) < kTieTol * (xn + cmax)) {
                    const unsigned int slot = atomicAdd(&st->amb_count, 1u);
                    worklist[slot] = (unsigned int)r;
                }
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// refine: one warp per near-tie row, exact float64 sum (x_i - c_ji)^2 over all k centroids.
__global__ void __launch_bounds__(256) refine_kernel(const float* __restrict__ X, int d,
                                                     const float* __restrict__ C, int k,
                                                     int32_t* __restrict__ labels, float* __restrict__ best,
                                                     IterState* st, const unsigned int* __restrict__ worklist,
                                                     int honor_done) {
    if (honor_done && st->done) return;
    const unsigned int count = *((volatile unsigned int*)&st->amb_count);
    const int lane = threadIdx.x & 31;
    const unsigned int warp = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const unsigned int nwarps = gridDim.x * (blockDim.x >> 5);
    for (unsigned int w = warp; w < count; w += nwarps) {
        const int64_t r = worklist[w];
        const float* x = X + r * d;
        double bd = DBL_MAX;
        int bi = INT_MAX;
        for (int j = lane; j < k; j += 32) {
            const float* c = C + (size_t)j * d;
            double s = 0.0;
            for (int i = 0; i < d; ++i) {
                const double df = (double)__ldg(x + i) - (double)c[i];
                s = fma(df, df, s);
            }
            if (s < bd) {
                bd = s;
                bi = j;
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const double od = __shfl_xor_sync(0xffffffffu, bd, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (od < bd || (od == bd && oi < bi)) {
                bd = od;
                bi = oi;
            }
        }
        if (lane == 0) {
            labels[r] = bi;
            best[r] = (float)bd;
        }
    }
    if (last_block_done(st)) {
        if (threadIdx.x == 0) st->amb_count = 0;
    }
}

// ------------------------------------------------------------------------------------------------
// update: persistent CTAs over 128-row tiles.  Rows of a tile are rank-sorted by label so that a
// lane owning a column adds consecutive equal-label rows in a register and flushes once per run.
template <bool SMEM_ACC>
__global__ void __launch_bounds__(256) update_kernel(const float* __restrict__ X, int64_t n, int d, int k,
                                                     const int32_t* __restrict__ labels,
                                                     const float* __restrict__ best, double* __restrict__ acc,
                                                     IterState* st, int honor_done, int64_t num_tiles) {
    extern __shared__ double sacc[];  // [k*d + k] when SMEM_ACC
    __shared__ int s_lab[128];
    __shared__ int s_sorted[128];
    __shared__ int s_perm[128];
    __shared__ double s_red[8];

    if (honor_done && st->done) return;

    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int warp = tid >> 5;
    const int acc_n = k * d + k;
    if (SMEM_ACC) {
        for (int i = tid; i < acc_n; i += 256) sacc[i] = 0.0;
        __syncthreads();
    }
    double* sums = SMEM_ACC ? sacc : acc;
    double* cnts = sums + (size_t)k * d;
    double inert = 0.0;

    for (int64_t tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int64_t row0 = tile * 128;
        const int rows = (int)((n - row0) < 128 ? (n - row0) : 128);
        int my = INT_MAX;
        if (tid < 128) {
            if (tid < rows) {
                my = labels[row0 + tid];
                inert += (double)best[row0 + tid];
            }
            s_lab[tid] = my;
        }
        __syncthreads();
        if (tid < 128) {
            int rank = 0;
#pragma unroll 8
            for (int q = 0; q < 128; ++q) {
                const int o = s_lab[q];
                rank += (o < my) || (o == my && q < tid);
            }
            s_sorted[rank] = my;
            s_perm[rank] = tid;
            if (tid < rows) atomicAdd(&cnts[my], 1.0);
        }
        __syncthreads();
        const int p0 = warp * 16;
        const int p1 = (p0 + 16 < rows) ? p0 + 16 : rows;
        if (p0 < p1) {
            for (int c = lane; c < d; c += 32) {
                float v[16];
#pragma unroll
                for (int q = 0; q < 16; ++q)
                    v[q] = (p0 + q < p1) ? __ldg(X + (row0 + s_perm[p0 + q]) * d + c) : 0.f;
                int cur = s_sorted[p0];
                float runsum = 0.f;
#pragma unroll
                for (int q = 0; q < 16; ++q) {
                    if (p0 + q < p1) {
                        const int l = s_sorted[p0 + q];
                        if (l != cur) {
                            atomicAdd(&sums[(size_t)cur * d + c], (double)runsum);
                            runsum = 0.f;
                            cur = l;
                        }
                        runsum += v[q];
                    }
                }
                atomicAdd(&sums[(size_t)cur * d + c], (double)runsum);
            }
        }
        __syncthreads();
    }

    if (SMEM_ACC) {
        for (int i = tid; i < acc_n; i += 256) {
            const double v = sacc[i];
            if (v != 0.0) atomicAdd(&acc[i], v);
        }
    }
    inert = warp_sum(inert);
    if (lane == 0) s_red[warp] = inert;
    __syncthreads();
    if (tid == 0) {
        double t = 0.0;
        for (int w = 0; w < 8; ++w) t += s_red[w];
        if (t != 0.0) atomicAdd(&acc[(size_t)k * d + k], t);
    }
}

__global__ void __launch_bounds__(256) sum_best_kernel(const float* __restrict__ best, int64_t n,
                                                       double* __restrict__ out) {
    __shared__ double s_red[8];
    double s = 0.0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        s += (double)best[i];
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < 8; ++w) t += s_red[w];
        atomicAdd(out, t);
    }
}

// ------------------------------------------------------------------------------------------------
// finalize: one block per centroid.  Reads the (all-reduced) accumulators, writes the new float32
// centroid (or keeps the old one when its count is 0), its squared norm and its squared shift, and
// zeroes the accumulators for the next pass.  The last block sums the per-centroid shifts in index
// order (every rank therefore takes the same decision from identical inputs), sets `done` when
// shift^2 < tol and counts the iteration.
__global__ void __launch_bounds__(128) finalize_kernel(float* __restrict__ C, int k, int d,
                                                       float* __restrict__ cnorm, double* __restrict__ acc,
                                                       double* __restrict__ shift_part, double* __restrict__ cnt_part,
                                                       double tol, IterState* st, IterState* snap) {
    __shared__ double red_s[4];
    __shared__ float red_c[4];
    if (st->done) return;
    const int j = blockIdx.x;
    const double cnt = acc[(size_t)k * d + j];
    double sh = 0.0;
    float cn = 0.f;
    for (int i = threadIdx.x; i < d; i += blockDim.x) {
        const size_t idx = (size_t)j * d + i;
        const float old = C[idx];
        float nw = old;
        if (cnt > 0.0) nw = (float)(acc[idx] / cnt);
        acc[idx] = 0.0;
        C[idx] = nw;
        const double df = (double)nw - (double)old;
        sh = fma(df, df, sh);
        cn = fmaf(nw, nw, cn);
    }
    sh = warp_sum(sh);
    cn = warp_sum(cn);
    if ((threadIdx.x & 31) == 0) {
        red_s[threadIdx.x >> 5] = sh;
        red_c[threadIdx.x >> 5] = cn;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        shift_part[j] = (red_s[0] + red_s[1]) + (red_s[2] + red_s[3]);
        cnt_part[j] = cnt;
        cnorm[j] = (red_c[0] + red_c[1]) + (red_c[2] + red_c[3]);
        acc[(size_t)k * d + j] = 0.0;
    }
    if (last_block_done(st)) {
        // fixed-order reduction: thread t sums indices t, t+128, ... then a fixed tree
        double s = 0.0;
        float m = 0.f;
        for (int q = threadIdx.x; q < k; q += blockDim.x) {
            s += ((volatile double*)shift_part)[q];
            m = fmaxf(m, ((volatile float*)cnorm)[q]);
        }
        s = warp_sum(s);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
        __syncthreads();
        if ((threadIdx.x & 31) == 0) {
            red_s[threadIdx.x >> 5] = s;
            red_c[threadIdx.x >> 5] = m;
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            const double shift2 = (red_s[0] + red_s[1]) + (red_s[2] + red_s[3]);
            st->cnorm_max = fmaxf(fmaxf(red_c[0], red_c[1]), fmaxf(red_c[2], red_c[3]));
            st->shift2 = shift2;
            st->n_iter += 1;
            acc[(size_t)k * d + k] = 0.0;
            if (shift2 < tol) st->done = 1;
        }
        if (k <= 32) {
            // update-warp ownership for the next pass: largest cluster first onto the least loaded warp with a
            // free slot (the per-tile update time follows the rows a warp owns).  Lane = cluster: the largest
            // remaining one by a warp max, the assignment by lane 0; everything stays in registers / shared.
            __shared__ signed char s_own[48];
            if (threadIdx.x < 32) {
                const int lane = threadIdx.x;
                s_own[lane] = (signed char)-1;
                if (lane < 16) s_own[32 + lane] = (signed char)-1;
                double mine = lane < k ? ((volatile double*)cnt_part)[lane] : -1.0;
                double load[kOwnBinsLater];
                int used[kOwnBinsLater];
#pragma unroll
                for (int u = 0; u < kOwnBinsLater; ++u) {
                    load[u] = 0.0;
                    used[u] = 0;
                }
                __syncwarp();
                for (int r = 0; r < k; ++r) {
                    // arg max over the lanes, ties to the lowest cluster index
                    double bc = mine;
                    int best = lane;
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) {
                        const double oc = __shfl_xor_sync(0xffffffffu, bc, o);
                        const int ob = __shfl_xor_sync(0xffffffffu, best, o);
                        if (oc > bc || (oc == bc && ob < best)) {
                            bc = oc;
                            best = ob;
                        }
                    }
                    if (lane == best) mine = -1.0;
                    int w = -1;
#pragma unroll
                    for (int u = 0; u < kOwnBinsLater; ++u)
                        if (used[u] < 4 && (w < 0 || load[u] < load[w])) w = u;
                    if (lane == 0) s_own[w * 4 + used[w]] = (signed char)best;
                    used[w] += 1;
                    load[w] += bc;
                }
                __syncwarp();
                st->own[lane] = s_own[lane];
                if (lane < 16) st->own[32 + lane] = s_own[32 + lane];
                if (lane == 0) st->own_bins = kOwnBinsLater;
            }
        }
        // a stable copy for the host: it is read asynchronously, while later iterations already change *st
        __syncthreads();
        if (threadIdx.x == 0 && snap) *snap = *st;
    }
}

// ------------------------------------------------------------------------------------------------
// Fused tail of an iteration for small k*d (the packed accumulator fits 64 KB): ONE CTA
//   1. (nranks > 1) one-shot all-reduce over NVLink: pushes this rank's accumulator into its slot of every
//      peer's exchange buffer (plain peer stores), raises its flag there, waits for all peers' flags in its own
//      buffer and adds the slots in RANK ORDER - every rank gets the same bits, whatever arrives first;
//   2. the centroid update of finalize_kernel (mean or keep, shift^2, next ||c||^2, stop test, iteration count),
//      the update-warp ownership for the next pass and the state snapshot.
// It replaces a 16 KB ncclAllReduce (~38 us of launch + protocol latency at 8 GPUs) plus a k-block finalize launch.
// xb: this rank's exchange buffer [2 parities][nranks][kXrDoubles]; flags [2][nranks] follow at kXrFlagOff bytes.
struct XrPeers {
    unsigned char* base[16];  // every rank's exchange buffer as mapped in THIS process (own rank: local pointer)
};

__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__global__ void __launch_bounds__(1024) finalize_fused_kernel(float* __restrict__ C, int k, int d, float* __restrict__ cnorm,
                                                              double* __restrict__ acc, double tol, IterState* st,
                                                              IterState* snap, XrPeers peers, int rank, int nranks,
                                                              unsigned int epoch) {
    __shared__ double s_shift[256];
    __shared__ float s_cn[256];
    __shared__ double s_cnt[256];
    if (st->done) return;  // identical on every rank: nobody pushes, nobody waits
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int acc_n = k * d + k;  // the inertia slot (acc_n) travels too
    const int len = acc_n + 1;
    if (nranks > 1) {
        const unsigned int par = epoch & 1u;
        // ---- push: my accumulator into slot `rank` of every peer (and of my own buffer) ----
        for (int r = 0; r < nranks; ++r) {
            double* dst = reinterpret_cast<double*>(peers.base[r]) + ((size_t)par * nranks + rank) * kXrDoubles;
            for (int i = tid; i < len; i += 1024) dst[i] = acc[i];
        }
        __threadfence_system();
        __syncthreads();
        if (tid < nranks) {
            unsigned int* flag = reinterpret_cast<unsigned int*>(peers.base[tid] + kXrFlagOff) + par * 16 + rank;
            st_release_sys(flag, epoch);
        }
        // ---- wait for every peer's slot in my own buffer ----
        if (tid < nranks) {
            const unsigned int* flag = reinterpret_cast<const unsigned int*>(peers.base[rank] + kXrFlagOff) + par * 16 + tid;
            long long spins = 0;
            while (ld_acquire_sys(flag) != epoch) {
                __nanosleep(100);
                if (++spins > 50000000ll) asm volatile("trap;");  // a peer never arrived (~5 s): fail instead of hanging
            }
        }
        __syncthreads();
        // ---- reduce in rank order ----
        const double* mine = reinterpret_cast<const double*>(peers.base[rank]) + (size_t)par * nranks * kXrDoubles;
        for (int i = tid; i < len; i += 1024) {
            double t = 0.0;
            for (int r = 0; r < nranks; ++r) t += mine[(size_t)r * kXrDoubles + i];
            acc[i] = t;
        }
        __syncthreads();
    }
    // ---- centroid update: warp per centroid ----
    for (int j = warp; j < k; j += 32) {
        const double cnt = acc[(size_t)k * d + j];
        double sh = 0.0;
        float cn = 0.f;
        for (int i = lane; i < d; i += 32) {
            const size_t idx = (size_t)j * d + i;
            const float old = C[idx];
            float nw = old;
            if (cnt > 0.0) nw = (float)(acc[idx] / cnt);
            C[idx] = nw;
            const double df = (double)nw - (double)old;
            sh = fma(df, df, sh);
            cn = fmaf(nw, nw, cn);
        }
        sh = warp_sum(sh);
        cn = warp_sum(cn);
        if (lane == 0) {
            s_shift[j] = sh;
            s_cn[j] = cn;
            s_cnt[j] = cnt;
            cnorm[j] = cn;
        }
    }
    __syncthreads();
    for (int i = tid; i < len; i += 1024) acc[i] = 0.0;  // ready for the next pass
    if (tid == 0) {
        double shift2 = 0.0;
        float m = 0.f;
        for (int j = 0; j < k; ++j) {  // index order: the same decision on every rank
            shift2 += s_shift[j];
            m = fmaxf(m, s_cn[j]);
        }
        st->cnorm_max = m;
        st->shift2 = shift2;
        st->n_iter += 1;
        if (shift2 < tol) st->done = 1;
    }
    if (k <= 32 && warp == 0) {
        // update-warp ownership for the next pass: clusters sorted by size (bitonic, descending; ties by index),
        // then placed greedily - the per-tile update time follows the rows a warp owns
        double key = lane < k ? s_cnt[lane] : -1.0;
        int idx = lane;
#pragma unroll
        for (int size = 2; size <= 32; size <<= 1) {
#pragma unroll
            for (int stride = size >> 1; stride > 0; stride >>= 1) {
                const double ok = __shfl_xor_sync(0xffffffffu, key, stride);
                const int oi = __shfl_xor_sync(0xffffffffu, idx, stride);
                const bool up = ((lane & size) == 0);          // this block sorts descending when `up`
                const bool lower = ((lane & stride) == 0);     // the lane keeps the "first" element of the pair
                const bool other_first = (ok > key) || (ok == key && oi < idx);  // descending order: larger first
                const bool take = (lower == up) ? other_first : !other_first;
                if (take) {
                    key = ok;
                    idx = oi;
                }
            }
        }
        // lane p now holds the p-th largest cluster: largest first onto the least loaded warp with a free slot
        // (lane w < 12 keeps warp w's load and slot count; the choice is a minimum over those lanes)
        st->own[lane] = (signed char)-1;
        if (lane < 16) st->own[32 + lane] = (signed char)-1;
        if (lane == 0) st->own_bins = kOwnBinsLater;
        __syncwarp();
        double load = 0.0;
        int used = 0;
        for (int pth = 0; pth < k; ++pth) {
            const double sz = __shfl_sync(0xffffffffu, key, pth);
            const int id = __shfl_sync(0xffffffffu, idx, pth);
            double cand = (lane < kOwnBinsLater && used < 4) ? load : 1e300;
            int who = lane;
#pragma unroll
            for (int o = 8; o > 0; o >>= 1) {
                const double oc = __shfl_xor_sync(0xffffffffu, cand, o);
                const int ow = __shfl_xor_sync(0xffffffffu, who, o);
                if (oc < cand || (oc == cand && ow < who)) {
                    cand = oc;
                    who = ow;
                }
            }
            if (lane < kOwnBinsLater && lane == who) {
                st->own[lane * 4 + used] = (signed char)id;
                used += 1;
                load += sz;
            }
        }
    }
    __syncthreads();
    if (tid == 0 && snap) *snap = *st;
}

}  // namespace

// ---- host launchers -----------------------------------------------------------------------------
void launch_prep(const float* C, int k, int d, float* cnorm, double* acc, int64_t acc_len, IterState* st,
                 bool reset_iter, cudaStream_t s) {
    if (acc) cudaMemsetAsync(acc, 0, sizeof(double) * (size_t)acc_len, s);
    cnorm_kernel<<<k, 128, 0, s>>>(C, k, d, cnorm, st, reset_iter ? 1 : 0);
    count_launch();
}

void launch_rownorm(const float* X, int64_t n, int d, float* xnorm, cudaStream_t s) {
    if (n <= 0) return;
    int64_t blocks = (n + 7) / 8;
    if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
    rownorm_kernel<<<(unsigned)blocks, 256, 0, s>>>(X, n, d, xnorm);
    count_launch();
}

void launch_assign_ffma(const float* X, int64_t n, int d, const float* C, int k, const float* cnorm,
                        const float* xnorm, int32_t* labels, float* best, IterState* st,
                        unsigned int* worklist, bool honor_done, cudaStream_t s) {
    if (n <= 0) return;
    const int vec_ok = ((d & 3) == 0) && ((reinterpret_cast<uintptr_t>(X) & 15) == 0) &&
                       ((reinterpret_cast<uintptr_t>(C) & 15) == 0);
    const unsigned blocks = (unsigned)((n + TM - 1) / TM);
    if (k <= 32)
        assign_ffma_kernel<2><<<blocks, 256, 0, s>>>(X, n, d, C, k, cnorm, xnorm, labels, best, st, worklist,
                                                     honor_done ? 1 : 0, vec_ok);
    else
        assign_ffma_kernel<4><<<blocks, 256, 0, s>>>(X, n, d, C, k, cnorm, xnorm, labels, best, st, worklist,
                                                     honor_done ? 1 : 0, vec_ok);
    count_launch();
}

void launch_refine(const float* X, int64_t n, int d, const float* C, int k, int32_t* labels, float* best,
                   IterState* st, const unsigned int* worklist, bool honor_done, cudaStream_t s) {
    if (n <= 0) return;
    int64_t blocks = (n + 127) / 128;  // no more warps than could ever be needed
    if (blocks > kNumSMs * 2) blocks = kNumSMs * 2;
    refine_kernel<<<(unsigned)blocks, 256, 0, s>>>(X, d, C, k, labels, best, st, worklist, honor_done ? 1 : 0);
    count_launch();
}

void launch_update(const float* X, int64_t n, int d, int k, const int32_t* labels, const float* best,
                   double* acc, IterState* st, bool honor_done, cudaStream_t s) {
    if (n <= 0) return;
    const int64_t tiles = (n + 127) / 128;
    const size_t smem = sizeof(double) * ((size_t)k * d + k);
    const bool smem_acc = smem <= 160 * 1024;
    int64_t blocks = smem_acc ? kNumSMs : kNumSMs * 4;
    if (blocks > tiles) blocks = tiles;
    if (smem_acc) {
        static PerDeviceOnce attr_set;  // the opt-in is a per-device attribute
        const int dev = PerDeviceOnce::current();
        if (dev < 0 || !attr_set.done[dev]) {
            cudaFuncSetAttribute(update_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
            if (dev >= 0) attr_set.done[dev] = true;
        }
        update_kernel<true><<<(unsigned)blocks, 256, smem, s>>>(X, n, d, k, labels, best, acc, st,
                                                               honor_done ? 1 : 0, tiles);
        count_launch();
    } else {
        update_kernel<false><<<(unsigned)blocks, 256, 0, s>>>(X, n, d, k, labels, best, acc, st,
                                                              honor_done ? 1 : 0, tiles);
        count_launch();
    }
}

void launch_sum_best(const float* best, int64_t n, double* out, cudaStream_t s) {
    if (n <= 0) return;
    int64_t blocks = (n + 255) / 256;
    if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
    sum_best_kernel<<<(unsigned)blocks, 256, 0, s>>>(best, n, out);
    count_launch();
}

bool finalize_fused_supported(int k, int d) { return k <= 256 && (size_t)k * d + k + 1 <= (size_t)kXrDoubles; }

void launch_finalize_fused(float* C, int k, int d, float* cnorm, double* acc, double tol, IterState* st, IterState* snap,
                           unsigned char* const* peer_bases, int rank, int nranks, unsigned int epoch, cudaStream_t s) {
    XrPeers peers = {};
    for (int r = 0; r < nranks && r < 16; ++r) peers.base[r] = peer_bases ? peer_bases[r] : nullptr;
    finalize_fused_kernel<<<1, 1024, 0, s>>>(C, k, d, cnorm, acc, tol, st, snap, peers, rank, nranks, epoch);
    count_launch();
}

void launch_finalize(float* C, int k, int d, float* cnorm, double* acc, double* shift_part, double* cnt_part,
                     double tol, IterState* st, IterState* snap, cudaStream_t s) {
    finalize_kernel<<<k, 128, 0, s>>>(C, k, d, cnorm, acc, shift_part, cnt_part, tol, st, snap);
    count_launch();
}

}  // namespace b200km
