// Per-cluster sums and counts for large k, without atomics on the data and with a fixed summation order.
//
// Used behind the tile-GEMM assignment (kernels_gemm.cu), where k*d float64 accumulators do not fit shared memory:
//   1. label_hist     every warp owns a contiguous chunk of rows and counts its labels in shared memory;
//   2. label_scan     per label, a prefix over the chunks (each chunk's first position inside the label's range),
//                     then an exclusive scan of the label totals: where every label's range starts;
//   3. label_scatter  the same walk again: row i goes to position start[label] + (rows of that label before i).
//                     The order inside a label is the row order - a stable counting sort, the same on every run;
//   4. segment_reduce the sorted order is cut into segments of S positions; a warp adds a segment's rows (gathered
//                     by index: whole rows, coalesced) into float64 registers, run by run.  A run that is a
//                     label's complete range is stored at once; a run that continues from the previous segment
//                     or into the next goes to that segment's head / tail slot;
//   5. label_combine  labels whose range spans segments add their slots in segment order.
// Every sum is therefore added in one fixed order: results are bitwise reproducible from run to run.
//
// Replaces cuML's reduce_rows_by_key behind reference python/src/spark_rapids_ml/clustering.py:367-370.
#include "common.cuh"
#include "tc.cuh"

#include <cstdio>

namespace b200km {

namespace {

thread_local char g_up_err[512] = "";

constexpr int kChunkRowsMin = 8192;   // rows per warp in the counting passes

// 1. one warp per chunk; shared memory: k counters
__global__ void __launch_bounds__(32) label_hist_kernel(const int32_t* __restrict__ labels, int64_t n, int k, int chunk_rows,
                                                        unsigned int* __restrict__ H, IterState* st, int honor_done) {
    extern __shared__ unsigned int s_cnt[];
    if (honor_done && st->done) return;
    const int lane = threadIdx.x;
    for (int j = lane; j < k; j += 32) s_cnt[j] = 0;
    __syncwarp();
    const int64_t r0 = (int64_t)blockIdx.x * chunk_rows;
    const int64_t r1 = r0 + chunk_rows < n ? r0 + chunk_rows : n;
    for (int64_t base = r0; base < r1; base += 32) {
        const int64_t r = base + lane;
        const int l = r < r1 ? labels[r] : -1;
        const unsigned peers = __match_any_sync(0xffffffffu, l);
        if (l >= 0 && (peers & ((1u << lane) - 1u)) == 0) s_cnt[l] += __popc(peers);  // lowest lane of each label
        __syncwarp();
    }
    unsigned int* out = H + (size_t)blockIdx.x * k;
    for (int j = lane; j < k; j += 32) out[j] = s_cnt[j];
}

// 2a. thread per label: H[c][j] <- rows of label j in the chunks before c; total[j]
__global__ void __launch_bounds__(256) label_colscan_kernel(unsigned int* __restrict__ H, int nc, int k,
                                                            unsigned int* __restrict__ total, IterState* st, int honor_done) {
    if (honor_done && st->done) return;
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= k) return;
    unsigned int run = 0;
    for (int c = 0; c < nc; ++c) {
        const unsigned int v = H[(size_t)c * k + j];
        H[(size_t)c * k + j] = run;
        run += v;
    }
    total[j] = run;
}

// 2b. one block: start[j] = exclusive scan of total (start[k] = n); counts go to the accumulator as float64
__global__ void __launch_bounds__(1024) label_offsets_kernel(const unsigned int* __restrict__ total, int k,
                                                             unsigned int* __restrict__ start, double* __restrict__ counts_out,
                                                             IterState* st, int honor_done) {
    __shared__ unsigned int s_warp[32];
    __shared__ unsigned int s_carry;
    if (honor_done && st->done) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_carry = 0;
    __syncthreads();
    for (int base = 0; base < k; base += 1024) {
        const int j = base + tid;
        const unsigned int v = j < k ? total[j] : 0u;
        unsigned int incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned int t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            unsigned int w = s_warp[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned int t = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += t;
            }
            s_warp[lane] = w;  // inclusive over warps
        }
        __syncthreads();
        const unsigned int before = s_carry + (warp ? s_warp[warp - 1] : 0u) + incl - v;
        if (j < k) {
            start[j] = before;
            if (counts_out) counts_out[j] += (double)v;
        }
        __syncthreads();
        if (tid == 1023) s_carry = before + v;
        __syncthreads();
    }
    if (tid == 0) start[k] = s_carry;
}

// 3. one warp per chunk; shared memory: the chunk's running position inside every label's range
__global__ void __launch_bounds__(32) label_scatter_kernel(const int32_t* __restrict__ labels, int64_t n, int k,
                                                           int chunk_rows, const unsigned int* __restrict__ H,
                                                           const unsigned int* __restrict__ start,
                                                           unsigned int* __restrict__ order, IterState* st, int honor_done) {
    extern __shared__ unsigned int s_pos[];
    if (honor_done && st->done) return;
    const int lane = threadIdx.x;
    const unsigned int* in = H + (size_t)blockIdx.x * k;
    for (int j = lane; j < k; j += 32) s_pos[j] = start[j] + in[j];
    __syncwarp();
    const int64_t r0 = (int64_t)blockIdx.x * chunk_rows;
    const int64_t r1 = r0 + chunk_rows < n ? r0 + chunk_rows : n;
    for (int64_t base = r0; base < r1; base += 32) {
        const int64_t r = base + lane;
        const int l = r < r1 ? labels[r] : -1;
        const unsigned peers = __match_any_sync(0xffffffffu, l);
        const unsigned before = peers & ((1u << lane) - 1u);
        unsigned int p0 = 0;
        if (l >= 0) p0 = s_pos[l];
        __syncwarp();
        if (l >= 0) {
            order[p0 + __popc(before)] = (unsigned int)r;
            if (before == 0) s_pos[l] = p0 + __popc(peers);
        }
        __syncwarp();
    }
}

// 4. one warp per (segment, column pass); lane = float4 column chunk, NC chunks of 32 float4 per pass
template <int NC>
__global__ void __launch_bounds__(128) segment_reduce_kernel(const float* __restrict__ X, int d, int k,
                                                             const int32_t* __restrict__ labels,
                                                             const unsigned int* __restrict__ order,
                                                             const unsigned int* __restrict__ start, int64_t n, int S,
                                                             int64_t nseg, int col_passes, double* __restrict__ sums,
                                                             double* __restrict__ head, double* __restrict__ tail,
                                                             IterState* st, int honor_done) {
    if (honor_done && st->done) return;
    const int lane = threadIdx.x & 31;
    const int64_t unit = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
    if (unit >= nseg * col_passes) return;
    const int64_t seg = unit / col_passes;
    const int cp = (int)(unit - seg * col_passes);
    const int64_t s0 = seg * S;
    const int64_t s1 = s0 + S < n ? s0 + S : n;
    const int nch = d >> 2;
    const int ch0 = cp * (NC * 32) + lane;  // this lane's first float4 chunk; then + 32, ...

    double acc[NC][4];
#pragma unroll
    for (int c = 0; c < NC; ++c) acc[c][0] = acc[c][1] = acc[c][2] = acc[c][3] = 0.0;

    auto flush = [&](int label, int64_t run_begin, int64_t run_end) {
        // [run_begin, run_end) is the part of the label's range [lb, le) inside this segment.  The complete range
        // is added to the label's sums at once; a part that continues from the previous segment goes to this
        // segment's head slot, a part that starts here and continues into the next one to its tail slot.
        const int64_t lb = start[label], le = start[label + 1];
        const bool complete = run_begin == lb && run_end == le;
        double* dst = complete ? sums + (size_t)label * d : ((lb < s0 ? head : tail) + (size_t)seg * d);
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            const int ch = ch0 + c * 32;
            if (ch < nch) {
                double* q = dst + ch * 4;
                if (complete) {
                    q[0] += acc[c][0];
                    q[1] += acc[c][1];
                    q[2] += acc[c][2];
                    q[3] += acc[c][3];
                } else {
                    q[0] = acc[c][0];
                    q[1] = acc[c][1];
                    q[2] = acc[c][2];
                    q[3] = acc[c][3];
                }
            }
            acc[c][0] = acc[c][1] = acc[c][2] = acc[c][3] = 0.0;
        }
    };

    constexpr int U = NC >= 8 ? 1 : (NC == 4 ? 2 : 4);  // rows whose loads are in flight together
    int cur = -1;
    int64_t run_begin = s0;
    for (int64_t base = s0; base < s1; base += 32) {
        // 32 sorted positions at a time: row indices and labels through registers
        const int64_t pos = base + lane;
        unsigned int row = 0;
        int lab = -1;
        if (pos < s1) {
            row = order[pos];
            lab = labels[row];
        }
        const int cnt = (int)((s1 - base) < 32 ? (s1 - base) : 32);
        for (int i0 = 0; i0 < cnt; i0 += U) {
            float4 v[U][NC];
            int ls[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int i = i0 + u < cnt ? i0 + u : cnt - 1;
                const unsigned int r = __shfl_sync(0xffffffffu, row, i);
                ls[u] = __shfl_sync(0xffffffffu, lab, i);
                const float4* xr = reinterpret_cast<const float4*>(X + (size_t)r * d);
#pragma unroll
                for (int c = 0; c < NC; ++c) {
                    const int ch = ch0 + c * 32;
                    v[u][c] = ch < nch ? __ldg(xr + ch) : make_float4(0.f, 0.f, 0.f, 0.f);
                }
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (i0 + u < cnt) {
                    if (ls[u] != cur) {
                        if (cur >= 0) flush(cur, run_begin, base + i0 + u);
                        cur = ls[u];
                        run_begin = base + i0 + u;
                    }
#pragma unroll
                    for (int c = 0; c < NC; ++c) {
                        acc[c][0] += (double)v[u][c].x;
                        acc[c][1] += (double)v[u][c].y;
                        acc[c][2] += (double)v[u][c].z;
                        acc[c][3] += (double)v[u][c].w;
                    }
                }
            }
        }
    }
    if (cur >= 0) flush(cur, run_begin, s1);
}

// 5. thread per (label, column) for the labels whose range spans segments: slots added in segment order
__global__ void __launch_bounds__(256) label_combine_kernel(int d, int k, const unsigned int* __restrict__ start, int S,
                                                            const double* __restrict__ head, const double* __restrict__ tail,
                                                            double* __restrict__ sums, IterState* st, int honor_done) {
    if (honor_done && st->done) return;
    const int label = blockIdx.x;
    const int64_t lb = start[label], le = start[label + 1];
    if (le <= lb) return;
    const int64_t sa = lb / S, sb = (le - 1) / S;
    if (sa == sb) return;  // inside one segment: stored by the reduce kernel
    for (int c = threadIdx.x; c < d; c += blockDim.x) {
        // first piece: starts at lb inside segment sa and runs to that segment's end -> its tail slot; every
        // later piece continues from the previous segment -> head slots
        double s = tail[(size_t)sa * d + c];
        for (int64_t sg = sa + 1; sg <= sb; ++sg) s += head[(size_t)sg * d + c];
        sums[(size_t)label * d + c] += s;
    }
}

template <typename T>
int grow_buf(T** p, size_t* cap, size_t need, const char* what) {
    if (need <= *cap && *p) return 0;
    if (*p) cudaFree(*p);
    *p = nullptr;
    *cap = 0;
    const size_t want = need + need / 8 + 64;
    if (cudaMalloc((void**)p, sizeof(T) * want) != cudaSuccess) {
        snprintf(g_up_err, sizeof(g_up_err), "cudaMalloc of %zu bytes for %s failed", sizeof(T) * want, what);
        return 1;
    }
    *cap = want;
    return 0;
}

}  // namespace

const char* sorted_update_last_error() { return g_up_err; }

bool sorted_update_supported(int64_t n, int d, int k) {
    return (d & 3) == 0 && k >= 1 && (size_t)k * 4 <= 200 * 1024 && n < ((int64_t)1 << 32);
}

void sorted_update_free(SortedUpdateWorkspace* ws) {
    for (void* q : {ws->d_H, ws->d_total, ws->d_start, ws->d_order, ws->d_head, ws->d_tail})
        if (q) cudaFree(q);
    *ws = SortedUpdateWorkspace();
}

int sorted_update(SortedUpdateWorkspace* ws, const float* X, int64_t n, int d, int k, const int32_t* labels, double* acc,
                  IterState* st, bool honor_done, cudaStream_t s) {
    if (n <= 0) return 0;
    if (!sorted_update_supported(n, d, k)) {
        snprintf(g_up_err, sizeof(g_up_err), "sorted update does not support n=%lld d=%d k=%d", (long long)n, d, k);
        return 1;
    }
    // chunk size of the counting passes: enough chunks to fill the machine, few enough that the [chunks x k] table stays small
    int64_t chunk_rows = kChunkRowsMin;
    while ((n + chunk_rows - 1) / chunk_rows * (int64_t)k > ((int64_t)64 << 20)) chunk_rows *= 2;
    const int nc = (int)((n + chunk_rows - 1) / chunk_rows);
    // segment length of the reduction: about 2^17 segments at most
    int S = 256;
    while ((n + S - 1) / S > (1 << 17)) S *= 2;
    const int64_t nseg = (n + S - 1) / S;
    if (grow_buf((unsigned int**)&ws->d_H, &ws->H_cap, (size_t)nc * k, "the label count table")) return 1;
    if (grow_buf((unsigned int**)&ws->d_total, &ws->total_cap, (size_t)k, "the label totals")) return 1;
    if (grow_buf((unsigned int**)&ws->d_start, &ws->start_cap, (size_t)k + 1, "the label offsets")) return 1;
    if (grow_buf((unsigned int**)&ws->d_order, &ws->order_cap, (size_t)n, "the sorted row order")) return 1;
    if (grow_buf((double**)&ws->d_head, &ws->head_cap, (size_t)nseg * d, "the segment head slots")) return 1;
    if (grow_buf((double**)&ws->d_tail, &ws->tail_cap, (size_t)nseg * d, "the segment tail slots")) return 1;

    static PerDeviceOnce attr_set;  // the opt-in is a per-device attribute
    const int dev = PerDeviceOnce::current();
    if (dev < 0 || !attr_set.done[dev]) {
        cudaFuncSetAttribute(label_hist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        cudaFuncSetAttribute(label_scatter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (dev >= 0) attr_set.done[dev] = true;
    }
    const int hd = honor_done ? 1 : 0;
    unsigned int* H = (unsigned int*)ws->d_H;
    unsigned int* total = (unsigned int*)ws->d_total;
    unsigned int* start = (unsigned int*)ws->d_start;
    unsigned int* order = (unsigned int*)ws->d_order;
    const size_t kd = (size_t)k * d;
    label_hist_kernel<<<nc, 32, (size_t)k * 4, s>>>(labels, n, k, (int)chunk_rows, H, st, hd);
    count_launch();
    label_colscan_kernel<<<(k + 255) / 256, 256, 0, s>>>(H, nc, k, total, st, hd);
    count_launch();
    label_offsets_kernel<<<1, 1024, 0, s>>>(total, k, start, acc + kd, st, hd);
    count_launch();
    label_scatter_kernel<<<nc, 32, (size_t)k * 4, s>>>(labels, n, k, (int)chunk_rows, H, start, order, st, hd);
    count_launch();
    const int nch = d >> 2;
    const int NC = nch <= 32 ? 1 : (nch <= 64 ? 2 : (nch <= 128 ? 4 : 8));
    const int col_passes = (nch + NC * 32 - 1) / (NC * 32);
    const int64_t units = nseg * col_passes;
    const unsigned blocks = (unsigned)((units + 3) / 4);
    double* head = (double*)ws->d_head;
    double* tail = (double*)ws->d_tail;
#define B200KM_SEG(NCV)                                                                                            \
    segment_reduce_kernel<NCV><<<blocks, 128, 0, s>>>(X, d, k, labels, order, start, n, S, nseg, col_passes, acc, head, \
                                                     tail, st, hd)
    if (NC == 1) B200KM_SEG(1);
    else if (NC == 2) B200KM_SEG(2);
    else if (NC == 4) B200KM_SEG(4);
    else B200KM_SEG(8);
#undef B200KM_SEG
    count_launch();
    label_combine_kernel<<<k, 256, 0, s>>>(d, k, start, S, head, tail, acc, st, hd);
    count_launch();
    cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess) {
        snprintf(g_up_err, sizeof(g_up_err), "sorted update launch failed: %s", cudaGetErrorString(err));
        return 1;
    }
    return 0;
}

}  // namespace b200km
