// tcgen05 / TMEM / TMA path: one persistent, warp-specialised kernel per Lloyd pass.
//
// For every 128-row tile of X the kernel computes  score_j = ||c_j||^2 - 2 x.c_j  on the 5th-gen
// tensor cores in split-TF32 (x = x_hi + x_lo, c = c_hi + c_lo, all four partial products, FP32
// accumulation in TMEM), takes the running (best, second-best, argmin) per row, re-decides near-ties
// in float64, and - in the same pass, while the tile is still in shared memory - adds the rows into
// per-cluster float64 sums / counts held in shared memory.  X is read from HBM exactly once per pass.
//
// Warp roles (512 threads, 1 CTA per SM, grid = #SMs; every scheduler hosts one warp of each role):
//   warp 0       TMA producer: split centroids once, then X tiles as K-atoms of 32 floats (SWIZZLE_128B)
//                plus a narrow tail atom (SWIZZLE_32B / 64B) when d % 32 is 4..16
//   warp 1       TMEM allocator + MMA issuer (warp-uniform loop, one elected lane issues
//                tcgen05.mma kind::tf32 with A from TMEM and B = [c_hi ; c_lo] through UMMA descriptors)
//   warps 4-7    transform: each thread owns one row; reads it from the swizzled tile and writes the
//                exact tf32 parts x_hi | x_lo into a TMEM ring slot (tcgen05.st); ||x||^2 on the first pass
//   warps 8-11   assign: tcgen05.ld the scores, argmin, float64 re-decision of near-ties, labels /
//                distances out, then a stable counting sort of the tile's rows by label
//   warps 12-15  update: label-sorted run accumulation of the tile into the float64 accumulators
//
// This kernel replaces cuML's fusedL2NN + reduce_rows_by_key pair that runs behind reference
// python/src/spark_rapids_ml/clustering.py:367-370 (fit) and :493-497 (predict).
#include "tc.cuh"
#include "tc_ptx.cuh"

#include <cuda.h>

#include <cfloat>
#include <climits>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace b200km {

namespace {

using namespace tcx;

thread_local char g_tc_err[512] = "";

constexpr int kAtomBytes = kTileM * 128;  // one full K-atom of an X tile: 128 rows x 32 floats
constexpr int kRing = 4;                  // TMEM ring slots for (x_hi | x_lo) atoms, 64 columns each
constexpr int kRingCol0 = 256;            // ring occupies TMEM columns [256, 512)
constexpr int kThreads = 768;
constexpr int kMaxSmem = 227 * 1024;
constexpr int kMaxStages = 4;
constexpr int kMaxD = 256;
constexpr int kMaxK = 256;
constexpr int kHist = kMaxK + 8;

struct TcParams {
    int64_t n;
    int64_t num_tiles;
    int d, k, k_pad;
    int n_full;     // number of 32-float SWIZZLE_128B atoms
    int tail_cols;  // 0, 8 (SWIZZLE_32B) or 16 (SWIZZLE_64B): width of the narrow last atom
    int num_stages, num_d;
    int stage_bytes;
    const float* C;
    const float* cnorm;
    float* xnorm;  // [n] ||x||^2: written when xn_compute, read otherwise
    int xn_compute;
    int32_t* labels;
    float* best;
    double* acc;
    IterState* st;
    int honor_done;
    float tie_tol;
    unsigned long long* trace;  // development aid (B200KM_TRACE): [256 tiles][16 events] clock64 of CTA 0
    // shared-memory offsets (bytes from the 1024-aligned base)
    int off_c, off_acc, off_cn, off_xn, off_bar;
};


// Byte offset of 16-byte chunk `chunk` (= float column / 4) of row r inside one tile stage.  Full atoms
// use the TMA 128-byte swizzle (chunk ^= row & 7); the narrow tail atom the 64-byte (chunk ^= (row>>1)&3)
// or 32-byte (chunk ^= (row>>2)&1) swizzle.
__device__ __forceinline__ uint32_t chunk_off(const TcParams& p, int r, int chunk) {
    const int a = chunk >> 3;
    if (a < p.n_full) return (uint32_t)(a * kAtomBytes + r * 128 + (((chunk & 7) ^ (r & 7)) << 4));
    const int c = chunk - p.n_full * 8;
    const uint32_t base = (uint32_t)(p.n_full * kAtomBytes);
    if (p.tail_cols == 16) return base + (uint32_t)(r * 64 + ((c ^ ((r >> 1) & 3)) << 4));
    return base + (uint32_t)(r * 32 + ((c ^ ((r >> 2) & 1)) << 4));
}


__device__ __forceinline__ void trace_ev(const TcParams& p, int64_t it, int ev) {
#ifdef B200KM_ENABLE_TRACE
    if (p.trace && blockIdx.x == 0 && it < 256) p.trace[it * 16 + ev] = (unsigned long long)clock64();
#endif
}

// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) split_centroids_kernel(const float* __restrict__ C, int k, int d, int k_pad,
                                                              int d_pad, float* __restrict__ out) {
    // out: [2*k_pad][d_pad]; rows [0,k_pad) = c_hi (tf32, round-to-nearest), rows [k_pad,2k_pad) = c_lo
    const int total = k_pad * d_pad;
    for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
        const int j = idx / d_pad, i = idx - j * d_pad;
        float hi = 0.f, lo = 0.f;
        if (j < k && i < d) {
            const float c = C[(size_t)j * d + i];
            hi = __uint_as_float(tf32_round(__float_as_uint(c)));
            lo = __uint_as_float(tf32_round(__float_as_uint(c - hi)));
        }
        out[idx] = hi;
        out[(size_t)total + idx] = lo;
    }
}

// One K-atom of one row -> exact tf32 parts in TMEM.  NCH = number of 16-byte chunks of the row in this
// atom (8 = full SWIZZLE_128B atom handled as two halves of 4, else the tail: 4 = SWIZZLE_64B, 2 = 32B).
template <bool XN>
__device__ __forceinline__ void split16(const uint8_t* rowp, int swz, int chunk0, uint32_t (&hi)[16],
                                        uint32_t (&lo)[16], float& xn) {
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        const float4 v = *reinterpret_cast<const float4*>(rowp + (((chunk0 + c) ^ swz) << 4));
        const float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const uint32_t h = tf32_trunc(__float_as_uint(e[t]));
            if (XN) xn = fmaf(e[t], e[t], xn);
            hi[c * 4 + t] = h;
            lo[c * 4 + t] = __float_as_uint(e[t] - __uint_as_float(h));  // exact; the MMA reads its top 19 bits
        }
    }
}

template <bool XN>
__device__ __forceinline__ float transform_tile(const TcParams& p, const uint8_t* stage, int row, uint32_t tmem_lane,
                                                uint32_t a_full0, uint32_t a_empty0, int64_t& ait, int lane) {
    float xn = 0.f;
    const int n_atoms = p.n_full + (p.tail_cols ? 1 : 0);
    for (int a = 0; a < n_atoms; ++a, ++ait) {
        const int slot = (int)(ait % kRing);
        const uint32_t taddr = tmem_lane + (uint32_t)(kRingCol0 + slot * 64);
        uint32_t hi[16], lo[16];
        if (a < p.n_full) {
            const uint8_t* rowp = stage + a * kAtomBytes + row * 128;
            split16<XN>(rowp, row & 7, 0, hi, lo, xn);
            mbar_wait(a_empty0 + 8u * slot, (uint32_t)(((ait / kRing) & 1) ^ 1));
            tc_fence_after();
            tmem_st16(taddr, hi);
            tmem_st16(taddr + 32, lo);
            split16<XN>(rowp, row & 7, 4, hi, lo, xn);
            tmem_st16(taddr + 16, hi);
            tmem_st16(taddr + 48, lo);
        } else if (p.tail_cols == 16) {
            const uint8_t* rowp = stage + p.n_full * kAtomBytes + row * 64;
            split16<XN>(rowp, (row >> 1) & 3, 0, hi, lo, xn);
            mbar_wait(a_empty0 + 8u * slot, (uint32_t)(((ait / kRing) & 1) ^ 1));
            tc_fence_after();
            tmem_st16(taddr, hi);
            tmem_st16(taddr + 32, lo);
        } else {
            const uint8_t* rowp = stage + p.n_full * kAtomBytes + row * 32;
            const int swz = (row >> 2) & 1;
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                const float4 v = *reinterpret_cast<const float4*>(rowp + ((c ^ swz) << 4));
                const float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    const uint32_t h = tf32_trunc(__float_as_uint(e[t]));
                    if (XN) xn = fmaf(e[t], e[t], xn);
                    hi[c * 4 + t] = h;
                    lo[c * 4 + t] = __float_as_uint(e[t] - __uint_as_float(h));
                }
            }
            mbar_wait(a_empty0 + 8u * slot, (uint32_t)(((ait / kRing) & 1) ^ 1));
            tc_fence_after();
            tmem_st8(taddr, hi);
            tmem_st8(taddr + 32, lo);
        }
        tmem_wait_st();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(a_full0 + 8u * slot);
    }
    return xn;
}

__global__ void __launch_bounds__(kThreads, 1)
    kmeans_fused_tc_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmXt,
                           const __grid_constant__ CUtensorMap tmC, const __grid_constant__ CUtensorMap tmCt,
                           const TcParams p) {
    if (p.honor_done && p.st->done) return;

    extern __shared__ uint8_t smem_raw[];
    __shared__ int s_hist[2][4][kHist];  // per assign group: per-warp label histograms, then offsets
    __shared__ int s_cand[2][kTileM][4];  // per assign group: candidate centroids of near-tie rows
    __shared__ int s_base[2][kHist];  // start of every label's run in the sorted order (double-buffered)
    __shared__ int s_perm[2][kTileM + 8];   // sorted position -> row offset in a full atom (r*128 | swizzle) | label<<16
    __shared__ int s_permT[2][kTileM + 8];  // same for the narrow tail atom (row pitch 64 / 32 bytes)
    // pointer arithmetic (not an integer round trip) keeps the shared address space visible to the compiler
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const uint32_t smem_base = smem_u32(smem);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int S = p.num_stages;
    const int ND = p.num_d;
    const int stage_bytes = p.stage_bytes;
    const int n_atoms = p.n_full + (p.tail_cols ? 1 : 0);
    const int c_block = 2 * p.k_pad * 128;  // bytes of one full-atom block [c_hi rows | c_lo rows]
    const bool stacked = p.k_pad <= 128;
    const int dcols = stacked ? 2 * p.k_pad : p.k_pad;

    // barrier table
    const uint32_t bar0 = smem_base + p.off_bar;
    const uint32_t x_full0 = bar0, x_empty0 = bar0 + 32, xn_full0 = bar0 + 64, a_full0 = bar0 + 96,
                   a_empty0 = bar0 + 128, d_full0 = bar0 + 160, d_empty0 = bar0 + 176, s_full0 = bar0 + 192,
                   s_empty0 = bar0 + 208, c_full = bar0 + 224;
    uint32_t* tmem_slot = (uint32_t*)(smem + p.off_bar + 232);

    if (threadIdx.x == 0) {
        for (int s = 0; s < kMaxStages; ++s) {
            mbar_init(x_full0 + 8u * s, 1);
            mbar_init(x_empty0 + 8u * s, p.acc ? 8 : 4);
            mbar_init(xn_full0 + 8u * s, 4);
        }
        for (int s = 0; s < kRing; ++s) {
            mbar_init(a_full0 + 8u * s, 4);
            mbar_init(a_empty0 + 8u * s, 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(d_full0 + 8u * s, 1);
            mbar_init(d_empty0 + 8u * s, 4);
            mbar_init(s_full0 + 8u * s, 4);
            mbar_init(s_empty0 + 8u * s, 8);
        }
        mbar_init(c_full, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    {
        // shared state every role relies on, written once by all threads before the first barrier
        float* cn_init = (float*)(smem + p.off_cn);
#pragma unroll 1
        for (int j = threadIdx.x; j < p.k_pad; j += kThreads) cn_init[j] = (j < p.k) ? p.cnorm[j] : FLT_MAX;
        if (p.acc) {
            double* sacc_init = (double*)(smem + p.off_acc);
#pragma unroll 1
            for (int i = threadIdx.x; i < p.k * p.d + p.k; i += kThreads) sacc_init[i] = 0.0;
        }
#pragma unroll 1
        for (int i = threadIdx.x; i < 2 * 4 * kHist; i += kThreads) (&s_hist[0][0][0])[i] = 0;
#pragma unroll 1
        for (int i = threadIdx.x; i < 2 * (kTileM + 8); i += kThreads) {
            (&s_perm[0][0])[i] = 0;
            (&s_permT[0][0])[i] = 0;
        }
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    const int64_t first_tile = blockIdx.x;
    const int64_t tile_step = gridDim.x;

    if (warp == 0) {
        // ===================================== TMA producer =====================================
        if (lane == 0) {
            mbar_arrive_expect_tx(c_full, (uint32_t)(2 * p.k_pad * (p.n_full * 128 + p.tail_cols * 4)));
            for (int a = 0; a < p.n_full; ++a)
                for (int part = 0; part < 2; ++part)  // atom block = [c_hi rows | c_lo rows], 2*k_pad rows
                    tma_load_2d(smem_base + p.off_c + a * c_block + part * p.k_pad * 128, &tmC, c_full, a * 32,
                                part * p.k_pad);
            if (p.tail_cols)
                for (int part = 0; part < 2; ++part)
                    tma_load_2d(smem_base + p.off_c + p.n_full * c_block + part * p.k_pad * p.tail_cols * 4, &tmCt,
                                c_full, p.n_full * 32, part * p.k_pad);
            int64_t it = 0;
            for (int64_t tile = first_tile; tile < p.num_tiles; tile += tile_step, ++it) {
                const int s = (int)(it % S);
                mbar_wait(x_empty0 + 8u * s, (uint32_t)(((it / S) & 1) ^ 1));
                trace_ev(p, it, 0);
                mbar_arrive_expect_tx(x_full0 + 8u * s, (uint32_t)stage_bytes);
                const uint32_t dst = smem_base + s * stage_bytes;
                for (int a = 0; a < p.n_full; ++a)
                    tma_load_2d(dst + a * kAtomBytes, &tmX, x_full0 + 8u * s, a * 32, (int)(tile * kTileM));
                if (p.tail_cols)
                    tma_load_2d(dst + p.n_full * kAtomBytes, &tmXt, x_full0 + 8u * s, p.n_full * 32,
                                (int)(tile * kTileM));
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ====================================== MMA issuer ======================================
        // The whole warp runs the loop and one elected lane issues (a tcgen05.mma issued from divergent
        // code costs ~100 cycles per instruction, ~40 from warp-uniform code: tools/mma_bench.cu).
        // k_pad <= 128: B is the stacked block [c_hi ; c_lo] (N = 2*k_pad), so a K-step is two instructions
        // of one shape (alternating shapes stalls the pipe):  D[:, 0:2k) (+)= x_hi.[c_hi;c_lo]^T  and
        // D[:, 0:2k) += x_lo.[c_hi;c_lo]^T, i.e. all four partial products of (x_hi+x_lo).(c_hi+c_lo).
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(kTileM >> 4) << 24) |
                               ((uint32_t)(dcols >> 3) << 17);
        const uint64_t desc_full = make_b_desc(smem_base + p.off_c, 128);
        const uint64_t desc_tail = make_b_desc(smem_base + p.off_c + p.n_full * c_block, p.tail_cols ? p.tail_cols * 4 : 128);
        mbar_wait(c_full, 0);
        int64_t it = 0;   // tile counter
        int64_t ait = 0;  // atom counter (ring position)
        for (int64_t tile = first_tile; tile < p.num_tiles; tile += tile_step, ++it) {
            const int db = (int)(it % ND);
            mbar_wait(d_empty0 + 8u * db, (uint32_t)(((it / ND) & 1) ^ 1));
            tc_fence_after();
            if (lane == 0) trace_ev(p, it, 8);
            const uint32_t d_tmem = tmem_base + (uint32_t)(db * dcols);
            for (int a = 0; a < n_atoms; ++a, ++ait) {
                const int slot = (int)(ait % kRing);
                mbar_wait(a_full0 + 8u * slot, (uint32_t)((ait / kRing) & 1));
                tc_fence_after();
                if (lane == 0 && a == 0) trace_ev(p, it, 9);
                if (lane == 0 && a == n_atoms - 1) trace_ev(p, it, 10);
                const bool full = a < p.n_full;
                const int ksteps = full ? 4 : p.tail_cols >> 3;
                const uint32_t a_hi = tmem_base + (uint32_t)(kRingCol0 + slot * 64);
                // descriptors differ between K-steps only in the start-address field: +32 B = +2 units
                const uint64_t dh0 = full ? desc_full + (uint64_t)(a * (c_block >> 4)) : desc_tail;
                const uint64_t dl0 = dh0 + (uint64_t)((p.k_pad * (full ? 128 : p.tail_cols * 4)) >> 4);
                if (elect_one()) {
#pragma unroll
                    for (int kk = 0; kk < 4; ++kk) {
                        if (kk < ksteps) {
                            if (stacked) {
                                mma_tf32_ts(d_tmem, a_hi + 8 * kk, dh0 + 2 * kk, idesc, (a | kk) != 0);
                                mma_tf32_ts(d_tmem, a_hi + 32 + 8 * kk, dh0 + 2 * kk, idesc, 1);
                            } else {
                                mma_tf32_ts(d_tmem, a_hi + 32 + 8 * kk, dl0 + 2 * kk, idesc, (a | kk) != 0);
                                mma_tf32_ts(d_tmem, a_hi + 32 + 8 * kk, dh0 + 2 * kk, idesc, 1);
                                mma_tf32_ts(d_tmem, a_hi + 8 * kk, dl0 + 2 * kk, idesc, 1);
                                mma_tf32_ts(d_tmem, a_hi + 8 * kk, dh0 + 2 * kk, idesc, 1);
                            }
                        }
                    }
                    tc_commit(a_empty0 + 8u * slot);
                    if (a == n_atoms - 1) {
                        tc_commit(d_full0 + 8u * db);
                        trace_ev(p, it, 3);
                    }
                }
                __syncwarp();
            }
        }
    } else if (warp >= 4 && warp < 8) {
        // ======================================= transform ======================================
        const int q = warp & 3;
        const int row = q * 32 + lane;
        const uint32_t tmem_lane = tmem_base + ((uint32_t)(q * 32) << 16);
        float* xn_s = (float*)(smem + p.off_xn);
        int64_t it = 0, ait = 0;
        for (int64_t tile = first_tile; tile < p.num_tiles; tile += tile_step, ++it) {
            const int s = (int)(it % S);
            const int64_t grow = tile * kTileM + row;
            float xn_g = 0.f;
            if (!p.xn_compute && grow < p.n) xn_g = p.xnorm[grow];  // issued early, consumed after the tile
            mbar_wait(x_full0 + 8u * s, (uint32_t)((it / S) & 1));
            if (threadIdx.x == 128) trace_ev(p, it, 1);
            const uint8_t* stage = smem + (size_t)s * stage_bytes;
            float xn;
            if (p.xn_compute) {
                xn = transform_tile<true>(p, stage, row, tmem_lane, a_full0, a_empty0, ait, lane);
                if (grow < p.n) p.xnorm[grow] = xn;
            } else {
                transform_tile<false>(p, stage, row, tmem_lane, a_full0, a_empty0, ait, lane);
                xn = xn_g;
            }
            xn_s[s * kTileM + row] = xn;
            __syncwarp();
            if (lane == 0) mbar_arrive(xn_full0 + 8u * s);
            if (threadIdx.x == 128) trace_ev(p, it, 2);
            if (lane == 0 && q < 2) trace_ev(p, it, 12 + q);
        }
    } else if (warp >= 8 && warp < 16) {
        // ======================================== assign ========================================
        // Two groups of four warps take alternate tiles (group g owns the D buffer and the sort buffer of
        // its parity), so the argmin / refine / sort latency of one tile overlaps the next tile's.
        const int g = (warp - 8) >> 2;
        if (g < ND) {
        const int q = warp & 3;
        const int row = q * 32 + lane;
        const int et = threadIdx.x - (8 + 4 * g) * 32;  // 0..127 within the group
        const int bar_id = 1 + g;
        float* cn_s = (float*)(smem + p.off_cn);
        float* xn_s = (float*)(smem + p.off_xn);
        const float cmax = p.st->cnorm_max;
        int (*hist)[kHist] = s_hist[g];

        double inert = 0.0;
        for (int64_t it = g;; it += ND) {
            const int64_t tile = first_tile + it * tile_step;
            if (tile >= p.num_tiles) break;
            const int s = (int)(it % S);
            const int db = g;
            const int sb = (int)(it & 1);
            const int64_t grow = tile * kTileM + row;
            const bool valid = grow < p.n;
            mbar_wait(d_full0 + 8u * db, (uint32_t)((it / ND) & 1));
            tc_fence_after();
            mbar_wait(x_full0 + 8u * s, (uint32_t)((it / S) & 1));
            mbar_wait(xn_full0 + 8u * s, (uint32_t)((it / S) & 1));
            if (et == 0) trace_ev(p, it, 4);

            // ---- scores -> running (best, second, argmin) ----
            float b1 = FLT_MAX, b2 = FLT_MAX;
            int i1 = 0;
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(db * dcols);
            for (int c0 = 0; c0 < p.k; c0 += 16) {
                uint32_t r[16], r2[16];
                tmem_ld16(taddr + c0, r);
                if (stacked) tmem_ld16(taddr + p.k_pad + c0, r2);
                tmem_wait_ld();
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    float dot = __uint_as_float(r[j]);
                    if (stacked) dot += __uint_as_float(r2[j]);
                    const float sc = fmaf(-2.f, dot, cn_s[c0 + j]);  // padded columns carry FLT_MAX
                    if (sc < b1) {
                        b2 = b1;
                        b1 = sc;
                        i1 = c0 + j;
                    } else {
                        b2 = fminf(b2, sc);
                    }
                }
            }
            const float xn = xn_s[s * kTileM + row];
            const float band = p.tie_tol * (xn + cmax);
            const bool amb = valid && p.k > 1 && (b2 - b1) < band;
            unsigned m = __ballot_sync(0xffffffffu, amb);
            // near-tie rows: a second look at the scores lists the candidates inside the band (first 4 kept)
            int ncand = 0;
            if (m) {
#pragma unroll 1
                for (int c0 = 0; c0 < p.k; c0 += 16) {
                    uint32_t r[16], r2[16];
                    tmem_ld16(taddr + c0, r);
                    if (stacked) tmem_ld16(taddr + p.k_pad + c0, r2);
                    tmem_wait_ld();
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        float dot = __uint_as_float(r[j]);
                        if (stacked) dot += __uint_as_float(r2[j]);
                        const float sc = fmaf(-2.f, dot, cn_s[c0 + j]);
                        if (amb && (c0 + j) < p.k && (sc - b1) < band) {
                            if (ncand < 4) s_cand[g][et][ncand] = c0 + j;
                            ++ncand;
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(d_empty0 + 8u * db);

            const uint8_t* stage = smem + (size_t)s * stage_bytes;
            int label = i1;
            float bestv = fmaxf(xn + b1, 0.f);
            // ---- float64 re-decision, warp-cooperative: lanes split the d dimension ----
            while (m) {
                const int src = __ffs(m) - 1;
                m &= m - 1;
                const int rr = q * 32 + src;
                const int nc = __shfl_sync(0xffffffffu, ncand, src);
                const int total = nc <= 4 ? nc : p.k;  // more than 4 candidates: scan every centroid
                double bd = DBL_MAX;
                int bi = 0;
                for (int t = 0; t < total; ++t) {
                    const int j = nc <= 4 ? s_cand[g][rr][t] : t;
                    const float* c = p.C + (size_t)j * p.d;
                    double sum = 0.0;
                    for (int i = lane; i < p.d; i += 32) {
                        const float xv = *reinterpret_cast<const float*>(stage + chunk_off(p, rr, i >> 2) + ((i & 3) << 2));
                        const double df = (double)xv - (double)__ldg(c + i);
                        sum = fma(df, df, sum);
                    }
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
                    if (sum < bd) {  // candidates come in ascending index order: ties keep the lowest
                        bd = sum;
                        bi = j;
                    }
                }
                if (lane == src) {
                    label = bi;
                    bestv = (float)bd;
                }
            }
            if (valid) {
                if (p.labels) p.labels[grow] = label;
                if (p.best) p.best[grow] = bestv;
                inert += (double)bestv;
            }
            if (et == 0) trace_ev(p, it, 5);

            if (p.acc) {
                // ---- stable counting sort of the tile's rows by label (bin k = rows past the end) ----
                mbar_wait(s_empty0 + 8u * sb, (uint32_t)(((it >> 1) & 1) ^ 1));
                const int bins = p.k + 1;
                const int bin = valid ? label : p.k;
                const unsigned peers = __match_any_sync(0xffffffffu, bin);
                const int rank_w = __popc(peers & ((1u << lane) - 1u));
                if (rank_w == 0) hist[q][bin] = __popc(peers);
                named_bar_sync(bar_id, 128);
                if (q == 0) {
                    // one warp turns the 4 per-warp histograms into offsets; lane owns `per` consecutive bins
                    const int per = (bins + 31) >> 5;
                    const int b0 = lane * per;
                    int lane_total = 0;
                    for (int bb = b0; bb < b0 + per && bb < bins; ++bb)
                        lane_total += hist[0][bb] + hist[1][bb] + hist[2][bb] + hist[3][bb];
                    int incl = lane_total;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) {
                        const int v = __shfl_up_sync(0xffffffffu, incl, o);
                        if (lane >= o) incl += v;
                    }
                    int run = incl - lane_total;
                    for (int bb = b0; bb < b0 + per && bb < bins; ++bb) {
                        const int h0 = hist[0][bb], h1 = hist[1][bb], h2 = hist[2][bb], h3 = hist[3][bb];
                        s_base[sb][bb] = run;
                        hist[0][bb] = run;
                        hist[1][bb] = run + h0;
                        hist[2][bb] = run + h0 + h1;
                        hist[3][bb] = run + h0 + h1 + h2;
                        run += h0 + h1 + h2 + h3;
                    }
                    if (lane == 31) s_base[sb][bins] = run;
                }
                named_bar_sync(bar_id, 128);
                {
                    // sorted position -> row offset inside an atom (r*pitch | swizzle bits), a last-of-run flag
                    // (bit 16) and the label (bits 17+)
                    const int pos = hist[q][bin] + rank_w;
                    const int meta = ((pos == s_base[sb][bin + 1] - 1) ? (1 << 16) : 0) | (bin << 17);
                    s_perm[sb][pos] = (row << 7) | ((row & 7) << 4) | meta;
                    if (p.tail_cols == 16) s_permT[sb][pos] = (row << 6) | (((row >> 1) & 3) << 4) | meta;
                    else if (p.tail_cols == 8) s_permT[sb][pos] = (row << 5) | (((row >> 2) & 1) << 4) | meta;
                }
                __syncwarp();
                for (int bb = lane; bb < bins; bb += 32) hist[q][bb] = 0;  // own row only: ready for the next tile
                __syncwarp();  // the warp's s_perm writes precede its arrival; the update group waits for all four
                if (et == 0) trace_ev(p, it, 6);
                if (lane == 0) mbar_arrive(s_full0 + 8u * sb);
            } else {
                __syncwarp();
                if (lane == 0) mbar_arrive(x_empty0 + 8u * s);
            }
        }
        if (p.acc) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) inert += __shfl_xor_sync(0xffffffffu, inert, o);
            if (lane == 0 && inert != 0.0) atomicAdd(&p.acc[p.k * p.d + p.k], inert);
        }
        }
    } else if (warp >= 16 && p.acc) {
        // ======================================== update ========================================
        // Eight warps; warp u owns the labels [u*k/8, (u+1)*k/8) for the whole kernel, i.e. one contiguous
        // range of each tile's label-sorted rows; lane = float4 column chunk.  No other thread ever touches
        // an owner's (label, column) sums, so no atomics are needed anywhere in the tile loop.
        const int u = warp - 16;
        const int ut = threadIdx.x - 16 * 32;
        double* sacc = (double*)(smem + p.off_acc);
        const int nchunks = p.d >> 2;
        const int l_lo = (u * p.k) >> 3, l_hi = ((u + 1) * p.k) >> 3;
        const bool reg_path = p.k <= 32 && nchunks <= 32;
        if (reg_path) {
            // k <= 32: at most 4 owned labels.  Their FP32 sums live in registers and are folded into the
            // float64 shared accumulators every 8 tiles.  The row loop is branch-free inside a label
            // (predicated adds): a single warp per scheduler pays ~20 cycles for every taken branch.
            const bool active = lane < nchunks;
            const int ch = active ? lane : 0;
            const int a = ch >> 3;
            const bool tail = a >= p.n_full;
            const int col_off = (tail ? p.n_full : a) * kAtomBytes;
            const int c16 = (tail ? (ch - p.n_full * 8) : (ch & 7)) << 4;
            float4 acc[4];
            int cnt[4];
#pragma unroll
            for (int li = 0; li < 4; ++li) {
                acc[li] = make_float4(0.f, 0.f, 0.f, 0.f);
                cnt[li] = 0;
            }
            auto fold = [&]() {
#pragma unroll
                for (int li = 0; li < 4; ++li) {
                    if (l_lo + li < l_hi) {
                        if (active) {
                            double2* dst = reinterpret_cast<double2*>(sacc + (size_t)(l_lo + li) * p.d + ch * 4);
                            double2 a0 = dst[0], a1 = dst[1];
                            a0.x += (double)acc[li].x;
                            a0.y += (double)acc[li].y;
                            a1.x += (double)acc[li].z;
                            a1.y += (double)acc[li].w;
                            dst[0] = a0;
                            dst[1] = a1;
                        }
                        if (lane == 0) sacc[p.k * p.d + l_lo + li] += (double)cnt[li];
                        acc[li] = make_float4(0.f, 0.f, 0.f, 0.f);
                        cnt[li] = 0;
                    }
                }
            };
            int64_t it = 0;
            for (int64_t tile = first_tile; tile < p.num_tiles; tile += tile_step, ++it) {
                const int s = (int)(it % S);
                const int sb = (int)(it & 1);
                mbar_wait(s_full0 + 8u * sb, (uint32_t)((it >> 1) & 1));
                mbar_wait(x_full0 + 8u * s, (uint32_t)((it / S) & 1));
                if (ut == 0) trace_ev(p, it, 11);
                const uint8_t* colp = smem + (size_t)s * stage_bytes + col_off;
                const int* perm = tail ? s_permT[sb] : s_perm[sb];
                int beg = s_base[sb][l_lo];
#pragma unroll
                for (int li = 0; li < 4; ++li) {
                    if (l_lo + li < l_hi) {
                        const int end = s_base[sb][l_lo + li + 1];
                        float4 r = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 1
                        for (int base = beg; base < end; base += 4) {
                            int w[4];
                            float4 v[4];
#pragma unroll
                            for (int j = 0; j < 4; ++j) w[j] = perm[base + j];  // past `end`: stale but valid offsets
#pragma unroll
                            for (int j = 0; j < 4; ++j)
                                v[j] = *reinterpret_cast<const float4*>(colp + ((w[j] & 0xFFFF) ^ c16));
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                if (base + j < end) {
                                    r.x += v[j].x;
                                    r.y += v[j].y;
                                    r.z += v[j].z;
                                    r.w += v[j].w;
                                }
                            }
                        }
                        acc[li].x += r.x;
                        acc[li].y += r.y;
                        acc[li].z += r.z;
                        acc[li].w += r.w;
                        cnt[li] += end - beg;
                        beg = end;
                    }
                }
                __syncwarp();
                if (lane == 0) {
                    mbar_arrive(s_empty0 + 8u * sb);
                    mbar_arrive(x_empty0 + 8u * s);
                }
                if (ut == 0) trace_ev(p, it, 7);
                if ((it & 7) == 7 || tile + tile_step >= p.num_tiles) fold();  // single (rare) fold site
            }
        } else {
            // general shapes: flat walk over the owned range, run sums flushed at the last row of each run
            int64_t it = 0;
            for (int64_t tile = first_tile; tile < p.num_tiles; tile += tile_step, ++it) {
                const int s = (int)(it % S);
                const int sb = (int)(it & 1);
                mbar_wait(s_full0 + 8u * sb, (uint32_t)((it >> 1) & 1));
                mbar_wait(x_full0 + 8u * s, (uint32_t)((it / S) & 1));
                const uint8_t* stage = smem + (size_t)s * stage_bytes;
                const int pb = s_base[sb][l_lo], pe = s_base[sb][l_hi];
                for (int ch = lane; ch < nchunks; ch += 32) {
                    const int a = ch >> 3;
                    const bool tail = a >= p.n_full;
                    const uint8_t* colp = stage + (tail ? p.n_full : a) * kAtomBytes;
                    const int c16 = (tail ? (ch - p.n_full * 8) : (ch & 7)) << 4;
                    const int* perm = tail ? s_permT[sb] : s_perm[sb];
                    float4 r = make_float4(0.f, 0.f, 0.f, 0.f);
                    for (int pos = pb; pos < pe; ++pos) {
                        const int w = perm[pos];
                        const float4 v = *reinterpret_cast<const float4*>(colp + ((w & 0xFFFF) ^ c16));
                        r.x += v.x;
                        r.y += v.y;
                        r.z += v.z;
                        r.w += v.w;
                        if (w & (1 << 16)) {  // last row of this label's run (warp-uniform)
                            double2* dst = reinterpret_cast<double2*>(sacc + (size_t)(w >> 17) * p.d + ch * 4);
                            double2 a0 = dst[0], a1 = dst[1];
                            a0.x += (double)r.x;
                            a0.y += (double)r.y;
                            a1.x += (double)r.z;
                            a1.y += (double)r.w;
                            dst[0] = a0;
                            dst[1] = a1;
                            r = make_float4(0.f, 0.f, 0.f, 0.f);
                        }
                    }
                }
                for (int l = l_lo + lane; l < l_hi; l += 32)
                    sacc[p.k * p.d + l] += (double)(s_base[sb][l + 1] - s_base[sb][l]);
                __syncwarp();
                if (lane == 0) {
                    mbar_arrive(s_empty0 + 8u * sb);
                    mbar_arrive(x_empty0 + 8u * s);
                }
            }
        }
        named_bar_sync(3, 256);
        const int acc_n = p.k * p.d + p.k;
#pragma unroll 1
        for (int i = ut; i < acc_n; i += 256) {
            const double v = sacc[i];
            if (v != 0.0) atomicAdd(&p.acc[i], v);
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
    }
}


constexpr int kStaticSmem = 17 * 1024;  // upper bound of the kernel's static shared arrays

struct Plan {
    int k_pad, n_full, tail_cols, d_pad, num_stages, num_d, stage_bytes;
    int off_c, off_acc, off_cn, off_xn, off_bar, total;
    bool ok;
};

Plan make_plan(int d, int k, bool with_acc) {
    Plan pl = {};
    pl.ok = false;
    if (d <= 0 || k <= 0 || d > kMaxD || k > kMaxK || (d & 3)) return pl;
    pl.k_pad = (k + 15) & ~15;
    // K-atoms: 32-float SWIZZLE_128B atoms, and a narrow SWIZZLE_32B / 64B tail when d % 32 is in 4..16
    const int rem = d & 31;
    pl.n_full = d >> 5;
    pl.tail_cols = 0;
    if (rem > 16) pl.n_full += 1;
    else if (rem > 8) pl.tail_cols = 16;
    else if (rem > 0) pl.tail_cols = 8;
    pl.d_pad = pl.n_full * 32 + pl.tail_cols;
    const int dcols = pl.k_pad <= 128 ? 2 * pl.k_pad : pl.k_pad;
    pl.num_d = dcols <= 128 ? 2 : 1;  // D buffers live in TMEM columns [0,256)
    pl.stage_bytes = pl.n_full * kAtomBytes + kTileM * pl.tail_cols * 4;
    const int c_bytes = 2 * pl.k_pad * pl.d_pad * 4;
    const int acc_bytes = with_acc ? (int)((sizeof(double) * ((size_t)k * d + k) + 15) & ~(size_t)15) : 0;
    const int fixed = c_bytes + acc_bytes + pl.k_pad * 4 + kMaxStages * kTileM * 4 + 256 /*barriers*/ + 1024 /*align*/ +
                      kStaticSmem;
    int S = (kMaxSmem - fixed) / pl.stage_bytes;
    if (S > kMaxStages) S = kMaxStages;
    if (S < 2) return pl;
    pl.num_stages = S;
    int off = S * pl.stage_bytes;
    off = (off + 1023) & ~1023;
    pl.off_c = off;
    off += c_bytes;
    pl.off_acc = off;
    off += acc_bytes;
    pl.off_cn = off;
    off += pl.k_pad * 4;
    pl.off_xn = off;
    off += kMaxStages * kTileM * 4;
    off = (off + 15) & ~15;
    pl.off_bar = off;
    off += 256;
    pl.total = off + 1024;
    pl.ok = pl.total + kStaticSmem <= kMaxSmem;
    return pl;
}

}  // namespace

bool tc_supported(const float* X, int d, int k, bool with_acc) {
    if ((reinterpret_cast<uintptr_t>(X) & 15) != 0) return false;
    return make_plan(d, k, with_acc).ok;
}
int tc_max_d() { return kMaxD; }
int tc_max_k() { return kMaxK; }
size_t tc_workspace_bytes(int d, int k) {
    Plan pl = make_plan(d, k, true);
    return pl.ok ? sizeof(float) * 2 * (size_t)pl.k_pad * pl.d_pad : 0;
}
const char* tc_last_error() { return g_tc_err; }

void tc_free(TcWorkspace* ws) {
    if (ws->d_buf) cudaFree(ws->d_buf);
    if (ws->d_pd) cudaFree(ws->d_pd);
    if (ws->d_bits) cudaFree(ws->d_bits);
    if (ws->d_slots) cudaFree(ws->d_slots);
    if (ws->d_open_rows) cudaFree(ws->d_open_rows);
    if (ws->d_open_info) cudaFree(ws->d_open_info);
    if (ws->d_open_blk) cudaFree(ws->d_open_blk);
    ws->d_open_rows = ws->d_open_info = ws->d_open_blk = nullptr;
    ws->open_cap = ws->open_blk_cap = 0;
    ws->d_buf = nullptr;
    ws->d_pd = nullptr;
    ws->d_bits = nullptr;
    ws->d_slots = nullptr;
    ws->bits_cap = 0;
    ws->slots_cap = 0;
    ws->cap = 0;
    ws->pd_cap = 0;
}

void launch_split_centroids(const float* C, int k, int d, int k_pad, int d_pad, float* out, cudaStream_t s) {
    const int total = k_pad * d_pad;
    int blocks = (total + 255) / 256;
    if (blocks > kNumSMs) blocks = kNumSMs;
    split_centroids_kernel<<<blocks, 256, 0, s>>>(C, k, d, k_pad, d_pad, out);
    count_launch();
}

int tc_pass(TcWorkspace* ws, const float* X, int64_t n, int d, int k, const float* C, const float* cnorm,
            float* xnorm, bool xn_compute, int32_t* labels, float* best, double* acc, IterState* st,
            bool honor_done, cudaStream_t s) {
    const Plan pl = make_plan(d, k, acc != nullptr);
    if (!pl.ok) {
        snprintf(g_tc_err, sizeof(g_tc_err), "shape d=%d k=%d does not fit the fused tcgen05 kernel", d, k);
        return 1;
    }
    const size_t need = sizeof(float) * 2 * (size_t)pl.k_pad * pl.d_pad;
    if (need > ws->cap) {
        if (ws->d_buf) cudaFree(ws->d_buf);
        ws->d_buf = nullptr;
        if (cudaMalloc(&ws->d_buf, need) != cudaSuccess) {
            snprintf(g_tc_err, sizeof(g_tc_err), "cudaMalloc of %zu bytes for split centroids failed", need);
            return 1;
        }
        ws->cap = need;
    }
    CUtensorMap tmX, tmXt, tmC, tmCt;
    const uint32_t full_box = pl.n_full ? 32 : (uint32_t)pl.tail_cols;   // d < 32 with a narrow tail only
    const uint32_t tail_box = pl.tail_cols ? (uint32_t)pl.tail_cols : full_box;
    if (encode_2d(g_tc_err, sizeof(g_tc_err), &tmX, X, (uint64_t)d, (uint64_t)n, (uint64_t)d * 4, full_box, kTileM)) return 1;
    if (encode_2d(g_tc_err, sizeof(g_tc_err), &tmXt, X, (uint64_t)d, (uint64_t)n, (uint64_t)d * 4, tail_box, kTileM)) return 1;
    if (encode_2d(g_tc_err, sizeof(g_tc_err), &tmC, ws->d_buf, (uint64_t)pl.d_pad, (uint64_t)2 * pl.k_pad, (uint64_t)pl.d_pad * 4, full_box,
                  (uint32_t)pl.k_pad))
        return 1;
    if (encode_2d(g_tc_err, sizeof(g_tc_err), &tmCt, ws->d_buf, (uint64_t)pl.d_pad, (uint64_t)2 * pl.k_pad, (uint64_t)pl.d_pad * 4, tail_box,
                  (uint32_t)pl.k_pad))
        return 1;

    static PerDeviceOnce attr_set;  // the opt-in is a per-device attribute
    const int dev = PerDeviceOnce::current();
    if (dev < 0 || !attr_set.done[dev]) {
        cudaFuncAttributes fa;
        if (cudaFuncGetAttributes(&fa, kmeans_fused_tc_kernel) != cudaSuccess ||
            (int)fa.sharedSizeBytes > kStaticSmem) {
            snprintf(g_tc_err, sizeof(g_tc_err), "unexpected static shared memory size of the fused kernel");
            return 1;
        }
        const int max_dyn = kMaxSmem - (int)fa.sharedSizeBytes;
        if (cudaFuncSetAttribute(kmeans_fused_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, max_dyn) !=
            cudaSuccess) {
            snprintf(g_tc_err, sizeof(g_tc_err), "cannot raise dynamic shared memory to %d bytes", max_dyn);
            return 1;
        }
        if (dev >= 0) attr_set.done[dev] = true;
    }
    launch_split_centroids(C, k, d, pl.k_pad, pl.d_pad, (float*)ws->d_buf, s);
    TcParams p;
    p.n = n;
    p.num_tiles = (n + kTileM - 1) / kTileM;
    p.d = d;
    p.k = k;
    p.k_pad = pl.k_pad;
    p.n_full = pl.n_full;
    p.tail_cols = pl.tail_cols;
    p.num_stages = pl.num_stages;
    p.num_d = pl.num_d;
    p.stage_bytes = pl.stage_bytes;
    p.C = C;
    p.cnorm = cnorm;
    p.xnorm = xnorm;
    p.xn_compute = xn_compute ? 1 : 0;
    p.labels = labels;
    p.best = best;
    p.acc = acc;
    p.st = st;
    p.honor_done = honor_done ? 1 : 0;
    // Near-tie band of the split-TF32 scores: the tensor core adds each MMA's products into the FP32
    // accumulator with truncation (<= 1 ulp of the running sum per MMA, 2*ceil(d/8) MMAs per row), on top
    // of the ~2^-20 relative error of the truncated operand splits.  Rows inside the band are re-decided
    // in float64, so the band only has to be an upper bound.
    p.tie_tol = (float)(2 * ((d + 7) / 8) + 4) * 2.384185791015625e-07f + 4e-6f;
    p.off_c = pl.off_c;
    p.off_acc = pl.off_acc;
    p.off_cn = pl.off_cn;
    p.off_xn = pl.off_xn;
    p.off_bar = pl.off_bar;
    p.trace = nullptr;
#ifdef B200KM_ENABLE_TRACE  // development builds only
    const char* trace_path = getenv("B200KM_TRACE");
    if (trace_path && acc) {
        cudaMalloc((void**)&p.trace, 256 * 16 * sizeof(unsigned long long));
        cudaMemsetAsync(p.trace, 0, 256 * 16 * sizeof(unsigned long long), s);
    }
#endif
    int64_t grid = p.num_tiles < kNumSMs ? p.num_tiles : kNumSMs;
    if (grid < 1) grid = 1;
    kmeans_fused_tc_kernel<<<(unsigned)grid, kThreads, pl.total, s>>>(tmX, tmXt, tmC, tmCt, p);
    count_launch();
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        snprintf(g_tc_err, sizeof(g_tc_err), "fused kernel launch failed: %s", cudaGetErrorString(e));
        return 1;
    }
#ifdef B200KM_ENABLE_TRACE
    if (p.trace) {  // development aid: synchronous dump of CTA 0's per-tile event clocks
        static unsigned long long host[256 * 16];
        cudaStreamSynchronize(s);
        cudaMemcpy(host, p.trace, sizeof(host), cudaMemcpyDeviceToHost);
        cudaFree(p.trace);
        if (FILE* f = fopen(trace_path, "w")) {
            for (int i = 0; i < 256; ++i) {
                for (int j = 0; j < 16; ++j) fprintf(f, "%llu ", host[i * 16 + j]);
                fprintf(f, "\n");
            }
            fclose(f);
        }
    }
#endif
    return 0;
}

}  // namespace b200km
