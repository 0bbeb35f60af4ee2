// tcgen05 / TMEM / TMA path: one persistent, warp-specialised kernel per Lloyd pass.
//
// For every 128-row tile of X the kernel computes  score_j = ||c_j||^2 - 2 x.c_j  on the 5th-gen
// tensor cores in 3xTF32 (x = x_hi + x_lo, c = c_hi + c_lo; products hi*hi + hi*lo + lo*hi with FP32
// accumulation in TMEM), takes the running (best, second-best, argmin) per row, re-decides near-ties
// in float64, and - in the same pass, while the tile is still in shared memory - adds the rows into
// per-cluster float64 sums / counts held in shared memory.  X is read from HBM exactly once per pass.
//
// Warp roles (320 threads, 1 CTA per SM, grid = #SMs):
//   warp 0      TMA producer: split centroids once, then X tiles [128 x 32-float atoms], SWIZZLE_128B
//   warp 1      TMEM allocator + MMA issuer (one elected lane issues tcgen05.mma kind::tf32, A from
//               TMEM, B = split centroids from shared memory through UMMA descriptors)
//   warps 2-5   transform: each thread owns one row; reads it from the swizzled tile, writes the exact
//               tf32 parts x_hi | x_lo into a TMEM ring slot (tcgen05.st) and accumulates ||x||^2
//   warps 6-9   epilogue: tcgen05.ld the scores, argmin, float64 refine of near-ties, labels / distances
//               out, then label-sorted run accumulation of the tile into the float64 accumulators
//
// This kernel replaces cuML's fusedL2NN + reduce_rows_by_key pair that runs behind reference
// python/src/spark_rapids_ml/clustering.py:367-370 (fit) and :493-497 (predict).
#include "tc.cuh"

#include <cuda.h>

#include <cfloat>
#include <climits>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace b200km {

namespace {

thread_local char g_tc_err[512] = "";

constexpr int kTileM = 128;
constexpr int kAtomBytes = kTileM * 128;  // one K-atom of an X tile: 128 rows x 32 floats
constexpr int kRing = 4;                  // TMEM ring slots for (x_hi | x_lo) atoms, 64 columns each
constexpr int kRingCol0 = 256;            // ring occupies TMEM columns [256, 512)
constexpr int kThreads = 320;
constexpr int kMaxSmem = 227 * 1024;
constexpr int kMaxStages = 4;
constexpr int kMaxD = 256;
constexpr int kMaxK = 256;

struct TcParams {
    int64_t n;
    int64_t num_tiles;
    int d, k, k_pad, n_atoms;
    int num_stages, num_d;
    const float* C;
    const float* cnorm;
    int32_t* labels;
    float* best;
    double* acc;
    IterState* st;
    int honor_done;
    float tie_tol;
    unsigned long long* trace;  // development aid (B200KM_TRACE): [256 tiles][8 events] clock64 of CTA 0
    // shared-memory offsets (bytes from the 1024-aligned base)
    int off_c, off_acc, off_cn, off_xn, off_sort, off_bar;
};

// ---- PTX wrappers ------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t"
        "}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0;
}
// Polls with a short sleep between probes: a warp that spins on try_wait competes for issue slots with
// the warps doing the work on its scheduler (measured: the spinning roles slowed the others ~2x).
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    if (mbar_try(bar, parity)) return;
    while (!mbar_try(bar, parity)) __nanosleep(40);
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
        "l"(tm), "r"(bar), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem desc], kind::tf32, M=128
__device__ __forceinline__ void mma_tf32_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc,
                                            uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t"
        "}" ::"r"(d_tmem),
        "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, "
        "%15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
        "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
        "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]),
        "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]),
        "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
        : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, "
        "[%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// K-major SWIZZLE_128B shared-memory matrix descriptor (sm_100 UMMA): start>>4 | LBO=1 | SBO=1024B |
// version 1 | layout type 2.
__device__ __forceinline__ uint64_t make_b_desc(uint32_t smem_addr) {
    uint64_t desc = 0;
    desc |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
    desc |= (uint64_t)1 << 16;
    desc |= (uint64_t)(1024 >> 4) << 32;
    desc |= (uint64_t)1 << 46;
    desc |= (uint64_t)2 << 61;
    return desc;
}

// byte offset of element (row r, float column i) inside one tile stage laid out as 32-float atoms with
// the TMA 128-byte swizzle (16-byte chunk index XOR (row & 7))
__device__ __forceinline__ uint32_t tile_off(int r, int i) {
    const int a = i >> 5, c = (i & 31) >> 2, e = i & 3;
    return (uint32_t)(a * kAtomBytes + r * 128 + (((c ^ (r & 7)) << 4) | (e << 2)));
}

__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "elect.sync _|p, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t"
        "}"
        : "=r"(pred));
    return pred != 0;
}

__device__ __forceinline__ uint32_t tf32_trunc(uint32_t bits) { return bits & 0xFFFFE000u; }
__device__ __forceinline__ uint32_t tf32_round(uint32_t bits) { return (bits + 0x1000u) & 0xFFFFE000u; }

struct Best {
    float b1, b2;
    int i1;
};

__device__ __forceinline__ void trace_ev(const TcParams& p, int64_t it, int ev) {
    if (p.trace && blockIdx.x == 0 && it < 256) p.trace[it * 8 + ev] = (unsigned long long)clock64();
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

__global__ void __launch_bounds__(kThreads, 1)
    kmeans_fused_tc_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmC,
                           const TcParams p) {
    if (p.honor_done && p.st->done) return;

    extern __shared__ uint8_t smem_raw[];
    __shared__ int s_hist[4][kMaxK + 8];  // per-epilogue-warp label histograms, then offsets
    __shared__ int s_base[kMaxK + 8];     // start of every label's run in the sorted order
    __shared__ int s_perm[kTileM];        // sorted position -> row of the tile
    // pointer arithmetic (not an integer round trip) keeps the shared address space visible to the compiler
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const uint32_t smem_base = smem_u32(smem);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int S = p.num_stages;
    const int ND = p.num_d;
    const int stage_bytes = p.n_atoms * kAtomBytes;
    const int c_block = p.k_pad * 128;  // bytes of one (part, atom) block of the split centroids

    // barrier table
    const uint32_t bar0 = smem_base + p.off_bar;
    auto x_full = [&](int s) { return bar0 + 8u * s; };
    auto x_empty = [&](int s) { return bar0 + 8u * (kMaxStages + s); };
    auto xn_full = [&](int s) { return bar0 + 8u * (2 * kMaxStages + s); };
    auto a_full = [&](int s) { return bar0 + 8u * (3 * kMaxStages + s); };
    auto a_empty = [&](int s) { return bar0 + 8u * (3 * kMaxStages + kRing + s); };
    auto d_full = [&](int s) { return bar0 + 8u * (3 * kMaxStages + 2 * kRing + s); };
    auto d_empty = [&](int s) { return bar0 + 8u * (3 * kMaxStages + 2 * kRing + 2 + s); };
    const uint32_t c_full = bar0 + 8u * (3 * kMaxStages + 2 * kRing + 4);
    uint32_t* tmem_slot = (uint32_t*)(smem + p.off_bar + 8 * (3 * kMaxStages + 2 * kRing + 5));

    if (threadIdx.x == 0) {
        for (int s = 0; s < kMaxStages; ++s) {
            mbar_init(x_full(s), 1);
            mbar_init(x_empty(s), 4);
            mbar_init(xn_full(s), 4);
        }
        for (int s = 0; s < kRing; ++s) {
            mbar_init(a_full(s), 4);
            mbar_init(a_empty(s), 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(d_full(s), 1);
            mbar_init(d_empty(s), 4);
        }
        mbar_init(c_full, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
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
            mbar_arrive_expect_tx(c_full, (uint32_t)(2 * p.n_atoms * c_block));
            for (int a = 0; a < p.n_atoms; ++a)
                for (int part = 0; part < 2; ++part)  // atom block = [c_hi rows | c_lo rows], 2*k_pad rows
                    tma_load_2d(smem_base + p.off_c + (2 * a + part) * c_block, &tmC, c_full, a * 32,
                                part * p.k_pad);
            int64_t it = 0;
            for (int64_t tile = first_tile; tile < p.num_tiles; tile += tile_step, ++it) {
                const int s = (int)(it % S);
                const uint32_t ph = (uint32_t)((it / S) & 1);
                mbar_wait(x_empty(s), ph ^ 1u);
                trace_ev(p, it, 0);
                mbar_arrive_expect_tx(x_full(s), (uint32_t)stage_bytes);
                for (int a = 0; a < p.n_atoms; ++a)
                    tma_load_2d(smem_base + s * stage_bytes + a * kAtomBytes, &tmX, x_full(s), a * 32,
                                (int)(tile * kTileM));
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ====================================== MMA issuer ======================================
        // The whole warp runs the loop and one elected lane issues (a tcgen05.mma issued from divergent
        // code costs ~100 cycles per instruction, ~40 from warp-uniform code: tools/mma_bench.cu).
        // k_pad <= 128: B is the stacked block [c_hi ; c_lo] (N = 2*k_pad), so a K-step is two instructions of
        // one shape:  D[:, 0:2k) (+)= x_hi.[c_hi;c_lo]^T   and   D[:, 0:2k) += x_lo.[c_hi;c_lo]^T  (all four
        // partial products of (x_hi+x_lo).(c_hi+c_lo)).
        const uint32_t idesc_base = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(kTileM >> 4) << 24);
        const uint32_t idesc_n = idesc_base | ((uint32_t)(p.k_pad >> 3) << 17);
        const uint32_t idesc_2n = idesc_base | ((uint32_t)((2 * p.k_pad) >> 3) << 17);
        const bool stacked = p.k_pad <= 128;
        const int dcols = stacked ? 2 * p.k_pad : p.k_pad;
        mbar_wait(c_full, 0);
        int64_t it = 0;   // tile counter
        int64_t ait = 0;  // atom counter (ring position)
        for (int64_t tile = first_tile; tile < p.num_tiles; tile += tile_step, ++it) {
            const int db = (int)(it % ND);
            mbar_wait(d_empty(db), (uint32_t)(((it / ND) & 1) ^ 1));
            tc_fence_after();
            const uint32_t d_tmem = tmem_base + (uint32_t)(db * dcols);
            for (int a = 0; a < p.n_atoms; ++a, ++ait) {
                const int slot = (int)(ait % kRing);
                mbar_wait(a_full(slot), (uint32_t)((ait / kRing) & 1));
                tc_fence_after();
                const int rem = p.d - a * 32;
                const int ksteps = rem >= 32 ? 4 : (rem + 7) >> 3;
                const uint32_t a_hi = tmem_base + (uint32_t)(kRingCol0 + slot * 64);
                const uint32_t a_lo = a_hi + 32;
                const uint32_t b_hi = smem_base + p.off_c + (2 * a) * c_block;
                const uint32_t b_lo = b_hi + c_block;
                for (int kk = 0; kk < ksteps; ++kk) {
                    const uint64_t dh = make_b_desc(b_hi + kk * 32);
                    if (stacked) {
                        // same shape and same B for both: D[:,0:k) += (x_hi+x_lo).c_hi, D[:,k:2k) += (x_hi+x_lo).c_lo
                        // (alternating instruction shapes between consecutive MMAs stalls the tensor pipe)
                        if (elect_one()) mma_tf32_ts(d_tmem, a_hi + 8 * kk, dh, idesc_2n, (a | kk) != 0);
                        if (elect_one()) mma_tf32_ts(d_tmem, a_lo + 8 * kk, dh, idesc_2n, 1);
                    } else {
                        const uint64_t dl = make_b_desc(b_lo + kk * 32);
                        if (elect_one()) mma_tf32_ts(d_tmem, a_lo + 8 * kk, dh, idesc_n, (a | kk) != 0);
                        if (elect_one()) mma_tf32_ts(d_tmem, a_hi + 8 * kk, dl, idesc_n, 1);
                        if (elect_one()) mma_tf32_ts(d_tmem, a_hi + 8 * kk, dh, idesc_n, 1);
                    }
                }
                if (elect_one()) tc_commit(a_empty(slot));
                if (a == p.n_atoms - 1) {
                    if (elect_one()) tc_commit(d_full(db));
                    if (lane == 0) trace_ev(p, it, 3);
                }
                __syncwarp();
            }
        }
    } else if (warp < 6) {
        // ======================================= transform ======================================
        const int q = warp & 3;
        const int row = q * 32 + lane;
        float* xn_s = (float*)(smem + p.off_xn);
        int64_t it = 0, ait = 0;
        for (int64_t tile = first_tile; tile < p.num_tiles; tile += tile_step, ++it) {
            const int s = (int)(it % S);
            mbar_wait(x_full(s), (uint32_t)((it / S) & 1));
            if (threadIdx.x == 64) trace_ev(p, it, 1);
            const uint8_t* stage = smem + (size_t)s * stage_bytes;
            float xn = 0.f;
            for (int a = 0; a < p.n_atoms; ++a, ++ait) {
                const int slot = (int)(ait % kRing);
                uint32_t hi[32], lo[32];
                const uint8_t* rowp = stage + a * kAtomBytes + row * 128;
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    const float4 v = *reinterpret_cast<const float4*>(rowp + ((c ^ (row & 7)) << 4));
                    const float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                    for (int t = 0; t < 4; ++t) {
                        const uint32_t h = tf32_trunc(__float_as_uint(e[t]));
                        xn = fmaf(e[t], e[t], xn);
                        hi[c * 4 + t] = h;
                        lo[c * 4 + t] = tf32_round(__float_as_uint(e[t] - __uint_as_float(h)));
                    }
                }
                mbar_wait(a_empty(slot), (uint32_t)(((ait / kRing) & 1) ^ 1));
                tc_fence_after();
                const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(kRingCol0 + slot * 64);
                tmem_st32(taddr, hi);
                tmem_st32(taddr + 32, lo);
                tmem_wait_st();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(a_full(slot));
            }
            xn_s[s * kTileM + row] = xn;
            __syncwarp();
            if (lane == 0) mbar_arrive(xn_full(s));
            if (threadIdx.x == 64) trace_ev(p, it, 2);
        }
    } else {
        // ======================================= epilogue =======================================
        const int q = warp & 3;
        const int row = q * 32 + lane;
        const int et = threadIdx.x - 6 * 32;  // 0..127 within the epilogue group
        float* cn_s = (float*)(smem + p.off_cn);
        float* xn_s = (float*)(smem + p.off_xn);
        double* sacc = (double*)(smem + p.off_acc);
        const int acc_n = p.k * p.d + p.k;
        for (int j = et; j < p.k_pad; j += 128) cn_s[j] = (j < p.k) ? p.cnorm[j] : FLT_MAX;
        if (p.acc)
            for (int i = et; i < acc_n; i += 128) sacc[i] = 0.0;
        for (int i = et; i < 4 * (kMaxK + 8); i += 128) (&s_hist[0][0])[i] = 0;
        const float cmax = p.st->cnorm_max;
        const bool stacked = p.k_pad <= 128;
        const int dcols = stacked ? 2 * p.k_pad : p.k_pad;
        named_bar_sync(1, 128);

        double inert = 0.0;
        int64_t it = 0;
        for (int64_t tile = first_tile; tile < p.num_tiles; tile += tile_step, ++it) {
            const int s = (int)(it % S);
            const int db = (int)(it % ND);
            const int64_t grow = tile * kTileM + row;
            const bool valid = grow < p.n;
            mbar_wait(d_full(db), (uint32_t)((it / ND) & 1));
            tc_fence_after();
            mbar_wait(x_full(s), (uint32_t)((it / S) & 1));
            mbar_wait(xn_full(s), (uint32_t)((it / S) & 1));
            if (et == 0) trace_ev(p, it, 4);

            // ---- scores -> running (best, second, argmin) ----
            Best b;
            b.b1 = FLT_MAX;
            b.b2 = FLT_MAX;
            b.i1 = 0;
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(db * dcols);
            for (int c0 = 0; c0 < p.k_pad; c0 += 16) {
                uint32_t r[16], r2[16];
                tmem_ld16(taddr + c0, r);
                if (stacked) tmem_ld16(taddr + p.k_pad + c0, r2);
                tmem_wait_ld();
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    float dot = __uint_as_float(r[j]);
                    if (stacked) dot += __uint_as_float(r2[j]);
                    const float sc = fmaf(-2.f, dot, cn_s[c0 + j]);
                    if (sc < b.b1) {
                        b.b2 = b.b1;
                        b.b1 = sc;
                        b.i1 = c0 + j;
                    } else {
                        b.b2 = fminf(b.b2, sc);
                    }
                }
            }
            const float xn = xn_s[s * kTileM + row];
            const float band = p.tie_tol * (xn + cmax);
            const bool amb = valid && p.k > 1 && (b.b2 - b.b1) < band;
            unsigned m = __ballot_sync(0xffffffffu, amb);
            // near-tie rows: second look at the scores to list the candidates inside the band (up to 4)
            int ncand = 0, cand0 = 0, cand1 = 0, cand2 = 0, cand3 = 0;
            if (m) {
                for (int c0 = 0; c0 < p.k_pad; c0 += 16) {
                    uint32_t r[16], r2[16];
                    tmem_ld16(taddr + c0, r);
                    if (stacked) tmem_ld16(taddr + p.k_pad + c0, r2);
                    tmem_wait_ld();
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        float dot = __uint_as_float(r[j]);
                        if (stacked) dot += __uint_as_float(r2[j]);
                        const float sc = fmaf(-2.f, dot, cn_s[c0 + j]);
                        if (amb && (c0 + j) < p.k && (sc - b.b1) < band) {
                            if (ncand == 0) cand0 = c0 + j;
                            else if (ncand == 1) cand1 = c0 + j;
                            else if (ncand == 2) cand2 = c0 + j;
                            else if (ncand == 3) cand3 = c0 + j;
                            ++ncand;
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(d_empty(db));

            const uint8_t* stage = smem + (size_t)s * stage_bytes;
            int label = b.i1;
            float bestv = fmaxf(xn + b.b1, 0.f);
            // ---- float64 re-decision, warp-cooperative: lanes split the d dimension ----
            while (m) {
                const int src = __ffs(m) - 1;
                m &= m - 1;
                const int rr = q * 32 + src;
                const int nc = __shfl_sync(0xffffffffu, ncand, src);
                const int c_0 = __shfl_sync(0xffffffffu, cand0, src), c_1 = __shfl_sync(0xffffffffu, cand1, src);
                const int c_2 = __shfl_sync(0xffffffffu, cand2, src), c_3 = __shfl_sync(0xffffffffu, cand3, src);
                const int total = nc <= 4 ? nc : p.k;  // more than 4 candidates: scan every centroid
                double bd = DBL_MAX;
                int bi = 0;
                for (int t = 0; t < total; ++t) {
                    const int j = nc <= 4 ? (t == 0 ? c_0 : t == 1 ? c_1 : t == 2 ? c_2 : c_3) : t;
                    const float* c = p.C + (size_t)j * p.d;
                    double sum = 0.0;
                    for (int i = lane; i < p.d; i += 32) {
                        const float xv = *reinterpret_cast<const float*>(stage + tile_off(rr, i));
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
                const int bins = p.k + 1;
                const int bin = valid ? label : p.k;
                const unsigned peers = __match_any_sync(0xffffffffu, bin);
                const int rank_w = __popc(peers & ((1u << lane) - 1u));
                if (rank_w == 0) s_hist[q][bin] = __popc(peers);
                named_bar_sync(1, 128);
                if (q == 0) {
                    // one warp turns the 4 per-warp histograms into offsets; lane owns `per` consecutive bins
                    const int per = (bins + 31) >> 5;
                    const int b0 = lane * per;
                    int lane_total = 0;
                    for (int bb = b0; bb < b0 + per && bb < bins; ++bb)
                        lane_total += s_hist[0][bb] + s_hist[1][bb] + s_hist[2][bb] + s_hist[3][bb];
                    int incl = lane_total;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) {
                        const int v = __shfl_up_sync(0xffffffffu, incl, o);
                        if (lane >= o) incl += v;
                    }
                    int run = incl - lane_total;
                    for (int bb = b0; bb < b0 + per && bb < bins; ++bb) {
                        const int h0 = s_hist[0][bb], h1 = s_hist[1][bb], h2 = s_hist[2][bb], h3 = s_hist[3][bb];
                        s_base[bb] = run;
                        s_hist[0][bb] = run;
                        s_hist[1][bb] = run + h0;
                        s_hist[2][bb] = run + h0 + h1;
                        s_hist[3][bb] = run + h0 + h1 + h2;
                        const int tot = h0 + h1 + h2 + h3;
                        if (bb < p.k && tot) sacc[p.k * p.d + bb] += (double)tot;  // sole owner of this count
                        run += tot;
                    }
                    if (lane == 31) s_base[bins] = run;
                }
                named_bar_sync(1, 128);
                s_perm[s_hist[q][bin] + rank_w] = row | (bin << 8);  // sorted position -> (row, label)
                __syncwarp();
                for (int bb = lane; bb < bins; bb += 32) s_hist[q][bb] = 0;  // own row only: ready for the next tile
                named_bar_sync(1, 128);
                if (et == 0) trace_ev(p, it, 6);
                // ---- accumulation: warp q owns the labels [q*k/4, (q+1)*k/4), i.e. one contiguous range of
                // sorted positions; lane = float4 column chunk.  Consecutive equal-label rows are summed in
                // FP32 registers and flushed once per run with a plain float64 read-modify-write (every
                // (label, column) has exactly one owner per tile, so no atomics are needed).
                const int l_lo = (q * p.k) >> 2, l_hi = ((q + 1) * p.k) >> 2;
                const int pb = s_base[l_lo], pe = s_base[l_hi];
                const int nchunks = p.d >> 2;
                for (int ch = lane; ch < nchunks; ch += 32) {
                    const uint8_t* colp = stage + (ch >> 3) * kAtomBytes;
                    const int c = ch & 7;
                    int cur = -1;
                    float4 run = make_float4(0.f, 0.f, 0.f, 0.f);
                    for (int base = pb; base < pe; base += 8) {
                        float4 v[8];
                        int pl[8];
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const int pos = base + j < pe ? base + j : pe - 1;
                            pl[j] = s_perm[pos];
                        }
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const int r = pl[j] & 127;
                            v[j] = *reinterpret_cast<const float4*>(colp + r * 128 + ((c ^ (r & 7)) << 4));
                        }
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            if (base + j < pe) {
                                const int l = pl[j] >> 8;
                                if (l != cur) {
                                    if (cur >= 0) {
                                        double2* dst = reinterpret_cast<double2*>(sacc + (size_t)cur * p.d + ch * 4);
                                        double2 a0 = dst[0], a1 = dst[1];
                                        a0.x += (double)run.x;
                                        a0.y += (double)run.y;
                                        a1.x += (double)run.z;
                                        a1.y += (double)run.w;
                                        dst[0] = a0;
                                        dst[1] = a1;
                                    }
                                    run = make_float4(0.f, 0.f, 0.f, 0.f);
                                    cur = l;
                                }
                                run.x += v[j].x;
                                run.y += v[j].y;
                                run.z += v[j].z;
                                run.w += v[j].w;
                            }
                        }
                    }
                    if (cur >= 0) {
                        double2* dst = reinterpret_cast<double2*>(sacc + (size_t)cur * p.d + ch * 4);
                        double2 a0 = dst[0], a1 = dst[1];
                        a0.x += (double)run.x;
                        a0.y += (double)run.y;
                        a1.x += (double)run.z;
                        a1.y += (double)run.w;
                        dst[0] = a0;
                        dst[1] = a1;
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(x_empty(s));
            if (et == 0) trace_ev(p, it, 7);
        }

        if (p.acc) {
            named_bar_sync(1, 128);
            for (int i = et; i < acc_n; i += 128) {
                const double v = sacc[i];
                if (v != 0.0) atomicAdd(&p.acc[i], v);
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) inert += __shfl_xor_sync(0xffffffffu, inert, o);
            if (lane == 0 && inert != 0.0) atomicAdd(&p.acc[acc_n], inert);
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
    }
}

// ---- host side ---------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode() {
    static EncodeTiledFn fn = nullptr;
    if (fn) return fn;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
        return nullptr;
    fn = (EncodeTiledFn)p;
    return fn;
}

int encode_2d(CUtensorMap* tm, const void* base, uint64_t cols, uint64_t rows, uint64_t row_pitch_bytes,
              uint32_t box_cols, uint32_t box_rows) {
    EncodeTiledFn enc = get_encode();
    if (!enc) {
        snprintf(g_tc_err, sizeof(g_tc_err), "cuTensorMapEncodeTiled not available from the driver");
        return 1;
    }
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {row_pitch_bytes};
    cuuint32_t box[2] = {box_cols, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        snprintf(g_tc_err, sizeof(g_tc_err), "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
        return 1;
    }
    return 0;
}

struct Plan {
    int k_pad, n_atoms, d_pad, num_stages, num_d;
    int off_c, off_acc, off_cn, off_xn, off_sort, off_bar, total;
    bool ok;
};

Plan make_plan(int d, int k, bool with_acc) {
    Plan pl = {};
    pl.ok = false;
    if (d <= 0 || k <= 0 || d > kMaxD || k > kMaxK || (d & 3)) return pl;
    pl.k_pad = (k + 15) & ~15;
    pl.n_atoms = (d + 31) / 32;
    pl.d_pad = pl.n_atoms * 32;
    pl.num_d = (pl.k_pad <= 128 ? 2 * pl.k_pad : pl.k_pad) <= 128 ? 2 : 1;  // D buffers in TMEM columns [0,256)
    const int stage = pl.n_atoms * kAtomBytes;
    const int c_bytes = 2 * pl.n_atoms * pl.k_pad * 128;
    int fixed = c_bytes;
    const int acc_bytes = with_acc ? (int)(sizeof(double) * ((size_t)k * d + k)) : 0;
    fixed += (acc_bytes + 15) & ~15;
    fixed += pl.k_pad * 4;            // cn
    fixed += kMaxStages * kTileM * 4;  // xn
    fixed += 3 * kTileM * 4;           // sort arrays
    fixed += 512;                      // barriers + tmem slot
    fixed += 1024;                     // alignment slack
    fixed += 6 * 1024;                 // static shared arrays of the kernel (histograms, offsets, permutation)
    int S = (kMaxSmem - fixed) / stage;
    if (S > kMaxStages) S = kMaxStages;
    if (S < 2) return pl;
    pl.num_stages = S;
    int off = S * stage;
    pl.off_c = off;
    off += c_bytes;
    pl.off_acc = off;
    off += (acc_bytes + 15) & ~15;
    pl.off_cn = off;
    off += pl.k_pad * 4;
    pl.off_xn = off;
    off += kMaxStages * kTileM * 4;
    pl.off_sort = off;
    off += 3 * kTileM * 4;
    off = (off + 15) & ~15;
    pl.off_bar = off;
    off += 512;
    pl.total = off + 1024;
    pl.ok = pl.total + 6 * 1024 <= kMaxSmem;
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
    ws->d_buf = nullptr;
    ws->cap = 0;
}

int tc_pass(TcWorkspace* ws, const float* X, int64_t n, int d, int k, const float* C, const float* cnorm,
            int32_t* labels, float* best, double* acc, IterState* st, bool honor_done, cudaStream_t s) {
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
    CUtensorMap tmX, tmC;
    if (encode_2d(&tmX, X, (uint64_t)d, (uint64_t)n, (uint64_t)d * 4, 32, kTileM)) return 1;
    if (encode_2d(&tmC, ws->d_buf, (uint64_t)pl.d_pad, (uint64_t)2 * pl.k_pad, (uint64_t)pl.d_pad * 4, 32,
                  (uint32_t)pl.k_pad))
        return 1;

    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncAttributes fa;
        if (cudaFuncGetAttributes(&fa, kmeans_fused_tc_kernel) != cudaSuccess || fa.sharedSizeBytes > 6 * 1024) {
            snprintf(g_tc_err, sizeof(g_tc_err), "unexpected static shared memory size of the fused kernel");
            return 1;
        }
        const int max_dyn = kMaxSmem - (int)fa.sharedSizeBytes;
        if (cudaFuncSetAttribute(kmeans_fused_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, max_dyn) !=
            cudaSuccess) {
            snprintf(g_tc_err, sizeof(g_tc_err), "cannot raise dynamic shared memory to %d bytes", max_dyn);
            return 1;
        }
        attr_set = true;
    }
    {
        const int total = pl.k_pad * pl.d_pad;
        int blocks = (total + 255) / 256;
        if (blocks > kNumSMs) blocks = kNumSMs;
        split_centroids_kernel<<<blocks, 256, 0, s>>>(C, k, d, pl.k_pad, pl.d_pad, (float*)ws->d_buf);
    }
    TcParams p;
    p.n = n;
    p.num_tiles = (n + kTileM - 1) / kTileM;
    p.d = d;
    p.k = k;
    p.k_pad = pl.k_pad;
    p.n_atoms = pl.n_atoms;
    p.num_stages = pl.num_stages;
    p.num_d = pl.num_d;
    p.C = C;
    p.cnorm = cnorm;
    p.labels = labels;
    p.best = best;
    p.acc = acc;
    p.st = st;
    p.honor_done = honor_done ? 1 : 0;
    // Near-tie band of the 3xTF32 scores: the tensor core adds each MMA's products into the FP32
    // accumulator with truncation (<= 1 ulp of the running sum per MMA, 3*ceil(d/8) MMAs per row), on
    // top of the ~2^-21 relative error of the split operands.  Rows inside the band are re-decided in
    // float64, so the band only has to be an upper bound.
    p.tie_tol = (float)(3 * ((d + 7) / 8)) * 2.384185791015625e-07f + 2e-6f;
    p.off_c = pl.off_c;
    p.off_acc = pl.off_acc;
    p.off_cn = pl.off_cn;
    p.off_xn = pl.off_xn;
    p.off_sort = pl.off_sort;
    p.off_bar = pl.off_bar;
    p.trace = nullptr;
    const char* trace_path = getenv("B200KM_TRACE");
    if (trace_path && acc) {
        cudaMalloc((void**)&p.trace, 256 * 8 * sizeof(unsigned long long));
        cudaMemsetAsync(p.trace, 0, 256 * 8 * sizeof(unsigned long long), s);
    }
    int64_t grid = p.num_tiles < kNumSMs ? p.num_tiles : kNumSMs;
    if (grid < 1) grid = 1;
    kmeans_fused_tc_kernel<<<(unsigned)grid, kThreads, pl.total, s>>>(tmX, tmC, p);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        snprintf(g_tc_err, sizeof(g_tc_err), "fused kernel launch failed: %s", cudaGetErrorString(e));
        return 1;
    }
    if (p.trace) {  // development aid: synchronous dump of CTA 0's per-tile event clocks
        static unsigned long long host[256 * 8];
        cudaStreamSynchronize(s);
        cudaMemcpy(host, p.trace, sizeof(host), cudaMemcpyDeviceToHost);
        cudaFree(p.trace);
        if (FILE* f = fopen(trace_path, "w")) {
            for (int i = 0; i < 256; ++i) {
                for (int j = 0; j < 8; ++j) fprintf(f, "%llu ", host[i * 8 + j]);
                fprintf(f, "\n");
            }
            fclose(f);
        }
    }
    return 0;
}

}  // namespace b200km
