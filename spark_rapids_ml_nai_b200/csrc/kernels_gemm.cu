// tcgen05 tile-GEMM path for k*d beyond shared memory: nearest centroid for any k and any d % 4 == 0.
//
// Two CTAs of a cluster (one SM pair) own 256 rows of X and walk all centroids in N-tiles of up to 256.  The
// scores  x^.c_j - 0.5 ||c_j||^2  are one `tcgen05.mma.cta_group::2.kind::tf32` GEMM  [256 x d] . [d x N]  with
//   A = the raw FP32 rows of X exactly as TMA wrote them (K-major, SWIZZLE_128B; the tensor core reads the top
//       19 bits of each word, i.e. x truncated to TF32 = x^); each CTA holds its 128 rows,
//   B = the TF32 split [c_hi ; c_lo] of the centroids (prepared per pass), streamed through a TMA ring in
//       sub-chunks of 16 floats (SWIZZLE_64B), each CTA loading its half of the N rows,
//   the norm folded in as one more K-step: A = a constant tile holding -0.5, B = the three-term TF32 split
//       of ||c_j||^2, so the accumulator is the score itself and the epilogue only compares,
//   D = FP32 accumulators in TMEM, two buffers of 256 columns: the epilogue of N-tile i overlaps the MMAs of i+1.
// A stays resident in shared memory across the N-tiles when d <= 256; otherwise it streams with B.
//
// Epilogue (8 warps per CTA, two per TMEM lane quarter, each half of the columns): tcgen05.ld 32 scores at a
// time, the centroid's index inside the 32 goes into the 5 low mantissa bits, a tournament yields the three
// largest, merged into the row's running triple.  After the last N-tile the row's best / second / third give
//   D_j = ||x||^2 - 2 score_j,
// and the same rigorous screen as kernels_screen.cu decides whether the TF32 view of x decided the row: the
// error of a margin D_a - D_b is 2 (x - x^).(c_a - c_b) + FP32 rounding <= 2 ||x - x^|| (r_a + r_b) + eps with
// r_j = ||x - c_j|| (triangle inequality: no k x k table needed).  Rows that fail are appended to a list and
// decided exactly (FP32 differences, float64 when those tie) by gemm_fixup_kernel; labels are exact for every
// row.  Sums are accumulated by a separate pass (this regime is tensor-bound: a second read of X is cheap).
//
// TMA loads here are plain (no multicast) and read only this GPU's memory.
//
// Replaces cuML's fusedL2NN behind reference python/src/spark_rapids_ml/clustering.py:367-370 (fit) and
// :493-497 (predict) for the shapes of BASELINE.json C3 / C5 and the reference's own k=1000 x d=3000 benchmark
// (python/benchmark/databricks/run_benchmark.sh:46-55).
#include "tc.cuh"
#include "tc_ptx.cuh"

#include <cfloat>
#include <climits>
#include <cstdio>
#include <cstdlib>

namespace b200km {

namespace {

using namespace tcx;

thread_local char g_gm_err[512] = "";

constexpr int kGmThreads = 384;  // warp 0 TMA, 1 MMA issue / relay, 2 TMEM, 3 spare, 4-11 epilogue
constexpr int kEpiWarp0 = 4;
constexpr int kRowsCta = 128;
constexpr int kRowsPair = 256;
constexpr int kASlotBytes = kRowsCta * 128;  // 128 rows x 32 floats, SWIZZLE_128B
constexpr int kMaxNA = 8, kMaxNB = 8;
constexpr int kConstABytes = kRowsCta * 32;  // the -0.5 tile: 128 rows x 8 floats, SWIZZLE_32B
constexpr int kScrBytes = 5 * kRowsCta * 4;  // hand-over of the column halves: (b1, cb1, b2, cb2, b3) per row
constexpr int kBarBytes = 512;
constexpr int kMaxSmem = 227 * 1024;
constexpr int kCntOpen = 0, kCntScan = 1, kCntCand = 2, kCntOverflow = 3;
constexpr float kCnPad = 1e30f;  // "norm" of the padded centroid columns: their score is -5e29

struct GmParams {
    int64_t n;
    int64_t num_tiles;  // 256-row tiles
    int d, k, k_pad;
    int n_sub;     // K sub-chunks of 16 floats
    int ntile;     // N of one MMA (multiple of 64, <= 256)
    int n_ntiles;  // k_pad / ntile
    int NA, NB, resident, rotate;
    int b_slot_bytes;
    const float* bt;   // pre-tiled B operand (gemm_prep_kernel)
    const float* cnt;  // pre-tiled norm operand
    const float* xnorm;
    const float* xlo;
    int32_t* labels;
    uint4* open;  // (row, runner-up | mode << 24, score threshold, -) of every row the TF32 scores did not decide
    unsigned int* counters;  // [kCntOpen, kCntScan, kCntCand, kCntOverflow]
    // second pass over the gathered multi-candidate rows (EMIT instantiation): n comes from counters[kCntScan]
    const float* thr;      // per gathered row: every centroid scoring above it is a candidate
    uint2* cand;           // (gathered row, centroid)
    unsigned int cand_cap, scan_cap;
    IterState* st;
    int honor_done;
    float tol_x, eps_rel, abs_rel;
    int off_a, off_b, off_ac, off_scr, off_bar;
    unsigned long long* trace;  // development aid (B200KM_ENABLE_TRACE builds): [16 roles][1024] clock64 of cluster 0
};

// ---- cluster / cta_group::2 PTX ------------------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// address of the same shared-memory offset in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t map_to_cta(uint32_t addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// arrives on the barrier at this offset in both CTAs of the pair once all MMAs issued so far have completed
__device__ __forceinline__ void tc_commit_pair(uint32_t bar) {
    asm volatile(
        "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
        "h"((unsigned short)3)
        : "memory");
}
// D[tmem, both CTAs] (+)= A[smem desc] * B[smem desc], kind::tf32, M = 256 across the CTA pair
__device__ __forceinline__ void mma_tf32_ss2(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                             uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void gtrace(const GmParams& p, int role, uint32_t idx) {
#ifdef B200KM_ENABLE_TRACE
    if (p.trace && blockIdx.x < 2 && idx < 1024u) p.trace[(role + (blockIdx.x ? 8 : 0)) * 1024 + idx] = (unsigned long long)clock64();
#endif
}
// Position in a ring of `n` slots, advanced without divisions: the single-thread roles of this kernel run at the
// latency of their dependent instructions, and one 32-bit division costs more than an MMA.
struct RingPos {
    uint32_t slot = 0, phase = 0;
    __device__ __forceinline__ void next(uint32_t n) {
        if (++slot == n) {
            slot = 0;
            phase ^= 1u;
        }
    }
};
// Inline wait for the single-thread roles (no call, no labels): try_wait blocks for the hardware's own time limit.
__device__ __forceinline__ void mbar_wait_inl(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile(
            "{\n\t"
            ".reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t"
            "}"
            : "=r"(ok)
            : "r"(bar), "r"(parity)
            : "memory");
    } while (!ok);
}
// 2-D tensor load into this CTA's shared memory whose completion bytes are counted on a barrier of the PAIR's
// leader CTA (`bar` is a shared::cluster address): the MMA issuer waits on one barrier for both halves of an
// operand.  Plain (non-multicast) load of this GPU's own memory.
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
        "l"(tm), "r"(bar), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}

// The three largest of 32 keys by a tournament of sorted lists (pairs -> triples -> ...); the third largest of
// two descending triples is max(a3, b3, min(a1,b2), min(a2,b1)).
__device__ __forceinline__ void top3max_of_32(const float (&key)[32], float& o1, float& o2, float& o3) {
    float m1[16], m2[16], m3[8];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        m1[i] = fmaxf(key[2 * i], key[2 * i + 1]);
        m2[i] = fminf(key[2 * i], key[2 * i + 1]);
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const float a1 = m1[2 * i], a2 = m2[2 * i], b1 = m1[2 * i + 1], b2 = m2[2 * i + 1];
        m1[i] = fmaxf(a1, b1);
        m2[i] = fmaxf(fminf(a1, b1), fmaxf(a2, b2));
        m3[i] = fmaxf(fminf(a1, b2), fminf(a2, b1));
    }
#pragma unroll
    for (int nn = 8; nn > 1; nn >>= 1) {
#pragma unroll
        for (int i = 0; i < nn / 2; ++i) {
            const float a1 = m1[2 * i], a2 = m2[2 * i], a3 = m3[2 * i];
            const float b1 = m1[2 * i + 1], b2 = m2[2 * i + 1], b3 = m3[2 * i + 1];
            m1[i] = fmaxf(a1, b1);
            m2[i] = fmaxf(fminf(a1, b1), fmaxf(a2, b2));
            m3[i] = fmaxf(fmaxf(a3, b3), fmaxf(fminf(a1, b2), fminf(a2, b1)));
        }
    }
    o1 = m1[0];
    o2 = m2[0];
    o3 = m3[0];
}

// Merges the descending triple (o1, o2, o3) (origins oc1, oc2) into the running one.
__device__ __forceinline__ void merge3(float& b1, int& cb1, float& b2, int& cb2, float& b3, float o1, int oc1, float o2,
                                       int oc2, float o3) {
    const bool w1 = o1 > b1;
    const float n1 = w1 ? o1 : b1;
    const int n1c = w1 ? oc1 : cb1;
    const float l1 = w1 ? b1 : o1;  // the loser of the heads
    const int l1c = w1 ? cb1 : oc1;
    const bool w2 = o2 > b2;
    const float m2 = w2 ? o2 : b2;
    const int m2c = w2 ? oc2 : cb2;
    const bool w3 = m2 > l1;
    const float n3 = fmaxf(fmaxf(b3, o3), fmaxf(fminf(b1, o2), fminf(b2, o1)));
    b1 = n1;
    cb1 = n1c;
    b2 = w3 ? m2 : l1;
    cb2 = w3 ? m2c : l1c;
    b3 = n3;
}

// ------------------------------------------------------------------------------------------------
// Operand preparation of a pass: the B operand in exactly the bytes the tensor core reads from shared memory, so
// that a ring slot is one contiguous bulk copy (a 2-D TMA box of 64-byte rows costs the TMA unit one request per
// row: measured ~1700 cycles per 16 KB slot, three times the slot's MMA time).
//   bt: blocks [(nt*2 + rank) * n_sub + sc] of N*16 floats = [hi: N/2 rows x 16 floats | lo: same], rows =
//       centroids nt*N + rank*N/2 + row, columns sc*16 .., SWIZZLE_64B (16-byte chunk ^= (row >> 1) & 3);
//       hi = centroid rounded to TF32, lo = the TF32-rounded remainder, zero padded.
//   cnt: blocks [nt*2 + rank] of N/2 rows x 8 floats, SWIZZLE_32B (chunk ^= (row >> 2) & 1): columns 0-2 = the
//       three-term TF32 split of ||c_j||^2 (kCnPad for padded rows), the rest zero.
__global__ void __launch_bounds__(256) gemm_prep_kernel(const float* __restrict__ C, const float* __restrict__ cnorm,
                                                        int k, int d, int k_pad, int n_sub, int halfN,
                                                        float* __restrict__ bt, float* __restrict__ cnt) {
    const size_t total = (size_t)k_pad * n_sub * 16;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    const int part = halfN * 16;  // floats of one block's hi part
    for (size_t o = (size_t)blockIdx.x * blockDim.x + threadIdx.x; o < total; o += stride) {
        const size_t b = o / part;
        const int r = (int)(o - b * part);
        const int row = r >> 4, pc = (r >> 2) & 3, w = r & 3;
        const int col = ((pc ^ ((row >> 1) & 3)) << 2) | w;
        const int sc = (int)(b % n_sub);
        const size_t j = (b / n_sub) * halfN + row;
        const int i = sc * 16 + col;
        float hi = 0.f, lo = 0.f;
        if (j < (size_t)k && i < d) {
            const float c = C[j * d + i];
            hi = __uint_as_float(tf32_round(__float_as_uint(c)));
            lo = __uint_as_float(tf32_round(__float_as_uint(c - hi)));
        }
        bt[b * 2 * part + r] = hi;
        bt[b * 2 * part + part + r] = lo;
    }
    for (size_t j = (size_t)blockIdx.x * blockDim.x + threadIdx.x; j < (size_t)k_pad; j += stride) {
        const float v = j < (size_t)k ? cnorm[j] : kCnPad;
        const float a = __uint_as_float(tf32_round(__float_as_uint(v)));
        const float b = __uint_as_float(tf32_round(__float_as_uint(v - a)));
        const float c = __uint_as_float(tf32_round(__float_as_uint((v - a) - b)));
        const int row = (int)(j % halfN);
        const int swz = (row >> 2) & 1;
        float4* o = reinterpret_cast<float4*>(cnt + j * 8);
        o[swz] = make_float4(a, b, c, 0.f);
        o[swz ^ 1] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
}

// ||x||^2 and ||x - trunc_tf32(x)||^2 per row: one warp per row (once per fit / predict call).
__global__ void __launch_bounds__(256) rownorm2_kernel(const float* __restrict__ X, int64_t n, int d,
                                                       float* __restrict__ xnorm, float* __restrict__ xlo) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
    const int nch = d >> 2;
    for (int64_t r = warp; r < n; r += nwarps) {
        const float4* row = reinterpret_cast<const float4*>(X + r * d);
        float s = 0.f, l = 0.f;
        for (int c = lane; c < nch; c += 32) {
            const float4 v = __ldg(row + c);
            const float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                s = fmaf(e[t], e[t], s);
                const float rr = e[t] - __uint_as_float(tf32_trunc(__float_as_uint(e[t])));  // exact
                l = fmaf(rr, rr, l);
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            s += __shfl_xor_sync(0xffffffffu, s, o);
            l += __shfl_xor_sync(0xffffffffu, l, o);
        }
        if (lane == 0) {
            xnorm[r] = s;
            xlo[r] = l;
        }
    }
}

// ---- exact decisions -------------------------------------------------------------------------------
__device__ __forceinline__ float sqdist_f32(const float4* __restrict__ x4, const float4* __restrict__ c4, int nchunks,
                                            int lane) {
    float s = 0.f;
    for (int ch = lane; ch < nchunks; ch += 32) {
        const float4 xv = __ldg(x4 + ch), cv = __ldg(c4 + ch);
        const float d0 = xv.x - cv.x, d1 = xv.y - cv.y, d2 = xv.z - cv.z, d3 = xv.w - cv.w;
        s += fmaf(d0, d0, d1 * d1) + fmaf(d2, d2, d3 * d3);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    return s;
}
__device__ __forceinline__ double sqdist_f64(const float* __restrict__ x, const float* __restrict__ c, int d, int lane) {
    double s = 0.0;
    for (int i = lane; i < d; i += 32) {
        const double df = (double)__ldg(x + i) - (double)__ldg(c + i);
        s = fma(df, df, s);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    return s;
}

// One warp per listed row.  mode 0: exactly the computed best (labels[row]) against the runner-up, FP32 squared
// differences (no cancellation), float64 when those tie; ties keep the lowest index.  mode 1 (more candidates):
// the row is copied into the gather buffer for the second GEMM pass, which lists its candidates.
__global__ void __launch_bounds__(256) gemm_fixup_kernel(const float* __restrict__ X, int d, int k,
                                                         const float* __restrict__ C, const uint4* __restrict__ open,
                                                         unsigned int* __restrict__ counters, int32_t* __restrict__ labels,
                                                         float* __restrict__ xs, unsigned int* __restrict__ scan_row,
                                                         float* __restrict__ scan_thr, unsigned long long* __restrict__ scan_key,
                                                         unsigned int scan_cap, IterState* st, int honor_done) {
    if (honor_done && st->done) return;
    const unsigned int count = counters[kCntOpen];
    const int lane = threadIdx.x & 31;
    const unsigned int warp0 = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const unsigned int nwarps = gridDim.x * (blockDim.x >> 5);
    const int nchunks = d >> 2;
    for (unsigned int w = warp0; w < count; w += nwarps) {
        const uint4 e = open[w];
        const int64_t row = e.x;
        const int i2 = (int)(e.y & 0xFFFFFFu);
        const int mode = (int)(e.y >> 24);
        const float* x = X + (size_t)row * d;
        const float4* x4 = reinterpret_cast<const float4*>(x);
        if (mode == 0) {
            const int i1 = labels[row];
            const int ja = min(i1, i2), jb = max(i1, i2);
            const float sa = sqdist_f32(x4, reinterpret_cast<const float4*>(C + (size_t)ja * d), nchunks, lane);
            const float sb = sqdist_f32(x4, reinterpret_cast<const float4*>(C + (size_t)jb * d), nchunks, lane);
            int bi = sb < sa ? jb : ja;
            if (fabsf(sa - sb) <= 4e-6f * (sa + sb)) {
                const double da = sqdist_f64(x, C + (size_t)ja * d, d, lane);
                const double db = sqdist_f64(x, C + (size_t)jb * d, d, lane);
                bi = db < da ? jb : ja;
            }
            if (bi != i1 && lane == 0) labels[row] = bi;
        } else {
            unsigned int pos = 0;
            if (lane == 0) pos = atomicAdd(counters + kCntScan, 1u);
            pos = __shfl_sync(0xffffffffu, pos, 0);
            if (pos >= scan_cap) {
                if (lane == 0) counters[kCntOverflow] = 1u;
                continue;
            }
            float4* dst = reinterpret_cast<float4*>(xs + (size_t)pos * d);
            for (int ch = lane; ch < nchunks; ch += 32) dst[ch] = __ldg(x4 + ch);
            if (lane == 0) {
                scan_row[pos] = (unsigned int)row;
                scan_thr[pos] = __uint_as_float(e.z);
                scan_key[pos] = ~0ull;
            }
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&st->screen_refined, count);
}

// One warp per candidate (gathered row, centroid): FP32 squared differences; the row keeps the smallest
// (distance, index) through a 64-bit atomic minimum (equal distances: the lowest index).
__global__ void __launch_bounds__(256) gemm_cand_kernel(const float* __restrict__ xs, int d, const float* __restrict__ C,
                                                        const uint2* __restrict__ cand,
                                                        const unsigned int* __restrict__ counters, unsigned int cand_cap,
                                                        unsigned long long* __restrict__ scan_key, IterState* st,
                                                        int honor_done) {
    if (honor_done && st->done) return;
    if (counters[kCntOverflow]) return;
    unsigned int count = counters[kCntCand];
    if (count > cand_cap) count = cand_cap;
    const int lane = threadIdx.x & 31;
    const unsigned int warp0 = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const unsigned int nwarps = gridDim.x * (blockDim.x >> 5);
    const int nchunks = d >> 2;
    for (unsigned int w = warp0; w < count; w += nwarps) {
        const uint2 e = cand[w];
        const float v = sqdist_f32(reinterpret_cast<const float4*>(xs + (size_t)e.x * d),
                                   reinterpret_cast<const float4*>(C + (size_t)e.y * d), nchunks, lane);
        if (lane == 0) atomicMin(scan_key + e.x, ((unsigned long long)__float_as_uint(v) << 32) | (unsigned long long)e.y);
    }
}

// Labels of the gathered rows from their (distance, index) minima - or, when the gather / candidate buffers
// overflowed (degenerate data: very many equidistant centroids), every centroid for every such row.
__global__ void __launch_bounds__(256) gemm_scan_finish_kernel(const float* __restrict__ X, int d, int k,
                                                               const float* __restrict__ C, const uint4* __restrict__ open,
                                                               const unsigned int* __restrict__ counters,
                                                               const unsigned int* __restrict__ scan_row,
                                                               const unsigned long long* __restrict__ scan_key,
                                                               int32_t* __restrict__ labels, IterState* st, int honor_done) {
    if (honor_done && st->done) return;
    if (!counters[kCntOverflow]) {
        const unsigned int m = counters[kCntScan];
        for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < m; i += gridDim.x * blockDim.x)
            labels[scan_row[i]] = (int32_t)(scan_key[i] & 0xFFFFFFFFull);
        return;
    }
    const unsigned int count = counters[kCntOpen];
    const int lane = threadIdx.x & 31;
    const unsigned int warp0 = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const unsigned int nwarps = gridDim.x * (blockDim.x >> 5);
    const int nchunks = d >> 2;
    for (unsigned int w = warp0; w < count; w += nwarps) {
        const uint4 e = open[w];
        if ((e.y >> 24) == 0) continue;
        const int64_t row = e.x;
        const float4* x4 = reinterpret_cast<const float4*>(X + (size_t)row * d);
        float f1 = FLT_MAX;
        int bi = 0;
        for (int j = 0; j < k; ++j) {
            const float sum = sqdist_f32(x4, reinterpret_cast<const float4*>(C + (size_t)j * d), nchunks, lane);
            if (sum < f1) {  // ascending index order: ties keep the lowest
                f1 = sum;
                bi = j;
            }
        }
        if (lane == 0) labels[row] = bi;
    }
}

// best[i] = ||x_i - c_label(i)||^2 from FP32 differences: one warp per row.
__global__ void __launch_bounds__(256) rowdist_kernel(const float* __restrict__ X, int64_t n, int d,
                                                      const float* __restrict__ C, const int32_t* __restrict__ labels,
                                                      float* __restrict__ best) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
    const int nch = d >> 2;
    for (int64_t r = warp; r < n; r += nwarps) {
        const float v = sqdist_f32(reinterpret_cast<const float4*>(X + r * d),
                                   reinterpret_cast<const float4*>(C + (size_t)labels[r] * d), nch, lane);
        if (lane == 0) best[r] = v;
    }
}

// ------------------------------------------------------------------------------------------------
// EMIT = false: the assignment pass (labels + open-row list).  EMIT = true: the same GEMM over the gathered rows
// that have more than two candidates; instead of ranking, the epilogue lists every centroid whose score reaches
// the row's threshold (the exact distances of those few are then compared by gemm_cand_kernel).
template <bool EMIT>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kGmThreads, 1)
    kmeans_gemm_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmB,
                       const __grid_constant__ CUtensorMap tmCn, GmParams p) {
    if (p.honor_done && p.st->done) return;  // both CTAs of a pair read the same flag
    if (EMIT) {
        if (p.counters[kCntOverflow]) return;  // too many such rows for the gather buffer: the brute-force kernel runs
        unsigned int m = p.counters[kCntScan];
        p.n = m;
        p.num_tiles = (m + kRowsPair - 1) / kRowsPair;
        if (p.num_tiles == 0) return;
    }

    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const uint32_t smem_base = smem_u32(smem);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const int64_t cluster_id = blockIdx.x >> 1;
    const int64_t n_clusters = gridDim.x >> 1;
    const int NA = p.NA, NB = p.NB;
    const int N = p.ntile;
    const int halfN = N >> 1;

    // barrier table (8 bytes each)
    const uint32_t bar0 = smem_base + p.off_bar;
    const uint32_t a_full0 = bar0, a_empty0 = bar0 + 64, b_full0 = bar0 + 128, b_empty0 = bar0 + 192,
                   d_full0 = bar0 + 256, d_empty0 = bar0 + 272;
    uint32_t* tmem_slot = (uint32_t*)(smem + p.off_bar + 288);

    if (threadIdx.x == 0) {
        // full barriers: used in the leader only; both CTAs' loads count their bytes there
        const uint32_t full_count = 1u;
        for (int s = 0; s < kMaxNA; ++s) {
            mbar_init(a_full0 + 8u * s, full_count);
            mbar_init(a_empty0 + 8u * s, 1);
        }
        for (int s = 0; s < kMaxNB; ++s) {
            mbar_init(b_full0 + 8u * s, full_count);
            mbar_init(b_empty0 + 8u * s, 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(d_full0 + 8u * s, 1);
            mbar_init(d_empty0 + 8u * s, 16);  // 8 epilogue warps of each CTA (used in the leader only)
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::);
    }
    {
        // the constant A tile of the norm K-step: columns 0-2 hold -0.5 (SWIZZLE_32B: chunk ^= (row >> 2) & 1)
        float* ac = (float*)(smem + p.off_ac);
#pragma unroll 1
        for (int i = threadIdx.x; i < kRowsCta * 8; i += kGmThreads) {
            const int row = i >> 3, col = i & 7;
            const int phys = (col >> 2) ^ ((row >> 2) & 1);
            ac[row * 8 + phys * 4 + (col & 3)] = col < 3 ? -0.5f : 0.f;
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // the tensor core reads it through the async proxy
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    const int n_sub = p.n_sub;
    const int n_nt = p.n_ntiles;
    const bool resident = p.resident != 0;
    // Every cluster walks the N-tiles and the K sub-chunks of B in its own rotation: with all clusters in step,
    // the whole chip would ask the same few L2 slices for the same B lines at the same time.
    // (Only while B is small: a B operand near the L2 capacity has to be walked in step, so that one cluster's
    // miss is the others' hit.)
    const int nt0 = p.rotate ? (int)(cluster_id % n_nt) : 0;
    const int sc0 = p.rotate ? 2 * (int)((cluster_id / n_nt) % ((n_sub + 1) / 2)) : 0;  // whole A chunks: even

    if (warp == 0) {
        // ===================================== TMA producer =====================================
        // Each CTA loads its half of every operand into its own shared memory; the bytes of both halves are
        // counted on the leader's barrier (the leader arms it with the total), so nobody relays.
        if (lane == 0) {
            RingPos ar, br;
            const uint32_t lead_a_full0 = map_to_cta(a_full0, 0), lead_b_full0 = map_to_cta(b_full0, 0);
            const int b_rows = N >> 4;       // a B block = N/16 rows of 256 floats in the pre-tiled buffer
            const int cn_rows = halfN >> 5;  // a norm block = N/64 rows of 256 floats
            for (int64_t tile = cluster_id; tile < p.num_tiles; tile += n_clusters) {
                const int row0 = (int)(tile * kRowsPair + rank * kRowsCta);
                int nt = nt0;
                for (int nti = 0; nti < n_nt; ++nti) {
                    const int blk0 = (nt * 2 + (int)rank) * n_sub;
                    int sc = sc0;
                    for (int sci = 0; sci < n_sub; ++sci) {
                        if ((sc & 1) == 0 && (!resident || nti == 0)) {  // first touch of this A chunk
                            mbar_wait_inl(a_empty0 + 8u * ar.slot, ar.phase ^ 1u);
                            if (rank == 0) mbar_arrive_expect_tx(a_full0 + 8u * ar.slot, 2u * (uint32_t)kASlotBytes);
                            tma_load_2d_pair(smem_base + p.off_a + ar.slot * kASlotBytes, &tmX, lead_a_full0 + 8u * ar.slot,
                                             (sc >> 1) * 32, row0);
                            ar.next((uint32_t)NA);
                        }
                        mbar_wait_inl(b_empty0 + 8u * br.slot, br.phase ^ 1u);
                        gtrace(p, 0, (uint32_t)sci);
                        if (rank == 0) mbar_arrive_expect_tx(b_full0 + 8u * br.slot, 2u * (uint32_t)p.b_slot_bytes);
                        tma_load_2d_pair(smem_base + p.off_b + br.slot * p.b_slot_bytes, &tmB, lead_b_full0 + 8u * br.slot, 0,
                                         (blk0 + sc) * b_rows);
                        br.next((uint32_t)NB);
                        if (++sc == n_sub) sc = 0;
                    }
                    {  // the norm K-step of this N-tile
                        mbar_wait_inl(b_empty0 + 8u * br.slot, br.phase ^ 1u);
                        if (rank == 0) mbar_arrive_expect_tx(b_full0 + 8u * br.slot, 2u * (uint32_t)(halfN * 32));
                        tma_load_2d_pair(smem_base + p.off_b + br.slot * p.b_slot_bytes, &tmCn, lead_b_full0 + 8u * br.slot, 0,
                                         (nt * 2 + (int)rank) * cn_rows);
                        br.next((uint32_t)NB);
                    }
                    if (++nt == n_nt) nt = 0;
                }
            }
        }
        __syncwarp();
    } else if (warp == 1 && rank == 0) {
        // ====================================== MMA issuer (leader CTA) ======================================
        // The whole warp runs the loop, one elected lane issues.
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(kRowsPair >> 4) << 24) |
                               ((uint32_t)(N >> 3) << 17);
        const uint64_t adesc0 = make_b_desc(smem_base + p.off_a, 128);
        const uint64_t bdesc0 = make_b_desc(smem_base + p.off_b, 64);
        const uint64_t cdesc0 = make_b_desc(smem_base + p.off_b, 32);
        const uint64_t acdesc = make_b_desc(smem_base + p.off_ac, 32);
        const uint64_t lo_off = (uint64_t)((halfN * 64) >> 4);
        const uint32_t b_slot16 = (uint32_t)p.b_slot_bytes >> 4;
        RingPos ar, br;
        uint32_t gnt = 0;
        for (int64_t tile = cluster_id; tile < p.num_tiles; tile += n_clusters) {
            for (int nti = 0; nti < n_nt; ++nti, ++gnt) {
                const uint32_t buf = gnt & 1u;
                const uint32_t d_tmem = tmem_base + buf * 256u;
                mbar_wait_inl(d_empty0 + 8u * buf, ((gnt >> 1) & 1u) ^ 1u);
                uint32_t aslot = 0, a_idx = 0;
                int sc = sc0;
                for (int sci = 0; sci < n_sub; ++sci) {
                    if ((sc & 1) == 0) {
                        if (!resident || nti == 0) {
                            aslot = ar.slot;
                            mbar_wait_inl(a_full0 + 8u * aslot, ar.phase);
                            ar.next((uint32_t)NA);
                        } else {
                            aslot = a_idx;  // resident: the chunks sit in the slots in this cluster's load order
                        }
                        ++a_idx;
                    }
                    const uint32_t bslot = br.slot;
                    mbar_wait_inl(b_full0 + 8u * bslot, br.phase);
                    br.next((uint32_t)NB);
                    tc_fence_after();
                    if (lane == 0) gtrace(p, 1, (uint32_t)sci);
                    if (elect_one()) {
                        // A: 64-byte half (sc & 1) of the 128-byte row, K-steps of 32 bytes
                        const uint64_t ad = adesc0 + (uint64_t)((aslot << 10) + (uint32_t)((sc & 1) * 4));
                        const uint64_t bd = bdesc0 + (uint64_t)(bslot * b_slot16);
#pragma unroll
                        for (int kk = 0; kk < 2; ++kk) {
                            mma_tf32_ss2(d_tmem, ad + 2 * kk, bd + 2 * kk, idesc, (sci | kk) != 0);
                            mma_tf32_ss2(d_tmem, ad + 2 * kk, bd + lo_off + 2 * kk, idesc, 1);
                        }
                        tc_commit_pair(b_empty0 + 8u * bslot);
                        const bool a_done = ((sc & 1) == 1 || sc == n_sub - 1) && (!resident || nti == n_nt - 1);
                        if (a_done) tc_commit_pair(a_empty0 + 8u * aslot);
                    }
                    __syncwarp();
                    if (++sc == n_sub) sc = 0;
                }
                {  // the norm K-step closes the N-tile
                    const uint32_t bslot = br.slot;
                    mbar_wait_inl(b_full0 + 8u * bslot, br.phase);
                    br.next((uint32_t)NB);
                    tc_fence_after();
                    if (elect_one()) {
                        mma_tf32_ss2(d_tmem, acdesc, cdesc0 + (uint64_t)(bslot * b_slot16), idesc, 1);
                        tc_commit_pair(b_empty0 + 8u * bslot);
                        tc_commit_pair(d_full0 + 8u * buf);
                    }
                    __syncwarp();
                }
            }
        }
    } else if (warp >= kEpiWarp0) {
        // ======================================= epilogue =======================================
        const int q = warp & 3;  // TMEM lane quarter of this warp
        const int half = (warp - kEpiWarp0) >> 2;
        const int row = q * 32 + lane;
        float* scr = (float*)(smem + p.off_scr);
        const float cmax = EMIT ? 0.f : p.st->cnorm_max;
        const uint32_t my_d_empty0 = map_to_cta(d_empty0, 0);
        uint32_t gnt = 0;
        for (int64_t tile = cluster_id; tile < p.num_tiles; tile += n_clusters) {
            const int64_t grow = tile * kRowsPair + rank * kRowsCta + row;
            const bool valid = grow < p.n;
            float xn = 0.f, xl = 0.f;
            if (!EMIT && valid && half == 0) {  // consumed after the last N-tile
                xn = p.xnorm[grow];
                xl = p.xlo[grow];
            }
            const float thr = (EMIT && valid) ? p.thr[grow] : FLT_MAX;
            float b1 = -FLT_MAX, b2 = -FLT_MAX, b3 = -FLT_MAX;
            int cb1 = 0, cb2 = 0;
            int nt = nt0;
            for (int nti = 0; nti < n_nt; ++nti, ++gnt) {
                const uint32_t buf = gnt & 1u;
                mbar_wait(d_full0 + 8u * buf, (gnt >> 1) & 1u);
                tc_fence_after();
                if (warp == kEpiWarp0 && lane == 0) gtrace(p, 4, gnt * 2);

                const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + buf * 256u + (uint32_t)(half * halfN);
#pragma unroll 1
                for (int c0 = 0; c0 < halfN; c0 += 32) {
                    uint32_t sc[32];
                    tmem_ld32(taddr + c0, sc);
                    tmem_wait_ld();
                    if (c0 + 32 >= halfN) {  // the scores of this N-tile are in registers: the buffer is free
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive_cluster(my_d_empty0 + 8u * buf);
                    }
                    const int col0 = nt * N + half * halfN + c0;
                    if constexpr (EMIT) {
                        unsigned hit = 0;
#pragma unroll
                        for (int j = 0; j < 32; ++j) hit |= (__uint_as_float(sc[j]) >= thr ? 1u : 0u) << j;
                        while (hit) {
                            const int j = __ffs(hit) - 1;
                            hit &= hit - 1;
                            const unsigned int pos = atomicAdd(p.counters + kCntCand, 1u);
                            if (pos < p.cand_cap) p.cand[pos] = make_uint2((unsigned int)grow, (unsigned int)(col0 + j));
                            else p.counters[kCntOverflow] = 1u;
                        }
                    } else {
                        float key[32];
#pragma unroll
                        for (int j = 0; j < 32; ++j) key[j] = __uint_as_float((sc[j] & ~31u) | (uint32_t)j);
                        float t1, t2, t3;
                        top3max_of_32(key, t1, t2, t3);
                        merge3(b1, cb1, b2, cb2, b3, t1, col0, t2, col0, t3);
                    }
                }
                if (warp == kEpiWarp0 && lane == 0) gtrace(p, 4, gnt * 2 + 1);
                if (++nt == n_nt) nt = 0;
            }
            if (EMIT) continue;
            // ---- the two column halves of a row meet; half 0 decides ----
            if (half == 1) {
                scr[0 * kRowsCta + row] = b1;
                scr[1 * kRowsCta + row] = __int_as_float(cb1);
                scr[2 * kRowsCta + row] = b2;
                scr[3 * kRowsCta + row] = __int_as_float(cb2);
                scr[4 * kRowsCta + row] = b3;
            }
            named_bar_sync(1, 256);
            if (half == 0) {
                merge3(b1, cb1, b2, cb2, b3, scr[0 * kRowsCta + row], __float_as_int(scr[1 * kRowsCta + row]),
                       scr[2 * kRowsCta + row], __float_as_int(scr[3 * kRowsCta + row]), scr[4 * kRowsCta + row]);
                const int i1 = cb1 + (int)(__float_as_uint(b1) & 31u);
                const int i2 = cb2 + (int)(__float_as_uint(b2) & 31u);
                // D_j = ||x||^2 - 2 score_j; every quantity of the test is an upper bound (file header)
                const float D1 = fmaf(-2.f, b1, xn), D2 = fmaf(-2.f, b2, xn), D3 = fmaf(-2.f, b3, xn);
                const float scale = xn + cmax;
                const float t = p.tol_x * sqrtf(xl);
                const float eps = p.eps_rel * scale;
                const float pad = p.abs_rel * scale;
                const float r1 = sqrtf(fmaxf(D1, 0.f) + pad);
                const float r2 = sqrtf(fmaxf(D2, 0.f) + pad);
                const float r3 = sqrtf(fmaxf(D3, 0.f) + pad);
                const bool open = valid && p.k > 1 && (D2 - D1) < 2.f * t * (r1 + r2) + eps;
                const bool third = open && p.k > 2 && (D3 - D1) < 2.f * t * (r1 + r3) + eps;
                // More than two candidates: every centroid j that can still be the nearest has D_j - D_1 below
                // 2 t (r_1 + r_j) + eps; with B = 4 t r_1 + eps that implies D_j < D_1 + 2 B (t << r_1), hence
                // r_j <= sqrt(D_1 + 2 B + pad) and, in score units, score_j > score_1 - band / 2.
                const float B = 4.f * t * r1 + eps;
                const float band = 2.f * t * (r1 + sqrtf(fmaxf(D1, 0.f) + 2.f * B + pad)) + 2.f * eps;
                const float thr_row = b1 - 0.5f * band;
                if (valid) p.labels[grow] = i1;
                const unsigned m = __ballot_sync(0xffffffffu, open);
                if (m) {
                    const int leader = __ffs(m) - 1;
                    unsigned int base = 0;
                    if (lane == leader) base = atomicAdd(p.counters + kCntOpen, (unsigned int)__popc(m));
                    base = __shfl_sync(0xffffffffu, base, leader);
                    if (open)
                        p.open[base + __popc(m & ((1u << lane) - 1u))] =
                            make_uint4((unsigned int)grow, (unsigned int)i2 | (third ? (1u << 24) : 0u),
                                       __float_as_uint(thr_row), 0u);
                }
            }
            named_bar_sync(2, 256);
        }
    }

    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    if (warp == 2) {
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
    }
}

// ---- host side ---------------------------------------------------------------------------------
struct Plan {
    int n_sub, n_chunks, d_pad, ntile, n_ntiles, k_pad, NA, NB, resident, b_slot_bytes;
    int off_a, off_b, off_ac, off_scr, off_bar, total;
    bool ok;
};

Plan make_plan(int d, int k) {
    Plan pl = {};
    pl.ok = false;
    if (d <= 0 || k <= 0 || (d & 3) || k > (1 << 24) - 256) return pl;
    pl.n_sub = (d + 15) / 16;
    pl.n_chunks = (pl.n_sub + 1) / 2;
    pl.d_pad = pl.n_sub * 16;
    const int tiles = (k + 255) / 256;
    const int per = (k + tiles - 1) / tiles;
    pl.ntile = ((per + 63) / 64) * 64;
    pl.n_ntiles = (k + pl.ntile - 1) / pl.ntile;
    pl.k_pad = pl.n_ntiles * pl.ntile;
    pl.b_slot_bytes = pl.ntile * 64;  // [hi: N/2 rows x 64 B | lo: N/2 rows x 64 B]
    const int fixed = kConstABytes + kScrBytes + kBarBytes + 1024 /*alignment*/;
    const int avail = kMaxSmem - fixed;
    if (pl.n_chunks <= kMaxNA && pl.n_chunks * kASlotBytes + 4 * pl.b_slot_bytes <= avail) {
        pl.resident = 1;
        pl.NA = pl.n_chunks;
    } else {
        pl.resident = 0;
        pl.NA = 4;
    }
    int nb = (avail - pl.NA * kASlotBytes) / pl.b_slot_bytes;
    if (nb > kMaxNB) nb = kMaxNB;
    if (nb < 2) return pl;
    pl.NB = nb;
    int off = 0;
    pl.off_a = off;
    off += pl.NA * kASlotBytes;
    pl.off_b = off;
    off += pl.NB * pl.b_slot_bytes;
    off = (off + 1023) & ~1023;
    pl.off_ac = off;
    off += kConstABytes;
    pl.off_scr = off;
    off += kScrBytes;
    pl.off_bar = off;
    off += kBarBytes;
    pl.total = off + 1024;
    pl.ok = pl.total <= kMaxSmem;
    return pl;
}

// Tensor map over a flat float buffer seen as [rows][256], box = box_rows full rows, no swizzle: the box lands in
// shared memory as box_rows * 1024 contiguous bytes.
int encode_flat(char* err, size_t err_len, CUtensorMap* tm, const void* base, uint64_t rows, uint32_t box_rows) {
    EncodeTiledFn enc = get_encode();
    if (!enc) {
        snprintf(err, err_len, "cuTensorMapEncodeTiled not available from the driver");
        return 1;
    }
    cuuint64_t dims[2] = {256, rows};
    cuuint64_t strides[1] = {1024};
    cuuint32_t box[2] = {256, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        snprintf(err, err_len, "cuTensorMapEncodeTiled (flat) failed with CUresult %d", (int)r);
        return 1;
    }
    return 0;
}

template <typename T>
int grow_buf(T** p, size_t* cap, size_t need, const char* what) {
    if (need <= *cap && *p) return 0;
    if (*p) cudaFree(*p);
    *p = nullptr;
    *cap = 0;
    const size_t want = need + need / 8 + 64;
    if (cudaMalloc((void**)p, sizeof(T) * want) != cudaSuccess) {
        snprintf(g_gm_err, sizeof(g_gm_err), "cudaMalloc of %zu bytes for %s failed", sizeof(T) * want, what);
        return 1;
    }
    *cap = want;
    return 0;
}

}  // namespace

bool gemm_supported(const float* X, int d, int k) {
    if ((reinterpret_cast<uintptr_t>(X) & 15) != 0) return false;
    return make_plan(d, k).ok;
}
size_t gemm_workspace_bytes(int64_t n, int d, int k) {
    const Plan pl = make_plan(d, k);
    if (!pl.ok) return 0;
    size_t scan = (size_t)n / 4 < 4096 ? 4096 : (size_t)n / 4;
    if (scan * (size_t)d * 4 > ((size_t)2 << 30)) scan = ((size_t)2 << 30) / ((size_t)d * 4);
    scan += scan / 8 + kRowsPair;
    return sizeof(float) * (2 * (size_t)pl.k_pad * pl.d_pad + 8 * (size_t)pl.k_pad) + 16 * ((size_t)n + (size_t)n / 8 + 64) + 64 +
           scan * ((size_t)d * 4 + 16 + 8 * 8);
}
const char* gemm_last_error() { return g_gm_err; }

void gemm_free(GemmWorkspace* ws) {
    if (ws->d_b) cudaFree(ws->d_b);
    if (ws->d_cn) cudaFree(ws->d_cn);
    if (ws->d_open) cudaFree(ws->d_open);
    if (ws->d_count) cudaFree(ws->d_count);
    for (void* q : {ws->d_xs, ws->d_scan_row, ws->d_scan_thr, ws->d_scan_key, ws->d_cand})
        if (q) cudaFree(q);
    *ws = GemmWorkspace();
}

void launch_rownorm2(const float* X, int64_t n, int d, float* xnorm, float* xlo, cudaStream_t s) {
    if (n <= 0) return;
    int64_t blocks = (n + 7) / 8;
    if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
    rownorm2_kernel<<<(unsigned)blocks, 256, 0, s>>>(X, n, d, xnorm, xlo);
    count_launch();
}

void launch_rowdist(const float* X, int64_t n, int d, const float* C, const int32_t* labels, float* best,
                    cudaStream_t s) {
    if (n <= 0) return;
    int64_t blocks = (n + 7) / 8;
    if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
    rowdist_kernel<<<(unsigned)blocks, 256, 0, s>>>(X, n, d, C, labels, best);
    count_launch();
}

int gemm_pass(GemmWorkspace* ws, const float* X, int64_t n, int d, int k, const float* C, const float* cnorm,
              const float* xnorm, const float* xlo, int32_t* labels, IterState* st, bool honor_done, cudaStream_t s) {
    const Plan pl = make_plan(d, k);
    if (!pl.ok) {
        snprintf(g_gm_err, sizeof(g_gm_err), "shape d=%d k=%d does not fit the tcgen05 tile-GEMM kernel (needs d %% 4 == 0)", d, k);
        return 1;
    }
    if (n <= 0) return 0;
    if (grow_buf((float**)&ws->d_b, &ws->b_cap, 2 * (size_t)pl.k_pad * pl.d_pad, "the TF32 centroids")) return 1;
    if (grow_buf((float**)&ws->d_cn, &ws->cn_cap, 8 * (size_t)pl.k_pad, "the centroid norms")) return 1;
    if (grow_buf((uint4**)&ws->d_open, &ws->open_cap, (size_t)n, "the open-row list")) return 1;
    if (!ws->d_count && cudaMalloc((void**)&ws->d_count, 64) != cudaSuccess) {
        snprintf(g_gm_err, sizeof(g_gm_err), "cudaMalloc of the open-row counter failed");
        return 1;
    }
    // second pass: gather buffer for the rows with more than two candidates (n / 4 rows, at most 2 GB) and the
    // candidate list; beyond that capacity the brute-force kernel takes over
    size_t scan_cap = (size_t)n / 4 < 4096 ? 4096 : (size_t)n / 4;
    if (scan_cap * (size_t)d * 4 > ((size_t)2 << 30)) scan_cap = ((size_t)2 << 30) / ((size_t)d * 4);
    scan_cap = (scan_cap + kRowsPair - 1) / kRowsPair * kRowsPair;
    const size_t cand_cap = 8 * scan_cap;
    if (ws->scan_cap < scan_cap || ws->scan_d < (size_t)d) {
        const size_t rows = scan_cap + scan_cap / 8;
        for (void** q : {&ws->d_xs, &ws->d_scan_row, &ws->d_scan_thr, &ws->d_scan_key, &ws->d_cand}) {
            if (*q) cudaFree(*q);
            *q = nullptr;
        }
        ws->scan_cap = 0;
        if (cudaMalloc(&ws->d_xs, rows * (size_t)d * 4) != cudaSuccess || cudaMalloc(&ws->d_scan_row, rows * 4) != cudaSuccess ||
            cudaMalloc(&ws->d_scan_thr, rows * 4) != cudaSuccess || cudaMalloc(&ws->d_scan_key, rows * 8) != cudaSuccess ||
            cudaMalloc(&ws->d_cand, 8 * rows * 8) != cudaSuccess) {
            snprintf(g_gm_err, sizeof(g_gm_err), "cudaMalloc of the second-pass buffers (%zu rows x %d) failed", rows, d);
            return 1;
        }
        ws->scan_cap = rows;
        ws->scan_d = (size_t)d;
    }
    // X: 128-row x 32-float boxes (SWIZZLE_128B).  The pre-tiled operands are read as rows of 256 floats without
    // swizzle, so that a ring slot (a contiguous block of the buffer) is one box.
    CUtensorMap tmX, tmXs, tmB, tmCn;
    char* e = g_gm_err;
    const size_t el = sizeof(g_gm_err);
    if (encode_2d(e, el, &tmX, X, (uint64_t)d, (uint64_t)n, (uint64_t)d * 4, 32, kRowsCta)) return 1;
    if (encode_2d(e, el, &tmXs, ws->d_xs, (uint64_t)d, (uint64_t)scan_cap, (uint64_t)d * 4, 32, kRowsCta)) return 1;
    if (encode_flat(e, el, &tmB, ws->d_b, 2 * (uint64_t)pl.k_pad * pl.d_pad / 256, (uint32_t)(pl.ntile / 16))) return 1;
    if (encode_flat(e, el, &tmCn, ws->d_cn, 8 * (uint64_t)pl.k_pad / 256, (uint32_t)(pl.ntile / 64))) return 1;

    static PerDeviceOnce attr_set;  // the opt-in is a per-device attribute
    const int dev = PerDeviceOnce::current();
    if (dev < 0 || !attr_set.done[dev]) {
        if (cudaFuncSetAttribute(kmeans_gemm_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxSmem) != cudaSuccess ||
            cudaFuncSetAttribute(kmeans_gemm_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxSmem) != cudaSuccess) {
            snprintf(g_gm_err, sizeof(g_gm_err), "cannot raise dynamic shared memory to %d bytes", kMaxSmem);
            return 1;
        }
        if (dev >= 0) attr_set.done[dev] = true;
    }
    {
        const size_t total = (size_t)pl.k_pad * pl.d_pad;
        size_t blocks = (total + 255) / 256;
        if (blocks > (size_t)kNumSMs * 8) blocks = (size_t)kNumSMs * 8;
        gemm_prep_kernel<<<(unsigned)blocks, 256, 0, s>>>(C, cnorm, k, d, pl.k_pad, pl.n_sub, pl.ntile / 2,
                                                         (float*)ws->d_b, (float*)ws->d_cn);
        count_launch();
    }
    cudaMemsetAsync(ws->d_count, 0, 64, s);

    GmParams p;
    p.n = n;
    p.num_tiles = (n + kRowsPair - 1) / kRowsPair;
    p.d = d;
    p.k = k;
    p.k_pad = pl.k_pad;
    p.n_sub = pl.n_sub;
    p.ntile = pl.ntile;
    p.n_ntiles = pl.n_ntiles;
    p.NA = pl.NA;
    p.NB = pl.NB;
    p.resident = pl.resident;
    p.rotate = (2 * (size_t)pl.k_pad * pl.d_pad * 4 <= ((size_t)16 << 20)) ? 1 : 0;
    p.b_slot_bytes = pl.b_slot_bytes;
    p.bt = (const float*)ws->d_b;
    p.cnt = (const float*)ws->d_cn;
    p.xnorm = xnorm;
    p.xlo = xlo;
    p.labels = labels;
    p.open = (uint4*)ws->d_open;
    p.counters = (unsigned int*)ws->d_count;
    p.thr = (const float*)ws->d_scan_thr;
    p.cand = (uint2*)ws->d_cand;
    p.cand_cap = (unsigned int)cand_cap;
    p.scan_cap = (unsigned int)scan_cap;
    p.st = st;
    p.honor_done = honor_done ? 1 : 0;
    // Error model (file header; the same as kernels_screen.cu).  x is read truncated to TF32, ||x - x^|| is the
    // row's own measured value (tol_x = slack of its FP32 evaluation) and <= 2^-10 ||x|| (abs_rel).  The rest is
    // FP32 rounding: the tensor core adds every MMA's products into the accumulator with truncation (<= 1 ulp of
    // the running sum per MMA, 2 ceil(d/8) + 1 MMAs), the c_hi / c_lo split leaves 2^-22 |c|, ||x||^2 rounds
    // once, and the index packed into the 5 low mantissa bits moves a score by < 2^-18 of itself.  With
    // 2 |x||c| <= |x|^2 + |c|^2 a margin carries at most eps_rel (||x||^2 + max ||c||^2) of all that.
    p.tol_x = 1.002f;
    p.eps_rel = 2.f * (float)(2 * ((d + 7) / 8) + 8) * 1.1920928955078125e-07f + 1e-6f + 1.6e-5f;
    p.abs_rel = 9.765625e-4f * 1.001f + p.eps_rel;
    p.off_a = pl.off_a;
    p.off_b = pl.off_b;
    p.off_ac = pl.off_ac;
    p.off_scr = pl.off_scr;
    p.off_bar = pl.off_bar;
    p.trace = nullptr;
#ifdef B200KM_ENABLE_TRACE
    const char* trace_path = getenv("B200KM_TRACE");
    if (trace_path) {
        cudaMalloc((void**)&p.trace, 16 * 1024 * sizeof(unsigned long long));
        cudaMemsetAsync(p.trace, 0, 16 * 1024 * sizeof(unsigned long long), s);
    }
#endif
    int64_t clusters = p.num_tiles < kNumSMs / 2 ? p.num_tiles : kNumSMs / 2;
    if (clusters < 1) clusters = 1;
    kmeans_gemm_kernel<false><<<(unsigned)(2 * clusters), kGmThreads, pl.total, s>>>(tmX, tmB, tmCn, p);
    count_launch();
    cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess) {
        snprintf(g_gm_err, sizeof(g_gm_err), "tile-GEMM kernel launch failed: %s", cudaGetErrorString(err));
        return 1;
    }
#ifdef B200KM_ENABLE_TRACE
    if (p.trace) {  // development aid: synchronous dump of cluster 0's event clocks, one line per index
        static unsigned long long host[16 * 1024];
        cudaStreamSynchronize(s);
        cudaMemcpy(host, p.trace, sizeof(host), cudaMemcpyDeviceToHost);
        cudaFree(p.trace);
        if (FILE* f = fopen(trace_path, "w")) {
            for (int i = 0; i < 1024; ++i) {
                for (int r = 0; r < 16; ++r) fprintf(f, "%llu ", host[r * 1024 + i]);
                fprintf(f, "\n");
            }
            fclose(f);
        }
    }
#endif
    {
        // exact decisions: pairs at once; rows with more candidates are gathered, scored again by the same GEMM
        // with the listing epilogue, and decided among the listed centroids
        const int hd = honor_done ? 1 : 0;
        int64_t blocks = (n + 7) / 8;
        if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
        unsigned int* counters = (unsigned int*)ws->d_count;
        gemm_fixup_kernel<<<(unsigned)blocks, 256, 0, s>>>(X, d, k, C, (const uint4*)ws->d_open, counters, labels,
                                                          (float*)ws->d_xs, (unsigned int*)ws->d_scan_row,
                                                          (float*)ws->d_scan_thr, (unsigned long long*)ws->d_scan_key,
                                                          (unsigned int)scan_cap, st, hd);
        count_launch();
        int64_t clusters2 = (int64_t)(scan_cap / kRowsPair) < kNumSMs / 2 ? (int64_t)(scan_cap / kRowsPair) : kNumSMs / 2;
        if (clusters2 < 1) clusters2 = 1;
        GmParams p2 = p;
        p2.trace = nullptr;
        kmeans_gemm_kernel<true><<<(unsigned)(2 * clusters2), kGmThreads, pl.total, s>>>(tmXs, tmB, tmCn, p2);
        count_launch();
        gemm_cand_kernel<<<kNumSMs * 8, 256, 0, s>>>((const float*)ws->d_xs, d, C, (const uint2*)ws->d_cand, counters,
                                                    (unsigned int)cand_cap, (unsigned long long*)ws->d_scan_key, st, hd);
        count_launch();
        gemm_scan_finish_kernel<<<kNumSMs * 8, 256, 0, s>>>(X, d, k, C, (const uint4*)ws->d_open, counters,
                                                           (const unsigned int*)ws->d_scan_row,
                                                           (const unsigned long long*)ws->d_scan_key, labels, st, hd);
        count_launch();
        err = cudaGetLastError();
        if (err != cudaSuccess) {
            snprintf(g_gm_err, sizeof(g_gm_err), "tile-GEMM fix-up launch failed: %s", cudaGetErrorString(err));
            return 1;
        }
    }
    return 0;
}

}  // namespace b200km
