// tcgen05 "screen" kernel: the Lloyd-iteration pass (and the labels-only predict pass).
//
// One persistent, warp-specialised kernel reads every 128-row tile of X once (TMA) and, while the tile sits
// in shared memory,
//   1. scores it against the centroids on the tensor cores straight from shared memory: A = the raw FP32
//      tile (the tensor core reads the top 19 bits of each word, i.e. x truncated to TF32), B = the stacked
//      TF32 split [c_hi ; c_lo] of the centroids, FP32 accumulation in TMEM; the assign warps order the centroids by
//      ||c_j||^2 - 2 x^.c_j and form D_j = ||x||^2 + ||c_j||^2 - 2 x^.c_j for the three nearest;
//   2. decides which rows the TF32 view of x decides.  With x^ = trunc(x) the error of a MARGIN D_a - D_b is
//      2 (x - x^).(c_a - c_b) plus FP32 rounding, hence bounded by 2 ||x - x^|| ||c_a - c_b|| + eps: it scales
//      with the distance between the two centroids (a k x k table), not with their norms, and ||x - x^|| is the
//      row's own value (computed with the row norms), not the 2^-10 ||x|| worst case.  A centroid j can truly
//      beat the computed best c_1 only if ||c_1 - c_j|| <= 2 r_1 (r_1 = distance to c_1); so a row is decided
//      when its runner-up margin exceeds the pair's bound and either no other centroid is that close to c_1
//      (each centroid's nearest-neighbour distances) or the third-best margin exceeds 4 ||x - x^|| r_1 + eps.
//      Every row leaves a verdict word; the open rows keep the computed best as a provisional label;
//   3. adds the decided rows (exact FP32 data, not the TF32 view) into per-cluster sums: every row ORs its bit into
//      its label's mask word (shared-memory atomic OR: order-free), the assign group turns the masks into the list of
//      the tile's rows ordered by (label, row) - no comparison sort: positions are popcounts -, and the update warps
//      add each label's run in that order into FP32 registers / float64 shared memory.
// The exact pass (kernels_fixup.cu) then decides the open rows exactly (FP32 squared differences, float64 when even
// those tie) and adds them - labels and sums are exact for every row.  Keeping that rarely executed code out of the
// persistent kernel matters: inside it, one such row cost 1400-2400 cycles of an assign warp (instruction fetch of
// cold code), 0.9 % of the rows cost 30 %.
// The pass produces no per-row distances; a caller that wants the inertia of the pass (b200km_lloyd_step) takes one
// more pass of exact per-row differences (rowdist_kernel) rather than the cancelling identity
// sum ||x||^2 - 2 sum c.S + sum n ||c||^2.
//
// When too many rows stay open (structureless data) b200km_fit switches to the split-TF32 kernel
// (kernels_tc.cu), which also splits x and therefore leaves nothing open beyond FP32 rounding.
//
// Warp roles (704 threads, 1 CTA per SM, grid = #SMs):
//   warp 0       TMA producer (centroids once, then X tiles: 32-float SWIZZLE_128B K-atoms + a narrow tail)
//   warp 1       TMEM allocator + MMA issuer (warp-uniform loop, one elected lane, A and B via UMMA descriptors);
//                k_pad <= 64: ONE instruction per K-step against [c_hi ; c_lo] stacked as 2 k_pad columns (the tile is
//                read from shared memory once), the assign warps add the two halves
//   warps 2-5    row norms ||x||^2 and ||x - x^||^2 on the first pass of a fit; later passes read them back and
//                these warps are four more update warps
//   warps 6-13   assign, two groups taking alternate tiles: tcgen05.ld of the real columns only (16 + 8 + 4 column
//                chunks), three smallest keys by a tournament, verdict, label masks, row list
//   warps 14-21  update: a warp owns <= 4 labels, lane = float4 column chunk; four list entries per broadcast load,
//                rows past a run's end are not loaded
// What the kernel runs out of (B200, d = 100, k = 20): a tile occupies its 52 KB stage for ~9000 cycles (three stages:
// 2 stages cost +32 % time), the shared-memory pipe is ~70 % busy (TMA writes + tensor-core operand reads + the
// update warps' row reads), and every role is a short dependent instruction stream.  What paid: fewer instructions
// on the assign warps (atomic-OR masks instead of a sort, only the real columns, keys without the row norm), fewer
// shared-memory wavefronts in the update warps, a small code footprint (the roles share a 32 KB instruction cache:
// an unrolled variant that doubled the code ran at 82 % hit rate and 1.7x the time).  What did not: starting the
// MMAs before the whole tile has landed, an L2 prefetch ahead of the TMA loads, both assign groups on one tile,
// two labels per update loop, inline waits everywhere.
//
// Replaces cuML's fusedL2NN + reduce_rows_by_key pair behind reference
// python/src/spark_rapids_ml/clustering.py:367-370 (fit) and :493-497 (predict).
#include "tc.cuh"
#include "tc_ptx.cuh"

#include <cfloat>
#include <climits>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace b200km {

namespace {

using namespace tcx;

thread_local char g_sc_err[512] = "";

constexpr int kAtomBytes = kTileM * 128;  // one full K-atom of an X tile: 128 rows x 32 floats
constexpr int kThreads = 704;
constexpr int kNormWarp0 = 2, kAssignWarp0 = 6, kUpdateWarp0 = 14;
constexpr int kMaxSmem = 227 * 1024;
constexpr int kMaxStages = 4;
constexpr int kMaxND = 4;
constexpr int kMaxD = 256;
constexpr int kMaxK = 128;    // labels-only passes; larger k runs the split-TF32 kernel
constexpr int kMaxKAcc = 32;  // passes with the update stage (register accumulators, lane = label in the sort)
#ifndef B200KM_UPD_BATCH
#define B200KM_UPD_BATCH 4
#endif
constexpr int kUpdBatch = B200KM_UPD_BATCH;  // rows of a label's run loaded together by an update warp

struct ScParams {
    int64_t n;
    int64_t num_tiles;
    int d, k, k_pad;
    int n_full;     // number of 32-float SWIZZLE_128B atoms
    int tail_cols;  // 0, 8 (SWIZZLE_32B) or 16 (SWIZZLE_64B): width of the narrow last atom
    int num_stages, num_d;
    int stacked;  // 1: x^.c_hi and x^.c_lo land in separate TMEM column ranges (one MMA per K-step)
    int stage_bytes;
    const float* X;
    int prefetch;  // tiles the producer asks into L2 ahead of its loads (0 = none)
    int skip;      // development builds: bit 0 = the update warps add nothing, bit 1 = the assign warps skip the top-3
    const float* C;
    const float* cnorm;
    float* xnorm;  // [n] ||x||^2: written when xn_compute, read otherwise
    float* xlo;    // [n] ||x - trunc_tf32(x)||^2: likewise
    int xn_compute;
    int32_t* labels;  // may be null
    unsigned int* info;  // [n] verdict word of every OPEN row (see kInfo*): read by screen_fixup_kernel
    unsigned int* open_bits;  // [num_tiles][4]: bit r of word q = row q*32 + r of the tile is open
    double* acc;      // [k*d | k | 1] or null (labels-only pass); only its non-nullness matters to the kernel
    double* slots;    // [grid][slot_stride]: every CTA's float64 partial sums | counts | sum ||x||^2
    int slot_stride;
    IterState* st;
    int honor_done;
    const float* pd;  // [k][k] upper bounds of ||c_a - c_b||
    float tol_x;      // slack on the FP32 value of ||x - trunc_tf32(x)||
    float eps_rel;    // FP32 rounding terms of a margin, relative to ||x||^2 + max ||c||^2
    float abs_rel;    // bound of the error of one D_j, relative to ||x||^2 + max ||c||^2
    unsigned long long* trace;  // development aid (B200KM_TRACE)
    int off_c, off_acc, off_cn, off_xn, off_pd, off_bar;
};

// D[tmem] (+)= A[smem desc] * B[smem desc], kind::tf32, M=128
__device__ __forceinline__ void mma_tf32_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                            uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}

// acc (two packed floats) += v: one FADD2 for two columns
__device__ __forceinline__ void add2(unsigned long long& acc, unsigned long long v) {
    asm("add.rn.f32x2 %0, %0, %1;" : "+l"(acc) : "l"(v));
}
__device__ __forceinline__ float2 unpack2(unsigned long long v) {
    float2 r;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(r.x), "=f"(r.y) : "l"(v));
    return r;
}


// The assign warps' waits: inline (a call would make the compiler wait for every load in flight first, and these
// warps keep the next tile's row norms in flight across their waits).
__device__ __forceinline__ void sc_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile(
            "{\n\t"
            ".reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t"
            "}"
            : "=r"(ok)
            : "r"(bar), "r"(parity), "r"(0x989680)
            : "memory");
    } while (!ok);
}
__device__ __forceinline__ float sqrt_up(float x) {  // >= sqrt(x): the approximate root (2 ulp) with slack
    float r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r * 1.000001f;
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}

// Entry of the sorted row list: byte offset of row r inside a full atom (r*128 | swizzle) in the low half, inside the
// narrow tail atom (row pitch 64 / 32 bytes) in the high half.
__device__ __forceinline__ unsigned int perm_entry(int r, int tail_cols) {
    const unsigned int full = (unsigned int)((r << 7) | ((r & 7) << 4));
    const unsigned int tl = tail_cols == 16 ? (unsigned int)((r << 6) | (((r >> 1) & 3) << 4))
                                            : (unsigned int)((r << 5) | (((r >> 2) & 1) << 4));
    return full | (tl << 16);
}

// Byte offset of 16-byte chunk `chunk` (= float column / 4) of row r inside one tile stage.  Full atoms
// use the TMA 128-byte swizzle (chunk ^= row & 7); the narrow tail atom the 64-byte (chunk ^= (row>>1)&3)
// or 32-byte (chunk ^= (row>>2)&1) swizzle.
__device__ __forceinline__ uint32_t chunk_off(const ScParams& p, int r, int chunk) {
    const int a = chunk >> 3;
    if (a < p.n_full) return (uint32_t)(a * kAtomBytes + r * 128 + (((chunk & 7) ^ (r & 7)) << 4));
    const int c = chunk - p.n_full * 8;
    const uint32_t base = (uint32_t)(p.n_full * kAtomBytes);
    if (p.tail_cols == 16) return base + (uint32_t)(r * 64 + ((c ^ ((r >> 1) & 3)) << 4));
    return base + (uint32_t)(r * 32 + ((c ^ ((r >> 2) & 1)) << 4));
}

__device__ __forceinline__ void trace_ev(const ScParams& p, int64_t it, int ev) {
#ifdef B200KM_ENABLE_TRACE
    if (p.trace && blockIdx.x == 0 && it < 256) p.trace[it * 16 + ev] = (unsigned long long)clock64();
#endif
}

// ------------------------------------------------------------------------------------------------
// Operand preparation of a pass, one launch.  Blocks [0, split_blocks): the B operand out[2*k_pad][d_pad] (rows
// [0,k_pad) = centroids rounded to TF32, rows [k_pad,2k_pad) = the TF32-rounded remainders, zero padded); the
// other blocks: pd[a][b] = an upper bound of ||c_a - c_b|| (FP32 sum inflated by its own rounding bound), one
// warp per pair.
__global__ void __launch_bounds__(128) screen_prep_kernel(const float* __restrict__ C, int k, int d, int k_pad,
                                                          int d_pad, int split_blocks, float* __restrict__ out,
                                                          float* __restrict__ pd) {
    if ((int)blockIdx.x < split_blocks) {
        const int total = k_pad * d_pad;
        for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += split_blocks * blockDim.x) {
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
        return;
    }
    const int pair = ((int)blockIdx.x - split_blocks) * 4 + (threadIdx.x >> 5);
    if (pair >= k * k) return;
    const int a = pair / k, b = pair - a * k;
    const int lane = threadIdx.x & 31;
    float s = 0.f;
    for (int i = lane; i < d; i += 32) {
        const float df = C[(size_t)a * d + i] - C[(size_t)b * d + i];
        s = fmaf(df, df, s);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) pd[pair] = sqrtf(s) * (1.f + 1e-5f) + 1e-30f;
}

// The three smallest of N (4, 8 or 16) keys by a tournament of sorted lists (independent operations, short dependency
// chains): pairs -> triples -> ...; the third smallest of two sorted triples is min(a3, b3, max(a1,b2), max(a2,b1)).
template <int N>
__device__ __forceinline__ void top3_of(const float (&key)[N], float& o1, float& o2, float& o3) {
    float m1[N / 2], m2[N / 2], m3[N / 4];
#pragma unroll
    for (int i = 0; i < N / 2; ++i) {
        m1[i] = fminf(key[2 * i], key[2 * i + 1]);
        m2[i] = fmaxf(key[2 * i], key[2 * i + 1]);
    }
#pragma unroll
    for (int i = 0; i < N / 4; ++i) {  // two sorted pairs -> sorted triple
        const float a1 = m1[2 * i], a2 = m2[2 * i], b1 = m1[2 * i + 1], b2 = m2[2 * i + 1];
        m1[i] = fminf(a1, b1);
        m2[i] = fminf(fmaxf(a1, b1), fminf(a2, b2));
        m3[i] = fminf(fmaxf(a1, b2), fmaxf(a2, b1));
    }
#pragma unroll
    for (int nn = N / 4; nn > 1; nn >>= 1) {
#pragma unroll
        for (int i = 0; i < nn / 2; ++i) {
            const float a1 = m1[2 * i], a2 = m2[2 * i], a3 = m3[2 * i];
            const float b1 = m1[2 * i + 1], b2 = m2[2 * i + 1], b3 = m3[2 * i + 1];
            m1[i] = fminf(a1, b1);
            m2[i] = fminf(fmaxf(a1, b1), fminf(a2, b2));
            m3[i] = fminf(fminf(a3, b3), fminf(fmaxf(a1, b2), fmaxf(a2, b1)));
        }
    }
    o1 = m1[0];
    o2 = m2[0];
    o3 = m3[0];
}
template <int N>
__device__ __forceinline__ void tmem_ldn(uint32_t taddr, uint32_t (&r)[N]);
template <>
__device__ __forceinline__ void tmem_ldn<16>(uint32_t taddr, uint32_t (&r)[16]) {
    tmem_ld16(taddr, r);
}
template <>
__device__ __forceinline__ void tmem_ldn<4>(uint32_t taddr, uint32_t (&r)[4]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(taddr)
                 : "memory");
}
template <>
__device__ __forceinline__ void tmem_ldn<8>(uint32_t taddr, uint32_t (&r)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr)
                 : "memory");
}
// N columns [c0, c0 + N) of this thread's row: scores from TMEM (the two column ranges of a stacked pass added),
// keys ||c_j||^2 - 2 x^.c_j with the centroid index in the low mantissa bits, their three smallest merged into
// the running sorted triple (b1, b2, b3).
template <int N>
__device__ __forceinline__ void score_chunk(uint32_t taddr, int c0, int lo_off, bool stacked, const float* cn_s,
                                            uint32_t idx_mask, float& b1, float& b2, float& b3) {
    uint32_t sc[N];
    tmem_ldn<N>(taddr + c0, sc);
    if (stacked) {
        uint32_t sl[N];
        tmem_ldn<N>(taddr + lo_off + c0, sl);
        tmem_wait_ld();
#pragma unroll
        for (int j = 0; j < N; ++j) sc[j] = __float_as_uint(__uint_as_float(sc[j]) + __uint_as_float(sl[j]));
    } else {
        tmem_wait_ld();
    }
    float key[N];
    const float4* cn4 = reinterpret_cast<const float4*>(cn_s + c0);
#pragma unroll
    for (int j4 = 0; j4 < N / 4; ++j4) {
        const float4 c = cn4[j4];
        const float cc[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const int j = j4 * 4 + t;
            const float dj = fmaf(-2.f, __uint_as_float(sc[j]), cc[t]);  // padded columns: FLT_MAX
            key[j] = __uint_as_float((__float_as_uint(dj) & ~idx_mask) | (uint32_t)(c0 + j));
        }
    }
    float t1, t2, t3;
    top3_of<N>(key, t1, t2, t3);
    const float n1 = fminf(b1, t1);
    const float n2 = fminf(fmaxf(b1, t1), fminf(b2, t2));
    const float n3 = fminf(fminf(b3, t3), fminf(fmaxf(b1, t2), fmaxf(b2, t1)));
    b1 = n1;
    b2 = n2;
    b3 = n3;
}

// ||x||^2 and ||x - trunc_tf32(x)||^2 of one row of the tile in `stage` (interleaved partial sums)
__device__ __forceinline__ void acc_norms(const float4 v, float (&s)[4], float (&l)[4]) {
    const float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int t = 0; t < 4; ++t) {
        s[t] = fmaf(e[t], e[t], s[t]);
        const float r = e[t] - __uint_as_float(tf32_trunc(__float_as_uint(e[t])));  // exact
        l[t] = fmaf(r, r, l[t]);
    }
}
__device__ __forceinline__ float2 row_norms(const ScParams& p, const uint8_t* stage, int row) {
    float s[4] = {0.f, 0.f, 0.f, 0.f}, l[4] = {0.f, 0.f, 0.f, 0.f};
    for (int a = 0; a < p.n_full; ++a) {
        const uint8_t* rowp = stage + a * kAtomBytes + row * 128;
#pragma unroll
        for (int c = 0; c < 8; ++c) acc_norms(*reinterpret_cast<const float4*>(rowp + ((c ^ (row & 7)) << 4)), s, l);
    }
    if (p.tail_cols) {
        const int nch = p.tail_cols >> 2;
        for (int c = 0; c < nch; ++c)
            acc_norms(*reinterpret_cast<const float4*>(stage + chunk_off(p, row, p.n_full * 8 + c)), s, l);
    }
    return make_float2((s[0] + s[1]) + (s[2] + s[3]), (l[0] + l[1]) + (l[2] + l[3]));
}

__global__ void __launch_bounds__(kThreads, 1)
    kmeans_screen_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmXt,
                         const __grid_constant__ CUtensorMap tmC, const __grid_constant__ CUtensorMap tmCt,
                         const ScParams p) {
    if (p.honor_done && p.st->done) return;

    extern __shared__ uint8_t smem_raw[];
    // per assign group (= tile parity): bit r of s_mask[g][label][q] = row q*32 + r of the tile goes to `label`
    __shared__ __align__(16) unsigned int s_mask[2][kMaxKAcc + 1][4];  // row kMaxKAcc stays zero (free slots)
    // per assign group: the tile's decided rows ordered by (label, row); low half of an entry = byte offset of the row
    // inside a full atom (r*128 | swizzle), high half = the same inside the narrow tail atom
    __shared__ __align__(16) unsigned int s_perm[2][kTileM + 4 * kMaxKAcc];  // runs start at multiples of the update batch
    __shared__ unsigned short s_base[2][kMaxKAcc];  // per label: start of its run in s_perm | rows << 8
    __shared__ unsigned int s_cntw[8][32];  // per assign warp: decided rows per label (lane = label)
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const uint32_t smem_base = smem_u32(smem);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int S = p.num_stages;
    constexpr int ND = kMaxND;  // score buffers (the plan always uses all four)
    // Ring positions are advanced by increments: the roles below are single instruction streams that run at the
    // latency of their dependent instructions, and a 64-bit division by a run-time stage count costs hundreds of
    // cycles on that path.
    const int stage_bytes = p.stage_bytes;
    const int c_block = 2 * p.k_pad * 128;  // bytes of one full-atom block [c_hi rows | c_lo rows]
    // TMEM columns of one score buffer.  stacked: one instruction per K-step against [c_hi ; c_lo] as 2*k_pad
    // columns (the tile is read from shared memory once, the assign warps add the two halves); otherwise two
    // instructions per K-step into the same k_pad columns.
    const bool stacked = p.stacked != 0;
    const int dcols = stacked ? 2 * p.k_pad : p.k_pad;   // N of the MMA
    const int dstride = (dcols + 31) & ~31;              // TMEM columns from one score buffer to the next
    const int n_groups = p.acc ? 2 : 4;  // assign groups: without an update stage its warps score tiles too
    // update warps: 14-21, and - on passes that read the row norms back - the four idle norm warps as well
    const int n_upd = p.xn_compute ? kOwnBinsFirst : kOwnBinsLater;

    // barrier table
    const uint32_t bar0 = smem_base + p.off_bar;
    const uint32_t x_full0 = bar0, x_empty0 = bar0 + 32, xn_full0 = bar0 + 64, d_full0 = bar0 + 96,
                   d_empty0 = bar0 + 128, s_full0 = bar0 + 160, s_empty0 = bar0 + 176, c_full = bar0 + 192;
    uint32_t* tmem_slot = (uint32_t*)(smem + p.off_bar + 200);

    if (threadIdx.x == 0) {
        for (int s = 0; s < kMaxStages; ++s) {
            mbar_init(x_full0 + 8u * s, 1);
            mbar_init(x_empty0 + 8u * s, p.acc ? n_upd : 4);
            mbar_init(xn_full0 + 8u * s, 4);
        }
        for (int s = 0; s < kMaxND; ++s) {
            mbar_init(d_full0 + 8u * s, 1);
            mbar_init(d_empty0 + 8u * s, 4);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(s_full0 + 8u * s, 4);
            mbar_init(s_empty0 + 8u * s, n_upd);
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
        if (p.k <= 32) {
            float* pd_init = (float*)(smem + p.off_pd);
#pragma unroll 1
            for (int i = threadIdx.x; i < 32 * 32; i += kThreads) {
                const int a = i >> 5, b = i & 31;
                pd_init[i] = (a < p.k && b < p.k) ? p.pd[a * p.k + b] : 0.f;
            }
        }
        if (p.acc) {
            double* sacc_init = (double*)(smem + p.off_acc);
#pragma unroll 1
            for (int i = threadIdx.x; i < p.k * p.d + p.k; i += kThreads) sacc_init[i] = 0.0;
        }
#pragma unroll 1
        for (int i = threadIdx.x; i < 2 * (kMaxKAcc + 1) * 4; i += kThreads) (&s_mask[0][0][0])[i] = 0;
#pragma unroll 1
        for (int i = threadIdx.x; i < 2 * (kTileM + 4 * kMaxKAcc); i += kThreads) (&s_perm[0][0])[i] = 0;
        if (threadIdx.x < 8 * 32) (&s_cntw[0][0])[threadIdx.x] = 0;
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    {
        // per centroid: its nearest other centroid and the distances to the nearest and the second nearest one
        // (the third-candidate test of the assign warps); written before any consumer can have a tile
        if (threadIdx.x < p.k) {
            const float* pd_init = (const float*)(smem + p.off_pd);
            float* nb = (float*)(smem + p.off_pd) + 32 * 32;
            const int i = threadIdx.x;
            float d1 = FLT_MAX, d2 = FLT_MAX;
            int n1 = -1;
            for (int j = 0; j < p.k; ++j) {
                if (j == i) continue;
                const float v = p.k <= 32 ? pd_init[i * 32 + j] : __ldg(p.pd + (size_t)i * p.k + j);
                if (v < d1) {
                    d2 = d1;
                    d1 = v;
                    n1 = j;
                } else if (v < d2) {
                    d2 = v;
                }
            }
            nb[i] = d1;
            nb[kMaxK + i] = d2;
            nb[2 * kMaxK + i] = __int_as_float(n1);
        }
        __syncthreads();
    }
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
            // L2 runs ahead of the shared-memory stages: a tile is one contiguous range of X
            const size_t tile_bytes = (size_t)kTileM * p.d * sizeof(float);
            const size_t x_bytes = (size_t)p.n * p.d * sizeof(float);
            auto prefetch_tile = [&](int64_t t) {
                if (t >= p.num_tiles) return;
                const size_t off = (size_t)t * tile_bytes;
                const size_t len = x_bytes - off < tile_bytes ? x_bytes - off : tile_bytes;
                prefetch_l2_bulk(reinterpret_cast<const char*>(p.X) + off, (uint32_t)len);
            };
            for (int a = 1; a < p.prefetch; ++a) prefetch_tile(first_tile + (int64_t)a * tile_step);
            int64_t it = 0;
            int s = 0;
            uint32_t sph = 0;  // phase of stage s
            for (int64_t tile = first_tile; tile < p.num_tiles; tile += tile_step, ++it) {
                if (p.prefetch) prefetch_tile(tile + (int64_t)p.prefetch * tile_step);
                mbar_wait(x_empty0 + 8u * s, sph ^ 1u);
                trace_ev(p, it, 0);
                mbar_arrive_expect_tx(x_full0 + 8u * s, (uint32_t)stage_bytes);
                const uint32_t dst = smem_base + s * stage_bytes;
                for (int a = 0; a < p.n_full; ++a)
                    tma_load_2d(dst + a * kAtomBytes, &tmX, x_full0 + 8u * s, a * 32, (int)(tile * kTileM));
                if (p.tail_cols)
                    tma_load_2d(dst + p.n_full * kAtomBytes, &tmXt, x_full0 + 8u * s, p.n_full * 32,
                                (int)(tile * kTileM));
                if (++s == S) {
                    s = 0;
                    sph ^= 1u;
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ====================================== MMA issuer ======================================
        // The whole warp runs the loop and one elected lane issues (a tcgen05.mma issued from divergent
        // code costs ~100 cycles per instruction, ~40 from warp-uniform code: tools/mma_bench.cu).  Both
        // operands come from shared memory: A = the X tile exactly as TMA wrote it (K-major, swizzled), B =
        // the TF32 centroids; consecutive K-steps differ only in the descriptors' start address (+32 B).
        // B holds [c_hi rows ; c_lo rows] per K-atom; a K-step is two instructions into the same k_pad columns:
        // x^.c_hi and x^.c_lo.
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(kTileM >> 4) << 24) |
                               ((uint32_t)(dcols >> 3) << 17);
        const int tail_row_bytes = p.tail_cols ? p.tail_cols * 4 : 128;
        const uint64_t bdesc_full = make_b_desc(smem_base + p.off_c, 128);
        const uint64_t bdesc_tail = make_b_desc(smem_base + p.off_c + p.n_full * c_block, tail_row_bytes);
        const uint64_t adesc_full = make_b_desc(smem_base, 128);
        const uint64_t adesc_tail = make_b_desc(smem_base + p.n_full * kAtomBytes, tail_row_bytes);
        const uint64_t lo_full = (uint64_t)((p.k_pad * 128) >> 4);  // c_lo rows inside an atom block
        const uint64_t lo_tail = (uint64_t)((p.k_pad * tail_row_bytes) >> 4);
        const int tail_steps = p.tail_cols >> 3;
        mbar_wait(c_full, 0);
        int64_t it = 0;
        int s = 0;
        uint32_t sph = 0;
        for (int64_t tile = first_tile; tile < p.num_tiles; tile += tile_step, ++it) {
            const int db = (int)(it & (ND - 1));
            mbar_wait(d_empty0 + 8u * db, (uint32_t)(((it / ND) & 1) ^ 1));
            mbar_wait(x_full0 + 8u * s, sph);
            tc_fence_after();
            if (lane == 0) trace_ev(p, it, 1);
            const uint32_t d_tmem = tmem_base + (uint32_t)(db * dstride);
            const uint64_t soff = (uint64_t)((s * stage_bytes) >> 4);
            if (elect_one()) {
                uint64_t ad = adesc_full + soff, bd = bdesc_full;
                if (stacked) {
#pragma unroll 1
                    for (int a = 0; a < p.n_full; ++a) {
#pragma unroll
                        for (int kk = 0; kk < 4; ++kk) mma_tf32_ss(d_tmem, ad + 2 * kk, bd + 2 * kk, idesc, (a | kk) != 0);
                        ad += (uint64_t)(kAtomBytes >> 4);
                        bd += (uint64_t)(c_block >> 4);
                    }
#pragma unroll 1
                    for (int kk = 0; kk < tail_steps; ++kk)
                        mma_tf32_ss(d_tmem, adesc_tail + soff + 2 * kk, bdesc_tail + 2 * kk, idesc, (p.n_full | kk) != 0);
                } else {
#pragma unroll 1
                    for (int a = 0; a < p.n_full; ++a) {
#pragma unroll
                        for (int kk = 0; kk < 4; ++kk) {
                            mma_tf32_ss(d_tmem, ad + 2 * kk, bd + 2 * kk, idesc, (a | kk) != 0);
                            mma_tf32_ss(d_tmem, ad + 2 * kk, bd + lo_full + 2 * kk, idesc, 1);
                        }
                        ad += (uint64_t)(kAtomBytes >> 4);
                        bd += (uint64_t)(c_block >> 4);
                    }
#pragma unroll 1
                    for (int kk = 0; kk < tail_steps; ++kk) {
                        mma_tf32_ss(d_tmem, adesc_tail + soff + 2 * kk, bdesc_tail + 2 * kk, idesc, (p.n_full | kk) != 0);
                        mma_tf32_ss(d_tmem, adesc_tail + soff + 2 * kk, bdesc_tail + lo_tail + 2 * kk, idesc, 1);
                    }
                }
                tc_commit(d_full0 + 8u * db);
                trace_ev(p, it, 2);
            }
            __syncwarp();
            if (++s == S) {
                s = 0;
                sph ^= 1u;
            }
        }
    } else if (warp >= kNormWarp0 && warp < kAssignWarp0 && p.xn_compute) {
        // ======================================= row norms ======================================
        {
            const int row = (warp - kNormWarp0) * 32 + lane;
            float* xn_s = (float*)(smem + p.off_xn);
            int s = 0;
            uint32_t sph = 0;
            for (int64_t tile = first_tile; tile < p.num_tiles; tile += tile_step) {
                mbar_wait(x_full0 + 8u * s, sph);
                const float2 nn = row_norms(p, smem + (size_t)s * stage_bytes, row);
                const int64_t grow = tile * kTileM + row;
                if (grow < p.n) {
                    p.xnorm[grow] = nn.x;
                    p.xlo[grow] = nn.y;
                }
                xn_s[s * kTileM + row] = nn.x;
                xn_s[(kMaxStages + s) * kTileM + row] = nn.y;
                __syncwarp();
                if (lane == 0) mbar_arrive(xn_full0 + 8u * s);
                if (++s == S) {
                    s = 0;
                    sph ^= 1u;
                }
            }
        }
    } else if (warp >= kAssignWarp0 && (warp < kUpdateWarp0 || !p.acc)) {
        // ======================================== assign ========================================
        // Groups of four warps take tiles in turn (two groups when there is an update stage - group g then owns
        // the sort buffer of its parity -, four when the update warps are free to score tiles too), so the
        // top-3 / verdict / sort latency of one tile overlaps the next tiles'.
        const int g = (warp - kAssignWarp0) >> 2;
        const int q = warp & 3;  // = the warp's TMEM lane quarter
        const int row = q * 32 + lane;
        const int et = threadIdx.x - (kAssignWarp0 + 4 * g) * 32;  // 0..127 within the group
        float* cn_s = (float*)(smem + p.off_cn);
        float* xn_s = (float*)(smem + p.off_xn);
        const float cmax = p.st->cnorm_max;
        const float* pd_s = (const float*)(smem + p.off_pd);        // [32][32] pair distances (k <= 32)
        const float* nb_s = pd_s + 32 * 32;                        // per centroid: nearest, second nearest, who
        const int sb = g;

        unsigned int cnt_lane = 0u;  // lane = label: rows this warp has decided for it
        // the centroid index rides in the low mantissa bits of a key: 5 bits (2^-18 of the key) for k <= 32, else 7
        const uint32_t idx_mask = p.k_pad <= 32 ? 31u : 127u;
        const float idx_eps = p.k_pad <= 32 ? 4e-6f : 1.6e-5f;
        int s = g % S;
        uint32_t sph = (uint32_t)((g / S) & 1);  // g < 4: cheap, once
        int64_t tile = first_tile + (int64_t)g * tile_step;
        const int64_t group_step = (int64_t)n_groups * tile_step;
        // the row norms of a pass that reads them back are fetched one tile ahead (HBM latency under load is
        // longer than a tile's turn)
        float xn_next = 0.f, xl_next = 0.f;
        if (!p.xn_compute && tile < p.num_tiles && tile * kTileM + row < p.n) {
            xn_next = p.xnorm[tile * kTileM + row];
            xl_next = p.xlo[tile * kTileM + row];
        }
        for (int64_t it = g; tile < p.num_tiles; it += n_groups, tile += group_step) {
            const int db = (int)(it & (ND - 1));
            const int64_t grow = tile * kTileM + row;
            const bool valid = grow < p.n;
            float xn = xn_next, xl = xl_next;
            {
                const int64_t grow_next = grow + group_step * kTileM;
                if (!p.xn_compute && grow_next < p.n) {
                    xn_next = p.xnorm[grow_next];
                    xl_next = p.xlo[grow_next];
                }
            }
            sc_wait(d_full0 + 8u * db, (uint32_t)((it / ND) & 1));
            tc_fence_after();
            if (et == 0) trace_ev(p, it, 4);

            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(db * dstride);
            int label = 0;
            unsigned int verdict = 0;  // this row's info word
            // ---- the scores a chunk of columns at a time into registers (score_chunk) ----
            if (p.xn_compute) {
                sc_wait(xn_full0 + 8u * s, sph);
                xn = xn_s[s * kTileM + row];
                xl = xn_s[(kMaxStages + s) * kTileM + row];
            }
            float b1 = FLT_MAX, b2 = FLT_MAX, b3 = FLT_MAX;
#ifdef B200KM_ENABLE_TRACE
            if (p.skip & 2) {
                b1 = __uint_as_float(0x3f800000u | (uint32_t)(lane % p.k));
                b2 = __uint_as_float(0x41800000u | (uint32_t)((lane + 1) % p.k));
                b3 = 100.f;
            } else
#endif
            {
                // the real columns only: 16 at a time, then 8 and 4 for the remainder (k = 20: 16 + 4, not 32)
                const int kcols = (p.k + 3) & ~3;
                int c0 = 0;
#pragma unroll 1
                for (; c0 + 16 <= kcols; c0 += 16) score_chunk<16>(taddr, c0, p.k_pad, stacked, cn_s, idx_mask, b1, b2, b3);
                if (c0 + 8 <= kcols) {
                    score_chunk<8>(taddr, c0, p.k_pad, stacked, cn_s, idx_mask, b1, b2, b3);
                    c0 += 8;
                }
                if (c0 < kcols) score_chunk<4>(taddr, c0, p.k_pad, stacked, cn_s, idx_mask, b1, b2, b3);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(d_empty0 + 8u * db);  // the score buffer is free again
            const int i1 = (int)(__float_as_uint(b1) & idx_mask), i2 = (int)(__float_as_uint(b2) & idx_mask);
            label = i1;
            // the keys are ||c_j||^2 - 2 x^.c_j: the row's own ||x||^2 does not change their order and is added to
            // the three that matter only.  The index bits perturb a key by less than idx_eps of its magnitude.
            const float kmag12 = fabsf(b1) + fabsf(b2), kmag3 = fabsf(b3);
            b1 += xn;
            b2 += xn;
            b3 += xn;
            // safety test (see the file header); every quantity is an upper bound
            const float scale = xn + cmax;
            const float t = p.tol_x * sqrt_up(xl);  // >= ||x - x^||
            const float eps = p.eps_rel * scale + idx_eps * kmag12;
            const float r1 = sqrt_up(fmaxf(b1, 0.f) + p.abs_rel * scale);
            const float safe = 4.f * t * r1 + eps;
            const float margin = b2 - b1;
            // runner-up: the pair bound.  Any other centroid j can truly beat c_1 only if ||c_1 - c_j|| <= 2 r_1:
            // impossible when even the nearest such centroid is farther than that; otherwise the third best
            // margin has to exceed `safe`.
            const float pd12 = p.k <= 32 ? pd_s[i1 * 32 + i2] : __ldg(p.pd + (size_t)i1 * p.k + min(i2, p.k - 1));
            const bool pair_ok = margin >= fminf(2.f * t * pd12 + eps, safe);
            const float others = (__float_as_int(nb_s[2 * kMaxK + i1]) == i2) ? nb_s[kMaxK + i1] : nb_s[i1];
            const bool third = valid && p.k > 2 && others <= 2.f * r1 &&
                               (b3 - b1) < safe + idx_eps * kmag3;  // a third candidate cannot be excluded
            const bool amb = (valid && p.k > 1 && !pair_ok) || third;
            verdict = (unsigned)i1 | ((unsigned)i2 << 8) | (amb ? kInfoUnsafe | (third ? kInfoReach : kInfoPair) : 0u);

            // open rows leave their verdict word and keep the computed best as a provisional label until the exact
            // pass has decided them (they are not part of this kernel's sums)
            if (valid) {
                if (amb) p.info[grow] = verdict;
                if (p.labels) p.labels[grow] = label;
            }
            {
                // open rows: one bitmap word per warp and tile (the exact pass finds them without reading a word per row)
                const unsigned bits = __ballot_sync(0xffffffffu, amb);
                if (lane == 0) p.open_bits[tile * 4 + q] = bits;
            }
            if (et == 0) trace_ev(p, it, 5);

            if (p.acc) {
                // ---- which rows go to which label: one mask word per (label, warp), then the row list ----
                // Rows that are open (the exact pass adds them) or past the end take no part.
                sc_wait(s_empty0 + 8u * sb, (uint32_t)(((it >> 1) & 1) ^ 1));
                const int bin = (valid && !amb) ? label : -1;
                // every row ORs its bit into its label's word (order does not matter for an OR)
                if (lane < p.k) s_mask[sb][lane][q] = 0u;
                __syncwarp();
                if (bin >= 0) atomicOr(&s_mask[sb][bin][q], 1u << lane);
                named_bar_sync(1 + g, 128);  // the group's four mask words of every label are complete
                // lane = label: where its rows start in the sorted order, and where this warp's rows of it start
                const uint4 m = *reinterpret_cast<const uint4*>(&s_mask[sb][lane][0]);  // rows >= k: zero
                const int c0 = __popc(m.x), c1 = __popc(m.y), c2 = __popc(m.z), c3 = __popc(m.w);
                const int tot = lane < p.k ? c0 + c1 + c2 + c3 : 0;
                const int tot_pad = (tot + kUpdBatch - 1) / kUpdBatch * kUpdBatch;  // runs start at multiples of the batch (one
                                                                                   // aligned load fetches its entries)
                int incl = tot_pad;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const int v = __shfl_up_sync(0xffffffffu, incl, o);
                    if (lane >= o) incl += v;
                }
                const int excl = incl - tot_pad;
                if (q == 0) s_base[sb][lane] = (unsigned short)(excl | (tot << 8));  // start | rows << 8
                const int mine = excl + (q > 0 ? c0 : 0) + (q > 1 ? c1 : 0) + (q > 2 ? c2 : 0);
                const unsigned int mq = q == 0 ? m.x : q == 1 ? m.y : q == 2 ? m.z : m.w;
                const int b = bin & 31;
                const int pos = __shfl_sync(0xffffffffu, mine, b) + __popc(__shfl_sync(0xffffffffu, mq, b) & ((1u << lane) - 1u));
                if (bin >= 0) s_perm[sb][pos] = perm_entry(row, p.tail_cols);
                if (lane < p.k) cnt_lane += (unsigned int)__popc(mq);
                __syncwarp();  // the warp's mask words precede its arrival; the update warps wait for all four
                if (et == 0) trace_ev(p, it, 6);
                if (lane == 0) mbar_arrive(s_full0 + 8u * sb);
            } else {
                __syncwarp();
                if (lane == 0) mbar_arrive(x_empty0 + 8u * s);
            }
            s += n_groups;  // this group's next tile: n_groups stages further round the ring
            while (s >= S) {
                s -= S;
                sph ^= 1u;
            }
        }
        if (p.acc) s_cntw[warp - kAssignWarp0][lane] = cnt_lane;
    } else if (p.acc && (warp >= kUpdateWarp0 || warp < kAssignWarp0)) {
        // ======================================== update ========================================
        // Eight warps; warp u owns the labels IterState::own[u*4..] (<= 4, balanced by cluster size after every
        // iteration) for the whole kernel; lane = float4 column chunk.  Their FP32 sums live in registers (two packed
        // pairs per label, FADD2) and are folded into the CTA's float64 shared accumulators every 8 tiles.  No other
        // thread touches an owner's sums: no atomics.  The code is kept small on purpose: the roles of this kernel
        // share a 32 KB instruction cache, and a miss costs an L2 round trip.
        const int u = warp >= kUpdateWarp0 ? warp - kUpdateWarp0 : 8 + warp - kNormWarp0;
        const int ut = u * 32 + lane;
        double* sacc = (double*)(smem + p.off_acc);
        const int nchunks = p.d >> 2;
        int lab[4];  // owned labels; kMaxKAcc = free slot (its mask row is all zero)
        {
            const bool table = p.st->own_bins == n_upd;  // else (never on b200km_fit's schedule): contiguous ranges
            const int lo = (u * p.k) / n_upd, hi = ((u + 1) * p.k) / n_upd;
#pragma unroll
            for (int li = 0; li < 4; ++li) {
                const int l = table ? (int)p.st->own[u * 4 + li] : (lo + li < hi ? lo + li : -1);
                lab[li] = l < 0 ? kMaxKAcc : l;
            }
        }
        const bool active = lane < nchunks;
        const int ch = active ? lane : 0;
        const int a = ch >> 3;
        const bool tail = a >= p.n_full;
        const int col_off = (tail ? p.n_full : a) * kAtomBytes;
        const int c16 = (tail ? (ch - p.n_full * 8) : (ch & 7)) << 4;
        // a list entry is a byte offset inside an atom (< 16 KB, 16-byte granular): XOR with this constant applies the
        // lane's swizzle and adds its atom's offset in one operation
        const unsigned int xk = (unsigned int)(col_off | c16);
        const int half_shift = tail ? 16 : 0;  // the half of a list entry this lane's column chunk needs
        unsigned long long acc[4][2];
#pragma unroll
        for (int li = 0; li < 4; ++li) acc[li][0] = acc[li][1] = 0ull;
        auto fold = [&]() {
#pragma unroll
            for (int li = 0; li < 4; ++li) {
                if (lab[li] < kMaxKAcc && active) {
                    double2* dst = reinterpret_cast<double2*>(sacc + (size_t)lab[li] * p.d + ch * 4);
                    const float2 lo = unpack2(acc[li][0]), hi = unpack2(acc[li][1]);
                    double2 a0 = dst[0], a1 = dst[1];
                    a0.x += (double)lo.x;
                    a0.y += (double)lo.y;
                    a1.x += (double)hi.x;
                    a1.y += (double)hi.y;
                    dst[0] = a0;
                    dst[1] = a1;
                }
                acc[li][0] = acc[li][1] = 0ull;
            }
        };
        int64_t it = 0;
        int s = 0;
        uint32_t sph = 0;
        for (int64_t tile = first_tile; tile < p.num_tiles; tile += tile_step, ++it) {
            const int sb = (int)(it & 1);
            mbar_wait(s_full0 + 8u * sb, (uint32_t)((it >> 1) & 1));
            if (ut == 0) trace_ev(p, it, 7);
            mbar_wait(x_full0 + 8u * s, sph);
            const uint8_t* stagep = smem + (size_t)s * stage_bytes;
            // Label by label, kUpdBatch list entries at a time; every label's run starts at a multiple of kUpdBatch.
#pragma unroll
            for (int li = 0; li < 4; ++li) {
                if (lab[li] == kMaxKAcc) continue;  // free slot (uniform)
#ifdef B200KM_ENABLE_TRACE
                if (p.skip & 1) continue;
#endif
                const unsigned int be = s_base[sb][lab[li]];  // start | rows << 8
                int left = (int)(be >> 8);
                // four list entries with one broadcast load (runs start at multiples of four entries); rows past the
                // end of the run are not loaded at all: shared-memory bandwidth is what the kernel runs out of
                const uint4* e4 = reinterpret_cast<const uint4*>(s_perm[sb] + (be & 255u));
                static_assert(kUpdBatch == 4, "the update warps fetch four list entries per load");
#pragma unroll 1
                for (; left > 0; left -= 4, ++e4) {
                    const uint4 ent = *e4;
                    const unsigned int w[4] = {ent.x, ent.y, ent.z, ent.w};
                    ulonglong2 v[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        v[j] = make_ulonglong2(0ull, 0ull);
                        if (j < left)
                            v[j] = *reinterpret_cast<const ulonglong2*>(stagep + (((w[j] >> half_shift) & 0xffffu) ^ xk));
                    }
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        add2(acc[li][0], v[j].x);
                        add2(acc[li][1], v[j].y);
                    }
                }
            }
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(s_empty0 + 8u * sb);
                mbar_arrive(x_empty0 + 8u * s);
            }
            if (lane == 0) trace_ev(p, it, 8 + u);
            if ((it & 7) == 7 || tile + tile_step >= p.num_tiles) fold();  // single (rare) fold site
            if (++s == S) {
                s = 0;
                sph ^= 1u;
            }
        }
        (void)ut;
    }

#ifdef B200KM_ENABLE_TRACE
    if (p.trace && threadIdx.x == 0) {  // per-CTA finish time (ns), slots past the per-tile events of CTA 0
        unsigned long long t;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
        p.trace[3900 + blockIdx.x] = t;
    }
#endif
    tc_fence_before();
    __syncthreads();
    if (p.acc) {
        // this CTA's partial sums go to its own slot: the slots are added in a fixed order afterwards
        // (slot_reduce_kernel), so the result does not depend on which CTA finishes first
        double* slot = p.slots + (size_t)blockIdx.x * p.slot_stride;
        const double* sacc = (const double*)(smem + p.off_acc);
        const int acc_n = p.k * p.d + p.k;
#pragma unroll 1
        for (int i = threadIdx.x; i < p.k * p.d; i += kThreads) slot[i] = sacc[i];
        if (threadIdx.x < p.k) {
            unsigned int c = 0;
            for (int w = 0; w < 8; ++w) c += s_cntw[w][threadIdx.x];
            slot[p.k * p.d + threadIdx.x] = (double)c;
        }
        if (threadIdx.x == 0) slot[acc_n] = 0.0;  // the inertia slot: this pass has no per-row distances
    }
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
    }
}

// ---- host side ---------------------------------------------------------------------------------
constexpr int kStaticSmem = 9 * 1024;  // upper bound of the kernel's static shared arrays

struct Plan {
    int k_pad, n_full, tail_cols, d_pad, num_stages, num_d, stage_bytes;
    int off_c, off_acc, off_cn, off_xn, off_pd, off_bar, total;
    bool ok;
};

Plan make_plan(int d, int k, bool with_acc) {
    Plan pl = {};
    pl.ok = false;
    if (d <= 0 || k <= 0 || d > kMaxD || k > kMaxK || (d & 3)) return pl;
    if (with_acc && (d > 128 || k > kMaxKAcc)) return pl;  // update warps: lane = float4 column chunk, <= 4 labels each
    pl.k_pad = (k + 15) & ~15;  // N of the MMA (M = 128: a multiple of 16); the assign warps read the real columns
    // K-atoms: 32-float SWIZZLE_128B atoms, and a narrow SWIZZLE_32B / 64B tail when d % 32 is in 4..16
    const int rem = d & 31;
    pl.n_full = d >> 5;
    pl.tail_cols = 0;
    if (rem > 16) pl.n_full += 1;
    else if (rem > 8) pl.tail_cols = 16;
    else if (rem > 0) pl.tail_cols = 8;
    pl.d_pad = pl.n_full * 32 + pl.tail_cols;
    pl.num_d = 4;  // score buffers of k_pad <= 128 TMEM columns each
    pl.stage_bytes = pl.n_full * kAtomBytes + kTileM * pl.tail_cols * 4;
    const int c_bytes = 2 * pl.k_pad * pl.d_pad * 4;
    const int pd_bytes = (32 * 32 + 3 * kMaxK) * 4;  // pair distances (k <= 32) + per-centroid neighbour data
    const int acc_bytes = with_acc ? (int)((sizeof(double) * ((size_t)k * d + k) + 15) & ~(size_t)15) : 0;
    const int fixed = c_bytes + pd_bytes + acc_bytes + pl.k_pad * 4 + 2 * kMaxStages * kTileM * 4 + 256 /*barriers*/ + 1024 /*align*/ +
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
    off += 2 * kMaxStages * kTileM * 4;  // ||x||^2 and ||x - x^||^2 per stage
    off = (off + 15) & ~15;
    pl.off_pd = off;
    off += pd_bytes;
    pl.off_bar = off;
    off += 256;
    pl.total = off + 1024;
    pl.ok = pl.total + kStaticSmem <= kMaxSmem;
    return pl;
}

}  // namespace

bool screen_supported(const float* X, int d, int k, bool with_acc) {
    if ((reinterpret_cast<uintptr_t>(X) & 15) != 0) return false;
    return make_plan(d, k, with_acc).ok;
}
size_t screen_workspace_bytes(int d, int k) {
    Plan pl = make_plan(d, k, true);
    return pl.ok ? sizeof(float) * (2 * (size_t)pl.k_pad * pl.d_pad + (size_t)k * k) : 0;
}
const char* screen_last_error() { return g_sc_err; }

#ifdef B200KM_ENABLE_TRACE
namespace {
struct KTime {  // development builds: B200KM_KTIME=1 prints the mean time of the pass's kernels at exit
    cudaEvent_t ev[64][4];
    int n = 0;
    bool on = getenv("B200KM_KTIME") != nullptr;
    ~KTime() {
        if (!on || n < 4) return;
        double t[3] = {0, 0, 0};
        int m = 0;
        for (int i = n / 2; i < n; ++i, ++m)
            for (int j = 0; j < 3; ++j) {
                float ms = 0;
                cudaEventElapsedTime(&ms, ev[i][j], ev[i][j + 1]);
                t[j] += ms;
            }
        fprintf(stderr, "KTIME over %d passes: prep %.4f  screen %.4f  fixup+reduce %.4f ms\n", m, t[0] / m, t[1] / m, t[2] / m);
    }
    void mark(int j, cudaStream_t s) {
        if (!on || n >= 64) return;
        if (j == 0) for (int q = 0; q < 4; ++q) cudaEventCreate(&ev[n][q]);
        cudaEventRecord(ev[n][j], s);
        if (j == 3) ++n;
    }
} g_ktime;
}
#define KT(j) g_ktime.mark(j, s)
#else
#define KT(j)
#endif

int screen_pass(TcWorkspace* ws, const float* X, int64_t n, int d, int k, const float* C, const float* cnorm,
                float* xnorm, float* xlo, unsigned int* info, bool xn_compute, int32_t* labels, double* acc,
                IterState* st, bool honor_done, cudaStream_t s) {
    const Plan pl = make_plan(d, k, acc != nullptr);
    if (!pl.ok) {
        snprintf(g_sc_err, sizeof(g_sc_err), "shape d=%d k=%d does not fit the tcgen05 screen kernel", d, k);
        return 1;
    }
    const size_t need = sizeof(float) * 2 * (size_t)pl.k_pad * pl.d_pad;
    if (need > ws->cap) {
        if (ws->d_buf) cudaFree(ws->d_buf);
        ws->d_buf = nullptr;
        ws->cap = 0;
        if (cudaMalloc(&ws->d_buf, need) != cudaSuccess) {
            snprintf(g_sc_err, sizeof(g_sc_err), "cudaMalloc of %zu bytes for the TF32 centroids failed", need);
            return 1;
        }
        ws->cap = need;
    }
    const size_t need_pd = sizeof(float) * (size_t)k * k;
    if (need_pd > ws->pd_cap) {
        if (ws->d_pd) cudaFree(ws->d_pd);
        ws->d_pd = nullptr;
        ws->pd_cap = 0;
        if (cudaMalloc(&ws->d_pd, need_pd) != cudaSuccess) {
            snprintf(g_sc_err, sizeof(g_sc_err), "cudaMalloc of %zu bytes for the centroid distance table failed", need_pd);
            return 1;
        }
        ws->pd_cap = need_pd;
    }
    const int64_t num_tiles = (n + kTileM - 1) / kTileM;
    const int acc_n = k * d + k;
    const int slot_stride = (acc_n + 1 + 7) & ~7;
    {
        const size_t need_bits = sizeof(unsigned int) * 4 * (size_t)(num_tiles + 1);
        if (need_bits > ws->bits_cap) {
            if (ws->d_bits) cudaFree(ws->d_bits);
            ws->d_bits = nullptr;
            ws->bits_cap = 0;
            if (cudaMalloc(&ws->d_bits, need_bits + need_bits / 8) != cudaSuccess) {
                snprintf(g_sc_err, sizeof(g_sc_err), "cudaMalloc of %zu bytes for the open-row bitmap failed", need_bits);
                return 1;
            }
            ws->bits_cap = need_bits + need_bits / 8;
        }
        const size_t need_slots = acc ? sizeof(double) * (size_t)slot_stride * (size_t)(kNumSMs + kNumSMs * kFixCtasPerSm) : 0;
        if (need_slots > ws->slots_cap) {
            if (ws->d_slots) cudaFree(ws->d_slots);
            ws->d_slots = nullptr;
            ws->slots_cap = 0;
            if (cudaMalloc(&ws->d_slots, need_slots) != cudaSuccess) {
                snprintf(g_sc_err, sizeof(g_sc_err), "cudaMalloc of %zu bytes for the partial-sum slots failed", need_slots);
                return 1;
            }
            ws->slots_cap = need_slots;
        }
    }
    CUtensorMap tmX, tmXt, tmC, tmCt;
    const uint32_t full_box = pl.n_full ? 32 : (uint32_t)pl.tail_cols;  // d < 32 with a narrow tail only
    const uint32_t tail_box = pl.tail_cols ? (uint32_t)pl.tail_cols : full_box;
    char* e = g_sc_err;
    const size_t el = sizeof(g_sc_err);
    if (encode_2d(e, el, &tmX, X, (uint64_t)d, (uint64_t)n, (uint64_t)d * 4, full_box, kTileM)) return 1;
    if (encode_2d(e, el, &tmXt, X, (uint64_t)d, (uint64_t)n, (uint64_t)d * 4, tail_box, kTileM)) return 1;
    if (encode_2d(e, el, &tmC, ws->d_buf, (uint64_t)pl.d_pad, (uint64_t)2 * pl.k_pad, (uint64_t)pl.d_pad * 4, full_box,
                  (uint32_t)pl.k_pad))
        return 1;
    if (encode_2d(e, el, &tmCt, ws->d_buf, (uint64_t)pl.d_pad, (uint64_t)2 * pl.k_pad, (uint64_t)pl.d_pad * 4, tail_box,
                  (uint32_t)pl.k_pad))
        return 1;

    static PerDeviceOnce attr_set;  // the opt-in is a per-device attribute
    const int dev = PerDeviceOnce::current();
    if (dev < 0 || !attr_set.done[dev]) {
        cudaFuncAttributes fa;
        if (cudaFuncGetAttributes(&fa, kmeans_screen_kernel) != cudaSuccess || (int)fa.sharedSizeBytes > kStaticSmem) {
            snprintf(g_sc_err, sizeof(g_sc_err), "unexpected static shared memory size of the screen kernel");
            return 1;
        }
        const int max_dyn = kMaxSmem - (int)fa.sharedSizeBytes;
        if (cudaFuncSetAttribute(kmeans_screen_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, max_dyn) !=
            cudaSuccess) {
            snprintf(g_sc_err, sizeof(g_sc_err), "cannot raise dynamic shared memory to %d bytes", max_dyn);
            return 1;
        }
        if (dev >= 0) attr_set.done[dev] = true;
    }
    {
        int split_blocks = (pl.k_pad * pl.d_pad + 127) / 128;
        if (split_blocks > kNumSMs) split_blocks = kNumSMs;
        KT(0);
        screen_prep_kernel<<<split_blocks + (k * k + 3) / 4, 128, 0, s>>>(C, k, d, pl.k_pad, pl.d_pad, split_blocks,
                                                                         (float*)ws->d_buf, (float*)ws->d_pd);
        count_launch();
    }
    ScParams p;
    p.n = n;
    p.num_tiles = num_tiles;
    p.d = d;
    p.k = k;
    p.k_pad = pl.k_pad;
    p.n_full = pl.n_full;
    p.tail_cols = pl.tail_cols;
    p.num_stages = pl.num_stages;
    p.num_d = pl.num_d;
    p.stacked = (2 * pl.k_pad * kMaxND <= 512) ? 1 : 0;
    p.stage_bytes = pl.stage_bytes;
    p.X = X;
    p.prefetch = 0;
    p.skip = 0;
#ifdef B200KM_ENABLE_TRACE
    if (const char* e = getenv("B200KM_PREFETCH")) p.prefetch = atoi(e);
    if (const char* e = getenv("B200KM_SKIP")) p.skip = atoi(e);
#endif
    p.C = C;
    p.cnorm = cnorm;
    p.xnorm = xnorm;
    p.xlo = xlo;
    p.xn_compute = xn_compute ? 1 : 0;
    p.labels = labels;
    p.info = info;
    p.open_bits = (unsigned int*)ws->d_bits;
    p.acc = acc;
    p.slots = (double*)ws->d_slots;
    p.slot_stride = slot_stride;
    p.st = st;
    p.honor_done = honor_done ? 1 : 0;
    // Error model (file header).  x is read truncated to TF32; ||x - x^||^2 is computed per row with the row
    // norms (tol_x = slack for its FP32 rounding and the square root), and ||x - x^|| <= 2^-10 ||x|| (abs_rel).  Everything else is FP32 rounding: the tensor core adds
    // each MMA's products into the FP32 accumulator with truncation (<= 1 ulp of the running sum per MMA,
    // ceil(d/8) MMAs), the c_hi/c_lo split leaves 2^-22 |c|, and ||c||^2, ||x||^2 and the final FMAs round
    // once each.  With 2|x||c| <= |x|^2+|c|^2 a margin (difference of two D_j) carries at most
    // eps_rel * (||x||^2 + max||c||^2) of it; one D_j alone is off by at most abs_rel * (same scale).
    p.pd = (const float*)ws->d_pd;
    p.tol_x = 1.002f;
    p.eps_rel = 2.f * (float)((d + 7) / 8 + 7) * 1.1920928955078125e-07f + 1e-6f;
    p.abs_rel = 9.765625e-4f * 1.001f + p.eps_rel;
    p.off_c = pl.off_c;
    p.off_acc = pl.off_acc;
    p.off_cn = pl.off_cn;
    p.off_xn = pl.off_xn;
    p.off_pd = pl.off_pd;
    p.off_bar = pl.off_bar;
    p.trace = nullptr;
#ifdef B200KM_ENABLE_TRACE  // development builds only (B200KM_ENABLE_TRACE=1 python -m ...build)
    const char* trace_path = getenv("B200KM_TRACE");
    if (getenv("B200KM_NO_STACK")) p.stacked = 0;
    if (trace_path && acc) {
        cudaMalloc((void**)&p.trace, 4096 * sizeof(unsigned long long));
        cudaMemsetAsync(p.trace, 0, 4096 * sizeof(unsigned long long), s);
    }
#endif
    int64_t grid = p.num_tiles < kNumSMs ? p.num_tiles : kNumSMs;
    if (grid < 1) grid = 1;
    KT(1);
    kmeans_screen_kernel<<<(unsigned)grid, kThreads, pl.total, s>>>(tmX, tmXt, tmC, tmCt, p);
    KT(2);
    count_launch();
    cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess) {
        snprintf(g_sc_err, sizeof(g_sc_err), "screen kernel launch failed: %s", cudaGetErrorString(err));
        return 1;
    }
    if (screen_fixup(ws, X, n, d, k, C, info, num_tiles, labels, acc, (int)grid, slot_stride, st, honor_done, s)) {
        snprintf(g_sc_err, sizeof(g_sc_err), "exact pass of the screen kernel failed: %s", fixup_last_error());
        return 1;
    }
    KT(3);
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
