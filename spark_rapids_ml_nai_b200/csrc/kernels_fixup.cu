// Exact pass of the tcgen05 screen kernel (kernels_screen.cu): decides the rows the TF32 scores left open - FP32
// squared differences, float64 when even those tie, lowest index on exact ties - and adds them to the sums.
//
// The open rows (C2: ~1 % of the rows, scattered) come out of the bitmap as ONE list in row order:
//   open_count  popcount of every block of 2048 bitmap words
//   open_list   every block adds up the counts before it (and the total), then writes (row index, verdict word) of
//               its open rows at their positions; beyond 4096 blocks open_scan does the prefix in a launch of its own
// and persistent CTAs walk the list 64 rows at a time in a fixed assignment (CTA c takes batches c, c + grid, ...):
// the batch's rows are fetched into shared memory all at once (one HBM latency for 64 scattered rows instead of one
// per row), decided there by groups of 16 lanes, and added - thread t owns column t, rows in list order - into the
// CTA's float64 partial, which goes to the CTA's slot.  No atomics touch the sums, nothing depends on the order in
// which anything finishes: the result is bitwise reproducible.  slot_reduce_kernel then adds all slots (the screen
// kernel's and these) per element in slot order.
//
// When more rows are open than the list holds (structureless data; b200km_fit then leaves the screen kernel anyway)
// the tile-range kernel below does the same work straight from the bitmap.
//
// Replaces, together with the screen kernel, cuML's fusedL2NN + reduce_rows_by_key pair behind reference
// python/src/spark_rapids_ml/clustering.py:367-370 (fit) and :493-497 (predict).
#include "tc.cuh"

#include <cfloat>
#include <climits>
#include <cstdio>

namespace b200km {

namespace {

thread_local char g_fx_err[256] = "";

// The exact decisions are made by groups of kSubW lanes (two rows per warp at a time: the loads of a row are
// independent, and the phase is bound by the latency of HBM, not by arithmetic).  sub = lane within the group,
// gmask = the group's lanes.
#ifndef B200KM_FIX_SUBW
#define B200KM_FIX_SUBW 16
#endif
constexpr int kSubW = B200KM_FIX_SUBW;
__device__ __forceinline__ float sqdist_f32(const float4* __restrict__ x4, const float4* __restrict__ c4, int nchunks,
                                            int sub, unsigned gmask) {
    float s = 0.f;
    for (int ch = sub; ch < nchunks; ch += kSubW) {
        const float4 xv = x4[ch], cv = c4[ch];
        const float d0 = xv.x - cv.x, d1 = xv.y - cv.y, d2 = xv.z - cv.z, d3 = xv.w - cv.w;
        s += fmaf(d0, d0, d1 * d1) + fmaf(d2, d2, d3 * d3);
    }
#pragma unroll
    for (int o = kSubW / 2; o > 0; o >>= 1) s += __shfl_xor_sync(gmask, s, o);
    return s;
}
__device__ __forceinline__ double sqdist_f64(const float* __restrict__ x, const float* __restrict__ c, int d, int sub,
                                             unsigned gmask) {
    double s = 0.0;
    for (int i = sub; i < d; i += kSubW) {
        const double df = (double)x[i] - (double)c[i];
        s = fma(df, df, s);
    }
#pragma unroll
    for (int o = kSubW / 2; o > 0; o >>= 1) s += __shfl_xor_sync(gmask, s, o);
    return s;
}

// Squared distances of one row to two centroids in one sweep (the row is read once).
__device__ __forceinline__ void sqdist2_f32(const float4* __restrict__ x4, const float4* __restrict__ a4,
                                            const float4* __restrict__ b4, int nchunks, int sub, unsigned gmask,
                                            float& sa, float& sb) {
    float ta = 0.f, tb = 0.f;
    for (int ch = sub; ch < nchunks; ch += kSubW) {
        const float4 xv = x4[ch], av = a4[ch], bv = b4[ch];
        const float a0 = xv.x - av.x, a1 = xv.y - av.y, a2 = xv.z - av.z, a3 = xv.w - av.w;
        const float b0 = xv.x - bv.x, b1 = xv.y - bv.y, b2 = xv.z - bv.z, b3 = xv.w - bv.w;
        ta += fmaf(a0, a0, a1 * a1) + fmaf(a2, a2, a3 * a3);
        tb += fmaf(b0, b0, b1 * b1) + fmaf(b2, b2, b3 * b3);
    }
#pragma unroll
    for (int o = kSubW / 2; o > 0; o >>= 1) {
        ta += __shfl_xor_sync(gmask, ta, o);
        tb += __shfl_xor_sync(gmask, tb, o);
    }
    sa = ta;
    sb = tb;
}

// Exact label of one open row (one group of kSubW lanes): FP32 squared differences, float64 when even those tie; ties
// keep the lowest index.  word = the row's verdict (computed best, runner-up, which centroids to look at).
__device__ __forceinline__ int exact_label(const float* __restrict__ x, unsigned int word, int d, int k,
                                           const float* __restrict__ C, const float* __restrict__ pd, int sub,
                                           unsigned gmask) {
    const int nchunks = d >> 2;
    const int i1 = (int)(word & 255u), i2 = (int)((word >> 8) & 255u);
    const unsigned int mode = word & (3u << 16);
    const float4* x4 = reinterpret_cast<const float4*>(x);
    if (mode == kInfoPair) {
        const int ja = min(i1, i2), jb = max(i1, i2);
        float sa, sb;
        sqdist2_f32(x4, reinterpret_cast<const float4*>(C + (size_t)ja * d), reinterpret_cast<const float4*>(C + (size_t)jb * d),
                    nchunks, sub, gmask, sa, sb);
        int bi = sb < sa ? jb : ja;
        if (fabsf(sa - sb) <= 4e-6f * (sa + sb)) {
            const double da = sqdist_f64(x, C + (size_t)ja * d, d, sub, gmask);
            const double db = sqdist_f64(x, C + (size_t)jb * d, d, sub, gmask);
            bi = db < da ? jb : ja;
        }
        return bi;
    }
    // every centroid, or those that can beat the computed best c_1 at all: ||c_1 - c_j|| <= 2 ||x - c_1||.  The
    // group's lanes test kSubW centroids at a time; only the candidates (usually one or two) are looked at.
    const float s1 = sqdist_f32(x4, reinterpret_cast<const float4*>(C + (size_t)i1 * d), nchunks, sub, gmask);
    const float reach = mode == kInfoReach ? 2.f * sqrtf(s1) * 1.0001f : FLT_MAX;
    const int gshift = (int)(threadIdx.x & 31) & ~(kSubW - 1);
    constexpr unsigned kSubMask = kSubW == 32 ? 0xffffffffu : (1u << (kSubW & 31)) - 1u;
    float f1 = s1, f2 = FLT_MAX;  // smallest (with its index bi: the lowest index among equals) and second smallest
    int bi = i1;
    for (int jb = 0; jb < k; jb += kSubW) {
        const int j = jb + sub;
        const bool cand = j < k && j != i1 && (reach == FLT_MAX || pd[(size_t)i1 * k + j] <= reach);
        unsigned int mask = (__ballot_sync(gmask, cand) >> gshift) & kSubMask;
        while (mask) {
            const int jj = jb + __ffs(mask) - 1;
            mask &= mask - 1;
            const float sum = sqdist_f32(x4, reinterpret_cast<const float4*>(C + (size_t)jj * d), nchunks, sub, gmask);
            if (sum < f1 || (sum == f1 && jj < bi)) {
                f2 = f1;
                f1 = sum;
                bi = jj;
            } else {
                f2 = fminf(f2, sum);
            }
        }
    }
    if (f2 < FLT_MAX && (f2 - f1) <= 4e-6f * (f1 + f2)) {
        double bd = sqdist_f64(x, C + (size_t)i1 * d, d, sub, gmask);
        bi = i1;
        for (int jb = 0; jb < k; jb += kSubW) {
            const int j = jb + sub;
            const bool cand = j < k && j != i1 && (reach == FLT_MAX || pd[(size_t)i1 * k + j] <= reach);
            unsigned int mask = (__ballot_sync(gmask, cand) >> gshift) & kSubMask;
            while (mask) {
                const int jj = jb + __ffs(mask) - 1;
                mask &= mask - 1;
                const double sum = sqdist_f64(x, C + (size_t)jj * d, d, sub, gmask);
                if (sum < bd || (sum == bd && jj < bi)) {
                    bd = sum;
                    bi = jj;
                }
            }
        }
    }
    return bi;
}

// ---- the open rows as a list ------------------------------------------------------------------------------------
constexpr int kWordsPerBlock = 2048;  // bitmap words per block of the list kernels (256 threads x 8 words)

__global__ void __launch_bounds__(256) open_count_kernel(const unsigned int* __restrict__ bits, int64_t n_words,
                                                         unsigned int* __restrict__ blk) {
    __shared__ unsigned int red[8];
    const int64_t w0 = (int64_t)blockIdx.x * kWordsPerBlock + (int64_t)threadIdx.x * 8;
    unsigned int c = 0;
    if (w0 + 8 <= n_words) {
        const uint4 a = *reinterpret_cast<const uint4*>(bits + w0), b = *reinterpret_cast<const uint4*>(bits + w0 + 4);
        c = __popc(a.x) + __popc(a.y) + __popc(a.z) + __popc(a.w) + __popc(b.x) + __popc(b.y) + __popc(b.z) + __popc(b.w);
    } else {
        for (int i = 0; i < 8; ++i)
            if (w0 + i < n_words) c += __popc(bits[w0 + i]);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) blk[blockIdx.x] = red[0] + red[1] + red[2] + red[3] + red[4] + red[5] + red[6] + red[7];
}

// blk[b] <- number of open rows before block b; *total <- all of them (also counted in the iteration state)
__global__ void __launch_bounds__(1024) open_scan_kernel(unsigned int* __restrict__ blk, int n_blk,
                                                         unsigned long long* __restrict__ total, IterState* st,
                                                         int honor_done) {
    __shared__ unsigned int s_warp[32];
    __shared__ unsigned long long s_carry;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (honor_done && st->done) {  // the screen kernel did not run: the bitmap is the previous pass's
        if (tid == 0) *total = 0ull;
        return;
    }
    if (tid == 0) s_carry = 0ull;
    __syncthreads();
    for (int base = 0; base < n_blk; base += 1024) {
        const int i = base + tid;
        const unsigned int v = i < n_blk ? blk[i] : 0u;
        unsigned int incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned int t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            unsigned int w = s_warp[lane], wi = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned int t = __shfl_up_sync(0xffffffffu, wi, o);
                if (lane >= o) wi += t;
            }
            s_warp[lane] = wi - w;  // exclusive
        }
        __syncthreads();
        const unsigned long long carry = s_carry;
        // the list positions are 32-bit: the list itself is capped far below that (open_cap)
        if (i < n_blk) blk[i] = (unsigned int)min(carry + s_warp[warp] + incl - v, 0xffffffffull);
        __syncthreads();
        if (tid == 1023) s_carry = carry + s_warp[warp] + incl;
        __syncthreads();
    }
    if (tid == 0) {
        *total = s_carry;
        st->screen_refined += (unsigned int)min(s_carry, 0xffffffffull);
    }
}

// scanned: blk holds the exclusive prefixes and *total is set (open_scan_kernel ran: very many blocks); otherwise blk
// holds the block counts and every block adds up its own prefix and the total (a few hundred values).
__global__ void __launch_bounds__(256) open_list_kernel(const unsigned int* __restrict__ bits, int64_t n_words,
                                                        const unsigned int* __restrict__ blk, int n_blk, int scanned,
                                                        const unsigned int* __restrict__ info,
                                                        unsigned long long* __restrict__ total, unsigned long long cap,
                                                        long long* __restrict__ rows, unsigned int* __restrict__ words,
                                                        IterState* st, int honor_done) {
    __shared__ unsigned int s_warp[8];
    __shared__ unsigned long long s_sum[2][8];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned long long before, all;
    if (scanned) {
        all = *total;
        before = blk[blockIdx.x];
    } else {
        if (honor_done && st->done) {  // the screen kernel did not run: the bitmap is the previous pass's
            if (blockIdx.x == 0 && tid == 0) *total = 0ull;
            return;
        }
        unsigned long long b = 0ull, a = 0ull;
        for (int j = tid; j < n_blk; j += 256) {
            const unsigned int v = blk[j];
            a += v;
            if (j < (int)blockIdx.x) b += v;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            a += __shfl_xor_sync(0xffffffffu, a, o);
            b += __shfl_xor_sync(0xffffffffu, b, o);
        }
        if (lane == 0) {
            s_sum[0][warp] = a;
            s_sum[1][warp] = b;
        }
        __syncthreads();
        all = 0ull;
        before = 0ull;
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            all += s_sum[0][q];
            before += s_sum[1][q];
        }
        if (blockIdx.x == 0 && tid == 0) {
            *total = all;
            st->screen_refined += (unsigned int)min(all, 0xffffffffull);
        }
    }
    if (all > cap) return;  // the tile-range kernel does the work
    const int64_t w0 = (int64_t)blockIdx.x * kWordsPerBlock + (int64_t)tid * 8;
    unsigned int w[8];
    unsigned int c = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        w[i] = w0 + i < n_words ? bits[w0 + i] : 0u;
        c += __popc(w[i]);
    }
    unsigned int incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    unsigned int off = (unsigned int)before + incl - c;
    for (int q = 0; q < warp; ++q) off += s_warp[q];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        unsigned int m = w[i];
        while (m) {
            const int b = __ffs(m) - 1;
            m &= m - 1;
            const long long row = (w0 + i) * 32 + b;  // 4 words of 32 rows per 128-row tile: bit index = row index
            rows[off] = row;
            words[off] = info[row];
            ++off;
        }
    }
}

// Persistent CTAs over the list, kFixRows rows at a time.
constexpr int kFixRows = 64;
#ifndef B200KM_FIX_MINB
#define B200KM_FIX_MINB 4
#endif
__global__ void __launch_bounds__(256, B200KM_FIX_MINB) fixup_list_kernel(const float* __restrict__ X, int d, int k, const float* C,
                                                         const float* pd, int stage_c,
                                                         const long long* __restrict__ rows,
                                                         const unsigned int* __restrict__ words,
                                                         const unsigned long long* __restrict__ total,
                                                         unsigned long long cap, int rows_per_batch,
                                                         int32_t* __restrict__ labels, double* __restrict__ slots,
                                                         int slot_stride, IterState* st, int honor_done) {
    extern __shared__ __align__(16) unsigned char f_smem[];
    __shared__ long long s_row[kFixRows];
    __shared__ unsigned int s_word[kFixRows];
    __shared__ unsigned char s_lab[kFixRows];
    if ((honor_done & 1) && st->done) return;
    const unsigned long long m_all = *total;
    if (m_all > cap) return;  // the tile-range kernel does the work
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int acc_n = k * d + k;
    float* xs = reinterpret_cast<float*>(f_smem);                                  // [rows_per_batch][d]
    double* f_acc = reinterpret_cast<double*>(f_smem + (size_t)rows_per_batch * d * sizeof(float));  // [k*d + k]
    if (slots)
        for (int i = tid; i < acc_n; i += 256) f_acc[i] = 0.0;
    if (stage_c) {  // the centroids and their distance table next to the partial sums (small k*d)
        float* sC = reinterpret_cast<float*>(f_acc + (slots ? ((acc_n + 1) & ~1) : 0));
        float* sP = sC + k * d;
        for (int i = tid; i < k * d; i += 256) sC[i] = C[i];
        for (int i = tid; i < k * k; i += 256) sP[i] = pd[i];
        C = sC;
        pd = sP;
    }
    const int nch = d >> 2;
    const long long n_batches = (long long)((m_all + rows_per_batch - 1) / rows_per_batch);
    for (long long b = blockIdx.x; b < n_batches; b += gridDim.x) {
        const long long e0 = b * rows_per_batch;
        const int m = (int)min((unsigned long long)rows_per_batch, m_all - (unsigned long long)e0);
        __syncthreads();  // the previous batch is no longer read
        if (tid < m) {
            s_row[tid] = rows[e0 + tid];
            s_word[tid] = words[e0 + tid];
        }
        __syncthreads();
        // ---- the batch's rows into shared memory, all loads in flight together ----
        // (a warp takes rows warp, warp + 8, ...: eight independent loads per lane before the first store)
#ifdef B200KM_ENABLE_TRACE
        if (!(honor_done & 8))
#endif
        for (int c = lane; c < nch; c += 32) {
            for (int r0 = warp; r0 < m; r0 += 64) {
                float4 v[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int r = r0 + 8 * j;
                    if (r < m) v[j] = __ldg(reinterpret_cast<const float4*>(X + (size_t)s_row[r] * d) + c);
                }
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int r = r0 + 8 * j;
                    if (r < m) reinterpret_cast<float4*>(xs + (size_t)r * d)[c] = v[j];
                }
            }
        }
        __syncthreads();
        // ---- exact labels: one group of kSubW lanes per row ----
#ifdef B200KM_ENABLE_TRACE
        if (honor_done & 2) {
            if (tid < m) s_lab[tid] = (unsigned char)(s_word[tid] & 255u);
        } else
#endif
        {
            const int sub = lane & (kSubW - 1);
            const unsigned gmask = kSubW == 32 ? 0xffffffffu : ((1u << (kSubW & 31)) - 1u) << (lane & ~(kSubW - 1));
            for (int i = tid / kSubW; i < m; i += 256 / kSubW) {
                const unsigned int e = s_word[i];
                const int bi = exact_label(xs + (size_t)i * d, e & 0x3FFFFu, d, k, C, pd, sub, gmask);
                if (sub == 0) {
                    s_lab[i] = (unsigned char)bi;
                    if (labels && bi != (int)(e & 255u)) labels[s_row[i]] = bi;
                }
            }
        }
        __syncthreads();
        // ---- into the partial sums, in list order: thread t owns column t (and t + 256), the last warp counts ----
#ifdef B200KM_ENABLE_TRACE
        if (honor_done & 4) continue;
#endif
        if (slots) {
            // thread t owns columns 2t, 2t + 1 (d is a multiple of 4); two rows at a time: different labels are
            // independent updates, equal labels add the two rows first
            for (int col = 2 * tid; col < d; col += 512) {
                int i = 0;
                for (; i + 1 < m; i += 2) {
                    const int la = s_lab[i], lb = s_lab[i + 1];
                    const float2 fa = *reinterpret_cast<const float2*>(xs + (size_t)i * d + col);
                    const float2 fb = *reinterpret_cast<const float2*>(xs + (size_t)(i + 1) * d + col);
                    double2* pa = reinterpret_cast<double2*>(f_acc + la * d + col);
                    double2* pb = reinterpret_cast<double2*>(f_acc + lb * d + col);
                    if (la == lb) {
                        double2 v = *pa;
                        v.x += (double)fa.x + (double)fb.x;
                        v.y += (double)fa.y + (double)fb.y;
                        *pa = v;
                    } else {
                        double2 va = *pa, vb = *pb;
                        va.x += (double)fa.x;
                        va.y += (double)fa.y;
                        vb.x += (double)fb.x;
                        vb.y += (double)fb.y;
                        *pa = va;
                        *pb = vb;
                    }
                }
                if (i < m) {
                    const float2 fa = *reinterpret_cast<const float2*>(xs + (size_t)i * d + col);
                    double2* pa = reinterpret_cast<double2*>(f_acc + (int)s_lab[i] * d + col);
                    double2 v = *pa;
                    v.x += (double)fa.x;
                    v.y += (double)fa.y;
                    *pa = v;
                }
            }
            if (tid >= 224 && tid - 224 < 32) {
                for (int j = tid - 224; j < k; j += 32) {
                    int c = 0;
                    for (int i = 0; i < m; ++i) c += (s_lab[i] == j);
                    f_acc[k * d + j] += (double)c;
                }
            }
        }
    }
    if (slots) {
        __syncthreads();
        double* slot = slots + (size_t)blockIdx.x * slot_stride;
        for (int i = tid; i < acc_n; i += 256) slot[i] = f_acc[i];
        if (tid == 0) slot[acc_n] = 0.0;
    }
}

// ---- the tile-range kernel (more open rows than the list holds) -------------------------------------------------
// Exact decision of the rows the screen left open, straight from HBM, and their contribution to the sums.
// Every CTA owns a contiguous range of tiles and walks it in order, kFixTiles tiles at a time: the open rows come
// from the bitmap (ascending row order), the warps decide them exactly, then thread t adds column t of every such
// row - in list order - into the CTA's own float64 partial, which goes to the CTA's slot.  No atomics: the sums
// do not depend on the order in which anything finishes.  These rows are few (C2: ~1 %), and keeping this code
// out of the persistent kernel keeps that kernel's instruction stream hot.
constexpr int kFixTiles = 32;
__global__ void __launch_bounds__(256) screen_fixup_kernel(const float* __restrict__ X, int64_t n, int d, int k,
                                                           const float* __restrict__ C, const float* __restrict__ pd,
                                                           const unsigned int* __restrict__ info,
                                                           const unsigned int* __restrict__ bits, int64_t num_tiles,
                                                           int32_t* __restrict__ labels, double* __restrict__ slots,
                                                           int slot_stride, IterState* st, int honor_done,
                                                           const unsigned long long* __restrict__ open_total,
                                                           unsigned long long open_cap) {
    extern __shared__ double f_acc[];  // [k*d + k] when slots != nullptr
    // per open row of the chunk: row offset inside the chunk (bits 18+) | the low 18 bits of its verdict word
    __shared__ unsigned int s_ent[kFixTiles * kTileM];
    __shared__ unsigned char s_lab[kFixTiles * kTileM];
    __shared__ int s_warp[4];
    if ((honor_done & 1) && st->done) return;
    if (*open_total <= open_cap) return;  // the list-based kernel has done the work
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int acc_n = k * d + k;
    if (slots)
        for (int i = tid; i < acc_n; i += 256) f_acc[i] = 0.0;
    const int64_t t0 = num_tiles * (int64_t)blockIdx.x / gridDim.x;
    const int64_t t1 = num_tiles * (int64_t)(blockIdx.x + 1) / gridDim.x;
    for (int64_t chunk = t0; chunk < t1; chunk += kFixTiles) {
        const int nt = (int)((t1 - chunk) < kFixTiles ? (t1 - chunk) : kFixTiles);
        const float* Xc = X + (size_t)chunk * kTileM * d;
        // ---- the chunk's open rows, in ascending order (their verdict words are fetched here, all at once) ----
        unsigned int w = 0;
        if (tid < nt * 4) w = bits[chunk * 4 + tid];
        int incl = __popc(w);
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        __syncthreads();  // the previous chunk's lists are no longer read
        if (warp < 4 && lane == 31) s_warp[warp] = incl;
        __syncthreads();
        int off = incl - __popc(w);
        for (int q = 0; q < warp && q < 4; ++q) off += s_warp[q];
        const int m = s_warp[0] + s_warp[1] + s_warp[2] + s_warp[3];
        while (w) {
            const int bpos = __ffs(w) - 1;
            w &= w - 1;
            const unsigned int roff = (unsigned int)(tid * 32 + bpos);
            s_ent[off++] = (roff << 18) | (info[(size_t)chunk * kTileM + roff] & 0x3FFFFu);
        }
        __syncthreads();
        // ---- exact labels: one group of kSubW lanes per row ----
        {
            const int sub = lane & (kSubW - 1);
            const unsigned gmask = kSubW == 32 ? 0xffffffffu : ((1u << (kSubW & 31)) - 1u) << (lane & ~(kSubW - 1));
            for (int i = tid / kSubW; i < m; i += 256 / kSubW) {
                const unsigned int e = s_ent[i];
                const unsigned int roff = e >> 18;
                const int bi = exact_label(Xc + (size_t)roff * d, e & 0x3FFFFu, d, k, C, pd, sub, gmask);
                if (sub == 0) {
                    s_lab[i] = (unsigned char)bi;
                    if (labels && bi != (int)(e & 255u)) labels[(size_t)chunk * kTileM + roff] = bi;
                }
            }
        }
        __syncthreads();
        // ---- their rows into the partial sums, in list order: thread t owns column t, threads 128.. count ----
        if (slots) {
            if (tid < d) {
                for (int i = 0; i < m; i += 8) {
                    float v[8];
                    int l[8];
#pragma unroll
                    for (int u = 0; u < 8; ++u) {
                        const int ii = i + u < m ? i + u : m - 1;
                        l[u] = s_lab[ii];
                        v[u] = __ldg(Xc + (size_t)(s_ent[ii] >> 18) * d + tid);
                    }
#pragma unroll
                    for (int u = 0; u < 8; ++u)
                        if (i + u < m) f_acc[l[u] * d + tid] += (double)v[u];
                }
            } else if (tid >= 128 && tid - 128 < k) {
                int c = 0;
                for (int i = 0; i < m; ++i) c += (s_lab[i] == tid - 128);
                f_acc[k * d + tid - 128] += (double)c;
            }
        }
    }
    if (slots) {
        __syncthreads();
        double* slot = slots + (size_t)blockIdx.x * slot_stride;
        for (int i = tid; i < acc_n; i += 256) slot[i] = f_acc[i];
        if (tid == 0) slot[acc_n] = 0.0;
    }
}

// acc[i] += sum over the slots of slot[i] in a fixed order.  A block takes 32 consecutive elements (a warp reads
// 256 contiguous bytes of a slot); warp g adds the slots g, g + 8, ... in ascending order and the eight partial sums
// of an element are added in warp order.
__global__ void __launch_bounds__(256) slot_reduce_kernel(const double* __restrict__ slots, int nslots, int slot_stride,
                                                          int len, double* __restrict__ acc, IterState* st, int honor_done) {
    __shared__ double part[8][32];
    if (honor_done && st->done) return;
    const int e = threadIdx.x & 31, g = threadIdx.x >> 5;
    const int i = blockIdx.x * 32 + e;
    double t = 0.0;
    if (i < len) {
        const double* src = slots + i;
#pragma unroll 8
        for (int sidx = g; sidx < nslots; sidx += 8) t += src[(size_t)sidx * slot_stride];
    }
    part[g][e] = t;
    __syncthreads();
    if (g == 0 && i < len) {
        double sum = part[0][e];
#pragma unroll
        for (int q = 1; q < 8; ++q) sum += part[q][e];
        acc[i] += sum;
    }
}

}  // namespace

const char* fixup_last_error() { return g_fx_err; }

int screen_fixup(TcWorkspace* ws, const float* X, int64_t n, int d, int k, const float* C, const unsigned int* info,
                 int64_t num_tiles, int32_t* labels, double* acc, int screen_slots, int slot_stride, IterState* st,
                 bool honor_done, cudaStream_t s) {
    const int64_t n_words = num_tiles * 4;
    const int n_blk = (int)((n_words + kWordsPerBlock - 1) / kWordsPerBlock);
    // the list holds up to 1/16 of the rows: beyond that the screen has nothing to work with anyway
    const size_t cap = (size_t)(n / 16) + 65536;
    if (cap > ws->open_cap) {
        if (ws->d_open_rows) cudaFree(ws->d_open_rows);
        if (ws->d_open_info) cudaFree(ws->d_open_info);
        ws->d_open_rows = ws->d_open_info = nullptr;
        ws->open_cap = 0;
        if (cudaMalloc(&ws->d_open_rows, cap * sizeof(long long)) != cudaSuccess ||
            cudaMalloc(&ws->d_open_info, cap * sizeof(unsigned int)) != cudaSuccess) {
            snprintf(g_fx_err, sizeof(g_fx_err), "cudaMalloc of %zu bytes for the list of open rows failed", cap * 12);
            return 1;
        }
        ws->open_cap = cap;
    }
    if ((size_t)n_blk + 2 > ws->open_blk_cap) {
        if (ws->d_open_blk) cudaFree(ws->d_open_blk);
        ws->d_open_blk = nullptr;
        ws->open_blk_cap = 0;
        // [n_blk] block counts / prefixes, then (8-byte aligned) the total
        const size_t bytes = (((size_t)n_blk + 2) * sizeof(unsigned int) + 15) / 16 * 16 + 16;
        if (cudaMalloc(&ws->d_open_blk, bytes) != cudaSuccess) {
            snprintf(g_fx_err, sizeof(g_fx_err), "cudaMalloc of %zu bytes for the open-row counts failed", bytes);
            return 1;
        }
        ws->open_blk_cap = (size_t)n_blk + 2;
    }
    unsigned int* blk = (unsigned int*)ws->d_open_blk;
    unsigned long long* total =
        (unsigned long long*)((char*)ws->d_open_blk + (((size_t)ws->open_blk_cap) * sizeof(unsigned int) + 15) / 16 * 16);
    const unsigned int* bits = (const unsigned int*)ws->d_bits;
    const float* pd = (const float*)ws->d_pd;
    int hd = honor_done ? 1 : 0;
#ifdef B200KM_ENABLE_TRACE
    if (const char* e = getenv("B200KM_FIXSKIP")) hd |= atoi(e) << 1;
#endif
    // CTAs (= partial-sum slots) of the exact pass: all that fit for a large shard, fewer for a small one (every slot
    // costs a write here and a read in the slot reduction; a small shard has few open rows)
    int fix_grid = (int)(n / 16384 < kNumSMs * kFixCtasPerSm ? n / 16384 : kNumSMs * kFixCtasPerSm);
    if (fix_grid < kNumSMs) fix_grid = kNumSMs;
    double* fix_slots = acc ? (double*)ws->d_slots + (size_t)screen_slots * slot_stride : nullptr;
    const int acc_n = k * d + k;

    open_count_kernel<<<n_blk, 256, 0, s>>>(bits, n_words, blk);
    count_launch();
    const int scanned = n_blk > 4096 ? 1 : 0;  // (more than 268M rows)
    if (scanned) {
        open_scan_kernel<<<1, 1024, 0, s>>>(blk, n_blk, total, st, hd & 1);
        count_launch();
    }
    open_list_kernel<<<n_blk, 256, 0, s>>>(bits, n_words, blk, n_blk, scanned, info, total, (unsigned long long)cap,
                                           (long long*)ws->d_open_rows, (unsigned int*)ws->d_open_info, st, hd & 1);
    count_launch();
    {
        // rows of a batch: 64, fewer when d is large (the batch sits in shared memory next to the float64 partial)
        int rpb = kFixRows;
        while (rpb > 8 && (size_t)rpb * d * sizeof(float) > 32 * 1024) rpb >>= 1;
        const size_t c_bytes = sizeof(float) * ((size_t)k * d + (size_t)k * k);
        const int stage_c = c_bytes <= 24 * 1024 ? 1 : 0;
        const size_t smem = (size_t)rpb * d * sizeof(float) + (acc ? sizeof(double) * (size_t)((acc_n + 1) & ~1) : 0) +
                            (stage_c ? c_bytes : 0);
        static PerDeviceOnce attr_set;
        const int dev = PerDeviceOnce::current();
        if (dev < 0 || !attr_set.done[dev]) {
            if (cudaFuncSetAttribute(fixup_list_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024) != cudaSuccess ||
                cudaFuncSetAttribute(fixup_list_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                     cudaSharedmemCarveoutMaxShared) != cudaSuccess) {
                snprintf(g_fx_err, sizeof(g_fx_err), "cannot raise the dynamic shared memory of the exact pass");
                return 1;
            }
            if (dev >= 0) attr_set.done[dev] = true;
        }
        fixup_list_kernel<<<fix_grid, 256, smem, s>>>(X, d, k, C, pd, stage_c, (const long long*)ws->d_open_rows,
                                                      (const unsigned int*)ws->d_open_info, total, (unsigned long long)cap,
                                                      rpb, labels, fix_slots, slot_stride, st, hd);
        count_launch();
    }
    {
        const size_t fsm = acc ? sizeof(double) * (size_t)acc_n : 0;
        screen_fixup_kernel<<<fix_grid, 256, fsm, s>>>(X, n, d, k, C, pd, info, bits, num_tiles, labels, fix_slots,
                                                       slot_stride, st, hd & 1, total, (unsigned long long)cap);
        count_launch();
    }
    if (acc) {
        slot_reduce_kernel<<<(acc_n + 1 + 31) / 32, 256, 0, s>>>((const double*)ws->d_slots, screen_slots + fix_grid,
                                                              slot_stride, acc_n + 1, acc, st, hd & 1);
        count_launch();
    }
    cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess) {
        snprintf(g_fx_err, sizeof(g_fx_err), "exact pass launch failed: %s", cudaGetErrorString(err));
        return 1;
    }
    return 0;
}

}  // namespace b200km
