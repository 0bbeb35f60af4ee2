// Interface of the tcgen05 / TMEM / TMA path (kernels_tc.cu).
#pragma once
#include "common.cuh"

namespace b200km {

struct TcWorkspace {
    void* d_buf = nullptr;   // split centroids (hi | lo), padded, in the layout the MMA reads
    size_t cap = 0;
    void* d_pd = nullptr;    // [k][k] upper bounds of the pairwise centroid distances (screen kernel)
    size_t pd_cap = 0;
    void* d_bits = nullptr;  // open-row bitmap of the screen kernel: 4 words per 128-row tile
    size_t bits_cap = 0;
    void* d_slots = nullptr;  // per-CTA float64 partial sums of the screen kernel and its exact pass
    size_t slots_cap = 0;
    // exact pass (kernels_fixup.cu): the open rows as one list in row order (row index, verdict word), the popcount
    // of every block of bitmap words and its prefix, the number of open rows
    void* d_open_rows = nullptr;
    void* d_open_info = nullptr;
    size_t open_cap = 0;  // entries
    void* d_open_blk = nullptr;
    size_t open_blk_cap = 0;  // blocks
};

constexpr int kTileM = 128;  // rows of a tile of the screen kernel = rows covered by 4 bitmap words
// Per-row verdict written by the screen kernel: bits 0-7 the computed best centroid (the row's provisional
// label, already counted in the sums), bits 8-15 the runner-up, bits 16-17 what the exact pass has to look at,
// bit 31 "not decided by the TF32 scores".
constexpr unsigned int kInfoUnsafe = 1u << 31;
constexpr unsigned int kInfoPair = 0u << 16;   // exactly the best and the runner-up
constexpr unsigned int kInfoReach = 1u << 16;  // every centroid within 2 r_1 of the computed best
[[maybe_unused]] constexpr unsigned int kInfoAll = 2u << 16;  // every centroid (kept for producers without a reach)

constexpr int kFixCtasPerSm = 4;  // CTAs (= partial-sum slots) of the exact pass per SM
// Exact decision of the rows the screen kernel left open and their contribution to the sums (kernels_fixup.cu): reads
// the bitmap ws->d_bits and the verdict words `info`, writes labels (where they change) and, when acc != nullptr, one
// float64 partial per CTA into ws->d_slots behind the screen kernel's `screen_slots` slots, then adds all slots into
// acc in slot order.
int screen_fixup(TcWorkspace* ws, const float* X, int64_t n, int d, int k, const float* C, const unsigned int* info,
                 int64_t num_tiles, int32_t* labels, double* acc, int screen_slots, int slot_stride, IterState* st,
                 bool honor_done, cudaStream_t s);
const char* fixup_last_error();


bool tc_supported(const float* X, int d, int k, bool with_acc);
int tc_max_d();
int tc_max_k();
size_t tc_workspace_bytes(int d, int k);
// One fused pass: labels + best for every row; when acc != nullptr also the float64 per-cluster
// sums / counts / inertia ([k*d | k | 1]) accumulated into acc.  Returns 0 on success.
// xnorm: device [n] row norms; written by the pass when xn_compute, otherwise read (a fit computes them
// in its first pass and reuses them in the later ones).
int tc_pass(TcWorkspace* ws, const float* X, int64_t n, int d, int k, const float* C, const float* cnorm,
            float* xnorm, bool xn_compute, int32_t* labels, float* best, double* acc, IterState* st,
            bool honor_done, cudaStream_t s);
const char* tc_last_error();
// out[2*k_pad][d_pad]: rows [0,k_pad) = centroids rounded to TF32, rows [k_pad,2k_pad) = the TF32-rounded
// remainders; zero padded.  The B operand of both tcgen05 kernels.
void launch_split_centroids(const float* C, int k, int d, int k_pad, int d_pad, float* out, cudaStream_t s);

// Screen kernel (kernels_screen.cu): the same pass with the TF32 view of X read by the tensor cores straight
// from shared memory, a rigorous error band and a float64 re-decision inside it.  Produces labels (optional)
// and, when acc != nullptr, sums / counts and sum_i ||x_i||^2 in the inertia slot.  No per-row distances.
bool screen_supported(const float* X, int d, int k, bool with_acc);
size_t screen_workspace_bytes(int d, int k);
int screen_pass(TcWorkspace* ws, const float* X, int64_t n, int d, int k, const float* C, const float* cnorm,
                float* xnorm, float* xlo, unsigned int* info, bool xn_compute, int32_t* labels, double* acc,
                IterState* st, bool honor_done, cudaStream_t s);
const char* screen_last_error();
void tc_free(TcWorkspace* ws);

// Tile-GEMM kernel (kernels_gemm.cu): nearest centroid for any k and any d % 4 == 0 on a CTA pair
// (tcgen05.mma.cta_group::2), centroids streamed in N-tiles, exact labels through a fix-up list.
struct GemmWorkspace {
    void* d_b = nullptr;  // [2*k_pad][d_pad] TF32 split of the centroids
    size_t b_cap = 0;
    void* d_cn = nullptr;  // [k_pad][8] three-term TF32 split of ||c||^2
    size_t cn_cap = 0;
    void* d_open = nullptr;  // (row, runner-up | mode) of the rows the TF32 scores did not decide
    size_t open_cap = 0;
    void* d_count = nullptr;
    // second pass over the rows with more than two candidates: gathered rows, their original index, score
    // threshold and (distance, index) minimum, and the candidate list
    void* d_xs = nullptr;
    void* d_scan_row = nullptr;
    void* d_scan_thr = nullptr;
    void* d_scan_key = nullptr;
    void* d_cand = nullptr;
    size_t scan_cap = 0, scan_d = 0;
};
bool gemm_supported(const float* X, int d, int k);
size_t gemm_workspace_bytes(int64_t n, int d, int k);
// labels[n] (required): exact nearest centroid of every row once the stream reaches this point.
// xnorm / xlo: ||x||^2 and ||x - trunc_tf32(x)||^2 per row (launch_rownorm2).
int gemm_pass(GemmWorkspace* ws, const float* X, int64_t n, int d, int k, const float* C, const float* cnorm,
              const float* xnorm, const float* xlo, int32_t* labels, IterState* st, bool honor_done, cudaStream_t s);
void launch_rownorm2(const float* X, int64_t n, int d, float* xnorm, float* xlo, cudaStream_t s);
// best[i] = ||x_i - c_labels[i]||^2 from FP32 differences
void launch_rowdist(const float* X, int64_t n, int d, const float* C, const int32_t* labels, float* best,
                    cudaStream_t s);
const char* gemm_last_error();
void gemm_free(GemmWorkspace* ws);

// Deterministic per-cluster sums / counts for large k (kernels_update.cu): stable counting sort of the rows by
// label + segmented float64 reduction in a fixed order; acc = [k*d sums | k counts | ...], added to.
struct SortedUpdateWorkspace {
    void* d_H = nullptr;
    size_t H_cap = 0;
    void* d_total = nullptr;
    size_t total_cap = 0;
    void* d_start = nullptr;
    size_t start_cap = 0;
    void* d_order = nullptr;
    size_t order_cap = 0;
    void* d_head = nullptr;
    size_t head_cap = 0;
    void* d_tail = nullptr;
    size_t tail_cap = 0;
};
bool sorted_update_supported(int64_t n, int d, int k);
int sorted_update(SortedUpdateWorkspace* ws, const float* X, int64_t n, int d, int k, const int32_t* labels, double* acc,
                  IterState* st, bool honor_done, cudaStream_t s);
const char* sorted_update_last_error();
void sorted_update_free(SortedUpdateWorkspace* ws);

}  // namespace b200km
