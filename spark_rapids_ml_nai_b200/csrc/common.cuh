// Shared declarations for the b200km kernels and the C-ABI layer.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200km {

// ---- device-resident iteration state (one per handle) -----------------------------------------
constexpr int kOwnBinsFirst = 8, kOwnBinsLater = 12;
struct IterState {
    int done;            // set by finalize when shift^2 < tol
    int n_iter;          // executed Lloyd iterations
    unsigned int ticket; // last-block election counter (reset by the elected block)
    unsigned int amb_count;  // entries in the near-tie work list of the current pass
    double shift2;       // sum_j ||c_new_j - c_old_j||^2 of the last executed iteration
    float cnorm_max;     // max_j ||c_j||^2 of the current centroids
    unsigned int screen_refined;  // rows the screen kernel re-decided in float64 since the last prep
    // k <= 32: labels owned by each update warp of the screen kernel (4 slots per warp, -1 = none), for own_bins
    // warps: 8 contiguous ranges after a centroid preparation (the first pass of a fit also computes the row norms
    // and has 8 update warps), 12 warps balanced by cluster size after every Lloyd iteration.
    signed char own[48];
    int own_bins;
};

// Near-tie band: a row whose best and second-best approximate scores differ by less than
// kTieTol * (||x||^2 + max_j ||c_j||^2) is re-decided in float64.  The bound covers the
// worst-case rounding of both arithmetic paths (FP32 FFMA chains; 3xTF32 with truncated splits).
constexpr float kTieTol = 4e-6f;

constexpr int kNumSMs = 148;

// Per-device "already done" flags for one-time set-up that CUDA keeps per device (function attributes such
// as the dynamic shared-memory opt-in): one process may drive several GPUs through several handles.
constexpr int kMaxTrackedDevices = 64;
struct PerDeviceOnce {
    bool done[kMaxTrackedDevices] = {};
    // index of the current device, or -1 when it cannot be tracked (the caller then repeats the set-up)
    static int current() {
        int dev = -1;
        if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxTrackedDevices) return -1;
        return dev;
    }
};

// Kernel launches issued by this thread through the library's host launchers (every `<<<>>>` site counts
// itself); the C-ABI layer reports the difference across a call as the handle's launch count.
extern thread_local unsigned long long g_launches;
inline void count_launch(int n = 1) { g_launches += (unsigned long long)n; }

// ---- generic (SIMT) path, kernels_generic.cu ---------------------------------------------------
void launch_prep(const float* C, int k, int d, float* cnorm, double* acc, int64_t acc_len,
                 IterState* st, bool reset_iter, cudaStream_t s);
void launch_rownorm(const float* X, int64_t n, int d, float* xnorm, cudaStream_t s);
void launch_assign_ffma(const float* X, int64_t n, int d, const float* C, int k, const float* cnorm,
                        const float* xnorm, int32_t* labels, float* best, IterState* st,
                        unsigned int* worklist, bool honor_done, cudaStream_t s);
void launch_refine(const float* X, int64_t n, int d, const float* C, int k, int32_t* labels,
                   float* best, IterState* st, const unsigned int* worklist, bool honor_done,
                   cudaStream_t s);
void launch_update(const float* X, int64_t n, int d, int k, const int32_t* labels, const float* best,
                   double* acc, IterState* st, bool honor_done, cudaStream_t s);
void launch_sum_best(const float* best, int64_t n, double* out, cudaStream_t s);
// snap: where the last block leaves a copy of *st once the iteration is complete (may be null)
void launch_finalize(float* C, int k, int d, float* cnorm, double* acc, double* shift_part, double* cnt_part,
                     double tol, IterState* st, IterState* snap, cudaStream_t s);

// Fused tail for small k*d (kernels_generic.cu): one-shot all-reduce over NVLink peer stores + centroid update in
// one CTA.  Exchange buffer of a rank: [2 parities][nranks <= 16][kXrDoubles] float64, then the arrival flags.
constexpr int kXrDoubles = 8192;
constexpr size_t kXrFlagOff = (size_t)2 * 16 * kXrDoubles * sizeof(double);
constexpr size_t kXrBytes = kXrFlagOff + 2 * 16 * sizeof(unsigned int);
bool finalize_fused_supported(int k, int d);
// peer_bases: every rank's exchange buffer as mapped in this process (nullptr when nranks == 1)
void launch_finalize_fused(float* C, int k, int d, float* cnorm, double* acc, double tol, IterState* st, IterState* snap,
                           unsigned char* const* peer_bases, int rank, int nranks, unsigned int epoch, cudaStream_t s);

// ---- shared-memory resident whole-fit kernel for small shards, kernels_small.cu -------------------------------
bool small_fit_supported(int64_t n, int d, int k);
size_t small_slots_bytes(int64_t n, int d, int k);
int small_fit(const float* X, int64_t n, int d, int k, const float* init, int max_iter, double tol, double* slots,
              float* C_out, int32_t* labels_out, double* inertia_out, IterState* st, cudaStream_t s);
const char* small_last_error();

// ---- k-means|| initialisation, kernels_init.cu ---------------------------------------------------
unsigned long long init_hash64(unsigned long long seed, unsigned long long stream, unsigned long long counter);
double init_hash_uniform(unsigned long long seed, unsigned long long stream, unsigned long long counter);
void launch_kpar_sample(const float* dist, int64_t n, long long row0, unsigned long long seed, int stream, double scale,
                        long long* list, unsigned int* count, unsigned int cap, cudaStream_t s);
void launch_min_inplace(float* dist, const float* other, int64_t n, cudaStream_t s);
void launch_det_sum(const float* x, int64_t n, double* partial, double* out, unsigned int* ticket, cudaStream_t s);
void launch_gather_rows(const float* X, int64_t n, int d, long long row0, const long long* ids, int m, float* out,
                        cudaStream_t s);
void launch_label_count(const int32_t* labels, int64_t n, unsigned long long* counts, cudaStream_t s);
size_t kpp_state_bytes();
void run_weighted_kmeanspp(const float* C, int m, int d, int k, const double* w, unsigned long long seed, int trials,
                           double* P, double* d2, double* Dt, unsigned char* taken, int* chosen, void* state, float* out,
                           cudaStream_t s);

}  // namespace b200km
