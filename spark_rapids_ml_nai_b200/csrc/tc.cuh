// Interface of the tcgen05 / TMEM / TMA path (kernels_tc.cu).
#pragma once
#include "common.cuh"

namespace b200km {

struct TcWorkspace {
    void* d_buf = nullptr;   // split centroids (hi | lo), padded, in the layout the MMA reads
    size_t cap = 0;
    void* tmap_cache = nullptr;  // host-side cache of encoded tensor maps
};

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
void tc_free(TcWorkspace* ws);

}  // namespace b200km
