// C-ABI layer of b200km (declared in include/b200km.h): handle, workspace, NCCL bring-up, the
// device-resident Lloyd loop and predict.  No C++ exception and no torch type crosses this file's
// extern "C" surface.
#include "../../include/b200km.h"
#include "common.cuh"
#include "tc.cuh"

#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <random>
#include <unordered_set>
#include <vector>

using namespace b200km;

namespace {

thread_local char g_err[1024] = "";

int fail(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return 1;
}

#define CU_TRY(expr)                                                                          \
    do {                                                                                      \
        cudaError_t _e = (expr);                                                              \
        if (_e != cudaSuccess)                                                                \
            return fail("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
    } while (0)

// ---- NCCL through dlopen: the library has no link-time dependency on it, a single-GPU process
// never loads it, and a process that already holds libnccl.so.2 (torch) shares that copy. --------
struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*CommAbort)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t,
                              cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
};
NcclApi g_nccl;

int load_nccl() {
    if (g_nccl.lib) return 0;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    void* lib = nullptr;
    for (const char* nm : names) {
        lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
        if (lib) break;
    }
    if (!lib) return fail("cannot dlopen libnccl.so.2: %s", dlerror());
#define SYM(field, name)                                                        \
    *(void**)(&g_nccl.field) = dlsym(lib, name);                                \
    if (!g_nccl.field) return fail("libnccl has no symbol %s", name);
    SYM(GetUniqueId, "ncclGetUniqueId")
    SYM(CommInitRank, "ncclCommInitRank")
    SYM(CommDestroy, "ncclCommDestroy")
    SYM(CommAbort, "ncclCommAbort")
    SYM(AllReduce, "ncclAllReduce")
    SYM(AllGather, "ncclAllGather")
    SYM(GetErrorString, "ncclGetErrorString")
#undef SYM
    g_nccl.lib = lib;
    return 0;
}

#define NCCL_TRY(expr)                                                                              \
    do {                                                                                            \
        ncclResult_t _r = (expr);                                                                   \
        if (_r != ncclSuccess)                                                                      \
            return fail("%s failed: %s (%s:%d)", #expr, g_nccl.GetErrorString(_r), __FILE__, __LINE__); \
    } while (0)

constexpr int kRing = 4;  // how many iterations the host may run ahead of the device

}  // namespace

struct b200km_handle {
    int device = 0, rank = 0, nranks = 1;
    ncclComm_t comm = nullptr;
    int path = B200KM_PATH_AUTO;
    const char* last_path = "none";
    // workspace
    IterState* d_state = nullptr;  // [1 + kRing]: the live state, then one snapshot per ring slot
    cudaStream_t side = nullptr;   // carries the snapshots to the host beside the compute stream
    cudaEvent_t evm[kRing] = {};   // "iteration finalized" on the compute stream
    IterState* h_state = nullptr;  // pinned ring [kRing] + 1 slot for the final read
    double* h_scalar = nullptr;    // pinned
    cudaEvent_t ev[kRing] = {};
    double* d_acc = nullptr;
    size_t acc_cap = 0;
    double* d_shift = nullptr;
    float* d_cnorm = nullptr;
    size_t k_cap = 0;
    float* d_C = nullptr;
    size_t kd_cap = 0;
    float* d_xnorm = nullptr;
    float* d_best = nullptr;
    int32_t* d_labels = nullptr;
    unsigned int* d_work = nullptr;
    size_t n_cap = 0;
    TcWorkspace tc;
    GemmWorkspace gm;
    SortedUpdateWorkspace su;
    bool split_only = false;  // this fit saw too many float64 re-decisions in the screen kernel
    int last_kernel = 0;      // 0 = FFMA kernels, 1 = split-TF32 fused kernel, 2 = screen kernel, 3 = tile GEMM
    // instrumentation (b200km_profile_*)
    bool prof = false;
    static constexpr int kProfMax = 256;
    cudaEvent_t pev[2 * kProfMax] = {};
    int prof_n = 0;
    unsigned long long launches = 0;
    int scope_depth = 0;  // nesting of the entry points that count launches (only the outermost adds)
    // one-shot all-reduce over NVLink for small accumulators (fused into finalize): this rank's exchange buffer
    // and every peer's, mapped through CUDA IPC at creation; NCCL does the large messages and the bootstrap
    double* d_red = nullptr;    // scratch of the fixed-order sums
    double* d_small = nullptr;  // slots of the shared-memory resident fit
    size_t small_cap = 0;
    unsigned char* xr_local = nullptr;
    unsigned char* xr_peer[16] = {};
    bool xr_ok = false;
    unsigned int fit_serial = 0;
    const char* last_reduce = "none";
};

namespace {

int free_ws(b200km_handle* h) {
    for (int r = 0; r < 16; ++r)
        if (h->xr_peer[r] && h->xr_peer[r] != h->xr_local) cudaIpcCloseMemHandle(h->xr_peer[r]);
    if (h->xr_local) cudaFree(h->xr_local);
    cudaFree(h->d_small);
    cudaFree(h->d_red);
    cudaFree(h->d_state);
    cudaFree(h->d_acc);
    cudaFree(h->d_shift);
    cudaFree(h->d_cnorm);
    cudaFree(h->d_C);
    cudaFree(h->d_xnorm);
    cudaFree(h->d_best);
    cudaFree(h->d_labels);
    cudaFree(h->d_work);
    tc_free(&h->tc);
    gemm_free(&h->gm);
    sorted_update_free(&h->su);
    if (h->h_state) cudaFreeHost(h->h_state);
    if (h->h_scalar) cudaFreeHost(h->h_scalar);
    for (auto& e : h->ev)
        if (e) cudaEventDestroy(e);
    for (auto& e : h->evm)
        if (e) cudaEventDestroy(e);
    if (h->side) cudaStreamDestroy(h->side);
    for (auto& e : h->pev)
        if (e) cudaEventDestroy(e);
    return 0;
}

template <typename T>
int grow(T** p, size_t* cap, size_t need) {
    if (need <= *cap && *p) return 0;
    if (*p) CU_TRY(cudaFree(*p));
    *p = nullptr;
    const size_t want = need + need / 8 + 64;
    CU_TRY(cudaMalloc((void**)p, sizeof(T) * want));
    *cap = want;
    return 0;
}

int ensure_ws(b200km_handle* h, int64_t n, int d, int k, bool need_row_bufs) {
    const size_t kd = (size_t)k * d;
    const size_t acc_len = kd + k + 1;
    if (acc_len > h->acc_cap || !h->d_acc) {
        if (h->d_acc) CU_TRY(cudaFree(h->d_acc));
        h->d_acc = nullptr;
        CU_TRY(cudaMalloc((void**)&h->d_acc, sizeof(double) * (acc_len + 64)));
        h->acc_cap = acc_len + 64;
    }
    if ((size_t)k > h->k_cap || !h->d_cnorm) {
        if (h->d_cnorm) CU_TRY(cudaFree(h->d_cnorm));
        if (h->d_shift) CU_TRY(cudaFree(h->d_shift));
        h->d_cnorm = nullptr;
        h->d_shift = nullptr;
        CU_TRY(cudaMalloc((void**)&h->d_cnorm, sizeof(float) * (k + 64)));
        CU_TRY(cudaMalloc((void**)&h->d_shift, sizeof(double) * 2 * (k + 64)));  // shifts | cluster sizes
        h->k_cap = k + 64;
    }
    if (grow(&h->d_C, &h->kd_cap, kd)) return 1;
    if (need_row_bufs && ((size_t)n > h->n_cap || !h->d_xnorm)) {
        const size_t want = (size_t)n + (size_t)n / 8 + 256;
        for (void** p : {(void**)&h->d_xnorm, (void**)&h->d_best, (void**)&h->d_labels, (void**)&h->d_work}) {
            if (*p) CU_TRY(cudaFree(*p));
            *p = nullptr;
            CU_TRY(cudaMalloc(p, 4 * want));
        }
        h->n_cap = want;
    }
    return 0;
}

// Maps every rank's exchange buffer into this process (CUDA IPC handles travel through one ncclAllGather).
// The ranks then agree (ncclAllReduce MIN) on whether all of them succeeded; if not, nobody uses the buffers.
void setup_exchange(b200km_handle* h) {
    int ok = 1;
    cudaIpcMemHandle_t mine;
    unsigned char* d_handles = nullptr;
    std::vector<cudaIpcMemHandle_t> all(h->nranks);
    if (cudaMalloc((void**)&h->xr_local, kXrBytes) != cudaSuccess || cudaMemset(h->xr_local, 0, kXrBytes) != cudaSuccess ||
        cudaIpcGetMemHandle(&mine, h->xr_local) != cudaSuccess ||
        cudaMalloc((void**)&d_handles, sizeof(cudaIpcMemHandle_t) * h->nranks + 16) != cudaSuccess)
        ok = 0;
    cudaGetLastError();
    if (d_handles) {
        // the collectives must run on every rank even when this one failed locally
        if (ok) cudaMemcpy(d_handles + sizeof(mine) * h->rank, &mine, sizeof(mine), cudaMemcpyHostToDevice);
        if (g_nccl.AllGather(d_handles + sizeof(mine) * h->rank, d_handles, sizeof(mine), ncclChar, h->comm, 0) != ncclSuccess) ok = 0;
        if (cudaStreamSynchronize(0) != cudaSuccess) ok = 0;
        if (ok) cudaMemcpy(all.data(), d_handles, sizeof(mine) * h->nranks, cudaMemcpyDeviceToHost);
    }
    if (ok) {
        for (int r = 0; r < h->nranks; ++r) {
            if (r == h->rank) {
                h->xr_peer[r] = h->xr_local;
                continue;
            }
            void* ptr = nullptr;
            if (cudaIpcOpenMemHandle(&ptr, all[r], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
                ok = 0;
                cudaGetLastError();
                break;
            }
            h->xr_peer[r] = (unsigned char*)ptr;
        }
    }
    if (d_handles) {
        int* d_ok = reinterpret_cast<int*>(d_handles + sizeof(mine) * h->nranks);
        cudaMemcpy(d_ok, &ok, sizeof(int), cudaMemcpyHostToDevice);
        int agreed = 0;
        if (g_nccl.AllReduce(d_ok, d_ok, 1, ncclInt32, ncclMin, h->comm, 0) == ncclSuccess &&
            cudaStreamSynchronize(0) == cudaSuccess && cudaMemcpy(&agreed, d_ok, sizeof(int), cudaMemcpyDeviceToHost) == cudaSuccess)
            h->xr_ok = agreed == 1;
        cudaFree(d_handles);
    }
    cudaGetLastError();
}

// Adds the kernel launches issued inside its scope (counted at the launch sites) to the handle's counter.
struct LaunchScope {
    b200km_handle* h;
    unsigned long long start;
    explicit LaunchScope(b200km_handle* hh) : h(hh), start(g_launches) {
        if (h) h->scope_depth++;
    }
    ~LaunchScope() {
        if (h && --h->scope_depth == 0) h->launches += g_launches - start;
    }
};

int check_shape(const b200km_handle* h, const void* X, int64_t n, int d, int k) {
    if (!h) return fail("null handle");
    if (!X && n > 0) return fail("X is null");
    if (n < 0 || n >= ((int64_t)1 << 31)) return fail("n=%lld out of range [0, 2^31)", (long long)n);
    if (d <= 0 || d > (1 << 20)) return fail("d=%d out of range", d);
    if (k <= 0 || k > (1 << 24)) return fail("k=%d out of range", k);
    return 0;
}

// Kernel family of a pass.
enum Kind { KIND_FFMA = 0, KIND_TC = 1, KIND_GEMM = 2 };

// Picks the kernel family for this shape, honouring a forced path (and failing if it cannot run).
int pick_path(b200km_handle* h, const float* X, int d, int k, bool with_acc, int* kind) {
    const bool can_tc = tc_supported(X, d, k, with_acc);
    const bool can_gemm = gemm_supported(X, d, k);
    if ((h->path == B200KM_PATH_TC || h->path == B200KM_PATH_TC_SPLIT) && !can_tc)
        return fail("tcgen05 path forced but not available for d=%d k=%d (needs d%%4==0, 16B-aligned X, "
                    "d<=%d, k<=%d and the split centroids%s to fit shared memory)", d, k, tc_max_d(), tc_max_k(),
                    with_acc ? " + float64 accumulators" : "");
    if (h->path == B200KM_PATH_GEMM && !can_gemm)
        return fail("tcgen05 tile-GEMM path forced but not available for d=%d k=%d (needs d%%4==0 and 16B-aligned X)", d, k);
    if (h->path == B200KM_PATH_TC || h->path == B200KM_PATH_TC_SPLIT) *kind = KIND_TC;
    else if (h->path == B200KM_PATH_GEMM) *kind = KIND_GEMM;
    else if (h->path == B200KM_PATH_FFMA) *kind = KIND_FFMA;
    else *kind = can_tc ? KIND_TC : (can_gemm ? KIND_GEMM : KIND_FFMA);
    h->last_path = *kind == KIND_TC ? "tc" : (*kind == KIND_GEMM ? "gemm" : "ffma");
    return 0;
}

// One assignment (+ optional accumulation) pass on the chosen path.
// labels / best: caller buffers, or nullptr when the caller does not need them (the FFMA kernels then use
// the workspace arrays).  acc == nullptr -> assignment only.  On the tcgen05 path a pass that needs no
// per-row distances runs the screen kernel, every other one the split-TF32 kernel.
// want_inertia: with acc, the inertia slot of acc has to be right (the tile-GEMM path then computes per-row distances).
int run_pass(b200km_handle* h, int kind, const float* X, int64_t n, int d, int k, const float* C,
             int32_t* labels, float* best, double* acc, bool honor_done, bool xn_compute, cudaStream_t s,
             bool want_inertia = false) {
    const bool use_tc = kind == KIND_TC;
    // the dominant kernel of the pass is bracketed by an event pair when profiling is on
    const bool rec = h->prof && h->prof_n < b200km_handle::kProfMax;
    if (rec) CU_TRY(cudaEventRecord(h->pev[2 * h->prof_n], s));
    if (use_tc) {
        const bool screen = h->path != B200KM_PATH_TC_SPLIT && !h->split_only && best == nullptr &&
                            screen_supported(X, d, k, acc != nullptr);
        if (screen) {
            if (screen_pass(&h->tc, X, n, d, k, C, h->d_cnorm, h->d_xnorm, (float*)h->d_work, (unsigned int*)h->d_best,
                            xn_compute, labels, acc, h->d_state,
                            honor_done, s))
                return fail("tcgen05 screen pass failed: %s", screen_last_error());
        } else if (tc_pass(&h->tc, X, n, d, k, C, h->d_cnorm, h->d_xnorm, xn_compute, labels, best, acc, h->d_state,
                           honor_done, s)) {
            return fail("tcgen05 pass failed: %s", tc_last_error());
        }
        h->last_kernel = screen ? 2 : 1;
        if (rec) CU_TRY(cudaEventRecord(h->pev[2 * h->prof_n + 1], s));
        if (rec) h->prof_n++;
        return 0;
    }
    if (kind == KIND_GEMM) {
        // tile GEMM on the tensor cores -> exact labels; distances and sums are separate (cheap) passes here
        int32_t* lab = labels ? labels : h->d_labels;
        float* xlo = reinterpret_cast<float*>(h->d_work);
        if (xn_compute) launch_rownorm2(X, n, d, h->d_xnorm, xlo, s);
        if (gemm_pass(&h->gm, X, n, d, k, C, h->d_cnorm, h->d_xnorm, xlo, lab, h->d_state, honor_done, s))
            return fail("tcgen05 tile-GEMM pass failed: %s", gemm_last_error());
        h->last_kernel = 3;
        if (rec) CU_TRY(cudaEventRecord(h->pev[2 * h->prof_n + 1], s));
        if (rec) h->prof_n++;
        float* bst = best ? best : h->d_best;
        if (best || (acc && want_inertia)) launch_rowdist(X, n, d, C, lab, bst, s);
        if (acc) {
            if (sorted_update_supported(n, d, k)) {
                // stable sort by label + segmented float64 sums in a fixed order (no atomics on the data)
                if (sorted_update(&h->su, X, n, d, k, lab, acc, h->d_state, honor_done, s))
                    return fail("sorted update failed: %s", sorted_update_last_error());
                if (want_inertia) launch_sum_best(bst, n, acc + (size_t)k * d + k, s);
            } else {
                if (!(best || want_inertia)) CU_TRY(cudaMemsetAsync(bst, 0, sizeof(float) * (size_t)n, s));
                launch_update(X, n, d, k, lab, bst, acc, h->d_state, honor_done, s);
            }
        }
        CU_TRY(cudaGetLastError());
        return 0;
    }
    if (!labels) labels = h->d_labels;
    if (!best) best = h->d_best;
    h->last_kernel = 0;
    launch_assign_ffma(X, n, d, C, k, h->d_cnorm, h->d_xnorm, labels, best, h->d_state, h->d_work, honor_done, s);
    if (rec) CU_TRY(cudaEventRecord(h->pev[2 * h->prof_n + 1], s));
    if (rec) h->prof_n++;
    launch_refine(X, n, d, C, k, labels, best, h->d_state, h->d_work, honor_done, s);
    if (acc) launch_update(X, n, d, k, labels, best, acc, h->d_state, honor_done, s);
    CU_TRY(cudaGetLastError());
    return 0;
}

// Shard sizes of all ranks (through the communicator), this rank's first global row and the total.
int gather_sizes(b200km_handle* h, int64_t n_local, int d, int k, std::vector<long long>* sizes, long long* my_off,
                 long long* total, cudaStream_t s) {
    sizes->assign(h->nranks, 0);
    (*sizes)[h->rank] = n_local;
    if (h->nranks > 1) {
        if (ensure_ws(h, 0, d, k, false)) return 1;
        long long* d_sizes = reinterpret_cast<long long*>(h->d_acc);
        CU_TRY(cudaMemcpyAsync(d_sizes + h->rank, &(*sizes)[h->rank], sizeof(long long), cudaMemcpyHostToDevice, s));
        NCCL_TRY(g_nccl.AllGather(d_sizes + h->rank, d_sizes, 1, ncclInt64, h->comm, s));
        CU_TRY(cudaMemcpyAsync(sizes->data(), d_sizes, sizeof(long long) * h->nranks, cudaMemcpyDeviceToHost, s));
        CU_TRY(cudaStreamSynchronize(s));
    }
    *total = 0;
    *my_off = 0;
    for (int r = 0; r < h->nranks; ++r) {
        if (r == h->rank) *my_off = *total;
        *total += (*sizes)[r];
    }
    return 0;
}

// Temporary device allocations of one initialisation call, released together.
struct DevPool {
    std::vector<void*> ptrs;
    ~DevPool() {
        for (void* q : ptrs) cudaFree(q);
    }
    template <typename T>
    T* get(size_t count) {
        void* q = nullptr;
        if (cudaMalloc(&q, sizeof(T) * (count ? count : 1)) != cudaSuccess) return nullptr;
        ptrs.push_back(q);
        return (T*)q;
    }
};

}  // namespace

extern "C" int b200km_predict(b200km_handle* h, const float* X, int64_t n, int d, const float* centroids, int k,
                              int32_t* labels_out, float* min_dist_out, void* stream);

namespace {

// k-means|| ("scalable k-means++"): see kernels_init.cu (the CPU checker restates the same hash keying and rules).
// Every rank ends with the same k rows in `out`.
int kmeans_parallel_init(b200km_handle* h, const float* X, int64_t n_local, int d, int k, unsigned long long seed,
                         double oversampling, float* out, cudaStream_t s) {
    std::vector<long long> sizes;
    long long my_off = 0, total = 0;
    if (gather_sizes(h, n_local, d, k, &sizes, &my_off, &total, s)) return 1;
    if (total < k) return fail("k=%d exceeds the number of rows %lld", k, total);
    const double ell = oversampling * k < 1.0 ? 1.0 : oversampling * k;
    const unsigned int cap = (unsigned int)(8.0 * ell) + 4096u;
    DevPool pool;
    float* dist = pool.get<float>((size_t)n_local);
    float* tmp = pool.get<float>((size_t)n_local);
    int32_t* lab = pool.get<int32_t>((size_t)n_local);
    long long* list = pool.get<long long>((size_t)h->nranks * (cap + 1));
    long long* all = pool.get<long long>((size_t)h->nranks * (cap + 1));
    unsigned int* cnt = pool.get<unsigned int>(4);
    double* red = pool.get<double>(1024 + 8);
    if (!dist || !tmp || !lab || !list || !all || !cnt || !red) return fail("k-means||: device allocation failed");
    CU_TRY(cudaMemsetAsync(cnt, 0, 16, s));

    std::vector<long long> cands;  // global row indices, in collection order
    float* cand = nullptr;         // their rows [cands.size(), d]
    size_t cand_cap = 0;

    auto phi_of = [&](double* phi) -> int {
        launch_det_sum(dist, n_local, red, red + 1024, cnt + 1, s);
        if (h->nranks > 1) NCCL_TRY(g_nccl.AllReduce(red + 1024, red + 1024, 1, ncclFloat64, ncclSum, h->comm, s));
        CU_TRY(cudaMemcpyAsync(phi, red + 1024, sizeof(double), cudaMemcpyDeviceToHost, s));
        CU_TRY(cudaStreamSynchronize(s));
        return 0;
    };
    // appends the rows `ids` (ascending global indices) to the candidate matrix and folds them into dist
    auto add = [&](const std::vector<long long>& ids, bool first) -> int {
        const size_t m0 = cands.size(), m1 = m0 + ids.size();
        if (m1 > cand_cap) {
            const size_t want = m1 * 2 + 64;
            float* bigger = pool.get<float>(want * d);
            if (!bigger) return fail("k-means||: device allocation of %zu candidate rows failed", want);
            if (cand) CU_TRY(cudaMemcpyAsync(bigger, cand, sizeof(float) * m0 * d, cudaMemcpyDeviceToDevice, s));
            cand = bigger;
            cand_cap = want;
        }
        CU_TRY(cudaMemcpyAsync(list, ids.data(), sizeof(long long) * ids.size(), cudaMemcpyHostToDevice, s));
        float* rows = cand + m0 * d;
        launch_gather_rows(X, n_local, d, my_off, list, (int)ids.size(), rows, s);
        if (h->nranks > 1)
            NCCL_TRY(g_nccl.AllReduce(rows, rows, ids.size() * (size_t)d, ncclFloat32, ncclSum, h->comm, s));
        cands.insert(cands.end(), ids.begin(), ids.end());
        if (n_local > 0) {
            if (b200km_predict(h, X, n_local, d, rows, (int)ids.size(), lab, first ? dist : tmp, s)) return 1;
            if (!first) launch_min_inplace(dist, tmp, n_local, s);
        }
        CU_TRY(cudaGetLastError());
        return 0;
    };

    const long long first_row = (long long)(init_hash64(seed, 0, 0) % (unsigned long long)total);
    if (add({first_row}, true)) return 1;
    double phi = 0.0;
    if (phi_of(&phi)) return 1;
    int rounds = (int)std::ceil(std::log(phi > 2.0 ? phi : 2.0));
    rounds = rounds < 1 ? 1 : (rounds > 8 ? 8 : rounds);
    std::vector<long long> host_list(cap), host_all((size_t)h->nranks * (cap + 1));
    for (int r = 0; r < rounds && phi > 0.0; ++r) {
        CU_TRY(cudaMemsetAsync(cnt, 0, sizeof(unsigned int), s));
        launch_kpar_sample(dist, n_local, my_off, seed, r + 1, ell / phi, list + 1, cnt, cap, s);
        unsigned int count = 0;
        CU_TRY(cudaMemcpyAsync(&count, cnt, sizeof(unsigned int), cudaMemcpyDeviceToHost, s));
        CU_TRY(cudaStreamSynchronize(s));
        if (count > cap) return fail("k-means||: a sampling round picked %u rows (capacity %u): oversampling factor too large", count, cap);
        CU_TRY(cudaMemcpyAsync(host_list.data(), list + 1, sizeof(long long) * count, cudaMemcpyDeviceToHost, s));
        CU_TRY(cudaStreamSynchronize(s));
        std::sort(host_list.begin(), host_list.begin() + count);  // the append order is arbitrary: ascending rows
        std::vector<long long> ids;
        if (h->nranks > 1) {
            // every rank contributes [count | ids ...]; ranks own ascending row ranges, so rank order is row order
            std::vector<long long> seg((size_t)cap + 1, 0);
            seg[0] = count;
            std::copy(host_list.begin(), host_list.begin() + count, seg.begin() + 1);
            CU_TRY(cudaMemcpyAsync(all + (size_t)h->rank * (cap + 1), seg.data(), sizeof(long long) * (cap + 1), cudaMemcpyHostToDevice, s));
            NCCL_TRY(g_nccl.AllGather(all + (size_t)h->rank * (cap + 1), all, (size_t)cap + 1, ncclInt64, h->comm, s));
            CU_TRY(cudaMemcpyAsync(host_all.data(), all, sizeof(long long) * host_all.size(), cudaMemcpyDeviceToHost, s));
            CU_TRY(cudaStreamSynchronize(s));
            for (int q = 0; q < h->nranks; ++q) {
                const long long* sg = host_all.data() + (size_t)q * (cap + 1);
                ids.insert(ids.end(), sg + 1, sg + 1 + sg[0]);
            }
        } else {
            ids.assign(host_list.begin(), host_list.begin() + count);
        }
        if (ids.empty()) continue;
        if (add(ids, false)) return 1;
        if (phi_of(&phi)) return 1;
    }

    const size_t m = cands.size();
    if (m < (size_t)k) {
        // tiny or degenerate data: further distinct rows, lowest hash first (the same rule on every rank)
        if (total > (1ll << 27)) return fail("k-means||: fewer candidates than k on %lld rows", total);
        std::unordered_set<long long> have(cands.begin(), cands.end());
        std::vector<std::pair<unsigned long long, long long>> order;
        order.reserve((size_t)total);
        for (long long g = 0; g < total; ++g)
            if (!have.count(g)) order.emplace_back(init_hash64(seed, 99, (unsigned long long)g), g);
        const size_t need = (size_t)k - m;
        std::partial_sort(order.begin(), order.begin() + need, order.end());
        // ties of the 64-bit hash are not expected; the (hash, row) pair order makes the rule total anyway
        for (size_t i = 0; i < need; ++i) {
            if (add({order[i].second}, false)) return 1;
        }
        CU_TRY(cudaMemcpyAsync(out, cand, sizeof(float) * (size_t)k * d, cudaMemcpyDeviceToDevice, s));
        CU_TRY(cudaStreamSynchronize(s));
        return 0;
    }
    if (m == (size_t)k) {
        CU_TRY(cudaMemcpyAsync(out, cand, sizeof(float) * (size_t)k * d, cudaMemcpyDeviceToDevice, s));
        CU_TRY(cudaStreamSynchronize(s));
        return 0;
    }
    // weights: rows nearest to each candidate
    unsigned long long* counts = pool.get<unsigned long long>(m);
    if (!counts) return fail("k-means||: device allocation failed");
    CU_TRY(cudaMemsetAsync(counts, 0, sizeof(unsigned long long) * m, s));
    if (n_local > 0) {
        if (b200km_predict(h, X, n_local, d, cand, (int)m, lab, nullptr, s)) return 1;
        launch_label_count(lab, n_local, counts, s);
    }
    if (h->nranks > 1) NCCL_TRY(g_nccl.AllReduce(counts, counts, m, ncclUint64, ncclSum, h->comm, s));
    std::vector<unsigned long long> hc(m);
    CU_TRY(cudaMemcpyAsync(hc.data(), counts, sizeof(unsigned long long) * m, cudaMemcpyDeviceToHost, s));
    CU_TRY(cudaStreamSynchronize(s));
    std::vector<double> hw(m);
    for (size_t j = 0; j < m; ++j) hw[j] = (double)hc[j];
    const int trials = 2 + (int)std::floor(std::log((double)k));
    double* w = pool.get<double>(m);
    double* P = pool.get<double>(m);
    double* d2 = pool.get<double>(m);
    double* Dt = pool.get<double>((size_t)trials * m);
    unsigned char* taken = pool.get<unsigned char>(m);
    int* chosen = pool.get<int>((size_t)k);
    void* state = pool.get<unsigned char>(kpp_state_bytes());
    if (!w || !P || !d2 || !Dt || !taken || !chosen || !state) return fail("k-means||: device allocation failed");
    CU_TRY(cudaMemcpyAsync(w, hw.data(), sizeof(double) * m, cudaMemcpyHostToDevice, s));
    run_weighted_kmeanspp(cand, (int)m, d, k, w, seed, trials > 16 ? 16 : trials, P, d2, Dt, taken, chosen, state, out, s);
    CU_TRY(cudaGetLastError());
    CU_TRY(cudaStreamSynchronize(s));
    return 0;
}

}  // namespace

extern "C" {

const char* b200km_version(void) { return "b200km 0.1.0 (sm_100a)"; }

const char* b200km_last_error(void) { return g_err; }

int b200km_comm_get_unique_id(void* out128) {
    if (!out128) return fail("out128 is null");
    if (load_nccl()) return 1;
    static_assert(sizeof(ncclUniqueId) == B200KM_UNIQUE_ID_BYTES, "NCCL unique id size");
    ncclUniqueId id;
    NCCL_TRY(g_nccl.GetUniqueId(&id));
    memcpy(out128, &id, sizeof(id));
    return 0;
}

int b200km_create(int device, int rank, int nranks, const void* uid128, b200km_handle** out) {
    if (!out) return fail("out is null");
    *out = nullptr;
    if (nranks < 1 || rank < 0 || rank >= nranks) return fail("bad rank %d / nranks %d", rank, nranks);
    int ndev = 0;
    CU_TRY(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return fail("device %d not present (%d visible)", device, ndev);
    CU_TRY(cudaSetDevice(device));
    cudaDeviceProp prop;
    CU_TRY(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10)
        return fail("b200km is built for sm_100a only; device %d is sm_%d%d", device, prop.major, prop.minor);
    b200km_handle* h = new b200km_handle();
    h->device = device;
    h->rank = rank;
    h->nranks = nranks;
    if (cudaMalloc((void**)&h->d_state, sizeof(IterState) * (1 + kRing)) != cudaSuccess ||
        cudaMemset(h->d_state, 0, sizeof(IterState) * (1 + kRing)) != cudaSuccess ||
        cudaStreamCreateWithFlags(&h->side, cudaStreamNonBlocking) != cudaSuccess ||
        cudaMallocHost((void**)&h->h_state, sizeof(IterState) * (kRing + 1)) != cudaSuccess ||
        cudaMallocHost((void**)&h->h_scalar, sizeof(double) * 8) != cudaSuccess) {
        free_ws(h);
        delete h;
        return fail("handle allocation failed: %s", cudaGetErrorString(cudaGetLastError()));
    }
    for (auto& e : h->evm) {
        if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) {
            free_ws(h);
            delete h;
            return fail("cudaEventCreate failed");
        }
    }
    for (auto& e : h->ev) {
        if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) {
            free_ws(h);
            delete h;
            return fail("cudaEventCreate failed");
        }
    }
    if (nranks > 1) {
        if (!uid128) {
            free_ws(h);
            delete h;
            return fail("uid128 is null with nranks=%d", nranks);
        }
        if (load_nccl()) {
            free_ws(h);
            delete h;
            return 1;
        }
        ncclUniqueId id;
        memcpy(&id, uid128, sizeof(id));
        ncclResult_t r = g_nccl.CommInitRank(&h->comm, nranks, id, rank);
        if (r != ncclSuccess) {
            free_ws(h);
            delete h;
            return fail("ncclCommInitRank failed: %s", g_nccl.GetErrorString(r));
        }
    }
    if (nranks > 1 && nranks <= 16) setup_exchange(h);  // failure leaves NCCL in charge of every reduction
    *out = h;
    return 0;
}

int b200km_destroy(b200km_handle* h) {
    if (!h) return 0;
    cudaSetDevice(h->device);
    if (h->comm) g_nccl.CommDestroy(h->comm);
    free_ws(h);
    delete h;
    return 0;
}

int b200km_abort(b200km_handle* h) {
    if (!h) return 0;
    cudaSetDevice(h->device);
    if (h->comm) g_nccl.CommAbort(h->comm);
    free_ws(h);
    delete h;
    return 0;
}

int b200km_set_path(b200km_handle* h, int path) {
    if (!h) return fail("null handle");
    if (path < B200KM_PATH_AUTO || path > B200KM_PATH_SMALL) return fail("unknown path %d", path);
    h->path = path;
    return 0;
}

const char* b200km_last_path(b200km_handle* h) { return h ? h->last_path : "none"; }

const char* b200km_last_reduce(b200km_handle* h) { return h ? h->last_reduce : "none"; }

int b200km_tc_supported(int d, int k, int with_update) {
    return tc_supported(nullptr, d, k, with_update != 0) ? 1 : 0;
}

int b200km_profile_enable(b200km_handle* h, int on) {
    if (!h) return fail("null handle");
    CU_TRY(cudaSetDevice(h->device));
    if (on && !h->pev[0]) {
        for (auto& e : h->pev) CU_TRY(cudaEventCreate(&e));
    }
    h->prof = on != 0;
    h->prof_n = 0;
    h->launches = 0;
    return 0;
}

int b200km_profile_read(b200km_handle* h, double* kernel_ms_total, int* kernel_launches, long long* all_launches) {
    if (!h) return fail("null handle");
    CU_TRY(cudaSetDevice(h->device));
    double total = 0.0;
    for (int i = 0; i < h->prof_n; ++i) {
        CU_TRY(cudaEventSynchronize(h->pev[2 * i + 1]));
        float ms = 0.f;
        CU_TRY(cudaEventElapsedTime(&ms, h->pev[2 * i], h->pev[2 * i + 1]));
        total += ms;
    }
    if (kernel_ms_total) *kernel_ms_total = total;
    if (kernel_launches) *kernel_launches = h->prof_n;
    if (all_launches) *all_launches = (long long)h->launches;
    h->prof_n = 0;
    h->launches = 0;
    return 0;
}

int b200km_screen_stats(b200km_handle* h, long long* refined_rows, int* last_kernel) {
    if (!h) return fail("null handle");
    CU_TRY(cudaSetDevice(h->device));
    IterState st;
    CU_TRY(cudaMemcpy(&st, h->d_state, sizeof(st), cudaMemcpyDeviceToHost));
    if (refined_rows) *refined_rows = (long long)st.screen_refined;
    if (last_kernel) *last_kernel = h->last_kernel;
    return 0;
}

int b200km_query_workspace(int64_t n, int d, int k, size_t* bytes_out) {
    if (!bytes_out) return fail("bytes_out is null");
    if (n < 0 || d <= 0 || k <= 0) return fail("bad shape");
    const size_t kd = (size_t)k * d;
    size_t b = sizeof(double) * (kd + k + 1 + 64) + sizeof(double) * (k + 64) + sizeof(float) * (k + 64) +
               sizeof(float) * (kd + kd / 8 + 64) + 16 * ((size_t)n + (size_t)n / 8 + 256) + sizeof(IterState);
    b += tc_workspace_bytes(d, k) > screen_workspace_bytes(d, k) ? tc_workspace_bytes(d, k) : screen_workspace_bytes(d, k);
    // fused small-k path: per-CTA partial-sum slots and the fixed part of the open-row list (its per-row part, the
    // bitmap and the block counts are within the per-row slack above)
    b += sizeof(double) * (kd + k + 9) * (size_t)(kNumSMs * (1 + kFixCtasPerSm)) + 12 * 65536 + 4096;
    if (!tc_supported(nullptr, d, k, true)) b += gemm_workspace_bytes(n, d, k);
    *bytes_out = b;
    return 0;
}

int b200km_fit(b200km_handle* h, const float* X, int64_t n_local, int d, int k, const float* init_centroids,
               int max_iter, double tol, float* centroids_out, double* inertia_out, int* n_iter_out,
               int32_t* labels_out, void* stream) {
    if (check_shape(h, X, n_local, d, k)) return 1;
    if (n_local == 0) return fail("this rank received no rows (every worker needs data)");
    if (!init_centroids || !centroids_out) return fail("init_centroids / centroids_out is null");
    if (max_iter < 0) return fail("max_iter=%d < 0", max_iter);
    cudaStream_t s = (cudaStream_t)stream;
    CU_TRY(cudaSetDevice(h->device));
    LaunchScope scope(h);
    const size_t kd = (size_t)k * d;
    const int64_t acc_len = (int64_t)kd + k + 1;
    {
        // latency regime: a shard that fits the shared memory of all SMs runs the whole fit in one cooperative kernel
        const bool can_small = h->nranks == 1 && small_fit_supported(n_local, d, k);
        if (h->path == B200KM_PATH_SMALL && !can_small)
            return fail("shared-memory resident fit forced but not available for n=%lld d=%d k=%d (one rank, k*d + k <= 2047, "
                        "the shard within the SMs' shared memory)", (long long)n_local, d, k);
        if (h->path == B200KM_PATH_SMALL || (h->path == B200KM_PATH_AUTO && can_small && kd <= 512)) {
            if (ensure_ws(h, n_local, d, k, false)) return 1;
            const size_t need = small_slots_bytes(n_local, d, k) / sizeof(double);
            if (grow(&h->d_small, &h->small_cap, need)) return 1;
            const bool rec = h->prof && h->prof_n < b200km_handle::kProfMax;
            if (rec) CU_TRY(cudaEventRecord(h->pev[2 * h->prof_n], s));
            double* d_inertia = h->d_acc + kd + k;
            if (small_fit(X, n_local, d, k, init_centroids, max_iter, tol, h->d_small, h->d_C, labels_out,
                          inertia_out ? d_inertia : nullptr, h->d_state, s))
                return fail("resident fit failed: %s", small_last_error());
            if (rec) CU_TRY(cudaEventRecord(h->pev[2 * h->prof_n + 1], s));
            if (rec) h->prof_n++;
            h->last_path = "small";
            h->last_kernel = 4;
            h->last_reduce = "none";
            if (inertia_out) CU_TRY(cudaMemcpyAsync(h->h_scalar, d_inertia, sizeof(double), cudaMemcpyDeviceToHost, s));
            CU_TRY(cudaMemcpyAsync(centroids_out, h->d_C, sizeof(float) * kd, cudaMemcpyDeviceToDevice, s));
            CU_TRY(cudaMemcpyAsync(&h->h_state[kRing], h->d_state, sizeof(IterState), cudaMemcpyDeviceToHost, s));
            CU_TRY(cudaStreamSynchronize(s));
            if (n_iter_out) *n_iter_out = h->h_state[kRing].n_iter;
            if (inertia_out) *inertia_out = h->h_scalar[0];
            return 0;
        }
    }
    int kind = KIND_FFMA;
    if (pick_path(h, X, d, k, true, &kind)) return 1;
    if (ensure_ws(h, n_local, d, k, true)) return 1;

    h->split_only = false;
    // small accumulators: one CTA does the all-reduce (NVLink peer stores, rank-ordered sum) and the centroid update
    const bool fused_tail = finalize_fused_supported(k, d) && (h->nranks == 1 || h->xr_ok) && max_iter < (1 << 20);
    const unsigned int epoch0 = (++h->fit_serial) << 20;
    h->last_reduce = h->nranks == 1 ? "none" : (fused_tail ? "nvlink-oneshot" : "nccl");
    CU_TRY(cudaMemcpyAsync(h->d_C, init_centroids, sizeof(float) * kd, cudaMemcpyDeviceToDevice, s));
    if (kind == KIND_FFMA) launch_rownorm(X, n_local, d, h->d_xnorm, s);
    launch_prep(h->d_C, k, d, h->d_cnorm, h->d_acc, acc_len, h->d_state, true, s);
    CU_TRY(cudaGetLastError());

    for (int it = 0; it < max_iter; ++it) {
        const int slot = it % kRing;
        if (it >= kRing) {
            CU_TRY(cudaEventSynchronize(h->ev[slot]));
            if (h->h_state[slot].done) break;
            // The screen kernel pays a float64 look for every row inside its (wide) TF32 band.  When the
            // passes seen so far re-decided more than 2% of their rows the data has no margin to screen
            // with: the remaining passes run the split-TF32 kernel, whose band is ~200x narrower.
            const double seen = (double)n_local * (double)(it - kRing + 1);
            if (!h->split_only && (double)h->h_state[slot].screen_refined > 0.02 * seen) h->split_only = true;
        }
        if (run_pass(h, kind, X, n_local, d, k, h->d_C, nullptr, nullptr, h->d_acc, true, it == 0, s)) return 1;
        if (fused_tail) {
            launch_finalize_fused(h->d_C, k, d, h->d_cnorm, h->d_acc, tol, h->d_state, h->d_state + 1 + slot,
                                  h->nranks > 1 ? h->xr_peer : nullptr, h->rank, h->nranks, epoch0 + (unsigned int)it + 1u, s);
        } else {
            if (h->nranks > 1)
                NCCL_TRY(g_nccl.AllReduce(h->d_acc, h->d_acc, (size_t)acc_len, ncclFloat64, ncclSum, h->comm, s));
            launch_finalize(h->d_C, k, d, h->d_cnorm, h->d_acc, h->d_shift, h->d_shift + h->k_cap, tol, h->d_state,
                            h->d_state + 1 + slot, s);
        }
        CU_TRY(cudaGetLastError());
        // the host looks at the state kRing iterations later; the copy rides a side stream so that the next
        // pass starts right behind finalize
        CU_TRY(cudaEventRecord(h->evm[slot], s));
        CU_TRY(cudaStreamWaitEvent(h->side, h->evm[slot], 0));
        CU_TRY(cudaMemcpyAsync(&h->h_state[slot], h->d_state + 1 + slot, sizeof(IterState), cudaMemcpyDeviceToHost,
                               h->side));
        CU_TRY(cudaEventRecord(h->ev[slot], h->side));
    }

    double* d_inertia = h->d_acc + kd + k;
    if (inertia_out || labels_out) {
        int32_t* lab = labels_out ? labels_out : h->d_labels;
        if (run_pass(h, kind, X, n_local, d, k, h->d_C, lab, h->d_best, nullptr, false, max_iter == 0, s)) return 1;
        CU_TRY(cudaMemsetAsync(d_inertia, 0, sizeof(double), s));
        launch_sum_best(h->d_best, n_local, d_inertia, s);
        CU_TRY(cudaGetLastError());
        if (h->nranks > 1)
            NCCL_TRY(g_nccl.AllReduce(d_inertia, d_inertia, 1, ncclFloat64, ncclSum, h->comm, s));
        CU_TRY(cudaMemcpyAsync(h->h_scalar, d_inertia, sizeof(double), cudaMemcpyDeviceToHost, s));
    }
    else if (fused_tail && h->nranks > 1 && max_iter > 0) {
        // no collective follows the loop: one tiny all-reduce keeps a fast rank's next fit from writing into an
        // exchange buffer a slow rank is still reading
        CU_TRY(cudaMemsetAsync(d_inertia, 0, sizeof(double), s));
        NCCL_TRY(g_nccl.AllReduce(d_inertia, d_inertia, 1, ncclFloat64, ncclSum, h->comm, s));
    }
    CU_TRY(cudaMemcpyAsync(centroids_out, h->d_C, sizeof(float) * kd, cudaMemcpyDeviceToDevice, s));
    CU_TRY(cudaMemcpyAsync(&h->h_state[kRing], h->d_state, sizeof(IterState), cudaMemcpyDeviceToHost, s));
    CU_TRY(cudaStreamSynchronize(s));
    if (n_iter_out) *n_iter_out = h->h_state[kRing].n_iter;
    if (inertia_out) *inertia_out = h->h_scalar[0];
    return 0;
}

int b200km_lloyd_step(b200km_handle* h, const float* X, int64_t n, int d, int k, const float* centroids,
                      double* sums_out, double* counts_out, double* inertia_out, int32_t* labels_out,
                      void* stream) {
    if (check_shape(h, X, n, d, k)) return 1;
    if (!centroids || !sums_out || !counts_out || !inertia_out) return fail("null output / centroids");
    cudaStream_t s = (cudaStream_t)stream;
    CU_TRY(cudaSetDevice(h->device));
    LaunchScope scope(h);
    int kind = KIND_FFMA;
    if (pick_path(h, X, d, k, true, &kind)) return 1;
    if (ensure_ws(h, n, d, k, true)) return 1;
    const size_t kd = (size_t)k * d;
    if (kind == KIND_FFMA) launch_rownorm(X, n, d, h->d_xnorm, s);
    launch_prep(centroids, k, d, h->d_cnorm, h->d_acc, (int64_t)kd + k + 1, h->d_state, true, s);
    CU_TRY(cudaGetLastError());
    h->split_only = false;
    if (n > 0) {
        int32_t* lab = labels_out ? labels_out : h->d_labels;
        if (run_pass(h, kind, X, n, d, k, centroids, lab, nullptr, h->d_acc, false, true, s, true)) return 1;
        if (h->last_kernel == 2) {
            // The screen kernel produces no per-row distances.  The inertia could be rebuilt from the sums
            // (sum ||x||^2 - 2 sum c.S + sum n ||c||^2), but that form cancels for data far from the origin; one more
            // pass takes exact FP32 differences per row instead, added in a fixed order.
            if (!h->d_red && cudaMalloc((void**)&h->d_red, sizeof(double) * 1040) != cudaSuccess) return fail("cudaMalloc failed");
            unsigned int* ticket = reinterpret_cast<unsigned int*>(h->d_red + 1032);
            CU_TRY(cudaMemsetAsync(ticket, 0, sizeof(unsigned int), s));
            launch_rowdist(X, n, d, centroids, lab, h->d_best, s);
            launch_det_sum(h->d_best, n, h->d_red, h->d_acc + kd + k, ticket, s);
        }
    }
    CU_TRY(cudaMemcpyAsync(sums_out, h->d_acc, sizeof(double) * kd, cudaMemcpyDeviceToDevice, s));
    CU_TRY(cudaMemcpyAsync(counts_out, h->d_acc + kd, sizeof(double) * k, cudaMemcpyDeviceToDevice, s));
    CU_TRY(cudaMemcpyAsync(inertia_out, h->d_acc + kd + k, sizeof(double), cudaMemcpyDeviceToDevice, s));
    return 0;
}

int b200km_predict(b200km_handle* h, const float* X, int64_t n, int d, const float* centroids, int k,
                   int32_t* labels_out, float* min_dist_out, void* stream) {
    if (check_shape(h, X, n, d, k)) return 1;
    if (!centroids) return fail("centroids is null");
    if (n == 0) return 0;
    if (!labels_out) return fail("labels_out is null");
    cudaStream_t s = (cudaStream_t)stream;
    CU_TRY(cudaSetDevice(h->device));
    LaunchScope scope(h);
    int kind = KIND_FFMA;
    if (pick_path(h, X, d, k, false, &kind)) return 1;
    if (ensure_ws(h, n, d, k, true)) return 1;
    if (kind == KIND_FFMA) launch_rownorm(X, n, d, h->d_xnorm, s);
    launch_prep(centroids, k, d, h->d_cnorm, nullptr, 0, h->d_state, false, s);
    CU_TRY(cudaGetLastError());
    h->split_only = false;
    return run_pass(h, kind, X, n, d, k, centroids, labels_out, min_dist_out, nullptr, false, true, s);
}

int b200km_init_centroids(b200km_handle* h, const float* X, int64_t n_local, int d, int k, int mode,
                          uint64_t seed, double oversampling, float* centroids_out, void* stream) {
    if (check_shape(h, X, n_local, d, k)) return 1;
    if (!centroids_out) return fail("centroids_out is null");
    cudaStream_t s = (cudaStream_t)stream;
    CU_TRY(cudaSetDevice(h->device));
    LaunchScope scope(h);
    if (mode == B200KM_INIT_KMEANS_PARALLEL)
        return kmeans_parallel_init(h, X, n_local, d, k, (unsigned long long)seed, oversampling > 0.0 ? oversampling : 2.0,
                                    centroids_out, s);
    if (mode != B200KM_INIT_RANDOM) return fail("unknown init mode %d", mode);
    // every rank learns all shard sizes, draws the same k distinct global rows, and contributes the
    // rows it owns; a float sum all-reduce of the otherwise-zero [k,d] buffer assembles the result.
    std::vector<long long> sizes(h->nranks, 0);
    sizes[h->rank] = n_local;
    if (h->nranks > 1) {
        if (ensure_ws(h, 0, d, k, false)) return 1;
        long long* d_sizes = reinterpret_cast<long long*>(h->d_acc);
        CU_TRY(cudaMemcpyAsync(d_sizes + h->rank, &sizes[h->rank], sizeof(long long), cudaMemcpyHostToDevice, s));
        NCCL_TRY(g_nccl.AllGather(d_sizes + h->rank, d_sizes, 1, ncclInt64, h->comm, s));
        CU_TRY(cudaMemcpyAsync(sizes.data(), d_sizes, sizeof(long long) * h->nranks, cudaMemcpyDeviceToHost, s));
        CU_TRY(cudaStreamSynchronize(s));
    }
    long long total = 0, my_off = 0;
    for (int r = 0; r < h->nranks; ++r) {
        if (r == h->rank) my_off = total;
        total += sizes[r];
    }
    if (total < k) return fail("k=%d exceeds the number of rows %lld", k, total);
    std::mt19937_64 rng(seed);
    std::unordered_set<long long> seen;
    std::vector<long long> pick;
    pick.reserve(k);
    for (long long j = total - k; j < total; ++j) {  // Floyd's distinct sampling
        std::uniform_int_distribution<long long> dist(0, j);
        long long t = dist(rng);
        if (!seen.insert(t).second) {
            seen.insert(j);
            t = j;
        }
        pick.push_back(t);
    }
    std::sort(pick.begin(), pick.end());
    CU_TRY(cudaMemsetAsync(centroids_out, 0, sizeof(float) * (size_t)k * d, s));
    for (int j = 0; j < k; ++j) {
        const long long g = pick[j];
        if (g >= my_off && g < my_off + n_local)
            CU_TRY(cudaMemcpyAsync(centroids_out + (size_t)j * d, X + (size_t)(g - my_off) * d, sizeof(float) * d,
                                   cudaMemcpyDeviceToDevice, s));
    }
    if (h->nranks > 1)
        NCCL_TRY(g_nccl.AllReduce(centroids_out, centroids_out, (size_t)k * d, ncclFloat32, ncclSum, h->comm, s));
    return 0;
}

}  // extern "C"
