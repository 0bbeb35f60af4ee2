/*
 * b200km.h - C ABI of the B200-native KMeans (Lloyd fit / predict) backend.
 *
 * This is the thin layer the Python plugin slots call in place of cuML's KMeansMG.  The
 * reference has no C interface for this path: its fit-function slot and transform slot are
 * Python closures that call into cuML (reference: python/src/spark_rapids_ml/clustering.py:348-386
 * and :466-499) and its communicator is created by python/src/spark_rapids_ml/common/
 * cuml_context.py:35-167.  Every entry point below names the reference call site it replaces.
 *
 * Conventions
 *  - every function returns 0 on success, non-zero on failure; b200km_last_error() then returns a
 *    thread-local, NUL-terminated message.  No C++ exception crosses this boundary.
 *  - handles are opaque, one per (process, GPU), not thread-safe.
 *  - all `const float* X`, centroid, label and accumulator pointers are DEVICE pointers owned by
 *    the caller (e.g. torch tensors' data_ptr()); scalars returned through `double*`/`int*`
 *    marked "host" are written to host memory after the stream has been synchronised.
 *  - `stream` is a cudaStream_t passed as void* (0 = the legacy default stream).  All work is
 *    enqueued on it.
 *  - X is row-major [n, d] float32, centroids row-major [k, d] float32, labels int32.
 *  - the library allocates only its own workspace (cudaMalloc at first use, grown on demand and
 *    released by b200km_destroy).
 */
#ifndef B200KM_H_
#define B200KM_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b200km_handle b200km_handle;

#define B200KM_UNIQUE_ID_BYTES 128

/* kernel path selectors for b200km_set_path (diagnostics / tests) */
#define B200KM_PATH_AUTO 0
#define B200KM_PATH_FFMA 1 /* generic SIMT FFMA tiles + separate update pass */
#define B200KM_PATH_TC 2   /* tcgen05/TMEM/TMA fused pass (where the shape allows): the TF32 screen kernel with
                              float64 re-decision for passes that need no per-row distance, else split-TF32 */
#define B200KM_PATH_TC_SPLIT 3 /* tcgen05 path, split-TF32 kernel for every pass */
#define B200KM_PATH_GEMM 4     /* tcgen05 tile GEMM on a CTA pair (any k, d % 4 == 0): centroids streamed in N-tiles,
                                  TF32 screen + exact fix-up, sums by a separate pass; AUTO takes it when k*d is
                                  beyond the fused kernels' shared memory */

#define B200KM_PATH_SMALL 5    /* b200km_fit only, one rank: the shard stays in the shared memory of all SMs and one
                                  cooperative kernel runs every Lloyd iteration (float64 distances, one grid barrier
                                  per iteration); AUTO takes it for small shards with k*d <= 512 */

/* init modes for b200km_init_centroids; reference: clustering.py:72 ("initMode" -> "init"),
 * values "random" and "k-means||" (clustering.py:185-186, tests/test_kmeans.py:84,93). */
#define B200KM_INIT_RANDOM 0
#define B200KM_INIT_KMEANS_PARALLEL 1

/* Library version string, e.g. "b200km 0.1.0 (sm_100a)". */
const char* b200km_version(void);

/* Message for the last failing call on this thread. */
const char* b200km_last_error(void);

/* Rank 0 creates the 128-byte NCCL unique id that every rank passes to b200km_create.
 * Replaces `nccl.get_unique_id()` in cuml_context.py:75-77. */
int b200km_comm_get_unique_id(void* out128);

/* Creates the per-GPU handle: selects `device`, and when nranks > 1 joins the NCCL communicator
 * (ncclCommInitRank) identified by uid128.  nranks == 1 needs no NCCL and uid128 may be NULL.
 * Replaces `Handle(n_streams=0)` + `nccl().init(nranks, uid, rank)` +
 * `inject_comms_on_handle_coll_only` in cuml_context.py:70,116-124. */
int b200km_create(int device, int rank, int nranks, const void* uid128, b200km_handle** out);

/* Orderly teardown (ncclCommDestroy + workspace free).  Replaces `nccl.destroy()` in
 * cuml_context.py:155-157. */
int b200km_destroy(b200km_handle* h);

/* Teardown after a failure on some rank (ncclCommAbort: does not wait for peers).  Replaces
 * `nccl.abort()` in cuml_context.py:158-159. */
int b200km_abort(b200km_handle* h);

/* Forces a kernel path (B200KM_PATH_*); AUTO picks per shape.  Asking for a path the shape does
 * not support makes the next compute call fail rather than fall back. */
int b200km_set_path(b200km_handle* h, int path);

/* Name of the path the last compute call on this handle used ("ffma" / "tc" / "gemm" / "small"). */
const char* b200km_last_path(b200km_handle* h);

/* How the per-iteration sums of the last b200km_fit on this handle crossed the GPUs: "none" (one rank),
 * "nvlink-oneshot" (small accumulators: every rank stores its [k*d | k | 1] float64 block into its slot of every
 * peer's exchange buffer - CUDA IPC mappings set up by b200km_create - and adds the slots in rank order inside the
 * fused finalize kernel) or "nccl" (ncclAllReduce: large accumulators, or when the peer mappings could not be made).
 * The reference's two FP32 all-reduces per iteration happen inside cuML; diagnostics only. */
const char* b200km_last_reduce(b200km_handle* h);

/* 1 when the fused tcgen05 kernel can run this (d, k): with_update != 0 asks for the fit pass
 * (assignment + float64 accumulation), 0 for the assignment-only pass.  Diagnostics; no counterpart
 * in the reference. */
int b200km_tc_supported(int d, int k, int with_update);

/* Instrumentation for the benchmark.  While enabled, every pass brackets its dominant kernel
 * (the fused tcgen05 kernel, or the FFMA assignment kernel) with a CUDA event pair on the caller's
 * stream and the handle counts its kernel launches.  b200km_profile_read waits for the recorded
 * events, returns the summed kernel time, the number of bracketed launches and the number of all
 * kernel launches since the last read, and resets the counters.  The reference has only wall-clock
 * prints for this path (python/benchmark/benchmark/bench_kmeans.py:166-174, utils.py:42-50). */
int b200km_profile_enable(b200km_handle* h, int on);
int b200km_profile_read(b200km_handle* h, double* kernel_ms_total, int* kernel_launches,
                        long long* all_launches);

/* Diagnostics of the tcgen05 path (synchronises the device): refined_rows = rows the screen kernel re-decided
 * in float64 since the last centroid preparation (one fit, or one lloyd_step / predict call); last_kernel =
 * kernel of the last pass (0 FFMA kernels, 1 split-TF32 fused kernel, 2 screen kernel).  No counterpart in
 * the reference. */
int b200km_screen_stats(b200km_handle* h, long long* refined_rows, int* last_kernel);

/* Bytes of library-owned device workspace a fit/predict of this shape needs. */
int b200km_query_workspace(int64_t n, int d, int k, size_t* bytes_out);

/* Lloyd fit from given initial centroids (identical on all ranks) over this rank's row shard.
 * Per iteration: nearest-centroid assignment, per-cluster float64 sums/counts, one NCCL sum
 * all-reduce of [k*d sums | k counts | inertia] when nranks > 1, mean update (empty clusters keep
 * their centroid), stop when sum_j ||c_new_j - c_old_j||^2 < tol or after max_iter iterations.
 * centroids_out: device [k,d].  labels_out: device [n_local] or NULL.  inertia_out / n_iter_out:
 * host scalars or NULL.  inertia is the float64 sum of squared distances of all ranks' rows to
 * the FINAL centroids (costs one extra assignment pass, skipped when both inertia_out and
 * labels_out are NULL).
 * Replaces `KMeansMG(handle, **cuml_params).fit(X, sample_weight=None)` and the read-back of
 * `cluster_centers_`, `n_iter_`, `inertia_` in clustering.py:355-384. */
int b200km_fit(b200km_handle* h, const float* X, int64_t n_local, int d, int k,
               const float* init_centroids, int max_iter, double tol, float* centroids_out,
               double* inertia_out, int* n_iter_out, int32_t* labels_out, void* stream);

/* One assignment + accumulation pass over this rank's rows (no all-reduce, no update):
 * sums_out device float64 [k*d], counts_out device float64 [k], inertia_out device float64 [1]
 * (sum of squared distances to `centroids`), labels_out device int32 [n] or NULL.
 * The inertia comes from per-row distances (FP32 differences on the fused small-k path, which costs it one more
 * pass over X), never from a cancelling sum-of-squares identity.
 * Building block of b200km_fit exposed for tests and the benchmark; it is the per-iteration
 * device work of `KMeansMG.fit` (clustering.py:367-370). */
int b200km_lloyd_step(b200km_handle* h, const float* X, int64_t n, int d, int k,
                      const float* centroids, double* sums_out, double* counts_out,
                      double* inertia_out, int32_t* labels_out, void* stream);

/* Nearest-centroid labels (int32) for n rows; min_dist_out (device float32 [n], squared distance
 * to the chosen centroid) may be NULL.  No communication.
 * Replaces `kmeans.predict(df, normalize_weights=False)` in clustering.py:493-497. */
int b200km_predict(b200km_handle* h, const float* X, int64_t n, int d, const float* centroids,
                   int k, int32_t* labels_out, float* min_dist_out, void* stream);

/* Chooses k initial centroids from the (row-sharded) data and leaves the same [k,d] array on
 * every rank in `centroids_out` (device).  Collective: every rank calls it with its shard.
 * mode B200KM_INIT_RANDOM: k distinct global rows (Floyd sampling, mt19937_64(seed)); `oversampling` is ignored.
 * mode B200KM_INIT_KMEANS_PARALLEL: k-means|| (scalable k-means++), all on the device: one random start row, then
 * up to 8 rounds that pick every row with probability oversampling*k*d^2(x)/phi (oversampling = cuML's
 * oversampling_factor, clustering.py:119; <= 0 means 2.0), the picks gathered over NCCL, the candidates weighted by
 * the rows they attract and reduced to k seeds by greedy weighted k-means++.  Its random numbers are a
 * counter-based hash of (seed, stream, GLOBAL row index): the result is the same for every sharding of the rows.
 * The returned centroids are rows of X.  Synchronises `stream` several times.
 * Replaces the `init=` handling inside `KMeansMG.fit` (clustering.py:72,117,355-359). */
int b200km_init_centroids(b200km_handle* h, const float* X, int64_t n_local, int d, int k,
                          int mode, uint64_t seed, double oversampling, float* centroids_out,
                          void* stream);

/* Nearest-centroid labels for callers that cannot allocate device memory: a transform task whose GPU
 * memory is already owned by another pool, a latency-bound serving thread (cudaMalloc may serialise the
 * device), or anything that must not fail for lack of memory.  It needs no handle and calls no
 * cudaMalloc / cudaFree or allocating library routine: its scratch (row and centroid norms, distances,
 * the near-tie work list) is a fixed array inside the loaded module, held by every GPU that loads the
 * library until the process exits (about 0.8 MB).  Limits: n <= B200KM_STATIC_MAX_ROWS rows and
 * k <= B200KM_STATIC_MAX_K centroids per call; larger sizes are refused with an error (split the batch).
 * X, centroids and labels_out are device pointers on the CURRENT device; work is enqueued on `stream`.
 * The first call on each device resolves the array's address (cudaGetSymbolAddress) - make that call
 * before any stream capture begins.  All calls share the one array, which keeps the last call's
 * contents: the entry point clears the part that has to start at zero, but two calls that overlap on the
 * GPU (different streams, or threads) race on it - serialise them.  Labels follow the same rule as
 * b200km_predict (FP32 scores, near-ties re-decided in float64, lowest index wins ties).
 * This entry point has no counterpart in the reference. */
#define B200KM_STATIC_MAX_ROWS 65536
#define B200KM_STATIC_MAX_K 1024
int b200km_predict_static(const float* X, int64_t n, int d, const float* centroids, int k,
                          int32_t* labels_out, void* stream);

/* Message for the last failing b200km_predict_static call on this thread.  No counterpart in the
 * reference. */
const char* b200km_static_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* B200KM_H_ */
