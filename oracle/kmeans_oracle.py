"""CPU oracle for the KMeans Lloyd fit / predict hot path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this module, and only as the checker or the
timed CPU baseline.  The product path (``spark_rapids_ml_nai_b200``) never imports it.

What it restates
----------------
The reference (``/root/reference``, spark-rapids-ml 24.06 fork) holds *no arithmetic* for
this path: ``KMeans._get_cuml_fit_func`` calls ``cuml.cluster.kmeans_mg.KMeansMG.fit``
(python/src/spark_rapids_ml/clustering.py:348-386) and ``KMeansModel._get_cuml_transform_func``
calls ``KMeansMG.predict`` (clustering.py:466-499).  The algorithm therefore lives in the
un-vendored dependency **cuML / RAFT 24.06** (``cuml-cu11~=24.6.0``; ci/Dockerfile:40-41,
docker/Dockerfile.pip:21,38-42, python/README.md:11-13), whose published Lloyd semantics
are restated here in NumPy float64:

* unit sample weights                      (clustering.py:367-370  ``sample_weight=None``)
* squared Euclidean distance only          (clustering.py:71  ``distanceMeasure: None``)
* float32 data and float32-valued centres  (core.py:463,481-486; clustering.py:383 ``dtype``)
* argmin with lowest-index tie-break       (RAFT fusedL2NN min-reduce, KVP ``<`` on value then key)
* mean update; a cluster that received no rows keeps its previous centre
* stop when  sum_j ||c_j_new - c_j_old||^2 < tol  (absolute, checked after the update) or at
  ``max_iter``; ``n_iter`` counts executed iterations
* ``tol == 0`` is mapped by the caller to ``np.finfo('float32').tiny``   (clustering.py:96-106)
* inertia = sum_i ||x_i - c[label_i]||^2 in float64 against the *final* centres, the
  reference's own definition in its benchmark scorer (python/benchmark/benchmark/bench_kmeans.py:59-113)
* labels are int32                         (clustering.py:459-461 ``"int"``)

Pinning status
--------------
The reference's tests hold three known answers for this path, all reproduced by
``tests/test_oracle.py``:  test_kmeans.py:154,177-183 (centres [1,1.5],[3.5,2.5]),
test_kmeans.py:351-356,393-409 (centres exactly [[0.5,0.5],[8.5,8.5]], paired labels) and the
docstring example clustering.py:224-254.  The slow 1000x20 blobs test (test_kmeans.py:268-330)
pins multi-GPU vs single-GPU *self-consistency* of cuML, not absolute values; it is restated
as "sharded fit == unsharded fit".  An independent implementation (scikit-learn's Lloyd with a
given init) cross-checks the restatement where no cluster empties.  The reference cannot be
imported or run here (pyspark, JVM and cuML are absent), so at every BASELINE.json shape the
oracle is **parity unpinned**: it is a restatement, checked only against the toy vectors above.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Tuple

import numpy as np

_ROW_BLOCK = 65536


@dataclass
class FitResult:
    centers: np.ndarray  # [k, d] float32
    labels: np.ndarray  # [n] int32, assignment against the final centres
    inertia: float  # float64 sum of squared distances against the final centres
    n_iter: int
    best: np.ndarray  # [n] float64 squared distance to the nearest final centre
    second: np.ndarray  # [n] float64 squared distance to the second nearest (inf when k == 1)


def _block_sqdist(Xb: np.ndarray, C: np.ndarray, cn: np.ndarray) -> np.ndarray:
    """||x||^2 - 2 x.c + ||c||^2 in float64 for one block of rows (clamped at 0)."""
    xn = np.einsum("ij,ij->i", Xb, Xb)
    D = Xb @ C.T
    D *= -2.0
    D += xn[:, None]
    D += cn[None, :]
    np.maximum(D, 0.0, out=D)
    return D


def assign(
    X: np.ndarray, C: np.ndarray, want_second: bool = True
) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """Nearest-centre assignment (the step behind KMeansMG.predict, clustering.py:493-497).

    Returns (labels int32, best float64, second float64).  Ties go to the lowest index
    (``np.argmin`` returns the first occurrence, the same rule as the backend's min-reduce).
    """
    X = np.asarray(X)
    C64 = np.asarray(C, dtype=np.float64)
    n = X.shape[0]
    k = C64.shape[0]
    cn = np.einsum("ij,ij->i", C64, C64)
    labels = np.empty(n, dtype=np.int32)
    best = np.empty(n, dtype=np.float64)
    second = np.full(n, np.inf, dtype=np.float64)
    for s in range(0, n, _ROW_BLOCK):
        e = min(n, s + _ROW_BLOCK)
        D = _block_sqdist(X[s:e].astype(np.float64), C64, cn)
        lab = np.argmin(D, axis=1)
        rows = np.arange(e - s)
        labels[s:e] = lab
        best[s:e] = D[rows, lab]
        if want_second and k > 1:
            D[rows, lab] = np.inf
            second[s:e] = D.min(axis=1)
    return labels, best, second


def predict(X: np.ndarray, C: np.ndarray) -> np.ndarray:
    """int32 labels, as ``KMeansModel.transform`` yields them (clustering.py:459-461,493-497)."""
    return assign(X, C, want_second=False)[0]


def inertia(X: np.ndarray, C: np.ndarray, labels: np.ndarray) -> float:
    """float64 sum ||x - c[label]||^2, the reference scorer (bench_kmeans.py:96-100)."""
    C64 = np.asarray(C, dtype=np.float64)
    total = 0.0
    for s in range(0, X.shape[0], _ROW_BLOCK):
        e = min(X.shape[0], s + _ROW_BLOCK)
        diff = X[s:e].astype(np.float64) - C64[labels[s:e]]
        total += float(np.sum(diff * diff))
    return total


def accumulate(
    X: np.ndarray, labels: np.ndarray, k: int
) -> Tuple[np.ndarray, np.ndarray]:
    """Per-cluster float64 sums [k,d] and counts [k] (the reduce-by-key step of one iteration)."""
    d = X.shape[1]
    sums = np.zeros((k, d), dtype=np.float64)
    counts = np.bincount(labels, minlength=k).astype(np.float64)
    order = np.argsort(labels, kind="stable")
    sorted_labels = labels[order]
    bounds = np.searchsorted(sorted_labels, np.arange(k + 1))
    for j in range(k):
        lo, hi = bounds[j], bounds[j + 1]
        if hi > lo:
            sums[j] = X[order[lo:hi]].astype(np.float64).sum(axis=0)
    return sums, counts


def lloyd_step(
    X: np.ndarray, C: np.ndarray
) -> Tuple[np.ndarray, np.ndarray, float, np.ndarray]:
    """One assign+accumulate pass: (sums f64 [k,d], counts f64 [k], inertia vs C, labels)."""
    labels, best, _ = assign(X, C, want_second=False)
    sums, counts = accumulate(X, labels, C.shape[0])
    return sums, counts, float(best.sum()), labels


def finalize(
    sums: np.ndarray, counts: np.ndarray, C_old: np.ndarray
) -> Tuple[np.ndarray, float]:
    """Mean update; zero-count clusters keep their previous centre; float32-valued result.

    Returns (new centres float32 [k,d], shift^2 = sum_j ||new_j - old_j||^2 in float64).
    """
    C_old32 = np.asarray(C_old, dtype=np.float32)
    newC = C_old32.astype(np.float64).copy()
    nz = counts > 0
    newC[nz] = sums[nz] / counts[nz, None]
    newC32 = newC.astype(np.float32)
    diff = newC32.astype(np.float64) - C_old32.astype(np.float64)
    return newC32, float(np.sum(diff * diff))


def lloyd_fit(
    X: np.ndarray,
    init_centers: np.ndarray,
    max_iter: int = 20,
    tol: float = 1e-4,
    shards: Optional[int] = None,
) -> FitResult:
    """Lloyd fit from given initial centres (the loop behind KMeansMG.fit, clustering.py:367-370).

    ``shards`` splits the rows into that many contiguous blocks whose partial sums are added
    (float64) before the update - the row-sharded multi-GPU schedule (core.py:651-652,700).
    """
    X = np.ascontiguousarray(X)
    C = np.asarray(init_centers, dtype=np.float32).copy()
    k = C.shape[0]
    n_iter = 0
    for _ in range(int(max_iter)):
        if shards and shards > 1:
            parts = np.array_split(np.arange(X.shape[0]), shards)
            sums = np.zeros((k, X.shape[1]), dtype=np.float64)
            counts = np.zeros(k, dtype=np.float64)
            for p in parts:
                s, c, _, _ = lloyd_step(X[p[0] : p[-1] + 1], C)
                sums += s
                counts += c
        else:
            sums, counts, _, _ = lloyd_step(X, C)
        C, shift2 = finalize(sums, counts, C)
        n_iter += 1
        if shift2 < tol:
            break
    labels, best, second = assign(X, C, want_second=True)
    return FitResult(C, labels, float(best.sum()), n_iter, best, second)


def ambiguous_rows(best: np.ndarray, second: np.ndarray, rel: float = 1e-5) -> np.ndarray:
    """Mask of rows whose best and second-best squared distances differ by < ``rel`` relative.

    BASELINE.json's parity rule exempts exactly these rows from bit-exact label comparison.
    """
    denom = np.maximum(best, np.finfo(np.float64).tiny)
    return (second - best) < rel * denom


def make_blobs(
    n: int, d: int, k: int, seed: int = 1, cluster_std: float = 1.0, box: float = 10.0
) -> Tuple[np.ndarray, np.ndarray]:
    """Gaussian blobs following the reference generator's recipe (python/benchmark/
    gen_data_distributed.py:123-137,152-157): centres ~ U(-box, box)^d, isotropic std, float32.
    NumPy's Generator replaces sklearn's RandomState, so values differ from sklearn's."""
    rng = np.random.default_rng(seed)
    centers = rng.uniform(-box, box, size=(k, d))
    which = rng.integers(0, k, size=n)
    X = centers[which] + cluster_std * rng.standard_normal(size=(n, d))
    return X.astype(np.float32), centers.astype(np.float32)


def random_init(X: np.ndarray, k: int, seed: int = 1) -> np.ndarray:
    """k distinct rows (``initMode="random"``, the reference benchmark setting run_benchmark.sh:164)."""
    rng = np.random.default_rng(seed)
    idx = rng.choice(X.shape[0], size=k, replace=False)
    return np.asarray(X[np.sort(idx)], dtype=np.float32).copy()


# ---- k-means|| ("scalable k-means++") initialisation -------------------------------------------------------------
# The reference hands initMode="k-means||" to the backend as init="scalable-k-means++" with oversampling_factor
# (python/src/spark_rapids_ml/clustering.py:72,117-119; the default init mode, :290).  The backend's algorithm is
# Bahmani et al.'s: oversample candidates in a few rounds with probability l*d^2(x)/phi, weight them by the rows
# they attract, reduce them to k seeds with weighted k-means++.  Its random stream is the backend's own and is
# not reproducible here, so the restatement fixes one: a counter-based hash keyed by (seed, stream, global row
# index) - the same keying the CUDA path uses, which makes the result independent of how rows are sharded.
_M64 = (1 << 64) - 1


def _mix64(z):
    with np.errstate(over="ignore"):
        z = np.asarray(z, dtype=np.uint64) + np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def hash64(seed: int, stream: int, counter) -> np.ndarray:
    """Counter-based 64-bit hash (splitmix64 finaliser, three rounds): the uniform of (seed, stream, counter)."""
    s = _mix64(np.uint64(seed & _M64))
    return _mix64(_mix64(s ^ np.uint64(stream)) ^ np.asarray(counter, dtype=np.uint64))


def hash_uniform(seed: int, stream: int, counter) -> np.ndarray:
    """float64 in [0, 1) from the top 53 bits of hash64."""
    return (hash64(seed, stream, counter) >> np.uint64(11)).astype(np.float64) * (2.0 ** -53)


def _sq_dists_to(X: np.ndarray, C: np.ndarray) -> np.ndarray:
    """min_j ||x_i - c_j||^2 in float64 (blocked)."""
    out = np.empty(X.shape[0], dtype=np.float64)
    C64 = C.astype(np.float64)
    cn = np.einsum("ij,ij->i", C64, C64)
    for s in range(0, X.shape[0], _ROW_BLOCK):
        e = min(X.shape[0], s + _ROW_BLOCK)
        out[s:e] = _block_sqdist(X[s:e].astype(np.float64), C64, cn).min(axis=1)
    return out


def kmeans_parallel_candidates(X: np.ndarray, k: int, seed: int, oversampling: float = 2.0, max_rounds: int = 8):
    """Global row indices of the k-means|| candidates, in the order they are collected (first the start row,
    then every round's picks in ascending row order)."""
    n = X.shape[0]
    first = int(hash64(seed, 0, 0) % np.uint64(n))
    cands = [first]
    dist = _sq_dists_to(X, X[first:first + 1])
    phi = float(dist.sum())
    ell = max(1.0, float(oversampling) * k)
    rounds = int(min(max_rounds, max(1, np.ceil(np.log(max(phi, 2.0))))))
    rows = np.arange(n, dtype=np.uint64)
    for r in range(rounds):
        if phi <= 0.0:
            break
        u = hash_uniform(seed, r + 1, rows)
        ids = np.nonzero(u < ell * dist / phi)[0]
        if ids.size == 0:
            continue
        cands.extend(int(i) for i in ids)
        dist = np.minimum(dist, _sq_dists_to(X, X[ids]))
        phi = float(dist.sum())
    return cands


def kmeans_parallel_init(X: np.ndarray, k: int, seed: int, oversampling: float = 2.0) -> np.ndarray:
    """k initial centres [k, d] float32 (rows of X) by k-means|| + greedy weighted k-means++ over the candidates."""
    n = X.shape[0]
    if n < k:
        raise ValueError(f"k={k} exceeds the number of rows {n}")
    cands = kmeans_parallel_candidates(X, k, seed, oversampling)
    m = len(cands)
    if m < k:  # tiny or degenerate data: top up with further distinct rows, lowest hash first
        have = set(cands)
        order = np.argsort(hash64(seed, 99, np.arange(n, dtype=np.uint64)), kind="stable")
        for i in order:
            if int(i) not in have:
                cands.append(int(i))
                have.add(int(i))
                if len(cands) == k:
                    break
        return np.ascontiguousarray(X[cands], dtype=np.float32)
    if m == k:
        return np.ascontiguousarray(X[cands], dtype=np.float32)
    C32 = np.ascontiguousarray(X[cands], dtype=np.float32)
    w = np.bincount(assign(X, C32, want_second=False)[0], minlength=m).astype(np.float64)
    C = C32.astype(np.float64)
    trials = 2 + int(np.log(k))

    def pick(P, u):
        return int(min(m - 1, np.searchsorted(P, u * P[-1], side="right")))

    idx = pick(np.cumsum(w), float(hash_uniform(seed, 200, 0)))
    chosen = [idx]
    d2 = ((C - C[idx]) ** 2).sum(axis=1)
    taken = np.zeros(m, dtype=bool)
    taken[idx] = True
    for s in range(1, k):
        P = np.cumsum(w * d2)
        if P[-1] <= 0.0:
            idx = int(np.nonzero(~taken)[0][0])
            dnew = ((C - C[idx]) ** 2).sum(axis=1)
        else:
            tries = [pick(P, float(hash_uniform(seed, 1000 + s, t))) for t in range(trials)]
            best_pot, idx, dnew = None, -1, None
            for t in tries:
                dt = ((C - C[t]) ** 2).sum(axis=1)
                pot = float((w * np.minimum(d2, dt)).sum())
                if best_pot is None or pot < best_pot:
                    best_pot, idx, dnew = pot, t, dt
        chosen.append(idx)
        taken[idx] = True
        d2 = np.minimum(d2, dnew)
    return np.ascontiguousarray(C32[chosen], dtype=np.float32)
