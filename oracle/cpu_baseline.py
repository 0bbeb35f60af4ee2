"""Timed CPU baseline for bench.py (`cpu_baseline` leg and `--impl reference`).  TEST/BENCH INFRASTRUCTURE.

The north star asks for pyspark.ml.clustering.KMeans in Spark local[N] beside the GPU number; pyspark
and a JVM are absent from this image (probed again at run time by bench.py), and the reference's own
code for this path is a call into the un-vendored cuML.  What is timed instead is this repository's
oracle port of the Lloyd iteration (oracle/kmeans_oracle.py), run over row blocks on all host cores.
"""
from __future__ import annotations

import os
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

from . import kmeans_oracle as ko


def host_threads() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def lloyd_iterations(X: np.ndarray, init: np.ndarray, iters: int, threads: int):
    """`iters` Lloyd iterations, rows split into `threads` blocks processed concurrently
    (NumPy releases the GIL inside BLAS and the large ufunc loops)."""
    try:
        from threadpoolctl import threadpool_limits
    except Exception:  # pragma: no cover
        threadpool_limits = None
    C = np.asarray(init, dtype=np.float32).copy()
    k = C.shape[0]
    bounds = np.linspace(0, X.shape[0], threads + 1).astype(np.int64)
    blocks = [(int(bounds[i]), int(bounds[i + 1])) for i in range(threads) if bounds[i + 1] > bounds[i]]

    def work(b):
        s, c, inertia, _ = ko.lloyd_step(X[b[0]:b[1]], C)
        return s, c, inertia

    ctx = threadpool_limits(limits=1) if threadpool_limits else None
    try:
        with ThreadPoolExecutor(max_workers=threads) as pool:
            for _ in range(iters):
                sums = np.zeros((k, X.shape[1]))
                counts = np.zeros(k)
                for s, c, _ in pool.map(work, blocks):
                    sums += s
                    counts += c
                C, _ = ko.finalize(sums, counts, C)
    finally:
        if ctx is not None:
            ctx.__exit__(None, None, None)
    return C


def time_fit(n: int, d: int, k: int, iters: int, seed: int = 1, threads: int | None = None):
    """Returns (rows*iters/s, seconds, threads) for `iters` iterations on an n-row blobs sample."""
    threads = threads or host_threads()
    X, _ = ko.make_blobs(n, d, k, seed=seed)
    init = ko.random_init(X, k, seed=seed)
    lloyd_iterations(X[: max(threads * 1024, 4096)], init, 1, threads)  # warm the pools
    t0 = time.perf_counter()
    lloyd_iterations(X, init, iters, threads)
    dt = time.perf_counter() - t0
    return n * iters / dt, dt, threads
