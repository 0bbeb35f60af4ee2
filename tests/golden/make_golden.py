"""Regenerates tests/golden/kmeans_golden.npz.

The reference package cannot be imported here (it needs pyspark + a JVM + cuML, all absent), so the
golden vectors are (1) the known answers written in the reference's own tests and docstrings, and
(2) oracle outputs on small seeded blobs, cross-checked at generation time against an independent
implementation (scikit-learn's Lloyd with the same initial centres).  Run from the repo root:

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
from oracle import kmeans_oracle as ko  # noqa: E402

out = {}

# --- known answers held by the reference ------------------------------------------------------
# python/tests/test_kmeans.py:154,177-183  (test_kmeans_basic)
out["basic_X"] = np.array([[1.0, 1.0], [1.0, 2.0], [3.0, 2.0], [4.0, 3.0]], dtype=np.float32)
out["basic_centers_sorted"] = np.array([[1.0, 1.5], [3.5, 2.5]], dtype=np.float64)
# python/tests/test_kmeans.py:351-356,393-409 (test_kmeans_spark_compat) and the class docstring
# python/src/spark_rapids_ml/clustering.py:224-254
out["compat_X"] = np.array([[0.0, 0.0], [1.0, 1.0], [9.0, 8.0], [8.0, 9.0]], dtype=np.float32)
out["compat_centers_sorted"] = np.array([[0.5, 0.5], [8.5, 8.5]], dtype=np.float64)
# python/tests/test_kmeans.py:236-242 (integer typed columns; must simply fit with k=2)
out["numeric_X"] = np.array(
    [[1, 4, 4, 4, 0], [2, 2, 2, 2, 1], [3, 3, 3, 2, 2], [3, 3, 3, 2, 3], [5, 2, 1, 3, 4]], dtype=np.float32
)

# --- oracle outputs on seeded blobs, cross-checked against scikit-learn -------------------------
from sklearn.cluster import KMeans as SkKMeans  # noqa: E402


def blob_case(tag, n, d, k, seed, max_iter):
    X, _ = ko.make_blobs(n, d, k, seed=seed)
    init = ko.random_init(X, k, seed=seed)
    res = ko.lloyd_fit(X, init, max_iter=max_iter, tol=1e-20)
    sk = SkKMeans(n_clusters=k, init=init.astype(np.float64), n_init=1, max_iter=res.n_iter, tol=0.0,
                  algorithm="lloyd").fit(X.astype(np.float64))
    # sklearn relocates empty clusters, so only compare when none emptied
    counts = np.bincount(res.labels, minlength=k)
    if counts.min() > 0:
        assert np.allclose(sk.cluster_centers_, res.centers, rtol=1e-4, atol=1e-5), tag
        amb = ko.ambiguous_rows(res.best, res.second)
        assert np.array_equal(sk.labels_[~amb], res.labels[~amb]), tag
        assert abs(sk.inertia_ - res.inertia) <= 1e-5 * res.inertia, tag
        print(tag, "cross-checked against sklearn", sk.n_iter_, res.n_iter)
    else:
        print(tag, "has an empty cluster; sklearn cross-check skipped")
    out[f"{tag}_X"] = X
    out[f"{tag}_init"] = init
    out[f"{tag}_centers"] = res.centers
    out[f"{tag}_labels"] = res.labels
    out[f"{tag}_inertia"] = np.float64(res.inertia)
    out[f"{tag}_n_iter"] = np.int32(res.n_iter)
    out[f"{tag}_best"] = res.best
    out[f"{tag}_second"] = res.second


blob_case("blobs1000x20k8", 1000, 20, 8, seed=0, max_iter=20)   # shape of test_kmeans.py:268-278
blob_case("blobs3000x100k20", 3000, 100, 20, seed=1, max_iter=10)  # C2's k and d, small n
blob_case("blobs2000x64k100", 2000, 64, 100, seed=2, max_iter=5)   # C4's k and d
blob_case("blobs777x37k5", 777, 37, 5, seed=3, max_iter=8)         # ragged: odd n, d % 4 != 0

np.savez_compressed(os.path.join(os.path.dirname(__file__), "kmeans_golden.npz"), **out)
print("written", sorted(out))
