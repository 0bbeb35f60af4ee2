"""Pins the CPU oracle against every known answer the reference's tests hold for the KMeans path,
against the committed golden vectors, and against an independent implementation (scikit-learn)."""
import numpy as np
import pytest

from oracle import kmeans_oracle as ko


def _sorted(c):
    return np.array(sorted(np.asarray(c, dtype=np.float64).tolist()))


def test_reference_known_answer_basic(golden):
    # python/tests/test_kmeans.py:154,177-183: sorted centres ~ [1,1.5],[3.5,2.5] rel 1e-3; labels pair up
    X = golden["basic_X"]
    for init in ([0, 2], [0, 3], [1, 2], [1, 3]):
        res = ko.lloyd_fit(X, X[init], max_iter=20, tol=float(np.finfo("float32").tiny))
        assert _sorted(res.centers) == pytest.approx(golden["basic_centers_sorted"], rel=1e-3)
        lab = res.labels
        assert lab[0] == lab[1] and lab[1] != lab[2] and lab[2] == lab[3]
        assert res.centers.dtype == np.float32 and res.labels.dtype == np.int32


def test_reference_known_answer_spark_compat(golden):
    # python/tests/test_kmeans.py:351-356,393-409 and clustering.py:224-254: exactly [[.5,.5],[8.5,8.5]]
    X = golden["compat_X"]
    for init in ([0, 2], [1, 3], [0, 3]):
        res = ko.lloyd_fit(X, X[init], max_iter=10, tol=1e-4)
        assert _sorted(res.centers).tolist() == golden["compat_centers_sorted"].tolist()
        assert res.labels[0] == res.labels[1] and res.labels[2] == res.labels[3]
        assert res.inertia == pytest.approx(2.0)  # 4 points, each 0.5 away in both coordinates


def test_reference_numeric_columns_fit(golden):
    # python/tests/test_kmeans.py:232-249: integer-typed columns only have to fit with k=2
    X = golden["numeric_X"]
    res = ko.lloyd_fit(X, X[[0, 4]], max_iter=20, tol=1e-4)
    assert res.centers.shape == (2, 5) and np.isfinite(res.centers).all()


@pytest.mark.parametrize("tag", ["blobs1000x20k8", "blobs3000x100k20", "blobs2000x64k100", "blobs777x37k5"])
def test_golden_vectors(golden, tag):
    X, init = golden[f"{tag}_X"], golden[f"{tag}_init"]
    res = ko.lloyd_fit(X, init, max_iter=int(golden[f"{tag}_n_iter"]), tol=1e-20)
    assert np.array_equal(res.centers, golden[f"{tag}_centers"])
    assert np.array_equal(res.labels, golden[f"{tag}_labels"])
    assert res.inertia == pytest.approx(float(golden[f"{tag}_inertia"]), rel=1e-12)
    # reference scorer definition (bench_kmeans.py:96-100) agrees with the expanded form
    assert ko.inertia(X, res.centers, res.labels) == pytest.approx(res.inertia, rel=1e-9)


def test_sharded_equals_unsharded(golden):
    # python/tests/test_kmeans.py:268-330 pins multi-GPU == single-GPU: same centres, identical labels
    X, init = golden["blobs1000x20k8_X"], golden["blobs1000x20k8_init"]
    a = ko.lloyd_fit(X, init, max_iter=20, tol=1e-20)
    for w in (2, 4):
        b = ko.lloyd_fit(X, init, max_iter=20, tol=1e-20, shards=w)
        assert np.allclose(a.centers, b.centers, rtol=1e-6, atol=0)
        assert np.array_equal(a.labels, b.labels)


def test_sklearn_cross_check():
    from sklearn.cluster import KMeans as SkKMeans

    X, _ = ko.make_blobs(5000, 16, 6, seed=11)
    init = X[:6].copy()
    res = ko.lloyd_fit(X, init, max_iter=7, tol=1e-20)
    if np.bincount(res.labels, minlength=6).min() == 0:
        pytest.skip("a cluster emptied; sklearn relocates those")
    sk = SkKMeans(n_clusters=6, init=init.astype(np.float64), n_init=1, max_iter=res.n_iter, tol=0.0,
                  algorithm="lloyd").fit(X.astype(np.float64))
    assert np.allclose(sk.cluster_centers_, res.centers, rtol=1e-4, atol=1e-5)
    amb = ko.ambiguous_rows(res.best, res.second)
    assert np.array_equal(sk.labels_[~amb], res.labels[~amb])
    assert sk.inertia_ == pytest.approx(res.inertia, rel=1e-5)


def test_empty_cluster_keeps_previous_centre():
    X = np.array([[0, 0], [0, 1], [1, 0], [1, 1]], dtype=np.float32)
    init = np.array([[0.5, 0.5], [100.0, 100.0]], dtype=np.float32)
    res = ko.lloyd_fit(X, init, max_iter=3, tol=1e-20)
    assert np.array_equal(res.centers[1], init[1])
    assert np.array_equal(res.centers[0], np.array([0.5, 0.5], dtype=np.float32))
    assert (res.labels == 0).all()


def test_tie_break_lowest_index_and_stop_rule():
    X = np.array([[0.0], [2.0]], dtype=np.float32)
    C = np.array([[1.0], [1.0], [1.0]], dtype=np.float32)
    labels, best, second = ko.assign(X, C)
    assert labels.tolist() == [0, 0] and best.tolist() == [1.0, 1.0] and second.tolist() == [1.0, 1.0]
    # stop rule: absolute shift^2 < tol, checked after the update; n_iter counts executed iterations
    Xb, _ = ko.make_blobs(400, 3, 3, seed=5)
    init = Xb[:3]
    r_all = ko.lloyd_fit(Xb, init, max_iter=50, tol=1e-20)
    r_tol = ko.lloyd_fit(Xb, init, max_iter=50, tol=1e9)
    assert r_tol.n_iter == 1 and 1 < r_all.n_iter <= 50
    assert ko.lloyd_fit(Xb, init, max_iter=2, tol=1e-20).n_iter == 2


def test_ambiguous_mask_and_single_cluster():
    best = np.array([1.0, 1.0, 0.0])
    second = np.array([1.0 + 5e-6, 1.1, 0.0])
    assert ko.ambiguous_rows(best, second).tolist() == [True, False, True]
    X = np.arange(12, dtype=np.float32).reshape(4, 3)
    res = ko.lloyd_fit(X, X[:1], max_iter=5, tol=1e-20)
    assert np.allclose(res.centers[0], X.mean(axis=0)) and np.isinf(res.second).all()


def test_kmeans_parallel_restatement_properties():
    """The oracle's k-means|| (initMode's default, python/src/spark_rapids_ml/clustering.py:290): k distinct rows of
    X, reproducible, one seed per blob on separated blobs, and independent of how the rows would be sharded (its
    random numbers are a function of the global row index)."""
    X, centers = ko.make_blobs(20000, 16, 10, seed=4)
    a = ko.kmeans_parallel_init(X, 10, seed=7)
    assert np.array_equal(a, ko.kmeans_parallel_init(X, 10, seed=7))
    assert not np.array_equal(a, ko.kmeans_parallel_init(X, 10, seed=8))
    rows = {tuple(r) for r in X}
    assert a.shape == (10, 16) and all(tuple(r) in rows for r in a) and len({tuple(r) for r in a}) == 10
    assert len(set(ko.predict(a, centers).tolist())) >= 9, "the seeds spread over (nearly) all blobs"
    u = ko.hash_uniform(5, 3, np.arange(100000, dtype=np.uint64))
    assert 0.0 <= u.min() and u.max() < 1.0 and abs(u.mean() - 0.5) < 0.01
    assert int(ko.hash64(7, 1, 0)) == 8075172986249684972  # pins the hash (the CUDA path must produce the same stream)
