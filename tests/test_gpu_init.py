"""GPU tests of the device-side k-means|| initialisation behind b200km_init_centroids (csrc/kernels_init.cu,
api.cu): the reference passes initMode="k-means||" (its default, python/src/spark_rapids_ml/clustering.py:290) to the
backend as init="scalable-k-means++" (:72,117-119).  The comparison partner is the oracle's restatement with the
same counter-based random keying: the seeds are rows of X, so they must agree exactly."""
import numpy as np
import pytest

from oracle import kmeans_oracle as ko

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def native():
    from spark_rapids_ml_nai_b200 import _native

    assert torch.cuda.is_available(), "the -m gpu tests need a CUDA device"
    _native.load()
    return _native


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.parametrize("n,d,k,seed", [(20000, 16, 10, 7), (30000, 100, 20, 1), (8000, 64, 100, 3), (5000, 7, 5, 11),
                                        (12000, 256, 300, 5)])
def test_seeds_equal_the_oracle(native, n, d, k, seed):
    X, _ = ko.make_blobs(n, d, max(2, k // 2), seed=seed)
    ref = ko.kmeans_parallel_init(X, k, seed=seed, oversampling=2.0)
    with native.Handle(device=0) as h:
        got = h.init_centroids(dev(X), k, mode=native.INIT_KMEANS_PARALLEL, seed=seed, oversampling=2.0).cpu().numpy()
        again = h.init_centroids(dev(X), k, mode=native.INIT_KMEANS_PARALLEL, seed=seed, oversampling=2.0).cpu().numpy()
    assert got.shape == (k, d) and np.array_equal(got, again)
    assert np.array_equal(got, ref), "the k seeds (rows of X) differ from the oracle's"
    assert len({tuple(r) for r in got}) == k


def test_tiny_and_degenerate_data(native):
    # fewer candidates than k: topped up with further distinct rows; duplicated rows: zero potential
    X = np.arange(24, dtype=np.float32).reshape(6, 4)
    with native.Handle(device=0) as h:
        got = h.init_centroids(dev(X), 5, mode=native.INIT_KMEANS_PARALLEL, seed=2).cpu().numpy()
    assert np.array_equal(got, ko.kmeans_parallel_init(X, 5, seed=2))
    Xd = np.repeat(np.array([[1.0, 2.0, 3.0, 4.0], [9.0, 9.0, 9.0, 9.0]], dtype=np.float32), 50, axis=0)
    with native.Handle(device=0) as h:
        got = h.init_centroids(dev(Xd), 2, mode=native.INIT_KMEANS_PARALLEL, seed=4).cpu().numpy()
    assert np.array_equal(got, ko.kmeans_parallel_init(Xd, 2, seed=4))
    assert {tuple(r) for r in got} == {(1.0, 2.0, 3.0, 4.0), (9.0, 9.0, 9.0, 9.0)}


def test_two_ranks_give_the_one_rank_seeds():
    """Row-sharded over two GPUs the initialisation picks the very same rows (its random numbers are keyed by the
    global row index): python/tests/test_kmeans.py:268-330 asks the multi-GPU fit to agree with the single-GPU one."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import pandas as pd

    from spark_rapids_ml_nai_b200.clustering import KMeans

    X, _ = ko.make_blobs(40000, 32, 12, seed=9)
    df = pd.DataFrame({"features": list(X)})
    one = KMeans(k=12, num_workers=1, seed=21).setMaxIter(0).fit(df)
    two = KMeans(k=12, num_workers=2, seed=21).setMaxIter(0).fit(df)
    assert np.array_equal(np.array(one.cluster_centers_), np.array(two.cluster_centers_))
    assert np.array_equal(np.array(one.cluster_centers_, dtype=np.float32), ko.kmeans_parallel_init(X, 12, seed=21))
