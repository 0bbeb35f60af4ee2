"""GPU parity tests of the tcgen05 tile-GEMM path (csrc/kernels_gemm.cu) through the C ABI: the regime where
k*d is beyond shared memory (BASELINE.json C3 / C5, the reference's own k=1000 x d=3000 benchmark,
python/benchmark/databricks/run_benchmark.sh:46-55).  Every shape is checked against the float64 oracle with
BASELINE.json's tolerances (labels exact outside the 1e-5 band, centroids / inertia within 1e-4)."""
import numpy as np
import pytest

from oracle import kmeans_oracle as ko
from tests import parity

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def native():
    from spark_rapids_ml_nai_b200 import _native

    assert torch.cuda.is_available(), "the -m gpu tests need a CUDA device"
    _native.load()
    return _native


@pytest.fixture(scope="module")
def handle(native):
    h = native.Handle(device=0)
    h.set_path(native.PATH_GEMM)
    yield h
    h.destroy()


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


# (n, d, k): one N-tile / several, resident A (d <= 256) / streamed A, ragged d and n, tiny k
PREDICT_SHAPES = [(1, 4, 1), (3, 8, 2), (300, 16, 5), (1000, 40, 7), (257, 64, 64), (2000, 100, 65), (3000, 64, 300),
                  (513, 260, 700), (2048, 256, 1000), (700, 1024, 300), (1500, 3000, 130), (5000, 32, 2500)]


@pytest.mark.parametrize("n,d,k", PREDICT_SHAPES)
def test_predict_matches_oracle(handle, n, d, k):
    rng = np.random.default_rng(n * 1000 + d * 10 + k)
    X = rng.standard_normal((n, d)).astype(np.float32) * 3
    C = rng.standard_normal((k, d)).astype(np.float32) * 3
    labels, dist = handle.predict(dev(X), dev(C), want_dist=True)
    assert handle.last_path() == "gemm"
    l_ref, best, second = ko.assign(X, C)
    parity.assert_labels_match(labels.cpu().numpy(), best, second, l_ref)
    assert np.allclose(dist.cpu().numpy(), best, rtol=1e-4, atol=1e-4 * max(1.0, best.max()))


@pytest.mark.parametrize("n,d,k,iters", [(6000, 64, 300, 4), (20000, 128, 500, 3), (50000, 256, 1000, 3),
                                         (20000, 1024, 4096, 2), (3000, 3000, 200, 3)])
def test_fit_matches_oracle(handle, n, d, k, iters):
    X, _ = ko.make_blobs(n, d, max(2, k // 3), seed=d + k)  # fewer blobs than centroids: close centroids inside blobs
    init = ko.random_init(X, k, seed=3)
    ref = ko.lloyd_fit(X, init, max_iter=iters, tol=1e-20)
    C, inertia, n_iter, labels = handle.fit(dev(X), dev(init), max_iter=iters, tol=1e-20, want_labels=True)
    assert handle.last_path() == "gemm" and n_iter == ref.n_iter
    Cn = C.cpu().numpy()
    parity.assert_centers_close(Cn, ref.centers)
    parity.assert_inertia_close(inertia, ref.inertia)
    l_ref, best, second = ko.assign(X, Cn)
    parity.assert_labels_match(labels.cpu().numpy(), best, second, l_ref)


def test_lloyd_step_sums_counts_inertia(handle):
    X, _ = ko.make_blobs(9000, 256, 40, seed=5)
    init = ko.random_init(X, 600, seed=5)
    sums, counts, inertia, labels = handle.lloyd_step(dev(X), dev(init), want_labels=True)
    s_ref, c_ref, i_ref, l_ref = ko.lloyd_step(X, init)
    _, best, second = ko.assign(X, init)
    lab = labels.cpu().numpy()
    parity.assert_labels_match(lab, best, second, l_ref)
    s_own, c_own = ko.accumulate(X, lab, 600)
    assert np.array_equal(counts.cpu().numpy(), c_own)
    assert np.allclose(sums.cpu().numpy(), s_own, rtol=1e-6, atol=1e-6 * np.abs(s_own).max())
    parity.assert_inertia_close(float(inertia.item()), i_ref)


def test_fit_is_bitwise_reproducible(handle):
    """Labels are exact and the sums are added in one fixed order (stable sort by label + segmented float64
    reduction): the same fit gives the same bits, run after run."""
    X, _ = ko.make_blobs(60000, 128, 50, seed=8)
    init = ko.random_init(X, 700, seed=8)
    Xd, Cd = dev(X), dev(init)
    a = handle.fit(Xd, Cd, max_iter=4, tol=0.0)
    b = handle.fit(Xd, Cd, max_iter=4, tol=0.0)
    assert torch.equal(a[0], b[0]) and a[1] == b[1]
    # a label with very many rows (its range spans many segments of the reduction) and empty labels
    C2 = init.copy()
    C2[5:] += 1000.0  # nearly everything goes to the first five centroids, most labels stay empty
    s1 = handle.lloyd_step(Xd, dev(C2), want_labels=True)
    s2 = handle.lloyd_step(Xd, dev(C2), want_labels=True)
    assert torch.equal(s1[0], s2[0]) and torch.equal(s1[1], s2[1])
    lab = s1[3].cpu().numpy()
    s_own, c_own = ko.accumulate(X, lab, 700)
    assert np.array_equal(s1[1].cpu().numpy(), c_own)
    assert np.allclose(s1[0].cpu().numpy(), s_own, rtol=1e-9, atol=1e-9 * np.abs(s_own).max())


def test_structureless_data_and_ties(handle):
    """Uniform data (margins of the order of the TF32 error: most rows go through the exact fix-up, many with
    more than two candidates), duplicated centroids (exact ties -> lowest index) and rows equal to a centroid."""
    rng = np.random.default_rng(9)
    X = rng.random((4000, 128), dtype=np.float32)
    C = rng.random((400, 128), dtype=np.float32)
    C[100] = C[7]  # exact duplicates
    C[399] = C[7]
    X[:50] = C[rng.integers(0, 400, 50)]
    labels, _ = handle.predict(dev(X), dev(C))
    lab = labels.cpu().numpy()
    l_ref, best, second = ko.assign(X, C)
    # exact ties (distance equal) must go to the lowest index: compare directly where the oracle's margin is 0
    tie = second == best
    assert np.array_equal(lab[tie], l_ref[tie])
    parity.assert_labels_match(lab, best, second, l_ref)


def test_far_from_origin(handle):
    """Close centroids far from the origin: the margin bound scales with the distance between centroids, so the
    screen still decides most rows, and the rest are exact."""
    X, _ = ko.make_blobs(8000, 64, 30, seed=6)
    X = (X * 0.01 + 1000.0).astype(np.float32)
    init = ko.random_init(X, 320, seed=6)
    labels, _ = handle.predict(dev(X), dev(init))
    l_ref, best, second = ko.assign(X, init)
    parity.assert_labels_match(labels.cpu().numpy(), best, second, l_ref)


def test_auto_picks_gemm_beyond_the_fused_kernels(native):
    with native.Handle(device=0) as h:
        X, _ = ko.make_blobs(4000, 256, 20, seed=2)
        init = ko.random_init(X, 512, seed=2)
        h.fit(dev(X), dev(init), max_iter=1, tol=0.0)
        assert h.last_path() == "gemm"
        h.fit(dev(X[:, :100].copy()), dev(init[:20, :100].copy()), max_iter=1, tol=0.0)
        assert h.last_path() == "tc"
        with pytest.raises(RuntimeError, match="tile-GEMM path forced but not available"):
            h.set_path(native.PATH_GEMM)
            h.predict(dev(X[:, :30].copy()), dev(init[:, :30].copy()))
