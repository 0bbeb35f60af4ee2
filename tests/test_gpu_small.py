"""GPU tests of the shared-memory resident whole-fit kernel (csrc/kernels_small.cu): BASELINE.json C1
(100k x 20, k = 8, 20 iterations) and its neighbourhood, against the float64 oracle.  The kernel works in float64
throughout, so its centroids equal the oracle's float32-rounded centroids bit for bit."""
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


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.parametrize("n,d,k,iters", [(100_000, 20, 8, 20), (1000, 20, 8, 30), (5000, 7, 3, 10), (37, 5, 4, 5),
                                         (200_000, 24, 12, 6), (3000, 100, 20, 5)])
def test_fit_matches_oracle(native, n, d, k, iters):
    X, _ = ko.make_blobs(n, d, k, seed=n % 97 + d)
    init = ko.random_init(X, k, seed=5)
    ref = ko.lloyd_fit(X, init, max_iter=iters, tol=1e-20)
    with native.Handle(device=0) as h:
        h.set_path(native.PATH_SMALL)
        C, inertia, n_iter, labels = h.fit(dev(X), dev(init), max_iter=iters, tol=1e-20, want_labels=True)
        assert h.last_path() == "small"
        again = h.fit(dev(X), dev(init), max_iter=iters, tol=1e-20)
    assert n_iter == ref.n_iter
    Cn = C.cpu().numpy()
    parity.assert_centers_close(Cn, ref.centers)
    parity.assert_inertia_close(inertia, ref.inertia)
    l_ref, best, second = ko.assign(X, Cn)
    parity.assert_labels_match(labels.cpu().numpy(), best, second, l_ref)
    assert torch.equal(C, again[0]) and inertia == again[1], "bitwise reproducible"


def test_auto_takes_it_for_c1_and_honours_the_stop_rule(native):
    X, _ = ko.make_blobs(100_000, 20, 8, seed=3)
    init = ko.random_init(X, 8, seed=3)
    ref = ko.lloyd_fit(X, init, max_iter=100, tol=1e-4)
    with native.Handle(device=0) as h:
        C, inertia, n_iter, _ = h.fit(dev(X), dev(init), max_iter=100, tol=1e-4)
        assert h.last_path() == "small"
        assert n_iter == ref.n_iter < 100
        parity.assert_centers_close(C.cpu().numpy(), ref.centers)
        # empty cluster keeps its centroid; max_iter = 0 returns the initial centroids with their inertia
        far = init.copy()
        far[7] = 1e4
        C2, _, _, _ = h.fit(dev(X), dev(far), max_iter=3, tol=0.0)
        assert np.array_equal(C2.cpu().numpy()[7], far[7])
        C0, in0, it0, _ = h.fit(dev(X), dev(init), max_iter=0, tol=0.0)
        assert it0 == 0 and np.array_equal(C0.cpu().numpy(), init)
        parity.assert_inertia_close(in0, ko.inertia(X, init, ko.predict(X, init)))
        h.set_path(native.PATH_SMALL)
        with pytest.raises(RuntimeError, match="resident fit forced but not available"):
            big = torch.zeros((4_000_000, 20), device="cuda")
            h.fit(big, big[:8].contiguous(), max_iter=1, tol=0.0)
