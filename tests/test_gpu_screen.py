"""GPU tests of the screen kernel + exact fix-up pair (run on the B200 box with -m gpu): rows the TF32 scores
cannot decide must come out exact, in the labels and in the sums, however many they are."""
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


@pytest.fixture()
def handle(native):
    h = native.Handle(device=0)
    h.set_path(native.PATH_TC)
    yield h
    h.destroy()


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _check_step(handle, X, C, expect_open):
    sums, counts, inertia, labels = handle.lloyd_step(dev(X), dev(C), want_labels=True)
    open_rows, kernel = handle.screen_stats()
    assert kernel == 2, "the Lloyd pass of the tcgen05 path is the screen kernel"
    assert (open_rows > 0) == expect_open
    l_ref, best, second = ko.assign(X, C)
    lab = labels.cpu().numpy()
    parity.assert_labels_match(lab, best, second, l_ref)
    # sums / counts belong to the FINAL labels (rows moved by the fix-up included)
    s_own, c_own = ko.accumulate(X, lab, C.shape[0])
    assert np.array_equal(counts.cpu().numpy(), c_own)
    assert np.allclose(sums.cpu().numpy(), s_own, rtol=1e-6, atol=1e-6 * max(1.0, np.abs(s_own).max()))
    parity.assert_inertia_close(float(inertia.item()), float(best.sum()))
    return open_rows


def test_separated_blobs_leave_nothing_open(handle):
    X, C = ko.make_blobs(20000, 100, 20, seed=5)  # the true centres: one centroid per blob
    _check_step(handle, X, C, expect_open=False)


@pytest.mark.parametrize("n,d,k", [(30000, 100, 20), (9000, 36, 7), (12000, 64, 32), (5000, 16, 3)])
def test_structureless_data_is_still_exact(handle, n, d, k):
    rng = np.random.default_rng(n + d + k)
    X = rng.uniform(-1.0, 1.0, size=(n, d)).astype(np.float32)
    C = X[rng.choice(n, size=k, replace=False)].copy()
    open_rows = _check_step(handle, X, C, expect_open=True)
    assert open_rows < n


def test_more_open_rows_than_the_list_holds(handle):
    """The exact pass walks a list of the open rows that holds n/16 + 65536 entries; beyond that (data the screen
    cannot decide: one blob shared by three nearly identical centroids) the tile-range kernel does the same work
    straight from the bitmap.  Labels, counts and sums are exact either way."""
    rng = np.random.default_rng(23)
    n, d = 400_000, 64
    centre = np.full(d, 40.0, dtype=np.float32)
    X = (centre + rng.standard_normal((n, d))).astype(np.float32)
    C = np.stack([centre, centre + np.float32(1e-3), centre - np.float32(1e-3), centre - 30.0]).astype(np.float32)
    open_rows = _check_step(handle, X, C, expect_open=True)
    assert open_rows > n // 16 + 65536
    labels, _ = handle.predict(dev(X), dev(C))  # the labels-only pass takes the same route
    l_ref, best, second = ko.assign(X, C)
    parity.assert_labels_match(labels.cpu().numpy(), best, second, l_ref)


def test_close_centroids_far_from_origin(handle):
    # three centroids 1e-3 apart inside one blob 40 units from the origin: every row of the blob stays open
    # (pair and third-candidate tests fail) and only the exact pass can separate them
    rng = np.random.default_rng(11)
    d = 64
    centre = np.full(d, 40.0, dtype=np.float32)
    X = (centre + rng.standard_normal((6000, d))).astype(np.float32)
    C = np.stack([centre, centre + np.float32(1e-3), centre - np.float32(1e-3), centre - 30.0]).astype(np.float32)
    open_rows = _check_step(handle, X, C, expect_open=True)
    assert open_rows > 1000


def test_predict_labels_only_uses_the_screen_and_is_exact(handle):
    rng = np.random.default_rng(3)
    X = rng.standard_normal((40000, 100)).astype(np.float32)
    C = X[:20].copy()
    labels, _ = handle.predict(dev(X), dev(C))
    _, kernel = handle.screen_stats()
    assert kernel == 2
    l_ref, best, second = ko.assign(X, C)
    parity.assert_labels_match(labels.cpu().numpy(), best, second, l_ref)
    labels2, dist = handle.predict(dev(X), dev(C), want_dist=True)  # distances: the split-TF32 kernel
    assert handle.screen_stats()[1] == 1
    assert torch.equal(labels, labels2)


@pytest.mark.parametrize("n,d,k", [(20000, 64, 100), (9000, 32, 128), (7000, 100, 33), (6000, 16, 65)])
def test_labels_only_predict_up_to_128_centroids_runs_the_screen(handle, n, d, k):
    # four assign groups, scores 32 at a time with a running top-3; blobs and structureless data
    for X in (ko.make_blobs(n, d, k, seed=n)[0],
              np.random.default_rng(k).uniform(-1.0, 1.0, size=(n, d)).astype(np.float32)):
        C = X[np.random.default_rng(d).choice(n, size=k, replace=False)].copy()
        labels, _ = handle.predict(dev(X), dev(C))
        assert handle.screen_stats()[1] == 2
        l_ref, best, second = ko.assign(X, C)
        parity.assert_labels_match(labels.cpu().numpy(), best, second, l_ref)


def test_fit_switches_to_the_split_kernel_when_the_screen_decides_too_little(handle, native):
    # one blob far from the origin shared by three centroids: a large share of its rows stays open in every pass
    rng = np.random.default_rng(17)
    d = 64
    centre = np.full(d, 40.0, dtype=np.float32)
    X = (centre + rng.standard_normal((200_000, d))).astype(np.float32)
    init = np.stack([centre, centre + np.float32(1e-3), centre - np.float32(1e-3), centre - 30.0]).astype(np.float32)
    C, _, it, _ = handle.fit(dev(X), dev(init), max_iter=7, tol=0.0, want_inertia=False)
    assert it == 7
    open_rows, kernel = handle.screen_stats()
    assert open_rows > 0.02 * len(X)          # far more open rows than the screen is worth
    assert kernel == 1                         # ... so the later iterations ran the split-TF32 kernel
    assert torch.isfinite(C).all()
    labels, _ = handle.predict(dev(X), C)      # and the centres it produced are usable
    l_ref, best, second = ko.assign(X, C.cpu().numpy())
    parity.assert_labels_match(labels.cpu().numpy(), best, second, l_ref)
    # and a well separated data set never leaves the screen
    Xb, _ = ko.make_blobs(100_000, 100, 20, seed=2)
    Cb, _, it, _ = handle.fit(dev(Xb), dev(ko.random_init(Xb, 20, seed=4)), max_iter=8, tol=0.0, want_inertia=False)
    assert handle.screen_stats()[1] == 2


def test_same_fit_twice_is_bitwise_identical(handle):
    """Determinism of the fused pass: every CTA adds its tiles in a fixed order into its own float64 partial, the
    open rows are added by the exact pass in row order, and the partials are combined in slot order - no atomics on
    the sums.  The same 4M-row fit therefore gives the same bits, run after run (labels are exact for every row)."""
    g = torch.Generator(device="cuda").manual_seed(5)
    n, d, k = 4_000_000, 100, 20
    centers = torch.rand((k, d), generator=g, device="cuda") * 20 - 10
    X = torch.randn((n, d), generator=g, device="cuda") * 3.0 + centers[torch.randint(0, k, (n,), generator=g, device="cuda")]
    init = X[torch.randperm(n, generator=g, device="cuda")[:k]].contiguous()
    runs = [handle.fit(X, init, max_iter=6, tol=0.0) for _ in range(3)]
    open_rows, kernel = handle.screen_stats()
    assert kernel in (1, 2) and open_rows > 0, "wide blobs: the exact pass has rows to decide"
    for r in runs[1:]:
        assert torch.equal(r[0], runs[0][0]), "centroids differ between identical runs"
        assert r[1] == runs[0][1], "inertia differs between identical runs"
    s1 = handle.lloyd_step(X, init)
    s2 = handle.lloyd_step(X, init)
    assert torch.equal(s1[0], s2[0]) and torch.equal(s1[1], s2[1]) and torch.equal(s1[2], s2[2])


def test_inertia_of_a_step_far_from_the_origin(handle):
    """Blobs shifted by 1e3: ||x||^2 is ~1e6 times the cluster distances, a sum-of-squares identity would lose the
    inertia to cancellation; the step's inertia comes from per-row differences and stays within 1e-4."""
    X, _ = ko.make_blobs(30000, 100, 20, seed=13)
    X = (X + 1000.0).astype(np.float32)
    C = ko.random_init(X, 20, seed=13)
    sums, counts, inertia, labels = handle.lloyd_step(dev(X), dev(C), want_labels=True)
    _, _, i_ref, l_ref = ko.lloyd_step(X, C)
    _, best, second = ko.assign(X, C)
    parity.assert_labels_match(labels.cpu().numpy(), best, second, l_ref)
    parity.assert_inertia_close(float(inertia.item()), i_ref)
