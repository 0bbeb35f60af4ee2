"""GPU parity tests (run on the B200 box with -m gpu): every compute call goes through the C ABI
(libb200km.so via ctypes) and is checked against the float64 oracle / the committed golden vectors."""
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


@pytest.fixture(scope="module", params=["auto", "ffma", "tc", "tc_split"])
def handle(request, native):
    h = native.Handle(device=0)
    h.set_path({"auto": native.PATH_AUTO, "ffma": native.PATH_FFMA, "tc": native.PATH_TC,
                "tc_split": native.PATH_TC_SPLIT}[request.param])
    h.forced = request.param
    yield h
    h.destroy()


def need(handle, native_mod, d, k, with_update=True):
    """A forced tcgen05 handle can only run shapes the fused kernel holds in shared memory; anything
    else must raise (never fall back), which test_forced_tc_rejects_unsupported_shapes checks."""
    if handle.forced in ("tc", "tc_split") and not native_mod.tc_supported(d, k, with_update):
        pytest.skip(f"fused tcgen05 kernel does not hold d={d} k={k}")


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


GOLDEN = ["blobs1000x20k8", "blobs3000x100k20", "blobs2000x64k100", "blobs777x37k5"]


@pytest.mark.parametrize("tag", GOLDEN)
def test_fit_matches_golden(handle, native, golden, tag):
    X, init = golden[f"{tag}_X"], golden[f"{tag}_init"]
    need(handle, native, X.shape[1], init.shape[0])
    n_iter = int(golden[f"{tag}_n_iter"])
    C, inertia, it, labels = handle.fit(dev(X), dev(init), max_iter=n_iter, tol=1e-20, want_labels=True)
    assert it == n_iter
    parity.assert_centers_close(C.cpu().numpy(), golden[f"{tag}_centers"])
    parity.assert_inertia_close(inertia, float(golden[f"{tag}_inertia"]))
    # labels are compared against an oracle assignment on the GPU's own final centres, so that a
    # centroid difference inside tolerance cannot move a boundary row
    lab_ref, best, second = ko.assign(X, C.cpu().numpy())
    parity.assert_labels_match(labels.cpu().numpy(), best, second, lab_ref)
    # and against the committed labels wherever the committed margins are unambiguous
    parity.assert_labels_match(labels.cpu().numpy(), golden[f"{tag}_best"], golden[f"{tag}_second"],
                               golden[f"{tag}_labels"])


@pytest.mark.parametrize("tag", GOLDEN)
def test_lloyd_step_matches_oracle(handle, native, golden, tag):
    X, init = golden[f"{tag}_X"], golden[f"{tag}_init"]
    need(handle, native, X.shape[1], init.shape[0])
    sums, counts, inertia, labels = handle.lloyd_step(dev(X), dev(init), want_labels=True)
    s_ref, c_ref, i_ref, l_ref = ko.lloyd_step(X, init)
    _, best, second = ko.assign(X, init)
    parity.assert_labels_match(labels.cpu().numpy(), best, second, l_ref)
    lab = labels.cpu().numpy()
    s_own, c_own = ko.accumulate(X, lab, init.shape[0])  # sums for the GPU's own labels must be exact-ish
    assert np.array_equal(counts.cpu().numpy(), c_own)
    assert np.allclose(sums.cpu().numpy(), s_own, rtol=1e-6, atol=1e-6 * np.abs(s_own).max())
    parity.assert_inertia_close(float(inertia.item()), i_ref)
    assert counts.sum().item() == X.shape[0]


def test_reference_known_answers(handle, native, golden):
    if handle.forced in ("tc", "tc_split"):
        pytest.skip("toy shapes have d % 4 != 0")
    # python/tests/test_kmeans.py:154,177-183 and :351-356,393-409
    X = golden["basic_X"]
    C, _, _, labels = handle.fit(dev(X), dev(X[[0, 3]]), max_iter=20, tol=float(np.finfo("float32").tiny),
                                 want_labels=True)
    got = np.array(sorted(C.cpu().numpy().tolist()))
    assert got == pytest.approx(golden["basic_centers_sorted"], rel=1e-3)
    lab = labels.cpu().numpy()
    assert lab[0] == lab[1] and lab[1] != lab[2] and lab[2] == lab[3]
    X = golden["compat_X"]
    C, inertia, _, labels = handle.fit(dev(X), dev(X[[0, 2]]), max_iter=10, tol=1e-4, want_labels=True)
    assert sorted(C.cpu().numpy().tolist()) == golden["compat_centers_sorted"].tolist()
    lab = labels.cpu().numpy()
    assert lab[0] == lab[1] and lab[2] == lab[3] and lab[0] != lab[2]
    assert inertia == pytest.approx(2.0)
    X = golden["numeric_X"]
    C, _, _, _ = handle.fit(dev(X), dev(X[[0, 4]]), max_iter=20, tol=1e-4)
    assert torch.isfinite(C).all()


@pytest.mark.parametrize("n,d,k", [(1, 1, 1), (1, 7, 3), (127, 4, 2), (128, 32, 32), (129, 33, 33), (257, 5, 65),
                                   (1000, 1, 4), (513, 260, 7), (300, 96, 130)])
def test_predict_ragged_shapes(handle, native, n, d, k):
    need(handle, native, d, k, with_update=False)
    rng = np.random.default_rng(n * 1000 + d * 10 + k)
    X = rng.standard_normal((n, d)).astype(np.float32) * 3
    C = rng.standard_normal((k, d)).astype(np.float32) * 3
    labels, dist = handle.predict(dev(X), dev(C), want_dist=True)
    l_ref, best, second = ko.assign(X, C)
    parity.assert_labels_match(labels.cpu().numpy(), best, second, l_ref)
    assert np.allclose(dist.cpu().numpy(), best, rtol=1e-4, atol=1e-4 * max(1.0, best.max()))


def test_forced_tc_rejects_unsupported_shapes(native):
    h = native.Handle(device=0)
    h.set_path(native.PATH_TC)
    try:
        with pytest.raises(RuntimeError, match="tcgen05 path forced but not available"):
            h.predict(torch.zeros((64, 7), device="cuda"), torch.zeros((3, 7), device="cuda"))  # d % 4 != 0
        with pytest.raises(RuntimeError, match="tcgen05 path forced but not available"):
            h.lloyd_step(torch.zeros((64, 256), device="cuda"), torch.zeros((256, 256), device="cuda"))
    finally:
        h.destroy()
    assert native.tc_supported(100, 20, True) and native.tc_supported(64, 100, False)


def test_auto_uses_tensor_cores_on_headline_shapes(native):
    h = native.Handle(device=0)
    try:
        X = torch.randn((4096, 100), device="cuda")
        h.lloyd_step(X, X[:20].contiguous())
        assert h.last_path() == "tc"
        h.predict(torch.randn((4096, 64), device="cuda"), torch.randn((100, 64), device="cuda"))
        assert h.last_path() == "tc"
        h.predict(torch.randn((512, 7), device="cuda"), torch.randn((3, 7), device="cuda"))
        assert h.last_path() == "ffma"
    finally:
        h.destroy()


def test_predict_empty_and_errors(handle, native):
    X = torch.empty((0, 8), dtype=torch.float32, device="cuda")
    C = torch.zeros((3, 8), dtype=torch.float32, device="cuda")
    labels, _ = handle.predict(X, C)
    assert labels.numel() == 0
    with pytest.raises(RuntimeError, match="no rows"):
        handle.fit(X, C, max_iter=1, tol=0.0)
    with pytest.raises(ValueError):
        handle.predict(torch.zeros((4, 7), device="cuda"), C)
    with pytest.raises(TypeError):
        handle.predict(torch.zeros((4, 8), dtype=torch.float64, device="cuda"), C)


def test_ties_go_to_lowest_index(handle, native):
    if handle.forced in ("tc", "tc_split"):
        handle.set_path(native.PATH_AUTO)  # second half uses d=2; first half (d=12) runs the fused kernel
    # duplicated centroids: every row is an exact tie between j and its copies
    rng = np.random.default_rng(3)
    X = rng.standard_normal((500, 12)).astype(np.float32)
    base = rng.standard_normal((5, 12)).astype(np.float32)
    C = np.concatenate([base, base, base[::-1]], axis=0)
    labels, _ = handle.predict(dev(X), dev(C))
    l_ref = ko.predict(X, C)
    assert np.array_equal(labels.cpu().numpy(), l_ref)
    assert labels.max().item() < 5
    # integer grid: exact ties between distinct centroids
    X = np.array([[0.0, 0.0], [2.0, 0.0], [1.0, 1.0], [1.0, -1.0]], dtype=np.float32)
    C = np.array([[1.0, 0.0], [1.0, 0.0], [0.0, 1.0], [2.0, 1.0]], dtype=np.float32)
    labels, _ = handle.predict(dev(X), dev(C))
    assert labels.cpu().numpy().tolist() == ko.predict(X, C).tolist()
    if handle.forced in ("tc", "tc_split"):
        handle.set_path(native.PATH_TC if handle.forced == "tc" else native.PATH_TC_SPLIT)


def test_empty_cluster_keeps_centroid_and_stop_rule(handle, native):
    X = np.array([[0, 0, 0, 0], [0, 1, 0, 0], [1, 0, 0, 0], [1, 1, 0, 0]], dtype=np.float32)
    init = np.array([[0.5, 0.5, 0, 0], [100.0, 100.0, 0, 0]], dtype=np.float32)
    C, inertia, it, labels = handle.fit(dev(X), dev(init), max_iter=3, tol=1e-20, want_labels=True)
    assert np.array_equal(C.cpu().numpy(), init)
    assert it == 1  # nothing moved: shift^2 == 0 < tol
    assert inertia == pytest.approx(2.0)
    Xb, _ = ko.make_blobs(4000, 8, 4, seed=5)
    init = Xb[:4].copy()
    ref = ko.lloyd_fit(Xb, init, max_iter=50, tol=1e-20)
    _, _, it_all, _ = handle.fit(dev(Xb), dev(init), max_iter=50, tol=1e-20)
    _, _, it_tol, _ = handle.fit(dev(Xb), dev(init), max_iter=50, tol=1e9)
    _, _, it_cap, _ = handle.fit(dev(Xb), dev(init), max_iter=2, tol=1e-20)
    assert it_tol == 1 and it_cap == 2 and it_all == ref.n_iter
    C0, _, it0, _ = handle.fit(dev(Xb), dev(init), max_iter=0, tol=1e-20)
    assert it0 == 0 and np.array_equal(C0.cpu().numpy(), init)


def test_near_ties_are_decided_in_float64(handle):
    # two centroids 1e-3 apart inside one blob far from the origin: FP32 expanded-form scores cannot
    # separate them, the float64 refine must
    rng = np.random.default_rng(9)
    d = 64
    centre = np.full(d, 40.0, dtype=np.float32)
    X = (centre + rng.standard_normal((20000, d))).astype(np.float32)
    C = np.stack([centre, centre + np.float32(1e-3), centre - 30.0]).astype(np.float32)
    labels, _ = handle.predict(dev(X), dev(C))
    l_ref, best, second = ko.assign(X, C)
    parity.assert_labels_match(labels.cpu().numpy(), best, second, l_ref)
    assert (parity.ko.ambiguous_rows(best, second) .sum()) < 0.5 * len(X)


@pytest.mark.parametrize("n,d,k,iters", [(200_000, 100, 20, 5), (100_000, 20, 8, 20), (50_000, 256, 300, 3)])
def test_fit_parity_medium(handle, native, n, d, k, iters):
    """Every iteration of the fit against the float64 oracle.  Iteration t of the CUDA fit is compared with one oracle
    step from the CUDA fit's own centroids of iteration t-1 (the fit is bitwise reproducible, so a shorter fit IS the
    prefix of a longer one): labels exact outside the 1e-5 band, centroids within 1e-4, at every step.  The
    free-running oracle trajectory is compared as well, as long as none of its rows sits inside the ambiguity band:
    such a row may legitimately take either label, and on this data (several centroids inside one blob) one flipped
    row moves a centroid by more than 1e-4 and the difference grows from iteration to iteration."""
    need(handle, native, d, k)
    X, _ = ko.make_blobs(n, d, k, seed=1)
    init = ko.random_init(X, k, seed=1)
    Xd = dev(X)
    prev_gpu = init
    prev_ref = init
    free_running = True
    for t in range(1, iters + 1):
        C, inertia, it, labels = handle.fit(Xd, dev(init), max_iter=t, tol=1e-20, want_labels=True)
        Cn = C.cpu().numpy()
        assert it == t
        # one oracle step from the centroids the CUDA fit had before this iteration
        sums, counts, _, _ = ko.lloyd_step(X, prev_gpu)
        parity.assert_centers_close(Cn, ko.finalize(sums, counts, prev_gpu)[0])
        if free_running:
            _, best, second = ko.assign(X, prev_ref)
            if ko.ambiguous_rows(best, second, parity.REL_AMBIG).any():
                free_running = False  # from here on the two trajectories may differ by tie rows
            else:
                sums, counts, _, _ = ko.lloyd_step(X, prev_ref)
                prev_ref = ko.finalize(sums, counts, prev_ref)[0]
                parity.assert_centers_close(Cn, prev_ref)
        prev_gpu = Cn
        if t == iters or t == 1:
            l_ref, best, second = ko.assign(X, Cn)
            parity.assert_labels_match(labels.cpu().numpy(), best, second, l_ref)
            parity.assert_inertia_close(inertia, float(best.sum()))


def test_size_independent_properties_large(handle):
    # 4M x 100 (1.6 GB): too large for the oracle in seconds; check invariants instead
    n, d, k = 4_000_000, 100, 20
    g = torch.Generator(device="cuda").manual_seed(1)
    centers = (torch.rand((k, d), generator=g, device="cuda") * 20 - 10)
    which = torch.randint(0, k, (n,), generator=g, device="cuda")
    X = centers[which] + torch.randn((n, d), generator=g, device="cuda")
    init = X[:: n // k][:k].contiguous()
    sums, counts, inertia, labels = handle.lloyd_step(X, init, want_labels=True)
    assert counts.sum().item() == n                                   # every row counted once
    col = X.double().sum(dim=0)
    assert torch.allclose(sums.sum(dim=0), col, rtol=1e-6, atol=1e-2)    # linearity: sum of sums = column sums
    assert torch.equal(torch.bincount(labels.long(), minlength=k).double(), counts)
    lab2, dist = handle.predict(X, init, want_dist=True)              # predict == assignment of the step
    assert torch.equal(lab2, labels)
    # the step's inertia comes from the float64 sums, predict's distances are FP32 per row (each carries
    # ~2^-24 of ||x||^2 + ||c||^2): agreement well inside the 1e-4 parity bar, not bit-for-bit
    assert dist.double().sum().item() == pytest.approx(inertia.item(), rel=2e-5)
    sub = torch.arange(0, n, 997, device="cuda")
    l_ref, best, second = ko.assign(X[sub].cpu().numpy(), init.cpu().numpy())
    parity.assert_labels_match(labels[sub].cpu().numpy(), best, second, l_ref)
    C, _, it, lab3 = handle.fit(X, init, max_iter=3, tol=0.0, want_labels=True)
    lab4, _ = handle.predict(X, C)                                    # idempotence of the final labels
    assert it == 3 and torch.equal(lab3, lab4)
