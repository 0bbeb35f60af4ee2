"""b200km_predict_static: the entry point for callers that cannot allocate device memory (a task whose GPU
memory belongs to another pool, a latency-bound serving thread).  Collected last on purpose: if it fails, a
run that stops at the first failure has already run every other test.  Called through the C ABI, two calls
in a row on the shared static scratch, compared with the existing b200km_predict entry point."""
import numpy as np
import pytest

from oracle import kmeans_oracle as ko
from tests import parity

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


def test_static_predict_matches_predict_twice_in_a_row():
    from spark_rapids_ml_nai_b200 import _native

    assert torch.cuda.is_available(), "the -m gpu tests need a CUDA device"
    torch.cuda.set_device(0)
    h = _native.Handle(device=0)
    try:
        # second call: fewer rows, other k and d, exact ties (duplicated centroids) -> stale scratch from
        # the first call must not leak into it
        for (n, d, k, seed) in [(50_000, 100, 20, 1), (7_777, 37, 65, 2)]:
            X, _ = ko.make_blobs(n, d, min(k, 32), seed=seed)
            C = ko.random_init(X, k, seed=seed)
            if seed == 2:
                C[5] = C[3]
            Xd, Cd = torch.from_numpy(X).cuda(), torch.from_numpy(C).cuda()
            got = _native.predict_static(Xd, Cd)
            torch.cuda.synchronize()
            # same arithmetic path in the existing entry point: bit-identical labels
            h.set_path(_native.PATH_FFMA)
            ref_ffma, _ = h.predict(Xd, Cd)
            assert torch.equal(got, ref_ffma)
            # whatever path AUTO picks: equal under the parity rule used by every other test
            h.set_path(_native.PATH_AUTO)
            ref_auto, _ = h.predict(Xd, Cd)
            l_ref, best, second = ko.assign(X, C)
            parity.assert_labels_match(got.cpu().numpy(), best, second, l_ref)
            parity.assert_labels_match(ref_auto.cpu().numpy(), best, second, l_ref)
    finally:
        h.destroy()


def test_static_predict_rejects_sizes_beyond_the_scratch():
    from spark_rapids_ml_nai_b200 import _native

    torch.cuda.set_device(0)
    C = torch.zeros((4, 8), device="cuda")
    with pytest.raises(RuntimeError, match="exceeds the static scratch"):
        _native.predict_static(torch.zeros((_native.STATIC_MAX_ROWS + 1, 8), device="cuda"), C)
    with pytest.raises(RuntimeError, match="exceeds the static scratch"):
        _native.predict_static(torch.zeros((16, 8), device="cuda"),
                               torch.zeros((_native.STATIC_MAX_K + 1, 8), device="cuda"))
    assert _native.predict_static(torch.zeros((0, 8), device="cuda"), C).numel() == 0
