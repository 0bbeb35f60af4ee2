"""One process driving two GPUs: handles on cuda:0 and cuda:1, fit + predict on each, and a two-partition
``KMeansModel.transform`` (the reference walks transform partitions with GPU = partition id modulo the GPU
count: python/src/spark_rapids_ml/core.py:358-363).  Catches per-device state kept process-wide, e.g. the
dynamic shared-memory opt-in of the tcgen05 kernels.  Skipped on a single-GPU box."""
import numpy as np
import pandas as pd
import pytest

from oracle import kmeans_oracle as ko
from tests import parity

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


@pytest.fixture(scope="module", autouse=True)
def _needs_two_gpus():
    assert torch.cuda.is_available(), "the -m gpu tests need a CUDA device"
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs in one process")


@pytest.mark.parametrize("shape", [(20000, 100, 20), (6000, 64, 100), (4000, 256, 600), (3000, 30, 7)])
def test_fit_and_predict_on_both_devices(shape):
    """The same fit on cuda:0 then cuda:1 (every kernel family: screen, split-TF32, tile GEMM, FFMA)."""
    from spark_rapids_ml_nai_b200 import _native

    n, d, k = shape
    X, _ = ko.make_blobs(n, d, k, seed=11)
    init = ko.random_init(X, k, seed=11)
    ref = ko.lloyd_fit(X, init, max_iter=4, tol=1e-20)
    results = []
    for dev in (0, 1):
        with torch.cuda.device(dev), _native.Handle(device=dev) as h:
            Xd = torch.from_numpy(X).to(f"cuda:{dev}")
            C, inertia, n_iter, labels = h.fit(Xd, torch.from_numpy(init).to(f"cuda:{dev}"), max_iter=4, tol=1e-20,
                                               want_labels=True)
            lab2, _ = h.predict(Xd, C)
            torch.cuda.synchronize(dev)
            results.append((C.cpu().numpy(), inertia, labels.cpu().numpy(), lab2.cpu().numpy()))
    for C, inertia, labels, lab2 in results:
        parity.assert_centers_close(C, ref.centers)
        parity.assert_inertia_close(inertia, ref.inertia)
        l_ref, best, second = ko.assign(X, C)
        parity.assert_labels_match(labels, best, second, l_ref)
        assert np.array_equal(labels, lab2)
    assert np.allclose(results[0][0], results[1][0], rtol=1e-6, atol=1e-7)


def test_two_partition_transform_uses_both_devices():
    from spark_rapids_ml_nai_b200.clustering import KMeans
    from spark_rapids_ml_nai_b200.runtime import LocalDataFrame

    X, _ = ko.make_blobs(30000, 100, 20, seed=12)
    pdf = pd.DataFrame({"features": list(X)})
    model = KMeans(k=20, num_workers=1, initMode="random", seed=2).setMaxIter(5).fit(pdf)
    ds = LocalDataFrame(pdf, num_partitions=2)
    out = model.transform(ds).toPandas()["prediction"].to_numpy()
    C = np.array(model.cluster_centers_, dtype=np.float32)
    l_ref, best, second = ko.assign(X, C)
    parity.assert_labels_match(out.astype(np.int32), best, second, l_ref)
