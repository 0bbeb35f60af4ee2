"""GPU tests of the drop-in KMeans / KMeansModel surface (fit -> model -> transform through the plugin
slots and the C ABI).  They re-express python/tests/test_kmeans.py:149-330,333-437 of the reference."""
import numpy as np
import pandas as pd
import pytest

from oracle import kmeans_oracle as ko
from tests import parity

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


@pytest.fixture(scope="module", autouse=True)
def _needs_gpu():
    assert torch.cuda.is_available(), "the -m gpu tests need a CUDA device"


def _sorted(c):
    return sorted([list(map(float, x)) for x in c])


def test_kmeans_basic(tmp_path, golden):
    # python/tests/test_kmeans.py:149-229
    from spark_rapids_ml_nai_b200.clustering import KMeans, KMeansModel

    data = golden["basic_X"].tolist()
    df = pd.DataFrame({"features": data})
    kmeans = KMeans(num_workers=1, n_clusters=2).setFeaturesCol("features").setSeed(0)
    kmeans.setTol(0.0)

    def check(model):
        assert len(model.cluster_centers_) == 2
        sc = _sorted(model.cluster_centers_)
        assert sc[0] == pytest.approx([1.0, 1.5], 0.001) and sc[1] == pytest.approx([3.5, 2.5], 0.001)
        assert model.dtype == "float32" and model.n_cols == 2

    model = kmeans.fit(df)
    check(model)
    path = str(tmp_path / "kmeans_model")
    model.write().overwrite().save(path)
    loaded = KMeansModel.load(path)
    check(loaded)
    out = model.transform(df).toPandas()
    assert sorted(out.columns) == ["features", "prediction"]
    labels = out[model.getPredictionCol()].tolist()
    assert len(labels) == 4 and labels[0] == labels[1] and labels[1] != labels[2] and labels[2] == labels[3]
    assert out["prediction"].dtype == np.int32
    model.predict([1.0, 1.0])  # single-vector predict does not raise
    assert loaded.transform(df).toPandas()["prediction"].tolist() == labels


def test_kmeans_spark_compat(tmp_path, golden):
    # python/tests/test_kmeans.py:333-437: centres exactly [[0.5,0.5],[8.5,8.5]], pairs share a label
    from spark_rapids_ml_nai_b200.clustering import KMeans, KMeansModel

    df = pd.DataFrame({"features": golden["compat_X"].tolist(), "weighCol": [2.0] * 4})
    kmeans = KMeans(k=2, solver="auto", maxBlockSizeInMB=0, num_workers=1)
    kmeans.setSeed(1)
    kmeans.setMaxIter(10)
    model = kmeans.fit(df)
    assert model.getDistanceMeasure() == "euclidean"
    model.setPredictionCol("newPrediction")
    assert _sorted(model.clusterCenters()) == [[0.5, 0.5], [8.5, 8.5]]
    rows = model.transform(df).toPandas()
    assert list(rows.columns) == ["features", "weighCol", "newPrediction"]
    p = rows["newPrediction"].tolist()
    assert p[0] == p[1] and p[2] == p[3] and p[0] != p[2]
    assert model.hasSummary is False
    kmeans.save(str(tmp_path / "kmeans"))
    assert KMeans.load(str(tmp_path / "kmeans")).getK() == 2
    model.save(str(tmp_path / "kmeans_model"))
    model2 = KMeansModel.load(str(tmp_path / "kmeans_model"))
    assert all(model.clusterCenters()[i].tolist() == model2.clusterCenters()[i].tolist() for i in range(2))
    assert model.transform(df).toPandas().iloc[0].tolist() == model2.transform(df).toPandas().iloc[0].tolist()


@pytest.mark.parametrize("data_type", ["int8", "int16", "int32", "int64"])
def test_kmeans_numeric_type(data_type, golden):
    # python/tests/test_kmeans.py:232-249: integer typed multi-column input just has to fit
    from spark_rapids_ml_nai_b200.clustering import KMeans

    cols = ["c1", "c2", "c3", "c4", "c5"]
    df = pd.DataFrame(golden["numeric_X"].astype(data_type), columns=cols)
    model = KMeans(num_workers=1, featuresCols=cols, n_clusters=2).fit(df)
    assert np.isfinite(np.array(model.cluster_centers_)).all() and model.n_cols == 5


@pytest.mark.parametrize("feature_type", ["array", "multi_cols"])
@pytest.mark.parametrize("data_type", [np.float32, np.float64])
@pytest.mark.parametrize("max_record_batch", [100, 10000])
def test_kmeans_matches_oracle(feature_type, data_type, max_record_batch, golden):
    # python/tests/test_kmeans.py:252-330 compares with single-GPU cuML (centres rel 1e-3, identical labels);
    # here the comparison partner is the float64 oracle started from the same initial centres
    from spark_rapids_ml_nai_b200 import _native
    from spark_rapids_ml_nai_b200.clustering import KMeans
    from spark_rapids_ml_nai_b200.runtime import LocalDataFrame

    X = golden["blobs1000x20k8_X"].astype(data_type)
    if feature_type == "array":
        pdf, est = pd.DataFrame({"features": list(X)}), KMeans(num_workers=1, n_clusters=8, initMode="random", seed=3)
    else:
        cols = [f"c{i}" for i in range(X.shape[1])]
        pdf, est = pd.DataFrame(X, columns=cols), KMeans(num_workers=1, n_clusters=8, initMode="random", seed=3,
                                                         featuresCols=cols)
    est.setTol(1e-20).setMaxIter(30)
    ds = LocalDataFrame(pdf, num_partitions=1, max_records_per_batch=max_record_batch)
    model = est.fit(ds)
    with _native.Handle(0) as h:
        init = h.init_centroids(torch.from_numpy(X.astype(np.float32)).cuda(), 8, seed=3).cpu().numpy()
    ref = ko.lloyd_fit(X.astype(np.float32), init, max_iter=30, tol=1e-20)
    got = np.array(model.cluster_centers_)
    assert np.allclose(np.array(_sorted(got)), np.array(_sorted(ref.centers)), rtol=1e-3, atol=1e-5)
    labels = model.transform(ds).toPandas()["prediction"].to_numpy()
    l_ref, best, second = ko.assign(X.astype(np.float32), got.astype(np.float32))
    parity.assert_labels_match(labels.astype(np.int32), best, second, l_ref)


def test_kmeans_parallel_init_quality_and_determinism():
    from spark_rapids_ml_nai_b200.clustering import KMeans

    X, true_centers = ko.make_blobs(20000, 16, 10, seed=4)
    df = pd.DataFrame({"features": list(X)})
    a = KMeans(k=10, num_workers=1, seed=7).fit(df)          # default initMode = "k-means||"
    b = KMeans(k=10, num_workers=1, seed=7).fit(df)
    assert a.cluster_centers_ == b.cluster_centers_        # same seed, same result
    inertia = ko.inertia(X, np.array(a.cluster_centers_), ko.predict(X, np.array(a.cluster_centers_)))
    ideal = ko.inertia(X, true_centers, ko.predict(X, true_centers))
    assert inertia <= 1.5 * ideal                           # k-means|| finds (nearly) every blob
    assert len({tuple(c) for c in a.cluster_centers_}) == 10


def test_two_gpu_fit_equals_one_gpu_fit(golden):
    # python/tests/test_kmeans.py:268-330: the multi-GPU fit agrees with the single-GPU one
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (the NCCL path is also exercised by bench.py --gpus N)")
    from spark_rapids_ml_nai_b200.clustering import KMeans

    X = golden["blobs3000x100k20_X"]
    df = pd.DataFrame({"features": list(X)})
    init = dict(n_clusters=20, initMode="random", seed=5)
    one = KMeans(num_workers=1, **init).setTol(1e-20).setMaxIter(10).fit(df)
    two = KMeans(num_workers=2, **init).setTol(1e-20).setMaxIter(10).fit(df)
    assert np.allclose(np.array(one.cluster_centers_), np.array(two.cluster_centers_), rtol=1e-5, atol=1e-6)
