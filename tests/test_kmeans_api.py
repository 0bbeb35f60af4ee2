"""CPU-side tests of the drop-in surface: parameter bridge, persistence, error behaviour.  They re-express
the assertions of the reference's python/tests/test_kmeans.py:66-146,333-457 without Spark or a GPU."""
import logging

import numpy as np
import pytest

from spark_rapids_ml_nai_b200.clustering import KMeans, KMeansModel


def test_backend_defaults_match_reference():
    # python/tests/test_kmeans.py:66-71 pins these against cuml.KMeans' signature (minus handle/output_type)
    assert KMeans()._get_cuml_params_default() == {
        "n_clusters": 8, "max_iter": 300, "tol": 0.0001, "verbose": False, "random_state": 1,
        "init": "scalable-k-means++", "n_init": 1, "oversampling_factor": 2.0, "max_samples_per_batch": 32768}


def _assert_params(est, spark, cuml):
    for k, v in spark.items():
        assert est.getOrDefault(k) == v, k
    for k, v in cuml.items():
        if k == "num_workers":
            continue
        assert est.cuml_params[k] == v, k


def test_kmeans_params(tmp_path, caplog):
    # python/tests/test_kmeans.py:79-146
    default_spark = {"initMode": "k-means||", "k": 2, "maxIter": 20}
    default_cuml = {"n_clusters": 2, "max_iter": 20, "tol": 0.0001, "verbose": False, "init": "k-means||",
                    "oversampling_factor": 2.0, "max_samples_per_batch": 32768}
    _assert_params(KMeans(), default_spark, default_cuml)

    spark_kmeans = KMeans(k=10, maxIter=100)
    _assert_params(spark_kmeans, {**default_spark, "k": 10, "maxIter": 100},
                   {**default_cuml, "n_clusters": 10, "max_iter": 100})

    cuml_params = {"n_clusters": 10, "max_iter": 100, "tol": 1e-1, "verbose": True, "random_state": 5,
                   "init": "k-means||", "oversampling_factor": 3, "max_samples_per_batch": 45678}
    cuml_kmeans = KMeans(**cuml_params)
    exp_spark = {**default_spark, "k": 10, "maxIter": 100}
    exp_cuml = {**default_cuml, **cuml_params}
    _assert_params(cuml_kmeans, exp_spark, exp_cuml)

    path = str(tmp_path / "kmeans")
    cuml_kmeans.write().overwrite().save(path)
    _assert_params(KMeans.load(path), exp_spark, exp_cuml)

    with pytest.raises(ValueError, match="set one or the other"):
        KMeans(k=2, n_clusters=10)

    kmeans_float32 = KMeans(float32_inputs=False)
    assert "float32_inputs to False" not in caplog.text
    assert not kmeans_float32._float32_inputs
    with pytest.raises(ValueError, match="Unsupported param"):
        KMeans(no_such_param=1)


def test_spark_compat_surface(caplog):
    # python/tests/test_kmeans.py:365-384: accepted-and-ignored params, unsupported ones, getters, clear()
    kmeans = KMeans(k=2, solver="auto", maxBlockSizeInMB=0)
    kmeans.setSeed(1)
    kmeans.setMaxIter(10)
    with pytest.raises(ValueError):
        kmeans.setWeightCol("weighCol")
    with pytest.raises(ValueError, match="not supported"):
        KMeans(distanceMeasure="cosine")
    assert (kmeans.getMaxIter(), kmeans.getK(), kmeans.getSeed()) == (10, 2, 1)
    assert kmeans.cuml_params["random_state"] == 1
    kmeans.clear(kmeans.maxIter)
    assert kmeans.getMaxIter() == 20 and kmeans.cuml_params["max_iter"] == 20
    with pytest.raises(ValueError, match="32-bit"):
        kmeans.setSeed(0x07FFFFFFF + 1)
    assert kmeans.getDistanceMeasure() == "euclidean"
    assert KMeans().getSeed() <= 0x07FFFFFFF
    with caplog.at_level(logging.WARNING):
        kmeans.setTol(0.0)
    assert "tol=0 is not supported in cuml yet" in caplog.text
    assert kmeans.cuml_params["tol"] == np.finfo("float32").tiny.item()
    km = KMeans().setFeaturesCol(["c1", "c2"])
    assert km.getFeaturesCol() == ["c1", "c2"] and km._get_input_columns() == (None, ["c1", "c2"])
    assert KMeans().setFeaturesCol("f").getFeaturesCol() == "f"


def test_fit_param_overrides_do_not_mutate_estimator():
    # python/tests/test_common_estimator.py:489-512
    est = KMeans(k=3)
    other = est.copy({est.k: 5, est.maxIter: 7})
    assert (est.getK(), est.cuml_params["n_clusters"]) == (3, 3)
    assert (other.getK(), other.cuml_params["n_clusters"], other.cuml_params["max_iter"]) == (5, 5, 7)
    with pytest.raises(TypeError):
        est.copy({"k": 5})


def test_model_surface_and_persistence(tmp_path):
    # python/tests/test_kmeans.py:393-437: centres, hasSummary False, save/load keeps centres exactly
    model = KMeansModel([[0.5, 0.5], [8.5, 8.5]], 2, "float32")
    model.setPredictionCol("newPrediction")
    assert model.getPredictionCol() == "newPrediction"
    assert model.hasSummary is False
    assert [c.tolist() for c in model.clusterCenters()] == [[0.5, 0.5], [8.5, 8.5]]
    assert model.predict([0.0, 0.0]) == 0 and model.predict(np.array([9.0, 8.0])) == 1
    path = str(tmp_path / "model")
    model.save(path)
    with pytest.raises(IOError):
        model.save(path)
    model.write().overwrite().save(path)
    m2 = KMeansModel.load(path)
    assert m2.hasSummary is False and m2.getPredictionCol() == "newPrediction"
    assert all(model.clusterCenters()[i].tolist() == m2.clusterCenters()[i].tolist() for i in range(2))
    assert (m2.n_cols, m2.dtype) == (2, "float32")
    # the data file is one JSON line {cluster_centers_, n_cols, dtype}, as the reference writes it (core.py:306-312)
    import json
    import os

    line = open(os.path.join(path, "data", "part-00000")).readline()
    assert set(json.loads(line)) == {"cluster_centers_", "n_cols", "dtype"}


def test_parameter_validation_messages():
    # python/tests/test_kmeans.py:440-457 (the JVM's IllegalArgumentException text)
    import pandas as pd

    df = pd.DataFrame({"features": [[1.0, 2.0], [3.0, 1.0]]})
    with pytest.raises(ValueError, match="k given invalid value -1"):
        KMeans(k=-1, num_workers=1)._validate_parameters()
    with pytest.raises(ValueError, match="maxIter given invalid value -1"):
        KMeans(num_workers=1).setMaxIter(-1)._validate_parameters()
    with pytest.raises(ValueError, match="feature column"):
        KMeans(num_workers=1).setFeaturesCol("nope")._pre_process_data(KMeans()._as_dataset(df))


def test_closures_do_not_capture_estimator_or_dataset():
    # python/tests/test_common_estimator.py:179-181: the fit function is shipped to executors
    import cloudpickle
    import pandas as pd

    from spark_rapids_ml_nai_b200.runtime import LocalDataFrame

    est = KMeans(k=2)
    ds = LocalDataFrame(pd.DataFrame({"features": [[0.0, 0.0], [1.0, 1.0]]}))
    blob = cloudpickle.dumps(est._get_cuml_fit_func(ds))
    assert b"LocalDataFrame" not in blob and est.uid.encode() not in blob
    model = KMeansModel([[0.0, 0.0]], 2, "float32")
    construct, transform, evaluate = model._get_cuml_transform_func(ds)
    assert evaluate is None and model.uid.encode() not in cloudpickle.dumps((construct, transform))
