"""KMeans / KMeansModel with the spark_rapids_ml.clustering surface, on the B200-native backend.

Mirrors reference python/src/spark_rapids_ml/clustering.py:67-499: the Spark<->backend parameter map
(:68-90), the tol=0 remap (:92-108), the backend defaults (:110-121), constructor defaults and setters
(:283-331), the fit-function slot (:336-386), the result schema (:388-397), the model with
``clusterCenters`` / ``hasSummary`` / ``predict`` (:403-457) and the transform slot (:466-499).  The cuML
calls inside the two slots are replaced by the C ABI of libb200km (include/b200km.h).
"""
from __future__ import annotations

import logging
from typing import Any, Callable, Dict, List, Optional, Tuple, Union

import numpy as np
import pandas as pd

from . import _native
from .core import (FitInputType, _CumlEstimator, _CumlModelWithPredictionCol, param_alias)
from .loader import partitions_to_device
from .params import Param, TypeConverters, _CumlClass, _CumlParams, keyword_only
from .runtime import LocalDataFrame
from .utils import get_logger

logger = logging.getLogger("spark_rapids_ml_nai_b200.clustering")


class KMeansClass(_CumlClass):
    @classmethod
    def _param_mapping(cls) -> Dict[str, Optional[str]]:
        # "" = accepted and ignored, None = raises when set (reference clustering.py:68-90)
        return {
            "distanceMeasure": None,
            "initMode": "init",
            "k": "n_clusters",
            "initSteps": "",
            "maxIter": "max_iter",
            "seed": "random_state",
            "tol": "tol",
            "weightCol": None,
            "solver": "",
            "maxBlockSizeInMB": "",
        }

    @classmethod
    def _param_value_mapping(cls) -> Dict[str, Callable[[Any], Union[None, str, float, int]]]:
        def tol_value_mapper(x: float) -> float:
            if x == 0.0:
                get_logger(cls).warning(
                    "tol=0 is not supported in cuml yet. "
                    + "It will be mapped to smallest positive float, i.e. numpy.finfo('float32').tiny."
                )
                return np.finfo("float32").tiny.item()
            return x

        return {"tol": lambda x: tol_value_mapper(x)}

    def _get_cuml_params_default(self) -> Dict[str, Any]:
        return {
            "n_clusters": 8,
            "max_iter": 300,
            "tol": 0.0001,
            "verbose": False,
            "random_state": 1,
            "init": "scalable-k-means++",
            "n_init": 1,
            "oversampling_factor": 2.0,
            "max_samples_per_batch": 32768,
        }


class _KMeansCumlParams(_CumlParams):
    """Spark Params shared by KMeans and KMeansModel (pyspark's _KMeansParams + HasFeaturesCols)."""

    def __init__(self) -> None:
        super().__init__()
        S, I, F = TypeConverters.toString, TypeConverters.toInt, TypeConverters.toFloat
        self._declare("featuresCol", "features column name.", S)
        self._declare("featuresCols", "features column names for multi-column input.", TypeConverters.toListString)
        self._declare("predictionCol", "prediction column name.", S)
        self._declare("k", "The number of clusters to create. Must be > 1.", I)
        self._declare("initMode", 'The initialization algorithm: "random" or "k-means||".', S)
        self._declare("initSteps", "The number of steps for k-means|| initialization mode. Must be > 0.", I)
        self._declare("maxIter", "max number of iterations (>= 0).", I)
        self._declare("tol", "the convergence tolerance for iterative algorithms (>= 0).", F)
        self._declare("seed", "random seed.", I)
        self._declare("distanceMeasure", "the distance measure: 'euclidean' or 'cosine'.", S)
        self._declare("weightCol", "weight column name.", S)
        self._declare("solver", 'The solver algorithm: "auto", "row" or "block".', S)
        self._declare("maxBlockSizeInMB", "maximum memory in MB for stacking input data into blocks.", F)
        self._setDefault(featuresCol="features", predictionCol="prediction", k=2, initMode="k-means||", initSteps=2,
                         tol=1e-4, maxIter=20, distanceMeasure="euclidean", solver="auto", maxBlockSizeInMB=0.0)
        # the default seed is restricted to a signed 32-bit value (reference clustering.py:134-135)
        self._setDefault(seed=hash(type(self).__name__) & 0x07FFFFFFF)

    # getters of pyspark's _KMeansParams
    def getK(self) -> int:
        return self.getOrDefault("k")

    def getInitMode(self) -> str:
        return self.getOrDefault("initMode")

    def getInitSteps(self) -> int:
        return self.getOrDefault("initSteps")

    def getMaxIter(self) -> int:
        return self.getOrDefault("maxIter")

    def getTol(self) -> float:
        return self.getOrDefault("tol")

    def getSeed(self) -> int:
        return self.getOrDefault("seed")

    def getDistanceMeasure(self) -> str:
        return self.getOrDefault("distanceMeasure")

    def getPredictionCol(self) -> str:
        return self.getOrDefault("predictionCol")

    def getFeaturesCols(self) -> List[str]:
        return self.getOrDefault("featuresCols")

    def getFeaturesCol(self) -> Union[str, List[str]]:
        if self.isDefined("featuresCols"):
            return self.getFeaturesCols()
        if self.isDefined("featuresCol"):
            return self.getOrDefault("featuresCol")
        raise RuntimeError("featuresCol is not set")

    def setFeaturesCol(self, value: Union[str, List[str]]) -> Any:
        return self._set_params(featuresCol=value) if isinstance(value, str) else self._set_params(featuresCols=value)

    def setFeaturesCols(self, value: List[str]) -> Any:
        return self._set_params(featuresCols=value)

    def setPredictionCol(self, value: str) -> Any:
        self._set_params(predictionCol=value)
        return self


# ---- executor-side bodies of the two slots ---------------------------------------------------------------
def _initial_centroids(handle: "_native.Handle", X: Any, cuml_params: Dict[str, Any]) -> Any:
    """``init`` may be an explicit [k,d] array (identical on all ranks), "random", or
    "k-means||" / "scalable-k-means++" (reference clustering.py:72,117; tests/test_kmeans.py:84,93)."""
    import torch

    k = int(cuml_params["n_clusters"])
    init = cuml_params.get("init", "scalable-k-means++")
    seed = int(cuml_params.get("random_state", 1) or 0)
    if not isinstance(init, str):
        arr = np.ascontiguousarray(np.asarray(init, dtype=np.float32))
        if arr.shape != (k, X.shape[1]):
            raise ValueError(f"init array has shape {arr.shape}, expected {(k, X.shape[1])}")
        return torch.from_numpy(arr).to(X.device)
    if init == "random":
        return handle.init_centroids(X, k, mode=_native.INIT_RANDOM, seed=seed)
    if init in ("k-means||", "scalable-k-means++"):
        return handle.init_centroids(X, k, mode=_native.INIT_KMEANS_PARALLEL, seed=seed,
                                     oversampling=float(cuml_params.get("oversampling_factor", 2.0)))
    raise ValueError(f"Value '{init}' for 'init' param is unsupported")


def kmeans_fit_partitions(parts: List, handle: "_native.Handle", cuml_params: Dict[str, Any], num_cols: int,
                          want_labels: bool = False) -> Dict[str, Any]:
    """Body of the fit-function slot (reference clustering.py:348-384): host batches -> device shard ->
    initial centroids -> device-resident Lloyd loop -> result dict in the ``_out_schema`` shape."""
    import torch

    device = torch.device("cuda", handle.device)
    X = partitions_to_device(parts, device, free_inputs=True)
    if X.shape[1] != num_cols:
        raise ValueError(f"partitions have {X.shape[1]} columns, expected {num_cols}")
    k = int(cuml_params["n_clusters"])
    if k < 1:
        raise ValueError(f"k given invalid value {k}")
    init = _initial_centroids(handle, X, cuml_params)
    centers, inertia, n_iter, labels = handle.fit(
        X, init, max_iter=int(cuml_params["max_iter"]), tol=float(cuml_params["tol"]), want_labels=want_labels)
    logger.info(f"iterations: {n_iter}, inertia: {inertia}")
    out = {
        "cluster_centers_": [centers.cpu().numpy().tolist()],
        "n_cols": int(num_cols),
        "dtype": "float32",
        "n_iter": int(n_iter),
        "inertia": float(inertia),
    }
    if want_labels:
        out["labels"] = labels
    return out


class _DevicePredictor:
    """Backend object of the transform slot: centroids resident on the device, one handle per task."""

    def __init__(self, centers: np.ndarray, device_index: int):
        import torch

        self.handle = _native.Handle(device=device_index)
        self.device = torch.device("cuda", device_index)
        self.centers = torch.from_numpy(np.ascontiguousarray(centers, dtype=np.float32)).to(self.device)

    def predict(self, data: np.ndarray) -> np.ndarray:
        import torch

        if data.shape[0] == 0:
            return np.empty(0, dtype=np.int32)
        X = partitions_to_device([data], self.device, free_inputs=False)
        labels, _ = self.handle.predict(X, self.centers)
        return labels.cpu().numpy()

    def predict_many(self, batches: List[np.ndarray]) -> np.ndarray:
        """Labels of several host batches at once: one packed host -> device copy, one pass, one read-back."""
        batches = [b for b in batches if b.shape[0]]
        if not batches:
            return np.empty(0, dtype=np.int32)
        X = partitions_to_device(list(batches), self.device, free_inputs=False)
        labels, _ = self.handle.predict(X, self.centers)
        return labels.cpu().numpy()

    def close(self) -> None:
        if self.handle is not None:
            self.handle.destroy()
            self.handle = None

    def __del__(self) -> None:
        try:
            self.close()
        except Exception:
            pass


class KMeans(KMeansClass, _CumlEstimator, _KMeansCumlParams):
    """Drop-in for ``spark_rapids_ml.clustering.KMeans``: Lloyd's algorithm from k initial centres
    ("k-means||" or "random"), one task per GPU, rows sharded across the tasks.

    >>> gpu_kmeans = KMeans(k=2).setFeaturesCol("features").setMaxIter(10)
    >>> gpu_model = gpu_kmeans.fit(df)
    >>> gpu_model.clusterCenters()
    [array([0.5, 0.5]), array([8.5, 8.5])]
    """

    @keyword_only
    def __init__(self, *, featuresCol: Union[str, List[str]] = "features", predictionCol: str = "prediction", k: int = 2,
                 initMode: str = "k-means||", tol: float = 0.0001, maxIter: int = 20, seed: Optional[int] = None,
                 num_workers: Optional[int] = None, verbose: Union[int, bool] = False, **kwargs: Any) -> None:
        super().__init__()
        kw = dict(self._input_kwargs)
        if kw.get("seed") is None:
            kw.pop("seed", None)
        if kw.get("num_workers") is None:
            kw.pop("num_workers", None)
        self._set_params(**kw)

    def setK(self, value: int) -> "KMeans":
        return self._set_params(k=value)

    def setMaxIter(self, value: int) -> "KMeans":
        return self._set_params(maxIter=value)

    def setSeed(self, value: int) -> "KMeans":
        if value > 0x07FFFFFFF:
            raise ValueError("cuML seed value must be a 32-bit integer.")
        return self._set_params(seed=value)

    def setTol(self, value: float) -> "KMeans":
        return self._set_params(tol=value)

    def setInitMode(self, value: str) -> "KMeans":
        return self._set_params(initMode=value)

    def setWeightCol(self, value: str) -> "KMeans":
        raise ValueError("'weightCol' is not supported by cuML.")

    def _validate_parameters(self) -> None:
        # text of the JVM's IllegalArgumentException (reference tests/test_kmeans.py:440-457)
        if self.getK() < 1:
            raise ValueError(f"k given invalid value {self.getK()}")
        if self.getMaxIter() < 0:
            raise ValueError(f"maxIter given invalid value {self.getMaxIter()}")
        if self.getTol() < 0:
            raise ValueError(f"tol given invalid value {self.getTol()}")
        if self.getInitMode() not in ("k-means||", "random", "scalable-k-means++"):
            raise ValueError(f"initMode given invalid value {self.getInitMode()}")

    def _fit_array_order(self) -> str:
        return "C"

    def _get_cuml_fit_func(self, dataset: LocalDataFrame, extra_params: Optional[List[Dict[str, Any]]] = None
                           ) -> Callable[[FitInputType, Dict[str, Any]], Dict[str, Any]]:
        def _cuml_fit(dfs: FitInputType, params: Dict[str, Any]) -> Dict[str, Any]:
            # heavy imports stay inside the closure: it runs on the executor (reference clustering.py:352-353)
            from spark_rapids_ml_nai_b200.clustering import kmeans_fit_partitions

            out = kmeans_fit_partitions([x for (x, _, _) in dfs], params[param_alias.handle],
                                        params[param_alias.cuml_init], params[param_alias.num_cols])
            return {"cluster_centers_": out["cluster_centers_"], "n_cols": out["n_cols"], "dtype": out["dtype"]}

        return _cuml_fit

    def _out_schema(self) -> str:
        return "cluster_centers_ array<array<double>>, n_cols int, dtype string"

    def _create_pyspark_model(self, result: Dict[str, Any]) -> "KMeansModel":
        return KMeansModel._from_row(result)


class KMeansModel(KMeansClass, _CumlModelWithPredictionCol, _KMeansCumlParams):
    """Fitted centres; ``transform`` appends the index of the nearest centre as an int column."""

    def __init__(self, cluster_centers_: List[List[float]], n_cols: int, dtype: str):
        super().__init__(dtype=dtype, n_cols=n_cols, cluster_centers_=[list(map(float, c)) for c in cluster_centers_])
        self.cluster_centers_ = self._model_attributes["cluster_centers_"]

    def clusterCenters(self) -> List[np.ndarray]:
        return [np.array(x) for x in self.cluster_centers_]

    @property
    def hasSummary(self) -> bool:
        return False

    def cpu(self) -> Any:
        """The reference builds a pyspark.ml KMeansModel through the JVM (clustering.py:422-442)."""
        try:
            import pyspark  # noqa: F401
        except ImportError as e:
            raise RuntimeError("cpu() needs pyspark and a JVM, which this environment does not have") from e
        raise RuntimeError("cpu() needs a running SparkSession")

    def predict(self, value: Any) -> int:
        """Single-vector prediction.  The reference delegates it to the CPU pyspark model because its GPU
        backend rejects one sample (clustering.py:453-457); a k x d distance needs no GPU either."""
        v = np.asarray(value.toArray() if hasattr(value, "toArray") else value, dtype=np.float64)
        c = np.asarray(self.cluster_centers_, dtype=np.float64)
        return int(np.argmin(((c - v[None, :]) ** 2).sum(axis=1)))

    def _out_schema(self, input_schema: Any = None) -> str:
        return "int"

    def _transform_array_order(self) -> str:
        return "C"

    def _get_cuml_transform_func(self, dataset: LocalDataFrame) -> Tuple[Callable, Callable, Optional[Callable]]:
        cluster_centers_ = self.cluster_centers_
        dtype = self.dtype

        def _construct_kmeans() -> Any:
            import torch

            from spark_rapids_ml_nai_b200.clustering import _DevicePredictor

            return _DevicePredictor(np.array(cluster_centers_).astype(dtype), torch.cuda.current_device())

        def _transform_internal(kmeans: Any, df: Union[pd.DataFrame, np.ndarray]) -> pd.Series:
            return pd.Series(kmeans.predict(np.asarray(df)))

        return _construct_kmeans, _transform_internal, None
