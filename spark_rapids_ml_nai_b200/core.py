"""Plugin core: the estimator / model base classes and the executor-side fit driver.

Mirrors the contract of the reference's ``core.py`` for the KMeans path
(python/src/spark_rapids_ml/core.py): ``_CumlCaller._call_cuml_fit_func`` / ``_train_udf`` (:622-787),
``_CumlEstimator._fit_internal`` (:1006-1046), ``_CumlModel`` / ``_CumlModelWithColumns._transform``
(:1138-1183, :1516-1578) and the persistence readers / writers (:249-336).  Spark itself is replaced by
``runtime`` (barrier stage, partitions, Arrow-sized batches); the backend handle is this package's C ABI.
"""
from __future__ import annotations

import json
import os
import time
from abc import abstractmethod
from collections import namedtuple
from typing import Any, Callable, Dict, Iterator, List, Optional, Tuple, Union

import numpy as np
import pandas as pd

from .params import Params, _CumlParams
from .runtime import BarrierTaskContext, LocalDataFrame, run_barrier_stage
from .utils import _get_gpu_id, get_logger

# keys of the params dict handed to the fit function (reference core.py:150-156)
param_alias = namedtuple("ParamAlias", ("cuml_init", "handle", "num_cols", "part_sizes", "loop", "fit_multiple_params"))(
    "cuml_init", "handle", "num_cols", "part_sizes", "loop", "fit_multiple_params")
# column aliases of the pre-processed dataset (reference core.py:143-147)
alias = namedtuple("Alias", ("data", "label", "row_number"))("cuml_values", "cuml_label", "unique_id")

FitInputType = List[Tuple[Any, Optional[Any], Optional[Any]]]
_TRANSFORM_COALESCE_ROWS = 1 << 20


def _batch_features(pdf: pd.DataFrame, input_col: Optional[str], multi_col_names: Optional[List[str]],
                    order: str, dtype: Any) -> np.ndarray:
    """One Arrow-sized batch -> dense [rows, d] array (reference core.py:727-733 and :1549-1551).

    Multi-column numeric input whose columns are the whole frame, in order, in one block of the right dtype is
    taken as a view (no copy: the loader copies it once, into pinned memory); an array column is stacked row
    by row, as the reference does."""
    if multi_col_names:
        if list(pdf.columns) == list(multi_col_names):
            a = pdf.to_numpy(copy=False)
            if a.dtype == dtype and (a.flags["C_CONTIGUOUS"] if order == "C" else a.flags["F_CONTIGUOUS"]):
                return a
        return np.array(pdf[multi_col_names], order=order, dtype=dtype)
    col = pdf[input_col]
    if len(col) and isinstance(col.iloc[0], np.ndarray) and col.iloc[0].dtype == dtype:
        return np.stack(col.to_numpy()).astype(dtype, order=order, copy=False)
    return np.array(list(col), order=order, dtype=dtype)


# ---- persistence: the on-disk layout of pyspark's DefaultParamsWriter + one JSON line of attributes ----
def _save_metadata(instance: "_CumlParams", path: str, overwrite: bool) -> None:
    if os.path.exists(path):
        if not overwrite:
            raise IOError(f"Path {path} already exists. To overwrite it, please use write.overwrite().save(path)")
        import shutil

        shutil.rmtree(path)
    os.makedirs(os.path.join(path, "metadata"))
    cls = type(instance)
    meta = {
        "class": f"{cls.__module__}.{cls.__name__}",
        "timestamp": int(time.time() * 1000),
        "sparkVersion": "none (spark_rapids_ml_nai_b200 local runtime)",
        "uid": instance.uid,
        "paramMap": {p.name: v for p, v in instance._paramMap.items()},
        "defaultParamMap": {p.name: v for p, v in instance._defaultParamMap.items()},
        "_cuml_params": instance._cuml_params,
        "_num_workers": instance._num_workers,
        "_float32_inputs": instance._float32_inputs,
    }
    with open(os.path.join(path, "metadata", "part-00000"), "w") as f:
        f.write(json.dumps(meta, separators=(",", ":")))
    open(os.path.join(path, "metadata", "_SUCCESS"), "w").close()


def _load_metadata(path: str) -> Dict[str, Any]:
    with open(os.path.join(path, "metadata", "part-00000")) as f:
        return json.loads(f.readline())


def _apply_metadata(instance: "_CumlParams", meta: Dict[str, Any]) -> None:
    instance._resetUid(meta["uid"])
    for name, value in meta.get("defaultParamMap", {}).items():
        if instance.hasParam(name):
            instance._setDefault(**{name: value})
    for name, value in meta["paramMap"].items():
        if instance.hasParam(name):
            instance._set(**{name: value})
    instance._cuml_params = meta["_cuml_params"]
    instance._num_workers = meta["_num_workers"]
    instance._float32_inputs = meta["_float32_inputs"]


class _Writer:
    def __init__(self, instance: Any) -> None:
        self.instance, self._overwrite = instance, False

    def overwrite(self) -> "_Writer":
        self._overwrite = True
        return self

    def save(self, path: str) -> None:
        self.instance._save_impl(path, self._overwrite)


class _CumlCommon:
    def write(self) -> _Writer:
        return _Writer(self)

    def save(self, path: str) -> None:
        self.write().save(path)

    @staticmethod
    def _set_gpu_device(context: Optional[BarrierTaskContext], is_transform: bool = False) -> int:
        """Selects the task's GPU before any CUDA work (reference core.py:343-388)."""
        import torch

        gpu_id = _get_gpu_id(context.partitionId() if context is not None else 0, is_transform)
        torch.cuda.set_device(gpu_id)
        return gpu_id


class _CumlCaller(_CumlParams, _CumlCommon):
    """Estimator-side driver: pre-processing, repartitioning, and the barrier fit stage."""

    def __init__(self) -> None:
        super().__init__()
        self._initialize_cuml_params()
        self.logger = get_logger(self.__class__)

    # -- hooks -----------------------------------------------------------------------------------
    @abstractmethod
    def _out_schema(self) -> Any:
        raise NotImplementedError()

    @abstractmethod
    def _get_cuml_fit_func(self, dataset: LocalDataFrame, extra_params: Optional[List[Dict[str, Any]]] = None
                           ) -> Callable[[FitInputType, Dict[str, Any]], Dict[str, Any]]:
        """Runs on the driver; the returned function runs on the executors and must not capture the
        estimator or the dataset (reference core.py:589-620, tests/test_common_estimator.py:179-181)."""
        raise NotImplementedError()

    def _require_nccl_ucx(self) -> Tuple[bool, bool]:
        return (True, False)

    def _fit_array_order(self) -> str:
        return "F"

    def _validate_parameters(self) -> None:
        """The reference validates by copying the params into the JVM estimator (core.py:564-587);
        without a JVM the same constraints are checked here, with the JVM's message text."""

    def _pre_process_data(self, dataset: LocalDataFrame) -> Tuple[Optional[str], Optional[List[str]], int, np.dtype]:
        """Resolves the feature column(s), the dimension (from the first row, reference core.py:523)
        and the compute dtype (float32 unless float32_inputs=False, core.py:463,481-486)."""
        input_col, input_cols = self._get_input_columns()
        first = dataset.first()
        if input_cols is not None:
            missing = [c for c in input_cols if c not in dataset.columns]
            if missing:
                raise ValueError(f"feature columns {missing} are not in the dataset")
            dimension = len(input_cols)
            src_dtype = np.result_type(*[np.asarray(first[c]).dtype for c in input_cols])
        else:
            if input_col not in dataset.columns:
                raise ValueError(f"feature column '{input_col}' is not in the dataset")
            v = np.asarray(first[input_col])
            if v.ndim != 1:
                raise ValueError("Unsupported feature column: expected an array / vector per row")
            dimension = int(v.shape[0])
            src_dtype = v.dtype
        dtype = np.dtype(np.float32) if (self._float32_inputs or src_dtype != np.float64) else np.dtype(np.float64)
        return input_col, input_cols, dimension, dtype

    def _call_cuml_fit_func(self, dataset: LocalDataFrame, partially_collect: bool = True) -> List[pd.DataFrame]:
        self._validate_parameters()
        cls = self.__class__
        input_col, multi_col_names, dimension, dtype = self._pre_process_data(dataset)
        num_workers = self.num_workers
        if dataset.getNumPartitions() != num_workers:
            dataset = dataset.repartition(num_workers)
        params: Dict[str, Any] = {param_alias.cuml_init: self.cuml_params, param_alias.fit_multiple_params: []}
        cuml_fit_func = self._get_cuml_fit_func(dataset, None)
        array_order = self._fit_array_order()
        (enable_nccl, require_ucx) = self._require_nccl_ucx()

        def _train_udf(pdf_iter: Iterator[pd.DataFrame]) -> Iterator[pd.DataFrame]:
            from .context import B200Context

            context = BarrierTaskContext.get()
            partition_id = context.partitionId()
            logger = get_logger(cls)
            gpu_id = _CumlCommon._set_gpu_device(context)
            logger.info("Loading data into python worker memory")
            inputs: FitInputType = []
            sizes: List[int] = []
            for pdf in pdf_iter:
                sizes.append(pdf.shape[0])
                inputs.append((_batch_features(pdf, input_col, multi_col_names, array_order, dtype), None, None))
            if len(sizes) == 0 or all(sz == 0 for sz in sizes):
                raise RuntimeError(
                    "A python worker received no data.  Please increase amount of data or use fewer workers."
                )
            logger.info("Initializing cuml context")
            with B200Context(partition_id, num_workers, context, enable_nccl, require_ucx, device=gpu_id) as cc:
                params[param_alias.handle] = cc.handle
                params[param_alias.part_sizes] = sizes
                params[param_alias.num_cols] = dimension
                params[param_alias.loop] = cc._loop
                logger.info("Invoking cuml fit")
                # the fit function may free its inputs: do not rely on them afterwards
                result = cuml_fit_func(inputs, params)
                logger.info("Cuml fit complete")
            if partially_collect:
                if enable_nccl:
                    context.barrier()
                if context.partitionId() == 0:
                    yield pd.DataFrame(data=result)
            else:
                yield pd.DataFrame(data=result)

        return run_barrier_stage(_train_udf, dataset, num_workers)


class _CumlEstimator(_CumlCaller):
    @abstractmethod
    def _create_pyspark_model(self, result: Dict[str, Any]) -> "_CumlModel":
        raise NotImplementedError()

    def _as_dataset(self, dataset: Any) -> LocalDataFrame:
        if isinstance(dataset, LocalDataFrame):
            return dataset
        input_col, _ = self._get_input_columns()
        return LocalDataFrame(dataset, num_partitions=1, features_col=input_col or "features")

    def fit(self, dataset: Any, params: Optional[Dict[Any, Any]] = None) -> "_CumlModel":
        """Fits a model; ``params`` overrides Spark params for this call only, leaving the estimator
        untouched (reference tests/test_common_estimator.py:489-512)."""
        est = self.copy(params) if params else self
        return est._fit(est._as_dataset(dataset))

    def _fit(self, dataset: LocalDataFrame) -> "_CumlModel":
        frames = self._call_cuml_fit_func(dataset, partially_collect=True)
        row = frames[0].iloc[0].to_dict()
        model = self._create_pyspark_model(row)
        model._num_workers = self._num_workers
        model._float32_inputs = self._float32_inputs
        self._copyValues(model)
        self._copy_cuml_params(model)
        return model

    def _save_impl(self, path: str, overwrite: bool) -> None:
        _save_metadata(self, path, overwrite)

    @classmethod
    def load(cls, path: str) -> "_CumlEstimator":
        meta = _load_metadata(path)
        est = cls()
        _apply_metadata(est, meta)
        return est


class _CumlModel(_CumlParams, _CumlCommon):
    """Fitted model: keeps its attributes (for persistence) and runs the transform stage."""

    def __init__(self, *, dtype: Optional[str] = None, n_cols: Optional[int] = None, **model_attributes: Any) -> None:
        super().__init__()
        self._initialize_cuml_params()
        self.logger = get_logger(self.__class__)
        self.dtype = dtype
        self.n_cols = n_cols
        self._model_attributes = model_attributes

    def _get_model_attributes(self) -> Dict[str, Any]:
        d = dict(self._model_attributes)
        d["dtype"], d["n_cols"] = self.dtype, self.n_cols
        return d

    @classmethod
    def _from_row(cls, model_attributes: Dict[str, Any]) -> "_CumlModel":
        return cls(**model_attributes)

    @abstractmethod
    def _get_cuml_transform_func(self, dataset: LocalDataFrame) -> Tuple[Callable, Callable, Optional[Callable]]:
        """(construct, transform, evaluate): construct builds the backend object once per task, transform
        maps one batch to a pandas Series (reference core.py:1147-1177)."""
        raise NotImplementedError()

    def _transform_array_order(self) -> str:
        return "F"

    def _out_schema(self, input_schema: Any = None) -> str:
        raise NotImplementedError()

    def transform(self, dataset: Any) -> LocalDataFrame:
        """Adds the prediction column, keeping every input column and the row order
        (reference core.py:1516-1578; tests/test_common_estimator.py:514-525)."""
        if not isinstance(dataset, LocalDataFrame):
            input_col, _ = self._get_input_columns()
            dataset = LocalDataFrame(dataset, num_partitions=1, features_col=input_col or "features")
        input_col, input_cols = self._get_input_columns()
        construct, transform_fn, _ = self._get_cuml_transform_func(dataset)
        array_order = self._transform_array_order()
        dtype = np.dtype(self.dtype or "float32")
        pred_col = self.getOrDefault("predictionCol")
        out_parts: List[pd.DataFrame] = []
        for pid in range(dataset.getNumPartitions()):
            # plain (non-barrier) task: GPU = partition id modulo the GPU count (reference core.py:358-363)
            _CumlCommon._set_gpu_device(BarrierTaskContext(pid, dataset.getNumPartitions(), False), is_transform=True)
            backend = construct()
            series: List[pd.Series] = []
            # Arrow-sized batches (10 000 rows by default) are far too small for one device round trip each:
            # a backend that offers `predict_many` gets them coalesced into >= _TRANSFORM_COALESCE_ROWS rows
            # (one packed host -> device copy, one kernel pass, one device -> host read per group)
            pending: List[np.ndarray] = []
            pending_rows = 0

            def drain() -> None:
                nonlocal pending, pending_rows
                if pending:
                    series.append(pd.Series(backend.predict_many(pending)))
                    pending, pending_rows = [], 0

            try:
                for pdf in dataset.batches(pid):
                    data = _batch_features(pdf, input_col, input_cols, array_order, dtype)
                    if hasattr(backend, "predict_many"):
                        pending.append(data)
                        pending_rows += data.shape[0]
                        if pending_rows >= _TRANSFORM_COALESCE_ROWS:
                            drain()
                    else:
                        series.append(transform_fn(backend, data))
                drain()
            finally:
                if hasattr(backend, "close"):
                    backend.close()  # the task owns the backend object: release its device handle now
            part = dataset.partitions[pid].copy()
            part[pred_col] = (pd.concat(series, ignore_index=True) if series else pd.Series([], dtype="int32")).to_numpy()
            out_parts.append(part)
        return dataset.withPartitions(out_parts)

    def _save_impl(self, path: str, overwrite: bool) -> None:
        _save_metadata(self, path, overwrite)
        os.makedirs(os.path.join(path, "data"))
        with open(os.path.join(path, "data", "part-00000"), "w") as f:
            f.write(json.dumps(self._get_model_attributes()) + "\n")
        open(os.path.join(path, "data", "_SUCCESS"), "w").close()

    @classmethod
    def load(cls, path: str) -> "_CumlModel":
        meta = _load_metadata(path)
        with open(os.path.join(path, "data", "part-00000")) as f:
            attrs = json.loads(f.readline())
        model = cls(**attrs)
        _apply_metadata(model, meta)
        return model


class _CumlModelWithPredictionCol(_CumlModel):
    def _is_single_pred(self) -> bool:
        return True
