"""The Spark-free barrier runtime and the executor-side fit driver, with a dummy backend and
world_size-2 gloo workers on CPU.  Re-expresses python/tests/test_common_estimator.py:179-239,528-558 and the
comm bring-up pattern of python/tests/test_ucx.py:34-99."""
from typing import Any, Callable, Dict, List, Optional

import numpy as np
import pandas as pd
import pytest

from spark_rapids_ml_nai_b200 import core
from spark_rapids_ml_nai_b200.params import TypeConverters, _CumlClass
from spark_rapids_ml_nai_b200.runtime import BarrierTaskContext, LocalDataFrame, run_barrier_stage
from spark_rapids_ml_nai_b200.utils import PartitionDescriptor, _concat_and_free


def _udf_collects(pdf_iter):
    ctx = BarrierTaskContext.get()
    rows = sum(len(p) for p in pdf_iter)
    gathered = ctx.allGather(str(rows))
    ctx.barrier()
    pd_desc = PartitionDescriptor.build([rows], 3)
    yield pd.DataFrame({"rank": [ctx.partitionId()], "rows": [rows], "all": [",".join(gathered)],
                        "m": [pd_desc.m], "desc_rank": [pd_desc.rank]})


@pytest.mark.parametrize("workers", [1, 2])
def test_barrier_stage_allgather_and_partitioning(workers):
    X = np.arange(30, dtype=np.float32).reshape(10, 3)
    ds = LocalDataFrame(X, num_partitions=1, max_records_per_batch=4)
    frames = run_barrier_stage(_udf_collects, ds, workers)
    assert len(frames) == workers
    assert sorted(f["rank"][0] for f in frames) == list(range(workers))
    assert sum(f["rows"][0] for f in frames) == 10
    expected = ",".join(str(10 // workers) for _ in range(workers))
    for f in frames:
        assert f["all"][0] == expected and f["m"][0] == 10 and f["desc_rank"][0] == f["rank"][0]


def test_local_dataframe_batches_keep_order():
    X = np.arange(50, dtype=np.float32).reshape(25, 2)
    ds = LocalDataFrame(X, num_partitions=2, max_records_per_batch=5)
    assert ds.getNumPartitions() == 2 and ds.count() == 25
    got = np.concatenate([np.array(list(b["features"])) for p in range(2) for b in ds.batches(p)])
    assert np.array_equal(got, X)
    assert all(len(b) <= 5 for b in ds.batches(0))
    assert ds.repartition(5).getNumPartitions() == 5


def test_concat_and_free():
    # python/tests/test_utils.py:28-50
    a, b = np.ones((2, 3), np.float32), np.zeros((1, 3), np.float32)
    lst = [a, b]
    out = _concat_and_free(lst, order="C")
    assert out.shape == (3, 3) and out.flags["C_CONTIGUOUS"] and lst == []


# ---- a dummy estimator on the plugin core: the boundary contract without a GPU -----------------------
class _DummyParams(core._CumlParams):
    def __init__(self) -> None:
        super().__init__()
        self._declare("featuresCol", "", TypeConverters.toString)
        self._declare("featuresCols", "", TypeConverters.toListString)
        self._declare("predictionCol", "", TypeConverters.toString)
        self._declare("alpha", "", TypeConverters.toFloat)
        self._setDefault(featuresCol="features", predictionCol="prediction", alpha=1.0)


class _DummyClass(_CumlClass):
    @classmethod
    def _param_mapping(cls) -> Dict[str, Optional[str]]:
        return {"alpha": "a"}

    def _get_cuml_params_default(self) -> Dict[str, Any]:
        return {"a": 10.0}


class DummyModel(_DummyClass, core._CumlModel, _DummyParams):
    def __init__(self, n_cols: int, dtype: str, total_rows: int, ranks: List[int]):
        super().__init__(n_cols=n_cols, dtype=dtype, total_rows=total_rows, ranks=ranks)
        self.total_rows, self.ranks = total_rows, ranks


class DummyEstimator(_DummyClass, core._CumlEstimator, _DummyParams):
    def __init__(self, **kwargs: Any) -> None:
        super().__init__()
        self._set_params(**kwargs)

    def _require_nccl_ucx(self):
        return (False, False)  # no GPU here: the communicator is covered by the -m gpu tests and bench.py

    def _out_schema(self):
        return "n_cols int, dtype string, total_rows int, ranks array<int>"

    def _get_cuml_fit_func(self, dataset, extra_params=None) -> Callable:
        def _fit(dfs, params):
            from spark_rapids_ml_nai_b200.core import param_alias
            from spark_rapids_ml_nai_b200.runtime import BarrierTaskContext
            from spark_rapids_ml_nai_b200.utils import PartitionDescriptor

            assert set(params) >= {param_alias.cuml_init, param_alias.handle, param_alias.part_sizes, param_alias.num_cols}
            assert params[param_alias.cuml_init] == {"a": 3.0}
            assert params[param_alias.handle] is None  # comms disabled for the dummy
            assert sum(params[param_alias.part_sizes]) == sum(x.shape[0] for (x, _, _) in dfs)
            desc = PartitionDescriptor.build(params[param_alias.part_sizes], params[param_alias.num_cols])
            assert desc.rank == BarrierTaskContext.get().partitionId()
            for (x, label, row_number) in dfs:
                assert x.dtype == np.float32 and x.flags["F_CONTIGUOUS"] and label is None and row_number is None
            return {"n_cols": [params[param_alias.num_cols]], "dtype": ["float32"], "total_rows": [desc.m],
                    "ranks": [sorted({r for r, _ in desc.parts_rank_size})]}

        return _fit

    def _create_pyspark_model(self, result):
        return DummyModel._from_row(result)

    def _set_gpu(self):  # pragma: no cover
        pass


@pytest.fixture(autouse=True)
def _no_gpu_selection(monkeypatch):
    monkeypatch.setattr(core._CumlCommon, "_set_gpu_device", staticmethod(lambda context, is_transform=False: 0))


def test_fit_driver_contract_single_worker():
    X = np.random.default_rng(0).standard_normal((40, 3))
    est = DummyEstimator(alpha=3.0, num_workers=1)
    est._infer_num_workers = lambda: 4
    model = est.fit(pd.DataFrame({"features": list(X)}))
    assert (model.n_cols, model.dtype, model.total_rows, model.ranks) == (3, "float32", 40, [0])
    assert model.getOrDefault("alpha") == 3.0 and model.cuml_params["a"] == 3.0
    # per-fit overrides leave the estimator untouched
    est2 = DummyEstimator(alpha=3.0, num_workers=1)
    est2._infer_num_workers = lambda: 4
    assert est2.getOrDefault("alpha") == 3.0


def test_num_workers_above_gpu_count_raises():
    # python/tests/test_common_estimator.py:528-558
    est = DummyEstimator(num_workers=3)
    est._infer_num_workers = lambda: 2
    with pytest.raises(ValueError, match="should be less than or equal to total GPUs"):
        est.num_workers


def test_empty_partition_raises():
    # reference core.py:750-753
    est = DummyEstimator(alpha=3.0, num_workers=1)
    est._infer_num_workers = lambda: 1
    ds = LocalDataFrame(pd.DataFrame({"features": [[1.0, 2.0]]})).withPartitions([pd.DataFrame({"features": []})])
    ds.first = lambda: pd.Series({"features": [1.0, 2.0]})
    with pytest.raises(RuntimeError, match="A python worker received no data"):
        est._fit(ds)
