"""Spark-free barrier runtime: the part of Spark the KMeans path needs, without a JVM.

The reference runs its fit as one barrier stage - ``mapInPandas(_train_udf).rdd.barrier().mapPartitions``
(python/src/spark_rapids_ml/core.py:781-785) - with one Python worker process per GPU whose
``BarrierTaskContext`` offers ``partitionId()``, ``allGather(str)`` and ``barrier()``
(core.py:699-700,772-774; common/cuml_context.py:75-81).  pyspark and a JVM are absent from this image,
so the same three calls are provided here over ``torch.distributed`` (gloo for the control plane; the
data plane is NCCL inside libb200km) or, for a single worker, in-process.  The executor-side functions
(`core.train_partition`, the fit / transform closures) only see this interface, so they run unchanged
under real Spark barrier tasks.
"""
from __future__ import annotations

import os
import pickle
import socket
import tempfile
from typing import Any, Callable, Iterator, List, Optional, Sequence

import numpy as np
import pandas as pd


class BarrierTaskContext:
    """partitionId / allGather / barrier, as pyspark.BarrierTaskContext offers them."""

    _current: Optional["BarrierTaskContext"] = None

    def __init__(self, rank: int, nranks: int, use_dist: bool):
        self._rank, self._nranks, self._use_dist = rank, nranks, use_dist

    @classmethod
    def get(cls) -> "BarrierTaskContext":
        if cls._current is None:
            raise RuntimeError("BarrierTaskContext.get() called outside a barrier task")
        return cls._current

    def partitionId(self) -> int:
        return self._rank

    def numPartitions(self) -> int:
        return self._nranks

    def allGather(self, message: str = "") -> List[str]:
        if not isinstance(message, str):
            raise TypeError("Argument `message` must be of type `str`")
        if not self._use_dist:
            return [message]
        import torch.distributed as dist

        out: List[Any] = [None] * self._nranks
        dist.all_gather_object(out, message)
        return [str(m) for m in out]

    def barrier(self) -> None:
        if self._use_dist:
            import torch.distributed as dist

            dist.barrier()


class LocalDataFrame:
    """A partitioned pandas DataFrame: what ``dataset.repartition(n)`` + Arrow batching hand to a task.

    Accepts a pandas DataFrame (feature column of per-row arrays, or numeric feature columns) or a 2-D
    array (becomes the column ``features``).  ``max_records_per_batch`` plays the role of
    ``spark.sql.execution.arrow.maxRecordsPerBatch`` (default 10000, reference core.py:1547-1557)."""

    def __init__(self, data: Any, num_partitions: int = 1, max_records_per_batch: int = 10000,
                 features_col: str = "features"):
        if isinstance(data, LocalDataFrame):
            parts = data.partitions
        elif isinstance(data, pd.DataFrame):
            parts = self._split(data, num_partitions)
        else:
            arr = np.asarray(data)
            if arr.ndim != 2:
                raise ValueError("expected a 2-D array or a pandas DataFrame")
            parts = self._split(pd.DataFrame({features_col: list(arr)}), num_partitions)
        self.partitions: List[pd.DataFrame] = parts
        self.max_records_per_batch = int(max_records_per_batch)

    @staticmethod
    def _split(pdf: pd.DataFrame, n: int) -> List[pd.DataFrame]:
        bounds = np.linspace(0, len(pdf), n + 1).astype(np.int64)
        return [pdf.iloc[bounds[i]:bounds[i + 1]].reset_index(drop=True) for i in range(n)]

    @property
    def columns(self) -> List[str]:
        return list(self.partitions[0].columns)

    def getNumPartitions(self) -> int:
        return len(self.partitions)

    def count(self) -> int:
        return sum(len(p) for p in self.partitions)

    def first(self) -> pd.Series:
        for p in self.partitions:
            if len(p):
                return p.iloc[0]
        raise ValueError("empty dataset")

    def repartition(self, n: int) -> "LocalDataFrame":
        out = LocalDataFrame.__new__(LocalDataFrame)
        out.partitions = self._split(self.toPandas(), n)
        out.max_records_per_batch = self.max_records_per_batch
        return out

    def toPandas(self) -> pd.DataFrame:
        return pd.concat(self.partitions, ignore_index=True)

    def batches(self, partition: int) -> Iterator[pd.DataFrame]:
        pdf = self.partitions[partition]
        step = self.max_records_per_batch
        if len(pdf) == 0:
            return
        for s in range(0, len(pdf), step):
            yield pdf.iloc[s:s + step]

    def withPartitions(self, parts: List[pd.DataFrame]) -> "LocalDataFrame":
        out = LocalDataFrame.__new__(LocalDataFrame)
        out.partitions = parts
        out.max_records_per_batch = self.max_records_per_batch
        return out


def _free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, nranks: int, port: int, payload_path: str, result_path: str) -> None:
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=nranks)
    try:
        with open(payload_path, "rb") as f:
            import cloudpickle  # the closures are shipped the way Spark ships them

            udf, dataset = cloudpickle.load(f)
        BarrierTaskContext._current = BarrierTaskContext(rank, nranks, use_dist=True)
        out = list(udf(dataset.batches(rank)))
        with open(f"{result_path}.{rank}", "wb") as f:
            pickle.dump(out, f)
    finally:
        BarrierTaskContext._current = None
        dist.destroy_process_group()


def run_barrier_stage(udf: Callable[[Iterator[pd.DataFrame]], Iterator[pd.DataFrame]], dataset: LocalDataFrame,
                      num_workers: int) -> List[pd.DataFrame]:
    """Runs ``udf`` once per partition as one all-or-nothing stage; returns the yielded frames in rank
    order.  One worker runs in-process; several are spawned as separate processes (one per GPU), the
    closure and the partitions shipped by cloudpickle, like Spark's local[N] barrier tasks."""
    if dataset.getNumPartitions() != num_workers:
        dataset = dataset.repartition(num_workers)
    if num_workers == 1:
        BarrierTaskContext._current = BarrierTaskContext(0, 1, use_dist=False)
        try:
            return list(udf(dataset.batches(0)))
        finally:
            BarrierTaskContext._current = None
    import cloudpickle
    import torch.multiprocessing as mp

    with tempfile.TemporaryDirectory() as tmp:
        payload = os.path.join(tmp, "payload.pkl")
        result = os.path.join(tmp, "result.pkl")
        with open(payload, "wb") as f:
            cloudpickle.dump((udf, dataset), f)
        mp.spawn(_worker, args=(num_workers, _free_port(), payload, result), nprocs=num_workers, join=True)
        frames: List[pd.DataFrame] = []
        for r in range(num_workers):
            with open(f"{result}.{r}", "rb") as f:
                frames.extend(pickle.load(f))
        return frames
