"""Helpers shared by the estimator code (counterparts of reference python/src/spark_rapids_ml/utils.py:74-308)."""
from __future__ import annotations

import logging
import os
import sys
from typing import Any, Callable, List, Tuple, Union

import numpy as np


def get_logger(cls_or_callable: Union[type, Callable, str], level: str = "INFO") -> logging.Logger:
    """Named stderr logger, configured once (reference utils.py:281-302)."""
    name = cls_or_callable if isinstance(cls_or_callable, str) else f"{cls_or_callable.__module__}.{cls_or_callable.__name__}"
    logger = logging.getLogger(name)
    logger.setLevel(level)
    if not logger.handlers and not logging.getLogger().handlers:
        handler = logging.StreamHandler(sys.stderr)
        handler.setFormatter(logging.Formatter("%(asctime)s - %(name)s - %(levelname)s - %(message)s"))
        logger.addHandler(handler)
    return logger


def _get_gpu_id(partition_id: int, is_transform: bool = False) -> int:
    """Local-mode device choice of the reference (core.py:343-388, utils.py:112-144): with
    CUDA_VISIBLE_DEVICES set the task uses entry 0; otherwise fit tasks use GPU = partition id and
    transform tasks partition id modulo the GPU count."""
    import torch

    if "CUDA_VISIBLE_DEVICES" in os.environ:
        if not os.environ["CUDA_VISIBLE_DEVICES"]:
            raise RuntimeError(
                "Couldn't get gpu id since CUDA_VISIBLE_DEVICES is set to an empty string.  Please check the GPU resource configuration."
            )
        if len(os.environ["CUDA_VISIBLE_DEVICES"].split(",")) == 1:
            return 0
    n = torch.cuda.device_count()
    if n == 0:
        raise RuntimeError("no CUDA device is visible to this task")
    if is_transform:
        return partition_id % n
    if partition_id >= n:
        raise RuntimeError(f"fit task {partition_id} has no GPU: only {n} visible (one task per GPU)")
    return partition_id


class PartitionDescriptor:
    """Total rows / cols and the (rank, rows) list of every partition (reference utils.py:173-210)."""

    def __init__(self, m: int, n: int, rank: int, parts_rank_size: List[Tuple[int, int]]) -> None:
        self.m, self.n, self.rank, self.parts_rank_size = m, n, rank, parts_rank_size

    @classmethod
    def build(cls, partition_rows: List[int], total_cols: int) -> "PartitionDescriptor":
        import json

        from .runtime import BarrierTaskContext

        context = BarrierTaskContext.get()
        rank = context.partitionId()
        messages = context.allGather(json.dumps([(rank, size) for size in partition_rows]))
        parts_rank_size = [tuple(item) for pair in messages for item in json.loads(pair)]
        return cls(sum(p[1] for p in parts_rank_size), total_cols, rank, parts_rank_size)


def _concat_and_free(array_list: List[np.ndarray], order: str = "F") -> np.ndarray:
    """Host-side concatenation into one preallocated array, emptying the input list
    (reference utils.py:213-252).  The fit path no longer needs it - `loader.partitions_to_device`
    copies the batches straight into the device shard - but the helper stays part of the surface."""
    rows = sum(a.shape[0] for a in array_list)
    shape: Tuple[int, ...] = (rows, array_list[0].shape[1]) if array_list[0].ndim > 1 else (rows,)
    out = np.empty(shape, order=order, dtype=array_list[0].dtype)
    np.concatenate(array_list, out=out)
    del array_list[:]
    return out


def dtype_to_pyspark_type(dtype: Any) -> str:
    dtype = np.dtype(dtype)
    table = {np.dtype(np.float32): "float", np.dtype(np.float64): "double", np.dtype(np.int32): "int",
             np.dtype(np.int16): "short"}
    if dtype not in table:
        raise RuntimeError("Unsupported dtype, found ", dtype)
    return table[dtype]
