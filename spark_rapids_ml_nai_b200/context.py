"""Communicator bootstrap for one barrier stage.

Mirrors the NCCL branch of the reference's ``CumlContext`` (python/src/spark_rapids_ml/common/
cuml_context.py:35-81,109-124,150-163): rank 0 creates the 128-byte NCCL unique id, it travels base64
through the barrier ``allGather`` (the control plane), every rank then joins the communicator - here
through ``b200km_create`` in the C ABI instead of raft_dask - and on exit the communicator is
destroyed, or aborted when an exception is in flight.  The UCX branch (``require_ucx``) belongs to
kNN / DBSCAN and is out of scope; KMeans never asks for it (core.py:555-562).
"""
from __future__ import annotations

import base64
from typing import Any, Optional

from . import _native


class B200Context:
    def __init__(self, rank: int, nranks: int, context: Any, enable: bool, require_ucx: bool = False,
                 device: Optional[int] = None) -> None:
        if require_ucx:
            raise NotImplementedError("UCX endpoints are only needed by kNN / DBSCAN, which this package does not cover")
        self.enable = enable
        self._handle: Optional[_native.Handle] = None
        self._loop = None  # kept for parity with the reference's params dict; always None without UCX
        if not enable:
            return
        self._rank, self._nranks = rank, nranks
        self._device = device if device is not None else 0
        uid = ""
        if nranks > 1 and context.partitionId() == 0:
            uid = base64.b64encode(_native.get_unique_id()).decode("utf-8")
        self._uid: Optional[bytes] = None
        if nranks > 1:
            uids = context.allGather(uid)
            self._uid = base64.b64decode(uids[0])

    @property
    def handle(self) -> Optional["_native.Handle"]:
        return self._handle

    def __enter__(self) -> "B200Context":
        if self.enable:
            # a GPU must be selected before the communicator is created (cuml_context.py:115)
            self._handle = _native.Handle(device=self._device, rank=self._rank, nranks=self._nranks, uid=self._uid)
        return self

    def __exit__(self, exc_type: Any, exc: Any, tb: Any) -> None:
        if not self.enable or self._handle is None:
            return
        h, self._handle = self._handle, None
        if exc_type is None:
            h.destroy()
        else:
            h.abort()
