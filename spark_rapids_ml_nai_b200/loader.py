"""Partition -> device loader.

Replaces the reference's host-side `np.array(list(...))` batches + `_concat_and_free` + implicit
H2D inside cuML (reference python/src/spark_rapids_ml/core.py:724-748, utils.py:213-252,
clustering.py:360-370): the per-batch host arrays are copied straight into one preallocated
row-major device tensor (no host-side concatenation), through a ring of pinned staging buffers when
the source memory is pageable, and the input list is emptied as the reference's helper does.
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import numpy as np

_STAGE_BYTES = 128 << 20
_staging = {}


def _staging_ring(device_index: int, nbuf: int = 2):
    import torch

    key = device_index
    if key not in _staging:
        bufs = [torch.empty(_STAGE_BYTES, dtype=torch.uint8, pin_memory=True) for _ in range(nbuf)]
        evs = [torch.cuda.Event() for _ in range(nbuf)]
        _staging[key] = (bufs, evs)
    return _staging[key]


def as_float32_matrix(a) -> np.ndarray:
    """C-ordered float32 [rows, d] view/copy of one host batch (ndarray or pandas DataFrame)."""
    if hasattr(a, "to_numpy"):
        a = a.to_numpy()
    a = np.asarray(a)
    if a.ndim == 1:
        a = a.reshape(-1, 1)
    if a.dtype != np.float32 or not a.flags["C_CONTIGUOUS"]:
        a = np.ascontiguousarray(a, dtype=np.float32)
    return a


def partitions_to_device(parts: List, device, free_inputs: bool = True, out=None):
    """Copies the host batches into one [sum(rows), d] float32 CUDA tensor on `device`.

    Returns the device tensor.  `parts` is emptied when `free_inputs` (the fit function may free its
    inputs: core.py:766-768).  Work is enqueued on the current stream; the result is safe to use on
    that stream without further synchronisation.
    """
    import torch

    device = torch.device(device)
    mats = [as_float32_matrix(p) for p in parts]
    if free_inputs:
        del parts[:]
    if not mats:
        raise RuntimeError("A python worker received no data.  Please increase amount of data or use fewer workers.")
    d = mats[0].shape[1]
    rows = sum(m.shape[0] for m in mats)
    for m in mats:
        if m.shape[1] != d:
            raise ValueError(f"feature dimension mismatch between batches: {m.shape[1]} vs {d}")
    if rows == 0:
        raise RuntimeError("A python worker received no data.  Please increase amount of data or use fewer workers.")
    X = out if out is not None else torch.empty((rows, d), dtype=torch.float32, device=device)
    flat = X.view(-1)
    off = 0
    with torch.cuda.device(device):
        stream = torch.cuda.current_stream()
        ring = None
        slot = 0
        for i, m in enumerate(mats):
            src = torch.from_numpy(m).view(-1)
            n_el = src.numel()
            if n_el == 0:
                continue
            if src.is_pinned():
                flat[off:off + n_el].copy_(src, non_blocking=True)
            else:
                if ring is None:
                    ring = _staging_ring(device.index or 0)
                bufs, evs = ring
                step = _STAGE_BYTES // 4
                for s in range(0, n_el, step):
                    e = min(n_el, s + step)
                    evs[slot].synchronize()  # the copy that last used this buffer has drained
                    stage = bufs[slot].view(torch.float32)[: e - s]
                    stage.copy_(src[s:e])
                    flat[off + s:off + e].copy_(stage, non_blocking=True)
                    evs[slot].record(stream)
                    slot = (slot + 1) % len(bufs)
            off += n_el
            mats[i] = None
    return X
