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

import os

_STAGE_BYTES = int(os.environ.get("B200KM_STAGE_MB", "128")) << 20  # one pinned staging slot
# worker threads of the pageable -> pinned copies: the staging copy is what a fit from host batches waits for
# (B200 box, 16 cores, 20 GB: 4 threads 1.04 s, 8 threads 0.86 s, 16 threads 0.77 s)
# The cores are shared by the ranks of a node (torchrun sets LOCAL_WORLD_SIZE).
_COPY_THREADS = int(os.environ.get("B200KM_COPY_THREADS", "0")) or max(
    4, min(16, (os.cpu_count() or 8) // max(1, int(os.environ.get("LOCAL_WORLD_SIZE", "1")))))
_staging = {}
_pool = None


def _copy_pool():
    """Worker threads for the pageable -> pinned copies (NumPy releases the GIL inside large copies, so the
    threads run in parallel: one thread moves ~5-8 GB/s, the host -> device link takes ~55 GB/s)."""
    global _pool
    if _pool is None:
        from concurrent.futures import ThreadPoolExecutor

        _pool = ThreadPoolExecutor(max_workers=_COPY_THREADS, thread_name_prefix="b200km-stage")
    return _pool


def _parallel_copy(dst: np.ndarray, src: np.ndarray) -> None:
    """dst[:] = src for flat arrays, split across the copy threads when it is worth it."""
    n = src.shape[0]
    if n < (4 << 20):
        np.copyto(dst, src)
        return
    pool = _copy_pool()
    step = (n + _COPY_THREADS - 1) // _COPY_THREADS
    futs = [pool.submit(np.copyto, dst[s:s + step], src[s:s + step]) for s in range(0, n, step)]
    for f in futs:
        f.result()


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
    with torch.cuda.device(device):
        stream = torch.cuda.current_stream()
        ring = None
        slot = 0
        off = 0          # elements of X already scheduled
        fill = 0         # elements assigned to the current slot, not yet sent
        cap = _STAGE_BYTES // 4
        pending = []     # copy tasks of the current slot: (dst view, src view)

        def flush():
            nonlocal fill, off, slot, pending
            if fill == 0:
                return
            bufs, evs = ring
            # the slot's batches are copied by the worker threads side by side (NumPy releases the GIL inside a
            # copy): Arrow-sized batches are a few MB each, far too small to split individually
            if len(pending) == 1:
                _parallel_copy(*pending[0])
            else:
                pool = _copy_pool()
                for f in [pool.submit(np.copyto, dst, src) for dst, src in pending]:
                    f.result()
            pending = []
            flat[off:off + fill].copy_(bufs[slot].view(torch.float32)[:fill], non_blocking=True)
            evs[slot].record(stream)
            off += fill
            fill = 0
            slot = (slot + 1) % len(bufs)
            evs[slot].synchronize()  # the copy that last used the next buffer has drained

        for i, m in enumerate(mats):
            n_el = m.size
            if n_el == 0:
                continue
            # only writable arrays can be views of pinned torch memory (pandas hands out read-only views)
            src_t = torch.from_numpy(m).view(-1) if m.flags.writeable else None
            if src_t is not None and src_t.is_pinned():
                if ring is not None:
                    flush()
                flat[off:off + n_el].copy_(src_t, non_blocking=True)
                off += n_el
                mats[i] = None
                continue
            # pageable memory: small (Arrow-sized) batches are packed into a 128 MB pinned slot and leave as one
            # host -> device copy; while that copy is in flight the next slot is filled
            if ring is None:
                ring = _staging_ring(device.index or 0)
                ring[1][slot].synchronize()
            src = m.reshape(-1)
            done = 0
            while done < n_el:
                take = min(n_el - done, cap - fill)
                stage = ring[0][slot].numpy().view(np.float32)
                pending.append((stage[fill:fill + take], src[done:done + take]))
                fill += take
                done += take
                if fill == cap:
                    flush()
            mats[i] = None
        if ring is not None:
            flush()
    return X
