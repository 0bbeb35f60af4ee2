"""ctypes binding of libb200km.so (the C ABI declared in include/b200km.h).

There is no CPU fallback: if the library is missing or a call fails this module raises.
The tensors passed in are torch CUDA tensors; only their ``data_ptr()`` crosses the boundary.
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional, Tuple

_HERE = os.path.dirname(os.path.abspath(__file__))
# B200KM_LIB: development override (same-box A/B timing of two builds, tools/ab_build.sh)
LIB_PATH = os.environ.get("B200KM_LIB") or os.path.join(_HERE, "lib", "libb200km.so")

PATH_AUTO, PATH_FFMA, PATH_TC, PATH_TC_SPLIT, PATH_GEMM, PATH_SMALL = 0, 1, 2, 3, 4, 5
INIT_RANDOM, INIT_KMEANS_PARALLEL = 0, 1
UNIQUE_ID_BYTES = 128

# every symbol include/b200km.h declares: (name, restype, argtypes)
_c = ctypes
_SIGNATURES = {
    "b200km_version": (_c.c_char_p, []),
    "b200km_last_error": (_c.c_char_p, []),
    "b200km_comm_get_unique_id": (_c.c_int, [_c.c_void_p]),
    "b200km_create": (_c.c_int, [_c.c_int, _c.c_int, _c.c_int, _c.c_void_p, _c.POINTER(_c.c_void_p)]),
    "b200km_destroy": (_c.c_int, [_c.c_void_p]),
    "b200km_abort": (_c.c_int, [_c.c_void_p]),
    "b200km_set_path": (_c.c_int, [_c.c_void_p, _c.c_int]),
    "b200km_last_path": (_c.c_char_p, [_c.c_void_p]),
    "b200km_last_reduce": (_c.c_char_p, [_c.c_void_p]),
    "b200km_tc_supported": (_c.c_int, [_c.c_int, _c.c_int, _c.c_int]),
    "b200km_profile_enable": (_c.c_int, [_c.c_void_p, _c.c_int]),
    "b200km_profile_read": (_c.c_int, [_c.c_void_p, _c.POINTER(_c.c_double), _c.POINTER(_c.c_int),
                                       _c.POINTER(_c.c_longlong)]),
    "b200km_screen_stats": (_c.c_int, [_c.c_void_p, _c.POINTER(_c.c_longlong), _c.POINTER(_c.c_int)]),
    "b200km_query_workspace": (_c.c_int, [_c.c_int64, _c.c_int, _c.c_int, _c.POINTER(_c.c_size_t)]),
    "b200km_fit": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_int64, _c.c_int, _c.c_int, _c.c_void_p, _c.c_int,
                              _c.c_double, _c.c_void_p, _c.POINTER(_c.c_double), _c.POINTER(_c.c_int),
                              _c.c_void_p, _c.c_void_p]),
    "b200km_lloyd_step": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_int64, _c.c_int, _c.c_int, _c.c_void_p,
                                     _c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_void_p]),
    "b200km_predict": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_int64, _c.c_int, _c.c_void_p, _c.c_int,
                                  _c.c_void_p, _c.c_void_p, _c.c_void_p]),
    "b200km_predict_static": (_c.c_int, [_c.c_void_p, _c.c_int64, _c.c_int, _c.c_void_p, _c.c_int, _c.c_void_p,
                                         _c.c_void_p]),
    "b200km_static_last_error": (_c.c_char_p, []),
    "b200km_init_centroids": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_int64, _c.c_int, _c.c_int, _c.c_int,
                                         _c.c_uint64, _c.c_double, _c.c_void_p, _c.c_void_p]),
}

_lib = None


def exported_symbols():
    return sorted(_SIGNATURES)


def load() -> ctypes.CDLL:
    """Loads the shared library and binds every declared entry point (raises if anything is missing)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -m spark_rapids_ml_nai_b200.build` "
            "(there is no CPU fallback for the KMeans path)"
        )
    lib = ctypes.CDLL(LIB_PATH)
    for name, (restype, argtypes) in _SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = restype
        fn.argtypes = argtypes
    _lib = lib
    return lib


def _check(rc: int) -> None:
    if rc != 0:
        raise RuntimeError("b200km: " + load().b200km_last_error().decode("utf-8", "replace"))


def version() -> str:
    return load().b200km_version().decode()


def get_unique_id() -> bytes:
    buf = ctypes.create_string_buffer(UNIQUE_ID_BYTES)
    _check(load().b200km_comm_get_unique_id(buf))
    return buf.raw


def _ptr(t) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _stream_ptr(stream=None) -> int:
    import torch

    s = stream if stream is not None else torch.cuda.current_stream()
    return s.cuda_stream


def _check_tensor(t, name, dtype, device_index):
    import torch

    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise TypeError(f"{name} must be a CUDA torch tensor")
    if t.dtype != dtype:
        raise TypeError(f"{name} must be {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise ValueError(f"{name} must be contiguous (row-major)")
    if t.device.index != device_index:
        raise ValueError(f"{name} lives on cuda:{t.device.index}, handle is on cuda:{device_index}")


class Handle:
    """One per (process, GPU).  ``nranks > 1`` joins the NCCL communicator named by ``uid``."""

    def __init__(self, device: int = 0, rank: int = 0, nranks: int = 1, uid: Optional[bytes] = None):
        lib = load()
        self._h = ctypes.c_void_p()
        self.device, self.rank, self.nranks = device, rank, nranks
        uid_buf = None
        if nranks > 1:
            if uid is None or len(uid) != UNIQUE_ID_BYTES:
                raise ValueError("uid must be the 128-byte id from get_unique_id()")
            uid_buf = ctypes.create_string_buffer(uid, UNIQUE_ID_BYTES)
        _check(lib.b200km_create(device, rank, nranks, uid_buf, ctypes.byref(self._h)))

    # -- lifetime ------------------------------------------------------------------------------
    def destroy(self) -> None:
        if self._h:
            h, self._h = self._h, ctypes.c_void_p()
            _check(load().b200km_destroy(h))

    def abort(self) -> None:
        if self._h:
            h, self._h = self._h, ctypes.c_void_p()
            _check(load().b200km_abort(h))

    def __enter__(self):
        return self

    def __exit__(self, exc_type, exc, tb):
        # same rule as the reference communicator: orderly destroy, abort when an exception is in flight
        if exc_type is None:
            self.destroy()
        else:
            self.abort()

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass

    # -- configuration ---------------------------------------------------------------------------
    def set_path(self, path: int) -> None:
        _check(load().b200km_set_path(self._h, path))

    def last_path(self) -> str:
        return load().b200km_last_path(self._h).decode()

    def last_reduce(self) -> str:
        """"none" / "nvlink-oneshot" / "nccl": how the last fit's sums crossed the GPUs."""
        return load().b200km_last_reduce(self._h).decode()

    def profile_enable(self, on: bool = True) -> None:
        _check(load().b200km_profile_enable(self._h, 1 if on else 0))

    def profile_read(self):
        """(summed dominant-kernel ms, bracketed launches, all kernel launches) since the last read."""
        ms, n, total = ctypes.c_double(0.0), ctypes.c_int(0), ctypes.c_longlong(0)
        _check(load().b200km_profile_read(self._h, ctypes.byref(ms), ctypes.byref(n), ctypes.byref(total)))
        return ms.value, n.value, total.value

    # -- compute ---------------------------------------------------------------------------------
    def fit(self, X, init_centroids, max_iter: int, tol: float, want_labels: bool = False,
            want_inertia: bool = True, stream=None):
        """Returns (centroids [k,d] float32 cuda, inertia or None, n_iter, labels int32 cuda or None)."""
        import torch

        _check_tensor(X, "X", torch.float32, self.device)
        _check_tensor(init_centroids, "init_centroids", torch.float32, self.device)
        n, d = X.shape
        k, d2 = init_centroids.shape
        if d2 != d:
            raise ValueError(f"init_centroids has {d2} columns, X has {d}")
        out = torch.empty_like(init_centroids)
        labels = torch.empty(n, dtype=torch.int32, device=X.device) if want_labels else None
        inertia = ctypes.c_double(0.0)
        n_iter = ctypes.c_int(0)
        _check(load().b200km_fit(self._h, X.data_ptr(), n, d, k, init_centroids.data_ptr(), int(max_iter),
                                 float(tol), out.data_ptr(), ctypes.byref(inertia) if want_inertia else None,
                                 ctypes.byref(n_iter), _ptr(labels), _stream_ptr(stream)))
        return out, (inertia.value if want_inertia else None), n_iter.value, labels

    def screen_stats(self) -> Tuple[int, int]:
        """(rows re-decided in float64 by the screen kernel since the last centroid preparation, kernel of the
        last pass: 0 FFMA, 1 split-TF32, 2 screen)."""
        r, kk = ctypes.c_longlong(0), ctypes.c_int(0)
        _check(load().b200km_screen_stats(self._h, ctypes.byref(r), ctypes.byref(kk)))
        return int(r.value), int(kk.value)

    def lloyd_step(self, X, centroids, want_labels: bool = False, stream=None):
        """One assign+accumulate pass: (sums f64 [k,d], counts f64 [k], inertia f64 [1], labels or None)."""
        import torch

        _check_tensor(X, "X", torch.float32, self.device)
        _check_tensor(centroids, "centroids", torch.float32, self.device)
        n, d = X.shape
        k = centroids.shape[0]
        sums = torch.empty((k, d), dtype=torch.float64, device=X.device)
        counts = torch.empty(k, dtype=torch.float64, device=X.device)
        inertia = torch.empty(1, dtype=torch.float64, device=X.device)
        labels = torch.empty(n, dtype=torch.int32, device=X.device) if want_labels else None
        _check(load().b200km_lloyd_step(self._h, X.data_ptr(), n, d, k, centroids.data_ptr(), sums.data_ptr(),
                                        counts.data_ptr(), inertia.data_ptr(), _ptr(labels), _stream_ptr(stream)))
        return sums, counts, inertia, labels

    def predict(self, X, centroids, want_dist: bool = False, out=None, stream=None) -> Tuple:
        import torch

        _check_tensor(X, "X", torch.float32, self.device)
        _check_tensor(centroids, "centroids", torch.float32, self.device)
        n, d = X.shape
        k, d2 = centroids.shape
        if d2 != d:
            raise ValueError(f"centroids have {d2} columns, X has {d}")
        labels = out if out is not None else torch.empty(n, dtype=torch.int32, device=X.device)
        _check_tensor(labels, "labels", torch.int32, self.device)
        dist = torch.empty(n, dtype=torch.float32, device=X.device) if want_dist else None
        _check(load().b200km_predict(self._h, X.data_ptr(), n, d, centroids.data_ptr(), k, labels.data_ptr(),
                                     _ptr(dist), _stream_ptr(stream)))
        return labels, dist

    def init_centroids(self, X, k: int, mode: int = INIT_RANDOM, seed: int = 1, oversampling: float = 2.0,
                       stream=None):
        import torch

        _check_tensor(X, "X", torch.float32, self.device)
        n, d = X.shape
        out = torch.empty((k, d), dtype=torch.float32, device=X.device)
        _check(load().b200km_init_centroids(self._h, X.data_ptr(), n, d, k, mode, seed, float(oversampling),
                                            out.data_ptr(), _stream_ptr(stream)))
        return out


STATIC_MAX_ROWS, STATIC_MAX_K = 65536, 1024


def predict_static(X, centroids, out=None, stream=None):
    """Labels through ``b200km_predict_static``: no handle, no allocation inside the library (the output
    tensor is the caller's).  For callers that cannot allocate device memory; at most STATIC_MAX_ROWS rows
    and STATIC_MAX_K centroids per call, calls must not overlap on the GPU.  The tensors must live on the
    current CUDA device."""
    import torch

    dev = torch.cuda.current_device()
    _check_tensor(X, "X", torch.float32, dev)
    _check_tensor(centroids, "centroids", torch.float32, dev)
    n, d = X.shape
    k, d2 = centroids.shape
    if d2 != d:
        raise ValueError(f"centroids have {d2} columns, X has {d}")
    labels = out if out is not None else torch.empty(n, dtype=torch.int32, device=X.device)
    _check_tensor(labels, "labels", torch.int32, dev)
    lib = load()
    rc = lib.b200km_predict_static(X.data_ptr(), n, d, centroids.data_ptr(), k, labels.data_ptr(), _stream_ptr(stream))
    if rc != 0:
        raise RuntimeError("b200km: " + lib.b200km_static_last_error().decode("utf-8", "replace"))
    return labels


def tc_supported(d: int, k: int, with_update: bool = True) -> bool:
    return bool(load().b200km_tc_supported(d, k, 1 if with_update else 0))


def query_workspace(n: int, d: int, k: int) -> int:
    out = ctypes.c_size_t(0)
    _check(load().b200km_query_workspace(n, d, k, ctypes.byref(out)))
    return out.value
