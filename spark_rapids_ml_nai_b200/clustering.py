"""KMeans fit / predict host layer (first slice: the executor-side functions the plugin slots call).

`kmeans_fit_partitions` is the body of the fit-function slot (reference
python/src/spark_rapids_ml/clustering.py:348-384): host batches -> device shard -> initial centroids ->
device-resident Lloyd loop through the C ABI -> result dict in the reference's `_out_schema` shape.
"""
from __future__ import annotations

import logging
from typing import Any, Dict, List

import numpy as np

from . import _native
from .loader import partitions_to_device

logger = logging.getLogger("spark_rapids_ml_nai_b200.clustering")


def _initial_centroids(handle: "_native.Handle", X, cuml_params: Dict[str, Any]):
    """`init` may be an explicit [k,d] array (identical on all ranks), "random", or
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
        from .init import kmeans_parallel_init

        return kmeans_parallel_init(handle, X, k, seed=seed,
                                    oversampling=float(cuml_params.get("oversampling_factor", 2.0)))
    raise ValueError(f"Value '{init}' for 'init' param is unsupported")


def kmeans_fit_partitions(parts: List, handle: "_native.Handle", cuml_params: Dict[str, Any], num_cols: int,
                          want_labels: bool = False) -> Dict[str, Any]:
    """Fits this rank's partitions.  Returns the reference's result dict plus `n_iter` / `inertia`."""
    import torch

    device = torch.device("cuda", handle.device)
    X = partitions_to_device(parts, device, free_inputs=True)
    if X.shape[1] != num_cols:
        raise ValueError(f"partitions have {X.shape[1]} columns, expected {num_cols}")
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
