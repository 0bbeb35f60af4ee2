"""Shared parity checks: the CUDA path (through the C ABI) against the float64 CPU oracle.

Tolerances are BASELINE.json's: labels bit-exact except rows whose best and second-best squared
distances differ by < 1e-5 relative; centroids and inertia within 1e-4 relative."""
import numpy as np

from oracle import kmeans_oracle as ko

REL_AMBIG = 1e-5
REL_TOL = 1e-4


def assert_labels_match(labels_gpu, res_best, res_second, labels_ref):
    labels_gpu = np.asarray(labels_gpu)
    assert labels_gpu.dtype == np.int32
    amb = ko.ambiguous_rows(res_best, res_second, REL_AMBIG)
    bad = np.nonzero((labels_gpu != labels_ref) & ~amb)[0]
    assert bad.size == 0, (
        f"{bad.size} rows differ outside the 1e-5 ambiguity band, first rows {bad[:5]}: "
        f"gpu {labels_gpu[bad[:5]]} ref {labels_ref[bad[:5]]} "
        f"best {res_best[bad[:5]]} second {res_second[bad[:5]]}"
    )


def assert_centers_close(c_gpu, c_ref):
    c_gpu = np.asarray(c_gpu, dtype=np.float64)
    c_ref = np.asarray(c_ref, dtype=np.float64)
    assert c_gpu.shape == c_ref.shape
    scale = np.maximum(np.linalg.norm(c_ref, axis=1, keepdims=True) / np.sqrt(c_ref.shape[1]), 1e-30)
    err = np.abs(c_gpu - c_ref) / np.maximum(np.abs(c_ref), scale)
    assert err.max() <= REL_TOL, f"centroid rel err {err.max():.3e} > {REL_TOL}"


def assert_inertia_close(i_gpu, i_ref):
    assert abs(i_gpu - i_ref) <= REL_TOL * max(abs(i_ref), 1e-30), (i_gpu, i_ref)
