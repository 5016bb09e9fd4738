"""Shared parity metrics of the GPU tests."""
import numpy as np


def rel(a, b):
    """max |a - b| / max |b| over the whole array (secondary check: insensitive to single small entries)."""
    a = np.asarray(a, dtype=float); b = np.asarray(b, dtype=float)
    assert a.shape == b.shape, (a.shape, b.shape)
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))


def rel_elem(a, b, floor_frac=1e-3):
    """Worst ELEMENT-WISE relative error.  Every entry is compared with its own reference value; entries whose
    reference is smaller than floor_frac x the median magnitude of the field (sums that cancel to ~0) are compared
    with that floor instead, because their rounding error scales with the summed terms, not with the result.
    Vector fields (N,3) are compared per particle against the particle's vector norm."""
    a = np.asarray(a, dtype=float); b = np.asarray(b, dtype=float)
    assert a.shape == b.shape, (a.shape, b.shape)
    if b.ndim > 1:
        mag = np.sqrt(np.sum(b * b, axis=-1)); err = np.sqrt(np.sum((a - b) ** 2, axis=-1))
    else:
        mag = np.abs(b); err = np.abs(a - b)
    fin = np.isfinite(mag)
    assert np.array_equal(np.isfinite(a).reshape(len(a), -1).all(axis=1), fin) or not fin.all(), "finiteness differs"
    if not fin.any():
        return 0.0
    nz = mag[fin & (mag > 0)]
    floor = floor_frac * (np.median(nz) if nz.size else 1.0)
    return float(np.max(err[fin] / np.maximum(mag[fin], max(floor, 1e-300))))
