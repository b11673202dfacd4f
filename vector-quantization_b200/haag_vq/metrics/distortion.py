"""Reconstruction distortion (mirrors src/haag_vq/metrics/distortion.py:4-7).

``compute_distortion`` = mean over rows of the squared L2 error between X and
``model.decompress(codes)``. The subtraction, squaring and sum run on the device
(``b200vq_sq_err_sum``: fp32 differences, fp64 accumulation); the decode is the model's own.
"""
from __future__ import annotations

import numpy as np

from b200vq import host, ops_sq


def sum_squared_error(X_original: np.ndarray, X_reconstructed: np.ndarray) -> float:
    a = host.as_f32_2d(X_original)
    b = host.as_f32_2d(np.asarray(X_reconstructed, dtype=np.float32))
    if a.shape != b.shape:
        raise ValueError(f"shape mismatch {a.shape} vs {b.shape}")
    if a.size == 0:
        return 0.0
    return float(ops_sq.sq_err_sum(host.to_device(a), host.to_device(b)).item())


def compute_distortion(X_original, X_compressed_codes, model) -> float:
    X_original = np.asarray(X_original)
    X_reconstructed = model.decompress(X_compressed_codes)
    n = X_original.shape[0]
    return sum_squared_error(X_original, X_reconstructed) / max(1, n)
