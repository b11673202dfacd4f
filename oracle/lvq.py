"""CPU restatement of the reference's LVQQuantizer. TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Follows src/haag_vq/methods/lvq_quantization.py:44-141 line by line (fit :44-46, compress :48-108,
decompress :110-141), in plain numpy with explicit float32 intermediates. PINNED:
oracle/make_golden_lvq.py runs the reference class itself (importable here: numpy only) on seeded
inputs, asserts that this restatement reproduces its bytes, and commits tests/golden/lvq_D*.npz.
"""
from __future__ import annotations

import math

import numpy as np


def lvq_fit(X: np.ndarray) -> np.ndarray:
    # lvq_quantization.py:46 — numpy adds the rows one after another in float32, then divides by N
    return np.asarray(X, dtype=np.float32).mean(axis=0).astype(np.float32)


def lvq_code_size(D: int, B: int) -> int:
    return math.ceil(D * B / 8) + 8


def lvq_compress(X: np.ndarray, mu: np.ndarray, B: int) -> np.ndarray:
    # lvq_quantization.py:58-108
    X = np.asarray(X, dtype=np.float32)
    N, D = X.shape
    levels = (1 << B) - 1
    index_bytes = math.ceil(D * B / 8)
    codes = np.empty((N, index_bytes + 8), dtype=np.uint8)
    R = X - mu.astype(np.float32)
    lo = R.min(axis=1) if D else np.zeros(N, np.float32)
    hi = R.max(axis=1) if D else np.zeros(N, np.float32)
    span = hi - lo
    eps = np.finfo(np.float32).tiny
    with np.errstate(divide="ignore", invalid="ignore"):
        delta = np.where(span == 0.0, eps, span / np.float32(levels)).astype(np.float32)
    idx = np.round((R - lo[:, None]) / delta[:, None]).astype(np.int32)
    idx = np.clip(idx, 0, levels)
    shifts = np.arange(B - 1, -1, -1, dtype=np.int32)
    bits = ((idx[:, :, None] >> shifts[None, None, :]) & 1).reshape(N, D * B).astype(np.uint8)
    codes[:, :index_bytes] = np.packbits(bits, axis=1, bitorder="big")
    codes[:, index_bytes:index_bytes + 4] = lo.astype(np.float32).view(np.uint8).reshape(N, 4)
    codes[:, index_bytes + 4:index_bytes + 8] = delta.view(np.uint8).reshape(N, 4)
    return codes


def lvq_decompress(codes: np.ndarray, mu: np.ndarray, B: int) -> np.ndarray:
    # lvq_quantization.py:117-141
    codes = np.asarray(codes, dtype=np.uint8)
    N, D = codes.shape[0], mu.shape[0]
    index_bytes = math.ceil(D * B / 8)
    lo = codes[:, index_bytes:index_bytes + 4].copy().view(np.float32).reshape(N)
    delta = codes[:, index_bytes + 4:index_bytes + 8].copy().view(np.float32).reshape(N)
    bits = np.unpackbits(codes[:, :index_bytes], axis=1, bitorder="big")[:, :D * B]
    shifts = np.arange(B - 1, -1, -1, dtype=np.int32)
    idx = (bits.reshape(N, D, B).astype(np.int32) * (1 << shifts)[None, None, :]).sum(axis=2)
    r_hat = lo[:, None] + idx.astype(np.float32) * delta[:, None]
    return r_hat + mu[None, :].astype(np.float32)
