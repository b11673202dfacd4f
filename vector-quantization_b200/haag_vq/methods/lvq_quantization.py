"""LVQQuantizer on the B200 kernels.

Mirrors src/haag_vq/methods/lvq_quantization.py:23-153 (single-level locally-adaptive vector
quantization, SVS style): fit = global mean; every row gets its own lo / delta over the residual,
B-bit indices bit-packed MSB first, and lo, delta appended as fp32, so ``decompress(codes[ids])``
is row-independent. Same constructor, attributes (num_bits, mu), errors and bytes as the
reference; the arithmetic runs in ``b200vq_lvq_encode`` / ``b200vq_lvq_decode`` (csrc/lvq.cu) and
the mean in ``b200vq_col_mean_seq`` (numpy's summation order, so ``mu`` is the reference's mu).
"""
from __future__ import annotations

import math

import numpy as np
import torch

from b200vq import host, ops_sq

from .base_quantizer import BaseQuantizer


class LVQQuantizer(BaseQuantizer):
    def __init__(self, num_bits: int = 8) -> None:
        if not (1 <= num_bits <= 8):
            raise ValueError(f"num_bits must be in [1, 8], got {num_bits}")
        self.num_bits: int = num_bits
        self.mu: np.ndarray | None = None   # global mean, shape (D,)
        self._mu_dev = None

    def __getstate__(self):
        st = dict(self.__dict__)
        st["_mu_dev"] = None
        return st

    def _mu(self) -> torch.Tensor:
        if self._mu_dev is None or self._mu_dev[0] is not self.mu:
            self._mu_dev = (self.mu, host.to_device(np.ascontiguousarray(self.mu, dtype=np.float32)))
        return self._mu_dev[1]

    def fit(self, X: np.ndarray) -> None:
        """Global mean of the training rows, X.mean(axis=0) in numpy's fp32 row order."""
        X = host.as_f32_2d(X)
        self.mu = host.to_host(ops_sq.col_mean_seq(host.to_device(X))).copy()
        self._mu_dev = None

    def compress(self, X: np.ndarray) -> np.ndarray:
        if self.mu is None:
            raise RuntimeError("LVQQuantizer.fit() must be called before compress/decompress.")
        X = host.as_f32_2d(X)
        n, D = X.shape
        if D != self.mu.shape[0]:
            raise ValueError(f"compress: got D={D}, fitted with D={self.mu.shape[0]}")
        mu = self._mu()
        out = torch.empty((n, ops_sq.lvq_code_size(D, self.num_bits)), dtype=torch.uint8, device=host.device())
        host.stream_rows(X, lambda xb, s, e: ops_sq.lvq_encode(xb, mu, self.num_bits, out=out[s:e]))
        return host.to_host(out)

    def decompress(self, codes: np.ndarray) -> np.ndarray:
        if self.mu is None:
            raise RuntimeError("LVQQuantizer.fit() must be called before compress/decompress.")
        codes = np.ascontiguousarray(codes, dtype=np.uint8)
        return host.to_host(ops_sq.lvq_decode(host.to_device(codes), self._mu(), self.num_bits))

    def get_compression_ratio(self, X: np.ndarray) -> float:
        D = X.shape[1]
        return D * 4 / (math.ceil(D * self.num_bits / 8) + 8)
