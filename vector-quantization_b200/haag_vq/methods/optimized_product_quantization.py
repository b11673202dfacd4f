"""OptimizedProductQuantizer on the B200 kernels.

Mirrors src/haag_vq/methods/optimized_product_quantization.py:8-46:
``opq = OPQMatrix(D, M, D).train(X)``; the PQ is trained on the rotated data;
``compress = pq.compute_codes(opq.apply(X))``; ``decompress =
opq.reverse_transform(pq.decode(codes))``.
"""
from __future__ import annotations

from typing import Optional

import numpy as np
import torch

from b200vq import host
from b200vq import pq as dpq

from .base_quantizer import BaseQuantizer
from .product_quantization import DevicePQ


class DeviceOPQMatrix:
    """Stand-in for faiss.OPQMatrix(d, M, d): orthonormal A with y = x A^T."""

    def __init__(self, d: int, M: int):
        self.d_in = self.d_out = int(d)
        self.M = int(M)
        self.A_dev: Optional[torch.Tensor] = None
        self.is_trained = False

    @property
    def A(self) -> np.ndarray:
        return host.to_host(self.A_dev).reshape(-1).copy()

    def train(self, X: np.ndarray, nbits: int = 8, seed: int = 123, niter: int = 50) -> None:
        X = host.as_f32_2d(X)
        ids = dpq.subsample_ids(X.shape[0], dpq.MAX_POINTS_PER_CENTROID * 256, seed)
        xt = host.to_device(X[ids] if ids.size != X.shape[0] else X)
        self.A_dev = dpq.opq_train(xt, self.M, nbits, niter=niter, seed=seed)
        self.is_trained = True

    def apply(self, X: np.ndarray) -> np.ndarray:
        X = host.as_f32_2d(X)
        out = np.empty_like(X)
        def f(xb, s, e):
            out[s:e] = host.to_host(dpq.rotate(xb, self.A_dev))
        host.stream_rows(X, f)
        return out

    def reverse_transform(self, Y: np.ndarray) -> np.ndarray:
        Y = host.as_f32_2d(Y)
        return host.to_host(dpq.rotate(host.to_device(Y), self.A_dev, transpose=False))


class OptimizedProductQuantizer(BaseQuantizer):
    def __init__(self, M: int, B: int = 8):
        self.M = M
        self.B = B
        self.opq: Optional[DeviceOPQMatrix] = None
        self.pq: Optional[DevicePQ] = None

    def fit(self, X: np.ndarray):
        X = np.ascontiguousarray(X, dtype=np.float32)
        N, D = X.shape
        assert D % self.M == 0, "D must be divisible by M"
        opq = DeviceOPQMatrix(D, self.M)
        opq.train(X, nbits=self.B)
        # train the PQ on the rotated data (tests/test_opq_trains_on_rotated.py:14-26)
        pq = DevicePQ(D, self.M, self.B)
        ids = dpq.subsample_ids(N, dpq.MAX_POINTS_PER_CENTROID * pq.ksub, 1234)
        xt = host.to_device(X[ids] if ids.size != N else X)
        pq.set_centroids(dpq.pq_train(dpq.rotate(xt, opq.A_dev), self.M, self.B, seed=1234))
        pq.codec.rotation = opq.A_dev
        self.opq, self.pq = opq, pq

    def compress(self, X: np.ndarray) -> np.ndarray:
        if self.pq is None:
            raise RuntimeError("OptimizedProductQuantizer must be fitted before compress().")
        # rotation is fused on the device: codes = assign(x A^T)
        return self.pq.compute_codes(np.asarray(X, dtype=np.float32))

    def decompress(self, compressed: np.ndarray) -> np.ndarray:
        if self.pq is None:
            raise RuntimeError("OptimizedProductQuantizer must be fitted before decompress().")
        return self.pq.decode(np.asarray(compressed))

    def get_compression_ratio(self, X: np.ndarray) -> float:
        D = int(X.shape[1])
        return float(D * 4 / int((self.M * self.B + 7) // 8))
