"""ProductQuantizer on the B200 kernels.

Mirrors src/haag_vq/methods/product_quantization.py:9-99 (a wrapper over
faiss.ProductQuantizer): same constructor aliases, attributes (``M, B,
num_chunks, num_clusters, codebooks, chunk_dim, pq``) and error behaviour. The
``pq`` attribute is a small state object exposing the faiss fields callers read
(``d, M, nbits, ksub, dsub, code_size, centroids``) plus ``compute_codes`` /
``decode``.
"""
from __future__ import annotations

from typing import List, Optional

import numpy as np
import torch

from b200vq import host, ops
from b200vq import pq as dpq

from .base_quantizer import BaseQuantizer


class DevicePQ:
    """Stand-in for faiss.ProductQuantizer: trained state + device codec."""

    def __init__(self, d: int, M: int, nbits: int):
        if nbits < 1 or nbits > 8:
            raise ValueError("b200vq ProductQuantizer supports 1 <= B <= 8")
        # nbits < 8: faiss packs the indices into a little-endian bit stream of ceil(M nbits / 8) bytes
        # (PQEncoderGeneric); the kernels work on one byte per index and b200vq_pq_pack_bits / _unpack_bits convert.
        # nbits > 8 (up to 65 536 centroids per sub-quantizer) is not built: the kernels keep ksub <= 256.
        self.d, self.M, self.nbits = int(d), int(M), int(nbits)
        self.ksub = 1 << self.nbits
        self.dsub = self.d // self.M
        self.code_size = (self.M * self.nbits + 7) // 8
        self.codec: Optional[dpq.PQCodec] = None
        self._centroids_host: Optional[np.ndarray] = None

    # -- state ---------------------------------------------------------------
    def set_centroids(self, centroids: torch.Tensor) -> None:
        self.codec = dpq.PQCodec(centroids)
        self._centroids_host = None

    @property
    def centroids(self) -> np.ndarray:
        """Flat (M*ksub*dsub,) float32 array, the layout of faiss' pq.centroids."""
        if self._centroids_host is None:
            self._centroids_host = host.to_host(self.codec.centroids).reshape(-1).copy()
        return self._centroids_host

    # -- faiss-like methods --------------------------------------------------
    def train(self, X: np.ndarray, seed: int = 1234) -> None:
        X = host.as_f32_2d(X)
        ksub = self.ksub
        ids = dpq.subsample_ids(X.shape[0], dpq.MAX_POINTS_PER_CENTROID * ksub, seed)
        xt = host.to_device(X[ids] if ids.size != X.shape[0] else X)
        # the subset is already drawn: train on it as given
        self.set_centroids(dpq.pq_train(xt, self.M, self.nbits, seed=seed))

    def compute_codes(self, X: np.ndarray) -> np.ndarray:
        X = host.as_f32_2d(X)
        if X.shape[1] != self.d:
            raise ValueError(f"compute_codes: got D={X.shape[1]}, trained with D={self.d}")
        n = X.shape[0]
        out = torch.empty((n, self.M), dtype=torch.uint8, device=host.device())
        host.stream_rows(X, lambda xb, s, e: self.codec.encode(xb, out=out[s:e]))
        if self.nbits != 8:
            out = ops.pq_pack_bits(out, self.nbits)
        return host.to_host(out)

    def unpack_codes(self, codes_dev: torch.Tensor) -> torch.Tensor:
        """(n, code_size) stored codes -> (n, M) one byte per index (identity for nbits = 8)."""
        return codes_dev if self.nbits == 8 else ops.pq_unpack_bits(codes_dev, self.M, self.nbits)

    def decode(self, codes: np.ndarray) -> np.ndarray:
        codes = np.ascontiguousarray(codes, dtype=np.uint8)
        if codes.ndim != 2 or codes.shape[1] != self.code_size:
            raise ValueError(f"decode: codes must be (n, {self.code_size})")
        return host.to_host(self.codec.decode(self.unpack_codes(host.to_device(codes))))


class ProductQuantizer(BaseQuantizer):
    def __init__(self, M: Optional[int] = None, B: int = 8, *, num_chunks: Optional[int] = None,
                 num_clusters: Optional[int] = None):
        if M is None and num_chunks is not None:
            M = int(num_chunks)
        if num_chunks is not None and M is not None and int(num_chunks) != int(M):
            raise ValueError("Conflicting values for M and num_chunks")
        if num_clusters is not None:
            if num_clusters <= 0:
                raise ValueError("num_clusters must be positive")
            log2_clusters = int(round(np.log2(num_clusters)))
            if 2 ** log2_clusters != int(num_clusters):
                raise ValueError("num_clusters must be a power of two")
            B = log2_clusters
        if M is None:
            M = 8
        self.M: int = int(M)
        self.B: int = int(B)
        self.num_chunks: int = self.M
        self.num_clusters: int = 2 ** self.B
        self.codebooks: List[np.ndarray] = []
        self.chunk_dim: Optional[int] = None
        self.pq: Optional[DevicePQ] = None

    def fit(self, X: np.ndarray) -> None:
        X = np.asarray(X, dtype=np.float32)
        N, D = X.shape
        if D % self.M != 0:
            raise AssertionError("D must be divisible by M (number of subquantizers)")
        self.chunk_dim = D // self.M
        pq = DevicePQ(D, self.M, self.B)
        pq.train(X)
        self.pq = pq
        centroids = pq.centroids.reshape(self.M, pq.ksub, pq.dsub)
        self.codebooks = [np.array(centroids[m], copy=True) for m in range(self.M)]

    def compress(self, X: np.ndarray) -> np.ndarray:
        if self.pq is None:
            raise RuntimeError("ProductQuantizer must be fitted before compress(). Call fit() first.")
        return self.pq.compute_codes(np.asarray(X, dtype=np.float32))

    def decompress(self, codes: np.ndarray) -> np.ndarray:
        if self.pq is None:
            raise RuntimeError("ProductQuantizer must be fitted before decompress(). Call fit() first.")
        return self.pq.decode(np.asarray(codes))

    def get_compression_ratio(self, X: np.ndarray) -> float:
        D = int(X.shape[1])
        if self.pq is not None:
            size = int(self.pq.code_size)
        else:
            size = int((self.M * self.B + 7) // 8)
        return float(D * 4 / size)

    # -- persistence helpers used by FlatQuantizedIndex.save/load ------------
    def _state(self) -> dict:
        return {'type': 'ProductQuantizer', 'M': self.M, 'B': self.B,
                'flat_centroids': self.pq.centroids.copy(), 'D': self.pq.d}

    @classmethod
    def _from_state(cls, qs: dict) -> "ProductQuantizer":
        q = cls(M=qs['M'], B=qs['B'])
        D = qs['D']
        q.chunk_dim = D // qs['M']
        q.pq = DevicePQ(D, qs['M'], qs['B'])
        flat = np.asarray(qs['flat_centroids'], dtype=np.float32)
        cent = flat.reshape(qs['M'], q.pq.ksub, q.pq.dsub)
        q.pq.set_centroids(host.to_device(cent))
        q.codebooks = [np.array(cent[m], copy=True) for m in range(qs['M'])]
        return q
