"""RaBitQIndex — 1-bit RaBitQ codes with the native popcount distance estimator.

Mirrors src/haag_vq/methods/search/rabitq_index.py:14-107 (a wrapper over
faiss.IndexRaBitQ): ``fit`` = train (centroid = mean of X) + add (encode);
``search_with_scores`` quantises each query residual to ``qb`` bits, transposes
it into bit planes and ranks codes by the unbiased RaBitQ estimate computed with
AND + popcount; ids uint32, distances float32. ``qb`` in [1, 8] (the reference's
``qb=0`` float path is not built).
"""
from __future__ import annotations

import pickle
from pathlib import Path
from typing import Literal, Optional, Tuple

import numpy as np
import torch

from b200vq import host, ops_sq

from haag_vq.methods.base_search_index import BaseSearchIndex

MAX_K = 128


class RaBitQIndex(BaseSearchIndex):
    def __init__(self, qb: int = 4) -> None:
        self._qb = int(qb)
        self._D: int = 0
        self._metric: Literal['l2', 'ip'] = 'l2'
        self._codes: Optional[np.ndarray] = None
        self._centroid: Optional[np.ndarray] = None
        self._codes_dev: Optional[torch.Tensor] = None
        self._centroid_dev: Optional[torch.Tensor] = None

    @property
    def ntotal(self) -> int:
        return 0 if self._codes is None else int(self._codes.shape[0])

    @property
    def code_size(self) -> int:
        return (self._D + 7) // 8 + 8

    def fit(self, X: np.ndarray, metric: Literal['l2', 'ip'] = 'l2') -> None:
        if not (1 <= self._qb <= 8):
            raise ValueError("RaBitQIndex on b200vq needs 1 <= qb <= 8")
        X = np.ascontiguousarray(X, dtype=np.float32)
        n, D = X.shape
        self._D = int(D)
        self._metric = metric
        acc = [torch.zeros(D, dtype=torch.float32, device=host.device())]
        def f(xb, s, e):
            acc[0] += xb.sum(0)
        host.stream_rows(X, f)
        c_dev = (acc[0] / float(n)).contiguous()
        out = torch.empty((n, self.code_size), dtype=torch.uint8, device=host.device())
        def g(xb, s, e):
            out[s:e] = ops_sq.rabitq1_encode(xb, c_dev, metric)
        host.stream_rows(X, g)
        self._centroid_dev, self._codes_dev = c_dev, out
        self._centroid = host.to_host(c_dev).copy()
        self._codes = host.to_host(out)

    def _require_fit(self):
        if self._codes is None:
            raise RuntimeError("RaBitQIndex must be fit() before use.")
        if self._codes_dev is None:
            self._codes_dev = host.to_device(self._codes)
            self._centroid_dev = host.to_device(self._centroid)
        return self._codes_dev

    def search(self, Q: np.ndarray, k: int) -> np.ndarray:
        ids, _ = self.search_with_scores(Q, k)
        return ids

    def search_with_scores(self, Q: np.ndarray, k: int) -> Tuple[np.ndarray, np.ndarray]:
        codes = self._require_fit()
        Q = np.ascontiguousarray(Q, dtype=np.float32)
        if k <= 0:
            raise ValueError(f"k must be positive, got {k}")
        # k > MAX_K: pages of MAX_K chained on the device (b200vq_rabitq1_scan_topk_after); beyond ntotal the ids are
        # 0xffffffff (faiss' -1) and the distances +-inf, as IndexRaBitQ.search returns them
        q_dev = host.to_device(Q)
        planes, qfac = ops_sq.rabitq1_prepare_queries(q_dev, self._centroid_dev, self._qb)
        vals, ids = ops_sq.rabitq1_scan_topk(codes, self._D, planes, qfac, self._qb, k, self._metric)
        return host.to_host(ids).view(np.uint32), host.to_host(vals)

    def memory_footprint(self) -> int:
        return self.ntotal * self.code_size

    def reconstruction_mse(self, X: np.ndarray, sample_ids: Optional[np.ndarray] = None) -> Optional[float]:
        if self._codes is None:
            return None
        self._require_fit()
        if sample_ids is None:
            sample_ids = np.arange(self.ntotal, dtype=np.int64)
        sample_ids = np.asarray(sample_ids, dtype=np.int64)
        x_hat = host.to_host(ops_sq.rabitq1_decode(host.to_device(self._codes[sample_ids]), self._D,
                                                   self._centroid_dev))
        return float(np.mean((X[sample_ids].astype(np.float32) - x_hat) ** 2))

    def save(self, path: str | Path) -> None:
        if self._codes is None:
            raise RuntimeError("RaBitQIndex.save() called before fit()")
        Path(path).parent.mkdir(parents=True, exist_ok=True)
        with open(path, 'wb') as f:
            pickle.dump({'format': 'b200vq.RaBitQIndex.v1', 'D': self._D, 'metric': self._metric,
                         'qb': self._qb, 'centroid': self._centroid, 'codes': self._codes}, f)

    def load(self, path: str | Path) -> None:
        with open(path, 'rb') as f:
            st = pickle.load(f)
        self._D, self._metric, self._qb = int(st['D']), st['metric'], int(st['qb'])
        self._centroid, self._codes = st['centroid'], st['codes']
        self._codes_dev = self._centroid_dev = None
