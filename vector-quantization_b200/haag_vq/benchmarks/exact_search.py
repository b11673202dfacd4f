"""Chunked exact-search core for the quantizer study, on the B200.

Mirrors src/haag_vq/benchmarks/exact_search.py:1-122. score(q, x_i) =
q . x_hat_i / ||x_i||_exact; reconstructions are scaled by 1/||x_i|| and ranked
by exact inner product. The reference builds a faiss.IndexFlatIP on the host;
here the "index" is a device matrix and ``search_index`` runs the exact flat
GEMM + top-k kernel.
"""
from __future__ import annotations

from typing import Callable, Dict, Iterator, Tuple

import numpy as np
import torch

from b200vq import host, ops

ReconstructFn = Callable[[np.ndarray], np.ndarray]  # ids (1D uint32) -> (m, D) float32


class DeviceFlatIP:
    """Stand-in for faiss.IndexFlatIP(d): rows appended with add(), exact IP search."""

    def __init__(self, d: int):
        self.d = int(d)
        self._chunks = []
        self._mat = None
        self.ntotal = 0

    def add(self, x: np.ndarray) -> None:
        x = np.ascontiguousarray(x, dtype=np.float32)
        self._chunks.append(host.to_device(x))
        self.ntotal += x.shape[0]
        self._mat = None

    def _matrix(self) -> torch.Tensor:
        if self._mat is None:
            self._mat = self._chunks[0] if len(self._chunks) == 1 else torch.cat(self._chunks, 0)
            self._chunks = [self._mat]
        return self._mat

    def search(self, Q: np.ndarray, k: int):
        Q = np.ascontiguousarray(Q, dtype=np.float32)
        if k > self.ntotal:
            raise AssertionError("faiss returned -1 sentinel; k likely > index.ntotal")
        vals, ids = ops.flat_topk(self._matrix(), host.to_device(Q), k, 'ip')
        return host.to_host(vals), host.to_host(ids).view(np.uint32).astype(np.int64)


def compute_exact_norms(X: np.ndarray, eps: float = 1e-12) -> np.ndarray:
    """Per-vector exact L2 norm, floored at eps to avoid division by zero."""
    X = host.as_f32_2d(X)
    out = torch.empty((X.shape[0],), dtype=torch.float32, device=host.device())
    def f(xb, s, e):
        out[s:e] = torch.sqrt(ops.row_sqnorms(xb))
    host.stream_rows(X, f)
    norms = np.maximum(host.to_host(out), eps)
    return norms.astype(np.float32)


def _chunks(n: int, chunk: int) -> Iterator[np.ndarray]:
    for start in range(0, n, chunk):
        yield np.arange(start, min(start + chunk, n), dtype=np.uint32)


def build_scaled_ip_index(reconstruct_fn: ReconstructFn, n: int, d: int, norms: np.ndarray,
                          chunk: int = 50_000) -> DeviceFlatIP:
    index = DeviceFlatIP(d)
    for ids in _chunks(n, chunk):
        x_hat = np.ascontiguousarray(reconstruct_fn(ids), dtype=np.float32)
        x_hat *= (1.0 / norms[ids])[:, None]
        index.add(x_hat)
    return index


def normalized_ground_truth(X: np.ndarray, Q: np.ndarray, k: int, norms: np.ndarray | None = None,
                            chunk: int = 50_000) -> np.ndarray:
    """Exact top-k under q . (x/||x||). Returns (nq, k) uint32 IDs."""
    X = np.asarray(X, dtype=np.float32)
    Q = np.ascontiguousarray(Q, dtype=np.float32)
    n, d = X.shape
    if norms is None:
        norms = compute_exact_norms(X)
    index = build_scaled_ip_index(lambda ids: X[ids], n, d, norms, chunk=chunk)
    _, ids = search_index(index, Q, k=k)
    return ids


def search_index(index, Q: np.ndarray, k: int) -> Tuple[np.ndarray, np.ndarray]:
    """Search an index. Returns (scores (nq,k) float32, ids (nq,k) uint32)."""
    Q = np.ascontiguousarray(Q, dtype=np.float32)
    scores, ids = index.search(Q, k)
    assert (ids >= 0).all(), "faiss returned -1 sentinel; k likely > index.ntotal"
    return scores.astype(np.float32), ids.astype(np.uint32)


def recall_at_ks(retrieved_ids: np.ndarray, gt_ids: np.ndarray,
                 ks: Tuple[int, ...] = (1, 10, 100)) -> Dict[int, float]:
    """Recall@k: mean over queries of |ret[:k] & gt[:k]| / min(k, |ret|, |gt|) (host-side metric)."""
    nq = retrieved_ids.shape[0]
    out: Dict[int, float] = {}
    for k in ks:
        kk_ret = min(k, retrieved_ids.shape[1])
        kk_gt = min(k, gt_ids.shape[1])
        denom = min(kk_ret, kk_gt)
        if denom == 0:
            out[k] = 0.0
            continue
        total = 0.0
        for i in range(nq):
            gt_set = set(gt_ids[i, :kk_gt].tolist())
            ret_set = set(retrieved_ids[i, :kk_ret].tolist())
            total += len(gt_set & ret_set) / denom
        out[k] = total / nq if nq else 0.0
    return out


def reconstruction_mse(X: np.ndarray, reconstruct_fn: ReconstructFn, sample_ids: np.ndarray,
                       chunk: int = 50_000) -> float:
    """Mean per-element squared error between X[sample] and its reconstruction."""
    X = np.asarray(X, dtype=np.float32)
    sample_ids = np.asarray(sample_ids, dtype=np.uint32)
    d = X.shape[1]
    sq_err = 0.0
    count = 0
    for start in range(0, sample_ids.size, chunk):
        block = sample_ids[start:start + chunk]
        x_hat = np.asarray(reconstruct_fn(block), dtype=np.float32)
        from b200vq import ops_sq
        sq_err += float(ops_sq.sq_err_sum(host.to_device(np.ascontiguousarray(X[block])),
                                          host.to_device(np.ascontiguousarray(x_hat))).item())
        count += block.size * d
    return sq_err / count if count else 0.0
