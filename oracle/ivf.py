"""CPU restatement (TEST INFRASTRUCTURE ONLY) of the reference's IVF shell after the clustering:

  src/haag_vq/methods/search/ivf_quantized_index.py:55-78    assignment, per-cluster residual quantizers
  src/haag_vq/methods/search/ivf_quantized_index.py:87-146   probe, per-query decompress + exact distance + top-k

The coarse centroids come from faiss.Kmeans in the reference (ivf_quantized_index.py:50-53); faiss is a third-party
dependency that is not vendored, so the centroids are an INPUT here ("parity unpinned" for the k-means itself,
pinned for everything computed from its output). Distances are evaluated in float64 so that the checker is not
itself a source of rounding ties.
"""
from __future__ import annotations

import numpy as np


def assign(X: np.ndarray, centroids: np.ndarray, metric: str):
    """km.index.search(X, 1): nearest centroid (IndexFlatL2) or largest inner product (spherical: IndexFlatIP).
    Returns (assignment, margin): margin = gap between the best and the second best score."""
    X64, C64 = X.astype(np.float64), centroids.astype(np.float64)
    if metric == 'ip':
        s = -(X64 @ C64.T)
    else:
        s = (X64 * X64).sum(1)[:, None] + (C64 * C64).sum(1)[None, :] - 2.0 * (X64 @ C64.T)
    a = np.argmin(s, axis=1)
    part = np.partition(s, 1, axis=1) if s.shape[1] > 1 else np.concatenate([s, s + 1.0], axis=1)
    return a.astype(np.int32), part[:, 1] - part[:, 0]


def build(X: np.ndarray, centroids: np.ndarray, assignments: np.ndarray, quantizer_factory):
    """Per-cluster residual quantizers exactly as the reference's fit loop; returns (quantizers, codes, ids)."""
    K = centroids.shape[0]
    qs, codes, ids = [], [], []
    for c in range(K):
        mask = assignments == c
        vid = np.where(mask)[0].astype(np.uint32)
        ids.append(vid)
        if not mask.any():
            qs.append(None)
            codes.append(np.array([], dtype=np.uint8))
            continue
        residuals = X[mask] - centroids[c]
        q = quantizer_factory()
        q.fit(residuals)
        qs.append(q)
        codes.append(q.compress(residuals))
    return qs, codes, ids


def search(Q: np.ndarray, k: int, centroids: np.ndarray, qs, codes, ids, nprobe: int, metric: str):
    """search_with_scores of the reference; returns (ids uint32, dists float64, probe lists)."""
    Q64, C64 = Q.astype(np.float64), centroids.astype(np.float64)
    if metric == 'ip':
        cs = -(Q64 @ C64.T)
    else:
        cs = (Q64 * Q64).sum(1)[:, None] + (C64 * C64).sum(1)[None, :] - 2.0 * (Q64 @ C64.T)
    npb = min(nprobe, centroids.shape[0])
    probe = np.argsort(cs, axis=1, kind='stable')[:, :npb]
    nq = Q.shape[0]
    out_ids = np.full((nq, k), np.iinfo(np.uint32).max, dtype=np.uint32)
    out_d = np.full((nq, k), np.inf, dtype=np.float64)
    for qi in range(nq):
        cid_list, vec_list = [], []
        for cid in probe[qi]:
            q = qs[cid]
            if q is None:
                continue
            recon = q.decompress(codes[cid]).astype(np.float32) + centroids[cid]
            cid_list.append(ids[cid])
            vec_list.append(recon)
        if not cid_list:
            continue
        vids = np.concatenate(cid_list)
        vecs = np.concatenate(vec_list, axis=0).astype(np.float64)
        d = ((Q64[qi] - vecs) ** 2).sum(1) if metric == 'l2' else -(vecs @ Q64[qi])
        kk = min(k, d.size)
        order = np.lexsort((vids, d))[:kk]          # (distance, id) order
        out_ids[qi, :kk] = vids[order]
        out_d[qi, :kk] = d[order]
    return out_ids, out_d, probe
