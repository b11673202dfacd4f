"""Recall of a quantizer's reconstructions (mirrors src/haag_vq/metrics/recall.py:6-42).

``evaluate_recall``: compress + decompress the corpus, rank the reconstructions for the first queries, compare
the top 10 / 100 with the ground truth. The reference materialises the (queries x N) distance matrix and fully
argsorts it; here, when the dataset's distance is Euclidean (the reference's default) or declares ``.metric``,
the ranking is the exact flat top-100 on the device. Any other callable is evaluated as the caller wrote it.
"""
from __future__ import annotations

import numpy as np

from b200vq import host, ops


def recall_at_k(true_nn: np.ndarray, retrieved: np.ndarray, k: int) -> float:
    """Mean over queries of |true[:k] ∩ retrieved[:k]| / k (recall.py:26-42)."""
    true_nn = np.asarray(true_nn)
    retrieved = np.asarray(retrieved)
    if len(true_nn) == 0:
        return 0.0
    hits = 0.0
    for gt, got in zip(true_nn, retrieved):
        hits += len(set(gt[:k].tolist()) & set(got[:k].tolist())) / k
    return hits / len(true_nn)


def evaluate_recall(data, model, num_queries: int = 100) -> dict:
    queries = np.ascontiguousarray(data.queries[:num_queries], dtype=np.float32)
    truth = data.ground_truth[:num_queries]
    recon = np.ascontiguousarray(model.decompress(model.compress(data.vectors)), dtype=np.float32)
    fn = data.distance_metric
    metric = getattr(fn, "metric", None)
    if metric is None and getattr(fn, "__name__", "") in ("pairwise_distances", "euclidean_distances"):
        metric = "l2"     # sklearn's default: Euclidean; squared L2 ranks identically
    if metric in ("l2", "ip"):
        k = min(100, recon.shape[0])
        _, ids = ops.flat_topk(host.to_device(recon), host.to_device(queries), k, metric)
        retrieved = host.to_host(ids).view(np.uint32).astype(np.int64)
    else:
        retrieved = np.asarray(fn(queries, recon)).argsort(axis=1)
    return {"recall@10": recall_at_k(truth, retrieved, k=10), "recall@100": recall_at_k(truth, retrieved, k=100)}
