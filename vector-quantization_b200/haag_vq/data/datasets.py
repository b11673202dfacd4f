"""Dataset container and exact ground truth on the B200.

Mirrors src/haag_vq/data/datasets.py:8-77: ``compute_ground_truth_faiss`` (faiss.IndexFlatL2 search in the
reference; the exact flat top-k kernel here) and ``Dataset`` (vectors, queries, ground truth, distance
callable). The HuggingFace / sentence-transformers loaders of the reference are outside the hot path.
"""
from __future__ import annotations

from typing import Callable, Optional

import numpy as np


def _euclidean(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    """Stand-in for sklearn.metrics.pairwise_distances (the reference's default callable): Euclidean."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    d2 = (a * a).sum(1)[:, None] + (b * b).sum(1)[None, :] - 2.0 * (a @ b.T)
    return np.sqrt(np.maximum(d2, 0.0))


_euclidean.metric = 'l2'   # lets evaluate_recall rank on the device


def compute_ground_truth_faiss(queries: np.ndarray, vectors: np.ndarray, k: int = 100) -> np.ndarray:
    """(num_queries, k) int64 exact L2 neighbours (datasets.py:8-34)."""
    from haag_vq.benchmarks.search_bench import compute_ground_truth
    return compute_ground_truth(vectors, queries, k=k, metric='l2')


class Dataset:
    def __init__(self, vectors: np.ndarray, queries: Optional[np.ndarray] = None,
                 ground_truth: Optional[np.ndarray] = None, num_queries: int = 100,
                 distance_metric: Callable[[np.ndarray, np.ndarray], np.ndarray] = _euclidean,
                 skip_ground_truth: bool = False):
        self.vectors = vectors
        if queries is None:
            self.queries = vectors[:num_queries]
        else:
            assert num_queries <= len(queries)
            self.queries = queries[:num_queries]
        if ground_truth is not None:
            assert num_queries <= len(ground_truth)
            self.ground_truth = ground_truth[:num_queries]
        elif skip_ground_truth:
            self.ground_truth = None
        else:
            self.ground_truth = compute_ground_truth_faiss(self.queries, self.vectors, k=100)
        self.distance_metric = distance_metric


def load_dummy_dataset(num_samples: int = 10000, dim: int = 1024, seed: int = 42) -> Dataset:
    np.random.seed(seed)
    return Dataset(np.random.randn(num_samples, dim).astype(np.float32))
