"""Precompute ground-truth nearest neighbours of a vectors file on the B200.

Mirrors src/haag_vq/benchmarks/precompute_ground_truth.py:14-135: load ``vectors_path`` (.npy), take the first
``num_queries`` rows as queries, exact L2 k-NN in query batches, save the (num_queries, k) int64 ids to
``output_path`` and the distances next to it (``.distances.npy``). The reference builds a faiss.IndexFlatL2
(optionally on a GPU through faiss); here the index is the device matrix and every batch is one call of the
exact flat top-k kernel (``b200vq_flat_topk_tc``). The typer CLI wrapper is not part of the hot path.
"""
from __future__ import annotations

from pathlib import Path
from time import perf_counter

import numpy as np

from b200vq import host, ops


def precompute_ground_truth(vectors_path: str, output_path: str, num_queries: int = 100, k: int = 100,
                            use_gpu: bool = True, batch_size: int = 1000) -> np.ndarray:
    vectors_path = Path(vectors_path)
    if not vectors_path.exists():
        raise FileNotFoundError(f"Vectors file not found: {vectors_path}")
    vectors = np.load(vectors_path)
    num_queries = min(num_queries, len(vectors))
    k = min(k, len(vectors))
    if k > 128:
        raise ValueError(f"k={k} exceeds the device top-k limit (128)")
    queries = np.ascontiguousarray(vectors[:num_queries], dtype=np.float32)
    t0 = perf_counter()
    xd = host.to_device(np.ascontiguousarray(vectors, dtype=np.float32))
    xn = ops.row_sqnorms(xd)
    ids = np.empty((num_queries, k), dtype=np.int64)
    dist = np.empty((num_queries, k), dtype=np.float32)
    for s in range(0, num_queries, batch_size):
        e = min(num_queries, s + batch_size)
        v, i = ops.flat_topk(xd, host.to_device(queries[s:e]), k, 'l2', xnorm=xn)
        ids[s:e] = host.to_host(i).view(np.uint32)
        dist[s:e] = host.to_host(v)
    elapsed = perf_counter() - t0
    output_path = Path(output_path)
    output_path.parent.mkdir(parents=True, exist_ok=True)
    np.save(output_path, ids)
    np.save(output_path.with_suffix('.distances.npy'), dist)
    print(f"ground truth {ids.shape} in {elapsed:.2f}s -> {output_path}")
    return ids
