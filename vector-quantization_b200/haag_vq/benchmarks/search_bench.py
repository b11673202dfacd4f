"""Benchmark harness for BaseSearchIndex implementations, ground truth on the B200.

Mirrors src/haag_vq/benchmarks/search_bench.py:25-144: ``compute_ground_truth`` (the reference builds a
faiss.IndexFlatL2 / IndexFlatIP and searches it; here one call of the exact flat top-k kernel,
``b200vq_flat_topk_tc`` = TF32 screen + exact fp32 finish), ``_compute_recall`` (set semantics) and
``benchmark_index`` (fit, best-of-repeats search timing, recall@k, footprint, reconstruction MSE).
"""
from __future__ import annotations

import time
from datetime import datetime, timezone

import numpy as np

from b200vq import host, ops
from haag_vq.methods.base_search_index import BaseSearchIndex

MAX_K = 128   # device top-k limit


def _utc_timestamp() -> str:
    return datetime.now(timezone.utc).strftime('%Y-%m-%dT%H:%M:%SZ')


def compute_ground_truth(X_train: np.ndarray, X_query: np.ndarray, k: int = 10, metric: str = 'l2',
                         query_batch: int = 4096) -> np.ndarray:
    """(nq, k) int64 ids of the exact nearest rows, sorted by distance (L2 ascending / IP descending)."""
    X_train = np.ascontiguousarray(X_train, dtype=np.float32)
    X_query = np.ascontiguousarray(X_query, dtype=np.float32)
    N = X_train.shape[0]
    k = min(k, N)
    if k > MAX_K:
        raise ValueError(f"k={k} exceeds the device top-k limit ({MAX_K})")
    if k <= 0 or X_query.shape[0] == 0:
        return np.empty((X_query.shape[0], max(k, 0)), dtype=np.int64)
    xd = host.to_device(X_train)
    xn = ops.row_sqnorms(xd)
    out = np.empty((X_query.shape[0], k), dtype=np.int64)
    for s in range(0, X_query.shape[0], query_batch):
        e = min(X_query.shape[0], s + query_batch)
        _, ids = ops.flat_topk(xd, host.to_device(X_query[s:e]), k, 'ip' if metric == 'ip' else 'l2', xnorm=xn)
        out[s:e] = host.to_host(ids).view(np.uint32)
    return out


def _compute_recall(ids: np.ndarray, ground_truth: np.ndarray, k: int) -> float:
    """Mean over queries of |returned[:k] ∩ truth[:k]| / min(k, k_gt) (search_bench.py:59-77)."""
    nq = ids.shape[0]
    k_gt = min(k, ground_truth.shape[1])
    k_ret = min(k, ids.shape[1])
    if nq == 0 or k_gt == 0:
        return 0.0
    hits = sum(len(set(ground_truth[i, :k_gt].tolist()) & set(ids[i, :k_ret].tolist())) for i in range(nq))
    return hits / (nq * k_gt)


def benchmark_index(index: BaseSearchIndex, X_train: np.ndarray, X_query: np.ndarray, gt_ids: np.ndarray, k: int = 10,
                    repeats: int = 3, mse_sample: int = 1000) -> dict:
    """Fit on X_train, search X_query (best of ``repeats``), report the reference's metric dict."""
    X_train = np.ascontiguousarray(X_train, dtype=np.float32)
    X_query = np.ascontiguousarray(X_query, dtype=np.float32)
    N, D = X_train.shape
    nq = X_query.shape[0]
    index.fit(X_train)
    best_ids, best_time = None, float('inf')
    for _ in range(repeats):
        t0 = time.perf_counter()
        cur = index.search(X_query, k)          # numpy result: the device work is complete when it returns
        dt = time.perf_counter() - t0
        if dt < best_time:
            best_time, best_ids = dt, cur
    mem = index.memory_footprint()
    if N > mse_sample:
        sample_ids = np.random.default_rng(0).choice(N, mse_sample, replace=False).astype(np.uint32)
    else:
        sample_ids = np.arange(N, dtype=np.uint32)
    return {
        'method': type(index).__name__,
        'recall_at_k': _compute_recall(best_ids, gt_ids, k),
        'qps': nq / best_time if best_time > 0 else float('inf'),
        'memory_bytes': mem,
        'compression_ratio': N * D * 4 / mem if mem > 0 else float('inf'),
        'mse': index.reconstruction_mse(X_train, sample_ids=sample_ids),
        'k': k, 'N': N, 'D': D,
    }


def sweep_bpd(index_factory, bpd_values, X_train: np.ndarray, X_query: np.ndarray, gt_ids: np.ndarray, k: int = 10):
    """One benchmark_index row per bits-per-dimension value (search_bench.py:147-190): a pandas DataFrame with
    the metric columns plus ``bpd`` and a shared UTC ``timestamp``."""
    import pandas as pd
    stamp = _utc_timestamp()
    rows = []
    for bpd in bpd_values:
        row = benchmark_index(index_factory(bpd), X_train, X_query, gt_ids, k=k)
        row['bpd'] = bpd
        row['timestamp'] = stamp
        rows.append(row)
    return pd.DataFrame(rows)


def compare_methods(method_configs: dict, X_train: np.ndarray, X_query: np.ndarray, gt_ids: np.ndarray, k: int = 10):
    """One benchmark_index row per named (unfitted) index (search_bench.py:192-218); ``method`` = the dict key."""
    import pandas as pd
    stamp = _utc_timestamp()
    rows = []
    for name, index in method_configs.items():
        row = benchmark_index(index, X_train, X_query, gt_ids, k=k)
        row['method'] = name
        row['timestamp'] = stamp
        rows.append(row)
    return pd.DataFrame(rows)
