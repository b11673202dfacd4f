"""Ground truth / recall harness rows of SURVEY 8a (a19, a20): haag_vq.benchmarks.search_bench,
haag_vq.data.datasets, haag_vq.metrics.recall against numpy brute force (reference
src/haag_vq/benchmarks/search_bench.py:25-144, data/datasets.py:8-77, metrics/recall.py:6-42)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _data(n, d, seed=0):
    return np.random.default_rng(seed).standard_normal((n, d)).astype(np.float32)


def test_ground_truth_matches_brute_force():
    from haag_vq.benchmarks.search_bench import _compute_recall, compute_ground_truth
    from haag_vq.data.datasets import Dataset, compute_ground_truth_faiss
    X, Q = _data(5000, 48, 1), _data(37, 48, 2)
    d2 = ((Q[:, None, :].astype(np.float64) - X[None].astype(np.float64)) ** 2).sum(-1)
    gt = compute_ground_truth(X, Q, k=10, metric='l2')
    assert gt.dtype == np.int64 and gt.shape == (37, 10)
    assert np.array_equal(gt, np.argsort(d2, axis=1, kind="stable")[:, :10])
    ip = Q.astype(np.float64) @ X.T.astype(np.float64)
    assert np.array_equal(compute_ground_truth(X, Q, k=5, metric='ip'), np.argsort(-ip, axis=1, kind="stable")[:, :5])
    assert compute_ground_truth(X[:3], Q, k=10).shape == (37, 3)                 # k clipped to N
    assert np.array_equal(compute_ground_truth_faiss(Q, X, k=10), gt)
    assert _compute_recall(gt, gt, 10) == 1.0
    assert _compute_recall(np.array([[0, 9, 2]]), np.array([[0, 1, 2]]), 3) == pytest.approx(2 / 3)
    ds = Dataset(X, num_queries=20)
    assert ds.ground_truth.shape == (20, 100) and np.array_equal(ds.ground_truth[:, 0], np.arange(20))


def test_benchmark_index_and_evaluate_recall():
    from haag_vq.benchmarks.search_bench import benchmark_index, compute_ground_truth, sweep_bpd
    from haag_vq.data.datasets import Dataset
    from haag_vq.methods.scalar_quantization import ScalarQuantizer
    from haag_vq.methods.search.flat_quantized_index import FlatQuantizedIndex
    from haag_vq.metrics.recall import evaluate_recall, recall_at_k
    X, Q = _data(4000, 32, 3), _data(50, 32, 4)
    gt = compute_ground_truth(X, Q, k=10)
    row = benchmark_index(FlatQuantizedIndex(ScalarQuantizer(8)), X, Q, gt, k=10, repeats=2)
    assert set(row) == {'method', 'recall_at_k', 'qps', 'memory_bytes', 'compression_ratio', 'mse', 'k', 'N', 'D'}
    assert row['method'] == 'FlatQuantizedIndex' and row['recall_at_k'] > 0.9 and row['qps'] > 0 and row['N'] == 4000
    df = sweep_bpd(lambda bpd: FlatQuantizedIndex(ScalarQuantizer(8 if bpd > 4.5 else 4)), [4.0, 8.0], X, Q, gt)
    assert list(df['bpd']) == [4.0, 8.0] and df['recall_at_k'][1] >= df['recall_at_k'][0]
    from haag_vq.benchmarks.search_bench import compare_methods
    from haag_vq.methods.search.rabitq_index import RaBitQIndex
    cm = compare_methods({"sq8": FlatQuantizedIndex(ScalarQuantizer(8)), "rabitq": RaBitQIndex()}, X, Q, gt)
    assert list(cm['method']) == ["sq8", "rabitq"] and cm['recall_at_k'][0] > cm['recall_at_k'][1] > 0.2
    ds = Dataset(X, queries=Q, num_queries=50)
    r = evaluate_recall(ds, _fitted(ScalarQuantizer(8), X), num_queries=50)
    assert r["recall@10"] > 0.9 and 0.9 < r["recall@100"] <= 1.0
    assert recall_at_k(np.array([[1, 2, 3]]), np.array([[3, 2, 9]]), 3) == pytest.approx(2 / 3)
    # a caller-supplied distance callable is used as written
    ds2 = Dataset(X, queries=Q, num_queries=50, ground_truth=ds.ground_truth,
                  distance_metric=lambda a, b: ((a[:, None, :] - b[None]) ** 2).sum(-1))
    assert evaluate_recall(ds2, _fitted(ScalarQuantizer(8), X), num_queries=50) == r


def _fitted(q, X):
    q.fit(X)
    return q


def test_precompute_ground_truth_files(tmp_path):
    from haag_vq.benchmarks.precompute_ground_truth import precompute_ground_truth
    X = _data(3000, 24, 9)
    np.save(tmp_path / "v.npy", X)
    ids = precompute_ground_truth(str(tmp_path / "v.npy"), str(tmp_path / "out" / "gt.npy"), num_queries=40, k=7,
                                  batch_size=16)
    d2 = ((X[:40, None, :].astype(np.float64) - X[None].astype(np.float64)) ** 2).sum(-1)
    assert np.array_equal(ids, np.argsort(d2, axis=1, kind="stable")[:, :7])
    assert np.array_equal(np.load(tmp_path / "out" / "gt.npy"), ids)
    dist = np.load(tmp_path / "out" / "gt.distances.npy")
    assert dist.shape == (40, 7) and np.allclose(dist, np.sort(d2, axis=1)[:, :7], rtol=1e-4, atol=1e-4)
    with pytest.raises(FileNotFoundError):
        precompute_ground_truth(str(tmp_path / "nope.npy"), str(tmp_path / "x.npy"))
