"""SURVEY 8f rows 1 and 3 on the GPU: streaming compression (reference
src/haag_vq/benchmarks/streaming_sweep.py:41-224) and codebook export / query
(src/haag_vq/utils/faiss_export.py:304-506; reference tests tests/test_faiss_export.py:40-110)."""
import numpy as np
import pytest

from oracle import pq as opq

pytestmark = pytest.mark.gpu


def _data(n, d, seed=0):
    return np.random.default_rng(seed).standard_normal((n, d)).astype(np.float32)


@pytest.mark.parametrize("method", ["pq", "opq", "sq", "rabitq"])
def test_stream_compress_matches_whole_array(method):
    from haag_vq.benchmarks.streaming_sweep import build_model, stream_compress
    from haag_vq.metrics.distortion import compute_distortion
    X = _data(3000, 64, 3)
    model, _ = build_model(method, pq_subquantizers=8, opq_quantizers=8)
    model.fit(X[:2048])
    got = []
    batches = [X[s:s + 700] for s in range(0, X.shape[0], 700)]      # ragged last batch
    m = stream_compress(model, batches, sink=lambda c, s: got.append((s, c)))
    assert m["total_vectors_compressed"] == X.shape[0] and m["num_batches"] == 5
    whole = model.compress(X)
    streamed = np.concatenate([c for _, c in got])
    if method == "opq":
        # the rotation GEMM picks its implementation by batch size (tensor-core 3xTF32 from 1024 rows,
        # library sgemm below): both are fp32-accurate, so only near-tie assignments may differ
        assert (streamed != whole).mean() < 5e-3
    else:
        assert np.array_equal(streamed, whole)
    assert [s for s, _ in got] == [0, 700, 1400, 2100, 2800]
    # OPQ: the few near-tie codes that differ between batch sizes move the MSE in the 4th digit
    tol = 2e-3 if method == "opq" else 1e-4
    ref_mse = compute_distortion(X, whole, model)                    # the reference's definition
    assert m["mse"] == pytest.approx(ref_mse, rel=tol)
    # host path (compress + compute_distortion per batch) gives the same number
    m2 = stream_compress(model, batches, device_path=False)
    assert m2["mse"] == pytest.approx(m["mse"], rel=tol)
    # numpy restatement of the metric
    R = model.decompress(whole)
    assert ref_mse == pytest.approx(float(np.mean(np.sum((X - R) ** 2, axis=1))), rel=1e-4)
    m3 = stream_compress(model, batches, max_batches=2)
    assert m3["num_batches"] == 2 and m3["total_vectors_compressed"] == 1400


def test_streaming_sweep_synthetic_small():
    from haag_vq.benchmarks.streaming_sweep import streaming_sweep
    out = streaming_sweep("pq", training_size=4096, batch_size=1000, n_total=5000, D=64, pq_subquantizers="8")
    mt = out["metrics"]
    assert mt["total_vectors_compressed"] == 5000 and mt["num_batches"] == 5
    assert mt["compression_ratio"] == pytest.approx(64 * 4 / 8)
    assert 0.0 < mt["mse"] < 1.0       # unit-norm rows
    with pytest.raises(ValueError):
        streaming_sweep("nope", training_size=16, n_total=16, D=8)


def test_query_codebook_scalar_and_pq(tmp_path):
    from haag_vq.methods.product_quantization import ProductQuantizer
    from haag_vq.methods.scalar_quantization import ScalarQuantizer
    from haag_vq.utils.faiss_export import (METRIC_INNER_PRODUCT, export_codebook, load_fvecs, load_ivecs,
                                            query_codebook)
    data = np.array([[0.0, 0.0], [1.0, 1.0], [2.0, 2.0]], dtype=np.float32)
    sq = ScalarQuantizer()
    sq.fit(data)
    res = export_codebook(sq, tmp_path / "sq", codes=sq.compress(data))
    assert np.array_equal(load_fvecs(res["codebook"]), np.stack([data.min(0), data.max(0)]))
    assert np.array_equal(load_ivecs(res["codes"]), sq.compress(data).astype(np.int32))
    d, i = query_codebook(np.array([[0.1, 0.1], [1.9, 1.9]], dtype=np.float32), codebook_path=res["codebook"])
    assert i[0, 0] == 0 and i[1, 0] == 1 and i.dtype == np.int64
    assert d[0, 0] == pytest.approx(0.02, rel=1e-5)

    X = _data(4096, 32, 5)
    pq = ProductQuantizer(M=4, B=8)
    pq.fit(X)
    res = export_codebook(pq, tmp_path / "pq")
    cb = res["codebook_vectors"]
    assert cb.shape == (4 * 256, 8)
    Q = X[:17]
    for metric, omet in ((1, opq.L2), (METRIC_INNER_PRODUCT, opq.IP)):
        dist, idx = query_codebook(Q, model=pq, codebook_vectors=cb, topk=3, metric=metric)
        assert dist.shape == idx.shape == (17, 12)
        for m in range(4):
            assert np.all((idx[:, 3 * m:3 * m + 3] >= 256 * m) & (idx[:, 3 * m:3 * m + 3] < 256 * (m + 1)))
            rv, ri = opq.flat_search(cb[256 * m:256 * (m + 1)], Q[:, 8 * m:8 * m + 8], 3, omet)
            assert np.array_equal(idx[:, 3 * m:3 * m + 3] - 256 * m, ri.astype(np.int64))
            assert np.allclose(dist[:, 3 * m:3 * m + 3], rv, rtol=1e-4, atol=1e-5)
    # top-1 under L2 is the PQ code itself
    dist, idx = query_codebook(Q, model=pq, codebook_vectors=cb, topk=1)
    assert np.array_equal(idx - 256 * np.arange(4), pq.compress(Q).astype(np.int64))
    with pytest.raises(ValueError):
        query_codebook(Q[:, :5], model=pq, codebook_vectors=cb)
    with pytest.raises(ValueError):
        query_codebook(Q, codebook_vectors=cb, topk=0)
