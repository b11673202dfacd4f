"""The reference's own contract tests, re-stated against the haag_vq drop-in on the GPU
(tests/test_method_registry.py, test_quantizer_adapters.py, test_exact_search.py,
test_flat_quantized.py, test_opq_trains_on_rotated.py, test_extended_rabitq.py of the reference)."""
import pickle

import numpy as np
import pytest

from oracle import pq as opq
from parity import assert_codes_parity, make_data

pytestmark = pytest.mark.gpu


def _data(n=256, d=48, seed=0):
    return np.random.default_rng(seed).standard_normal((n, d)).astype(np.float32)


@pytest.mark.parametrize("method,bpd", [("pq", 1.0), ("pq", 2.0), ("opq", 1.0), ("sq", 4.0), ("sq", 8.0),
                                        ("sq", 16.0), ("rabitq", 1.0), ("rabitq", 4.0)])
def test_registry_methods_fit_and_reconstruct(method, bpd):
    from haag_vq.benchmarks.method_registry import build_quantizer
    from haag_vq.benchmarks.quantizer_adapters import Quantizer
    X = _data()
    q = build_quantizer(method, bpd, X.shape[1])
    assert isinstance(q, Quantizer)
    with pytest.raises(RuntimeError):
        q.reconstruct(np.arange(3))
    q.fit(X)
    R = q.reconstruct(np.arange(X.shape[0], dtype=np.uint32))
    assert R.shape == X.shape and R.dtype == np.float32 and np.isfinite(R).all()
    assert q.code_bytes() > 0
    ids = np.array([7, 3, 200])
    assert np.allclose(q.reconstruct(ids), R[ids], rtol=1e-5, atol=1e-6)
    with pytest.raises(ValueError):
        q.reconstruct(np.array([-1]))


def test_registry_errors_and_constants():
    from haag_vq.benchmarks import method_registry as mr
    assert mr.FAISS_METHODS == ("pq", "opq", "sq") and mr.PQ_BITS_PER_SUB == 8
    assert set(mr.ALL_METHODS) == set(mr.FAISS_METHODS) | set(mr.SAQ_METHODS)
    with pytest.raises(ValueError):
        mr.build_quantizer("nope", 4.0, 48)
    with pytest.raises(NotImplementedError):
        mr.build_quantizer("perdim_mse_exact", 4.0, 48)          # optimal-DP codebooks: not built
    from haag_vq.benchmarks.quantizer_adapters import SaqEngineAdapter
    assert isinstance(mr.build_quantizer("saq_paper", 4.0, 48), SaqEngineAdapter)


def test_adapter_code_bytes_and_sq_mse():
    from haag_vq.benchmarks.quantizer_adapters import FaissQuantizerAdapter
    from haag_vq.methods.product_quantization import ProductQuantizer
    from haag_vq.methods.scalar_quantization import ScalarQuantizer
    X = _data(512, 64)
    a = FaissQuantizerAdapter(ProductQuantizer(M=8, B=8))
    a.fit(X)
    assert a.code_bytes() == 512 * 8 + 512 * 4
    s = FaissQuantizerAdapter(ScalarQuantizer(8))
    s.fit(X)
    assert float(np.mean((s.reconstruct(np.arange(512)) - X) ** 2)) < 0.01


def test_product_quantizer_surface_and_parity_with_oracle():
    from haag_vq.methods.product_quantization import ProductQuantizer
    X = make_data(5000, 64, 3)
    pq = ProductQuantizer(M=8, B=8)
    with pytest.raises(RuntimeError):
        pq.compress(X)
    with pytest.raises(AssertionError):
        ProductQuantizer(M=7).fit(X)
    with pytest.raises(ValueError):
        ProductQuantizer(M=8, num_chunks=4)
    assert ProductQuantizer(num_chunks=4, num_clusters=256).B == 8
    pq.fit(X)
    assert len(pq.codebooks) == 8 and pq.codebooks[0].shape == (256, 8) and pq.chunk_dim == 8
    assert pq.pq.d == 64 and pq.pq.ksub == 256 and pq.pq.dsub == 8 and pq.pq.code_size == 8
    C = pq.pq.centroids.reshape(8, 256, 8)
    codes = pq.compress(X)
    assert codes.dtype == np.uint8 and codes.shape == (5000, 8)
    ref, d1, d2 = opq.pq_assign(X, C, return_gap=True)  # parity conditional on shared trained state
    assert_codes_parity(codes, ref, d1, d2, "ProductQuantizer.compress")
    assert np.array_equal(pq.decompress(codes), opq.pq_decode(codes, C))
    assert pq.get_compression_ratio(X) == 64 * 4 / 8
    # trained state is close to the oracle's (same seeds): objective within 1%
    Co = opq.pq_train(X, 8, 8)
    so = opq.pq_sse(X, opq.pq_assign(X, Co), Co)
    sg = opq.pq_sse(X, ref, C)
    assert abs(sg - so) < 0.01 * so


def test_opq_trains_on_rotated():
    from haag_vq.methods.optimized_product_quantization import OptimizedProductQuantizer
    from haag_vq.methods.product_quantization import ProductQuantizer
    rng = np.random.default_rng(0)
    D = 64
    Z = rng.standard_normal((4000, D)).astype(np.float32) * np.linspace(4, 0.2, D, dtype=np.float32)
    X = np.ascontiguousarray((Z @ rng.standard_normal((D, D)).astype(np.float32)) / np.float32(np.sqrt(D)))
    pq = ProductQuantizer(M=8, B=8)
    pq.fit(X)
    mse_pq = float(np.mean((X - pq.decompress(pq.compress(X))) ** 2))
    o = OptimizedProductQuantizer(M=8, B=8)
    o.fit(X)
    codes = o.compress(X)
    assert codes.shape == (4000, 8)
    mse_opq = float(np.mean((X - o.decompress(codes)) ** 2))
    assert mse_opq <= 1.02 * mse_pq
    A = o.opq.A.reshape(D, D)
    assert np.allclose(A @ A.T, np.eye(D), atol=1e-3)
    Y = o.opq.apply(X[:10])
    assert np.allclose(Y, X[:10] @ A.T, rtol=1e-4, atol=1e-5)


def test_exact_search_helpers():
    from haag_vq.benchmarks import exact_search as es
    X = _data(600, 32, 1)
    Q = _data(9, 32, 2)
    norms = es.compute_exact_norms(X)
    assert np.allclose(norms, np.linalg.norm(X, axis=1), rtol=1e-5)
    Z = X.copy()
    Z[3] = 0
    assert es.compute_exact_norms(Z)[3] > 0
    gt = es.normalized_ground_truth(X, Q, k=10)
    assert gt.dtype == np.uint32 and gt.shape == (9, 10)
    cos = Q @ (X / np.linalg.norm(X, axis=1, keepdims=True)).T
    brute = np.argsort(-cos, axis=1)[:, :10]
    for i in range(9):
        assert set(gt[i].tolist()) == set(brute[i].tolist()) and gt[i, 0] == brute[i, 0]
    idx = es.build_scaled_ip_index(lambda ids: X[ids], 600, 32, norms)
    _, ids = es.search_index(idx, Q, 10)
    assert es.recall_at_ks(ids, gt, ks=(1, 10))[10] == 1.0
    assert es.recall_at_ks(np.array([[0, 1, 2]]), np.array([[0, 9, 2]]), ks=(1, 3)) == {1: 1.0, 3: 2 / 3}
    assert es.reconstruction_mse(X, lambda ids: X[ids], np.arange(50)) == 0.0
    assert es.reconstruction_mse(X, lambda ids: X[ids] + 1.0, np.arange(50)) == pytest.approx(1.0, rel=1e-5)


@pytest.mark.parametrize("kind", ["pq", "sq", "erabitq", "opq", "saq"])
@pytest.mark.parametrize("metric", ["l2", "ip"])
def test_flat_quantized_index_contract(kind, metric, tmp_path):
    from haag_vq.methods.search import FlatQuantizedIndex
    from haag_vq.methods.extended_rabitq import ExtendedRaBitQuantizer
    from haag_vq.methods.optimized_product_quantization import OptimizedProductQuantizer
    from haag_vq.methods.product_quantization import ProductQuantizer
    from haag_vq.methods.saq import SAQ
    from haag_vq.methods.scalar_quantization import ScalarQuantizer
    X = _data(2000, 32, 4)
    X /= np.linalg.norm(X, axis=1, keepdims=True)
    Q = _data(13, 32, 5)
    mk = {"pq": lambda: ProductQuantizer(M=8, B=8), "sq": lambda: ScalarQuantizer(8),
          "erabitq": lambda: ExtendedRaBitQuantizer(4), "opq": lambda: OptimizedProductQuantizer(M=4, B=8),
          "saq": lambda: SAQ(total_bits=32 * 4, n_segments=4)}[kind]
    idx = FlatQuantizedIndex(mk())
    idx.fit(X, metric=metric)
    ids, d = idx.search_with_scores(Q, 10)
    assert ids.shape == (13, 10) and ids.dtype == np.uint32 and d.dtype == np.float32
    assert idx.memory_footprint() > 0 and np.isfinite(idx.reconstruction_mse(X))
    mono = np.diff(d, axis=1)
    assert np.all(mono >= 0) if metric == "l2" else np.all(mono <= 0)
    # == the reference algorithm (decompress everything, exact distance, top-k) on the same codes
    Xhat = idx._quantizer.decompress(idx._codes).astype(np.float32)
    _, ref_ids = opq.decompress_search(Xhat, Q, 10, 0 if metric == "l2" else 1)
    agree = np.mean([len(set(ids[i].tolist()) & set(ref_ids[i].tolist())) / 10 for i in range(13)])
    assert agree >= 0.98
    e_ids, e_d = idx.search_with_scores(Q, 0)
    assert e_ids.shape == (13, 0) and e_d.shape == (13, 0)
    p = tmp_path / f"{kind}.pkl"
    idx.save(p)
    st = pickle.load(open(p, "rb"))
    assert set(st) == {"quantizer_state", "codes", "metric", "N", "D"}
    if kind == "pq":
        assert set(st["quantizer_state"]) == {"type", "M", "B", "flat_centroids", "D"}
    idx2 = FlatQuantizedIndex(mk())
    idx2.load(p)
    assert np.array_equal(idx2.search(Q, 10), ids)


def test_scalar_and_saq_quantizer_surface():
    from haag_vq.methods.saq import SAQ
    from haag_vq.methods.scalar_quantization import ScalarQuantizer
    X = _data(500, 33, 7)
    with pytest.raises(ValueError):
        ScalarQuantizer(5)
    for bits, shape, dt in ((4, (500, 17), np.uint8), (8, (500, 33), np.uint8), (16, (500, 33), np.uint16)):
        sq = ScalarQuantizer(bits)
        sq.fit(X)
        assert np.array_equal(sq.min, X.min(0)) and np.array_equal(sq.max, X.max(0))
        c = sq.compress(X)
        assert c.shape == shape and c.dtype == dt
        assert sq.decompress(c).shape == X.shape
    mses = []
    for b in (2, 4, 8):
        s = SAQ(num_bits=b)
        s.fit(X)
        c = s.compress(X)
        assert c.shape == (500, (33 * b + 7) // 8)
        mses.append(float(np.mean((X - s.decompress(c)) ** 2)))
    assert mses[0] > mses[1] > mses[2]
    s = SAQ(total_bits=33 * 4, allowed_bits=(0, 2, 4, 6, 8), n_segments=4)
    s.fit(X)
    # 132 bits round up to 17 bytes per row: ratio = 4*33/17
    assert int(s.bit_offset[-1]) <= 33 * 4 and s.get_compression_ratio(X) >= 4 * 33 / 17 - 1e-9


@pytest.mark.parametrize("B,M", [(4, 8), (6, 5), (1, 16), (7, 3)])
def test_sub_byte_pq_codes_use_the_faiss_bit_stream_layout(B, M):
    """ProductQuantizer(M, B < 8): faiss packs the M indices as a little-endian bit stream of ceil(M B / 8) bytes
    (the reference forwards any B: product_quantization.py:67-73); indices, decode and search agree with the oracle."""
    from haag_vq.methods.product_quantization import ProductQuantizer
    from haag_vq.methods.search.flat_quantized_index import FlatQuantizedIndex
    rng = np.random.default_rng(B * 10 + M)
    D = M * 6
    X = rng.standard_normal((1500, D)).astype(np.float32)
    pq = ProductQuantizer(M=M, B=B)
    pq.fit(X)
    assert pq.num_clusters == 1 << B and pq.pq.code_size == (M * B + 7) // 8
    codes = pq.compress(X)
    assert codes.shape == (1500, (M * B + 7) // 8) and codes.dtype == np.uint8
    C = pq.pq.centroids.reshape(M, 1 << B, D // M)
    ref, d1, d2 = opq.pq_assign(X, C, return_gap=True)
    # unpack the little-endian bit stream on the host: index m = bits [m B, (m + 1) B)
    bits = np.unpackbits(codes, axis=1, bitorder="little")[:, :M * B].reshape(1500, M, B)
    idx = (bits.astype(np.int64) << np.arange(B)).sum(-1)
    bad = (idx != ref) & ((d2 - d1) > 1e-6 * np.maximum(d1, 1e-30))
    assert not bad.any()
    assert np.array_equal(pq.decompress(codes), opq.pq_decode(idx.astype(np.uint8), C))
    assert pq.get_compression_ratio(X) == pytest.approx(D * 4 / ((M * B + 7) // 8))
    index = FlatQuantizedIndex(ProductQuantizer(M=M, B=B))
    index.fit(X, metric="l2")
    ids, d = index.search_with_scores(X[:5], 3)
    assert ids.shape == (5, 3) and np.all(np.diff(d, axis=1) >= 0)
    Xhat = index._quantizer.decompress(index._codes)
    true = ((X[:5, None, :] - Xhat[None]) ** 2).sum(-1)
    assert np.allclose(d[:, 0], true.min(1), rtol=1e-4, atol=1e-5)
