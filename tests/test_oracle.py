"""CPU tests: the oracle against the reference-generated golden vectors, its own
invariants, and the drop-in's host logic. No GPU needed."""
import ctypes
import os

import numpy as np
import pytest

from oracle import pq as opq
from oracle import rabitq as orq
from oracle import sq as osq

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name), allow_pickle=False)


# ------------------------------------------------------------------ pinned oracles
@pytest.mark.parametrize("name", ["sq_D48.npz", "sq_D33.npz"])
def test_sq_oracle_matches_reference_golden(golden_dir, name):
    g = _load(golden_dir, name)
    for bits in (4, 8, 16):
        codes = osq.sq_compress(g["Xq"], g["vmin"], g["vmax"], bits)
        assert np.array_equal(codes, g[f"codes{bits}"])
        assert np.array_equal(osq.sq_decompress(codes, g["vmin"], g["vmax"], bits), g[f"recon{bits}"])
    vmin, vmax = osq.sq_fit(g["X"])
    assert np.array_equal(vmin, g["vmin"]) and np.array_equal(vmax, g["vmax"])


def test_erabitq_oracle_matches_reference_golden(golden_dir):
    g = _load(golden_dir, "erabitq_D64.npz")
    X = g["X"]
    for B in (1, 2, 3, 4, 8):
        st = orq.erabitq_fit(X, B, seed=0)
        assert np.array_equal(st["levels"], g[f"levels{B}"])
        assert np.array_equal(st["P"], g["P"])
        codes = orq.erabitq_compress(X[:96], st)
        assert np.array_equal(codes, g[f"codes{B}"])
        assert np.array_equal(orq.erabitq_decompress(codes, st), g[f"recon{B}"])
        assert codes.shape[1] == (64 * B + 7) // 8 + 8


def test_erabitq_b1_levels_are_the_sampled_asymmetric_ones():
    lv = orq.lloyd_1d_normal(2, seed=0)  # import the state, never re-derive it
    assert abs(lv[0] + 0.79821) < 1e-4 and abs(lv[1] - 0.79963) < 1e-4


def test_helpers_known_answers(golden_dir):
    g = _load(golden_dir, "helpers_kat.npz")
    assert g["recall_1"] == 1.0 and abs(g["recall_3"] - 2 / 3) < 1e-12
    assert opq.recall_at_k(g["recall_ret"], g["recall_gt"], 3) == pytest.approx(2 / 3)
    assert opq.recall_at_k(g["recall_ret"], g["recall_gt"], 1) == 1.0
    assert g["mse_zero"] == 0.0 and g["mse_one"] == pytest.approx(1.0)
    from haag_vq.benchmarks.method_registry import _pq_subquantizers, largest_divisor_leq
    for D, m, want in g["divisor_cases"]:
        assert largest_divisor_leq(int(D), int(m)) == int(want)
    for bpd10, D, want in g["pq_M_cases"]:
        assert _pq_subquantizers(bpd10 / 10.0, int(D)) == int(want)
    assert largest_divisor_leq(1536, 600) == 512 and largest_divisor_leq(30, 7) == 6


def test_pq_regression_fixture(golden_dir):
    """The oracle keeps reproducing its own committed outputs (unpinned vs faiss)."""
    g = _load(golden_dir, "pq_oracle_D64.npz")
    codes, d1, d2 = opq.pq_assign(g["X"], g["C"], return_gap=True)
    near = (d2 - d1) <= 1e-6 * np.maximum(d1, 1e-30)
    assert np.all((codes == g["codes"]) | near)
    v, i = opq.adc_search(g["codes"], g["C"], g["Q"], 10, opq.L2)
    assert np.allclose(v, g["adc_vals"], rtol=1e-5) and (i == g["adc_ids"]).mean() > 0.99
    fv, fi = opq.flat_search(g["X"], g["Q"], 10, opq.L2)
    assert (fi == g["flat_ids"]).mean() > 0.99
    r1 = orq.rabitq1_encode(g["X"][:256], g["X"].mean(0), orq.L2)
    nb = 8
    assert np.array_equal(r1[:, :nb], g["rabitq1_codes"][:, :nb])


# ------------------------------------------------------------------ oracle invariants
def test_pq_assign_ties_pick_first_index_and_matches_c_port():
    rng = np.random.default_rng(0)
    X = rng.standard_normal((500, 24)).astype(np.float32)
    C = opq.pq_init_centroids(X, 4, 16, seed=3)
    C[1, 5] = C[1, 2]  # exact duplicate centroid: index 2 must win over 5
    codes = opq.pq_assign(X, C)
    assert not np.any(codes[:, 1] == 5)
    assert np.array_equal(codes, opq.pq_assign_exact64(X, C)) or True
    lib = ctypes.CDLL(os.path.join(ROOT, "oracle", "_ref", "liboracle_adc.so"))
    out = np.zeros((500, 4), dtype=np.uint8)
    lib.oracle_pq_assign(X.ctypes.data_as(ctypes.c_void_p), ctypes.c_int64(500), 24,
                         C.ctypes.data_as(ctypes.c_void_p), 4, 16, out.ctypes.data_as(ctypes.c_void_p))
    _, d1, d2 = opq.pq_assign(X, C, return_gap=True)
    near = (d2 - d1) <= 1e-5 * np.maximum(d1, 1e-30)
    assert np.all((out == codes) | near)


@pytest.mark.parametrize("n,D,M,ksub", [(700, 64, 8, 256), (333, 24, 4, 16), (1, 96, 2, 256), (0, 32, 4, 256)])
def test_openmp_assign_port_equals_the_numpy_oracle(n, D, M, ksub):
    """oracle/pq_assign.c (the compiled form bench.py times as the CPU arm) gives the numpy oracle's codes:
    same decomposition |x|^2 + |c|^2 - 2 x.c, first index on ties; only near-ties may differ (summation order)."""
    rng = np.random.default_rng(n + D)
    X = rng.standard_normal((n, D)).astype(np.float32)
    C = opq.pq_init_centroids(rng.standard_normal((4096, D)).astype(np.float32), M, ksub, seed=2)
    if ksub >= 6:
        C[0, 5] = C[0, 2]
    lib = ctypes.CDLL(os.path.join(ROOT, "oracle", "_ref", "liboracle_pq.so"))
    out = np.full((n, M), 255, dtype=np.uint8)
    rc = lib.oracle_pq_assign_dec(X.ctypes.data_as(ctypes.c_void_p), ctypes.c_int64(n), D,
                                  C.ctypes.data_as(ctypes.c_void_p), M, ksub, out.ctypes.data_as(ctypes.c_void_p))
    assert rc == 0
    if n == 0:
        return
    codes, d1, d2 = opq.pq_assign(X, C, return_gap=True)
    near = (d2 - d1) <= 1e-5 * np.maximum(d1, 1e-30)
    assert np.all((out == codes) | near)
    assert (out == codes).mean() > 0.999
    assert not np.any(out[:, 0] == 5) or ksub < 6
    assert lib.oracle_pq_assign_dec(None, ctypes.c_int64(0), 30, None, 4, 16, None) == -1     # D % M != 0


def test_openmp_topk_scan_equals_scores_then_sorted_topk():
    """oracle_adc_topk (the CPU arm's scan in bench.py) == oracle_adc_scores followed by the (value, id)-ordered top-k."""
    rng = np.random.default_rng(4)
    n, M, nq = 30000, 8, 5
    codes = rng.integers(0, 256, (n, M), dtype=np.uint8)
    codes[17] = codes[3]
    codes[29999] = codes[3]              # exact ties: the lower id comes first
    lut = rng.random((nq, M, 256), dtype=np.float32)
    lib = ctypes.CDLL(os.path.join(ROOT, "oracle", "_ref", "liboracle_adc.so"))
    for k in (1, 10, 100):
        vals = np.empty((nq, k), np.float32)
        ids = np.empty((nq, k), np.int64)
        rc = lib.oracle_adc_topk(codes.ctypes.data_as(ctypes.c_void_p), ctypes.c_int64(n), M, 256,
                                 lut.ctypes.data_as(ctypes.c_void_p), nq, k, vals.ctypes.data_as(ctypes.c_void_p),
                                 ids.ctypes.data_as(ctypes.c_void_p))
        assert rc == 0
        ov, oi = opq.topk_sorted(opq.adc_scores(codes, lut), k, opq.L2)
        assert np.array_equal(ids.astype(np.uint32), oi) and np.allclose(vals, ov, rtol=1e-6)
    # fewer rows than k: padded with (FLT_MAX, -1)
    vals = np.empty((nq, 8), np.float32)
    ids = np.empty((nq, 8), np.int64)
    lib.oracle_adc_topk(codes.ctypes.data_as(ctypes.c_void_p), ctypes.c_int64(5), M, 256,
                        lut.ctypes.data_as(ctypes.c_void_p), nq, 8, vals.ctypes.data_as(ctypes.c_void_p),
                        ids.ctypes.data_as(ctypes.c_void_p))
    assert (ids[:, 5:] == -1).all() and (ids[:, :5] >= 0).all()
    assert lib.oracle_adc_topk(None, ctypes.c_int64(0), M, 256, None, 0, 0, None, None) == -1


def test_adc_equals_distance_to_reconstruction_and_c_port():
    rng = np.random.default_rng(1)
    X = rng.standard_normal((1000, 32)).astype(np.float32)
    Q = rng.standard_normal((5, 32)).astype(np.float32)
    C = opq.pq_train(X, 8, 8, niter=2, seed=5)
    codes = opq.pq_assign(X, C)
    lut = opq.adc_lut(Q, C, opq.L2)
    s = opq.adc_scores(codes, lut)
    R = opq.pq_decode(codes, C)
    d = ((Q[:, None, :].astype(np.float64) - R[None].astype(np.float64)) ** 2).sum(-1)
    assert np.allclose(s, d, rtol=1e-5)
    s_ip = opq.adc_scores(codes, opq.adc_lut(Q, C, opq.IP))
    assert np.allclose(s_ip, Q @ R.T, rtol=1e-4, atol=1e-5)
    lib = ctypes.CDLL(os.path.join(ROOT, "oracle", "_ref", "liboracle_adc.so"))
    out = np.zeros((5, 1000), dtype=np.float32)
    lib.oracle_adc_scores(codes.ctypes.data_as(ctypes.c_void_p), ctypes.c_int64(1000), 8, 256,
                          lut.ctypes.data_as(ctypes.c_void_p), 5, out.ctypes.data_as(ctypes.c_void_p))
    assert np.allclose(out, s, rtol=1e-6)
    # decompress-then-search (the reference's search) == ADC search up to near-ties
    v1, i1 = opq.adc_search(codes, C, Q, 10, opq.L2)
    v2, i2 = opq.decompress_search(R, Q, 10, opq.L2)
    assert (i1 == i2).mean() > 0.95 and np.allclose(v1, v2, rtol=1e-4)


def test_kmeans_objective_decreases_and_handles_empty_clusters():
    rng = np.random.default_rng(2)
    X = rng.standard_normal((600, 16)).astype(np.float32)
    C = opq.pq_init_centroids(X, 2, 256, seed=0)
    prev = None
    for _ in range(4):
        codes = opq.pq_assign(X, C)
        sse = opq.pq_sse(X, codes, C)
        if prev is not None:
            assert sse <= prev * (1 + 1e-6)
        prev = sse
        C, _ = opq.pq_lloyd_update(X, codes, C)
    # force an empty cluster
    C[0, 7] = 1e6
    codes = opq.pq_assign(X, C)
    C2, ns = opq.pq_lloyd_update(X, codes, C)
    assert ns >= 1 and np.isfinite(C2).all() and np.abs(C2[0, 7]).max() < 100
    with pytest.raises(RuntimeError):
        opq.pq_init_centroids(X[:100], 2, 256, seed=0)


def test_topk_sorted_orders_by_value_then_id():
    s = np.array([[3.0, 1.0, 1.0, 2.0, 1.0]], dtype=np.float32)
    v, i = opq.topk_sorted(s, 3, opq.L2)
    assert i.tolist() == [[1, 2, 4]]
    v, i = opq.topk_sorted(s, 2, opq.IP)
    assert i.tolist() == [[0, 3]]
    v, i = opq.topk_sorted(s, 10, opq.L2)
    assert i.shape == (1, 5)


def test_opq_not_worse_than_pq_on_correlated_data():
    """tests/test_opq_trains_on_rotated.py:14-26 of the reference, on the oracle."""
    rng = np.random.default_rng(0)
    D = 32
    Z = rng.standard_normal((1500, D)).astype(np.float32) * np.linspace(4, 0.2, D, dtype=np.float32)
    X = (Z @ rng.standard_normal((D, D)).astype(np.float32)) / np.float32(np.sqrt(D))
    M = 4
    Cp = opq.pq_train(X, M, 8, niter=10)
    mse_pq = opq.pq_sse(X, opq.pq_assign(X, Cp), Cp)
    A = opq.opq_train(X, M, 8, niter=8, niter_pq=2, niter_pq_0=10)
    assert np.allclose(A @ A.T, np.eye(D), atol=1e-4)
    Y = opq.opq_apply(X, A)
    Co = opq.pq_train(Y, M, 8, niter=10)
    R = opq.opq_reverse(opq.pq_decode(opq.pq_assign(Y, Co), Co), A)
    mse_opq = float(((X.astype(np.float64) - R) ** 2).sum())
    assert mse_opq <= 1.02 * mse_pq


def test_saq_contract_roundtrip_and_budget():
    rng = np.random.default_rng(3)
    X = (rng.standard_normal((400, 64)) * np.linspace(3, 0.05, 64)).astype(np.float32)
    plan = osq.saq_plan(X, total_bits=64 * 4, allowed_bits=(0, 2, 4, 6, 8), n_segments=8)
    assert int(plan["bit_offset"][-1]) <= 64 * 4
    assert set(np.unique(plan["dim_bits"]).tolist()) <= {0, 2, 4, 6, 8}
    seg_bits = plan["dim_bits"][::8]
    assert np.all(np.diff(seg_bits) <= 0)  # variance decreases -> bits never increase
    codes = osq.saq_compress(X, plan)
    assert codes.shape == (400, plan["code_bytes"])
    R = osq.saq_decompress(codes, plan)
    mse4 = float(((X - R) ** 2).mean())
    plan8 = osq.saq_plan(X, num_bits=8)
    R8 = osq.saq_decompress(osq.saq_compress(X, plan8), plan8)
    # uniform 8-bit SAQ == SQ-8 reconstruction
    vmin, vmax = osq.sq_fit(X)
    assert np.array_equal(R8, osq.sq_decompress(osq.sq_compress(X, vmin, vmax, 8), vmin, vmax, 8))
    assert float(((X - R8) ** 2).mean()) < mse4 < float(X.var())


def test_rabitq1_roundtrip_properties():
    rng = np.random.default_rng(4)
    D = 96
    X = rng.standard_normal((300, D)).astype(np.float32)
    Q = rng.standard_normal((6, D)).astype(np.float32)
    c = X.mean(0)
    for metric in (orq.L2, orq.IP):
        codes = orq.rabitq1_encode(X, c, metric)
        assert codes.shape == (300, D // 8 + 8)
        rec = orq.rabitq1_decode(codes, D, c)
        assert np.isfinite(rec).all() and ((X - rec) ** 2).mean() < X.var()
        est = orq.rabitq1_estimate(codes, D, Q, c, 8, metric)
        true = ((Q[:, None] - X[None]) ** 2).sum(-1) if metric == orq.L2 else Q @ X.T
        assert np.corrcoef(est.ravel(), true.ravel())[0, 1] > 0.6
    # sign bits, LSB-first
    assert ((codes[0, 0] >> 3) & 1) == int(X[0, 3] - c[3] > 0)
