"""CPU tests of the SAQ-engine oracle (oracle/caq.c, oracle/caq.py) and of the host-side plan DP that
ships in libb200vq.so (a host function: no device work, so it runs without a GPU).

The engine has no golden codes in the reference tree; what its tests pin are properties
(tests/test_saq_engine_adapter.py:14-35: shape, dtype, MSE(8 bit) < MSE(2 bit); the engine's own
saq_quantizer_test.cpp metric ranges). Those are asserted here on the oracle, together with
C-vs-pure-Python agreement of the two restatements of each algorithm."""
import numpy as np
import pytest

from oracle import caq


def _rows(n, ln, seed):
    rng = np.random.default_rng(seed)
    return (rng.standard_normal((n, ln)) * np.linspace(2.0, 0.1, ln)).astype(np.float32)


@pytest.mark.parametrize("bits", [1, 2, 3, 4, 8, 13])
def test_caq_c_and_python_twins_agree(bits):
    o = _rows(12, 64, 100 + bits)
    enc = caq.caq_encode(o, bits)
    for i in range(o.shape[0]):
        code, ip, oa2 = caq.caq_encode_row_py(o[i], bits)
        assert list(enc["codes"][i]) == code
        cos = ip / np.sqrt(float(np.float32((o[i] * o[i]).cumsum(dtype=np.float32)[-1]))) / np.sqrt(oa2)
        if bits > 1:
            assert abs(cos - enc["cos"][i]) < 1e-6


def test_caq_adjustment_improves_the_cosine_and_codes_stay_in_range():
    o = _rows(200, 128, 7)
    for bits in (1, 2, 4, 6):
        adj = caq.caq_encode(o, bits, rd_lmt=6)
        raw = caq.caq_encode(o, bits, rd_lmt=0)
        assert adj["codes"].max() <= (1 << bits) - 1
        # cosine between o and the half-step reconstruction, recomputed independently in fp64
        def cosine(codes):
            vmx = np.abs(o).max(1).astype(np.float64)
            delta = 2 * vmx / (1 << bits)
            oa = (codes.astype(np.float64) + 0.5) * delta[:, None] - vmx[:, None]
            return (oa * o).sum(1) / np.linalg.norm(oa, axis=1) / np.linalg.norm(o.astype(np.float64), axis=1)
        assert np.all(cosine(adj["codes"]) >= cosine(raw["codes"]) - 1e-12)
        if bits >= 2:   # at 1 bit the sign code is already the optimum
            assert cosine(adj["codes"]).mean() > cosine(raw["codes"]).mean()


def test_caq_factors_and_edge_rows():
    o = _rows(5, 64, 3)
    o[1] = 0.0                     # all-zero row: delta = 0, codes 0, rescale 0
    o[2] = 0.25                    # constant row
    o[3, :] = 0.0
    o[3, 5] = -3.0                 # one-hot row
    enc = caq.caq_encode(o, 4)
    assert np.all(enc["codes"][1] == 0) and enc["rescale"][1] == 0.0 and enc["l2norm"][1] == 0.0
    assert np.allclose(enc["l2norm"], np.linalg.norm(o, axis=1), rtol=1e-6)
    # rescale * <q, oa_unit> reproduces <q, o> for q = o:  rescale = |o|^2 / <o, oa_unit>
    delta = 2.0 / 16
    oa_unit = (enc["codes"].astype(np.float64) + 0.5) * delta - 1.0
    ip = (oa_unit * o).sum(1)
    ok = ip != 0
    assert np.allclose(enc["rescale"][ok] * ip[ok], (o.astype(np.float64) ** 2).sum(1)[ok], rtol=1e-5)
    z = caq.caq_encode(o, 0)
    assert np.all(z["codes"] == 0) and np.all(z["rescale"] == 0) and np.all(z["error"] == 0)
    assert np.allclose(z["mse"], (o ** 2).mean(1), rtol=1e-6)


@pytest.mark.parametrize("avg_bits", [0.5, 1.0, 2.0, 4.0, 8.0])
def test_plan_dp_oracle_twins_and_product_agree(avg_bits):
    import b200vq  # noqa: F401  (host function of the product library; no device needed)
    from b200vq import saq_engine
    rng = np.random.default_rng(11)
    for d_pad in (64, 192, 320):
        var = (np.sort(rng.random(d_pad))[::-1] * np.linspace(3, 0.01, d_pad)).astype(np.float32)
        a = caq.plan_dp(var, avg_bits)
        b = caq.plan_dp_py(var, avg_bits)
        c = saq_engine.plan_dp(var, avg_bits)
        assert a == b == c, (d_pad, avg_bits, a, b, c)
        assert sum(ln for ln, _ in a) == d_pad
        used = sum(ln * bits + (64 if bits else 0) for ln, bits in a)
        assert used <= int(np.float32(avg_bits) * np.float32(d_pad) + np.float32(64))
        assert all(ln % 64 == 0 and 0 <= bits <= 13 for ln, bits in a)
        # high-variance dimensions never get fewer bits than low-variance ones
        assert [bits for _, bits in a] == sorted([bits for _, bits in a], reverse=True)


def test_plan_dp_flat_variance_is_uniform_and_equal_segmentation():
    var = np.ones(256, np.float32)
    plan = caq.plan_dp(var, 4.0)
    assert sum(ln for ln, _ in plan) == 256
    # budget 4*256 + 64 = 1088: one segment of 256 dims at 4 bits uses 1088 exactly
    assert plan == [(256, 4)]
    assert caq.equal_segmentation(256, 4.0, 4) == [(64, 4)] * 4
    assert caq.equal_segmentation(320, 3.0, 2) == [(192, 3), (128, 3)]
    from b200vq import saq_engine
    assert saq_engine.equal_segmentation(320, 3.0, 2) == [(192, 3), (128, 3)]


def test_mt19937_known_answer_and_centroid_row():
    # std::mt19937 default-seeded 5489: first output 3499211612 (ISO C++ [rand.predef] gives the
    # 10000th; the first is the widely published value)
    assert caq.mt19937_first(5489) == 3499211612
    from b200vq import saq_engine
    for seed in (0, 1, 42):
        assert saq_engine.mt19937_first(seed) == caq.mt19937_first(seed)
        assert saq_engine.kmeans1_row(100000, seed) == caq.kmeans1_row(100000, seed)
        assert 0 <= caq.kmeans1_row(100000, seed) < 100000


def test_engine_oracle_fit_decompress_properties():
    """tests/test_saq_engine_adapter.py:14-35 on the oracle: shape/dtype, MSE(8) < MSE(2)."""
    rng = np.random.default_rng(0)
    X = (rng.standard_normal((600, 128)) * np.linspace(3, 0.2, 128)).astype(np.float32)
    mses = {}
    for ab in (2.0, 8.0):
        st = caq.fit(X, ab)
        ids = np.arange(0, 600, 7)
        R = caq.decompress(st, ids)
        assert R.shape == (ids.size, 128) and R.dtype == np.float32
        mses[ab] = float(((R - X[ids]) ** 2).mean())
        # the accurate estimator tracks the true squared distance
        Q = X[:5] + 0.1 * rng.standard_normal((5, 128)).astype(np.float32)
        est = caq.accurate_dist(st, Q, ids)
        true = ((Q[:, None, :] - X[ids][None]) ** 2).sum(-1)
        rel = np.abs(est - true) / np.maximum(true, 1e-6)
        assert np.median(rel) < (0.25 if ab == 2.0 else 0.01)
    assert mses[8.0] < mses[2.0]
