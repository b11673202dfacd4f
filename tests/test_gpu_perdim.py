"""RankAwareQuantizer on the GPU (csrc/perdim.cu, csrc/ffd.cu) against the reference's golden vectors (shared
trained state) and the reference's own property tests (tests/test_rank_aware_quantization.py:13-86)."""
import os

import numpy as np
import pytest

from oracle import perdim as opd

pytestmark = pytest.mark.gpu


def _data(seed=0, n=3000, d=64):
    rng = np.random.default_rng(seed)
    A = np.diag(np.linspace(4.0, 0.2, d)) @ rng.standard_normal((d, d))
    return (rng.standard_normal((n, d)) @ A).astype(np.float32)


def _Q(**kw):
    from haag_vq.methods.rank_aware_quantization import RankAwareQuantizer
    return RankAwareQuantizer(**kw)


def test_codes_match_reference_given_shared_state(golden_dir):
    g = np.load(os.path.join(golden_dir, "perdim_D24.npz"))
    cb, o = [], 0
    for n in g["cb_len"]:
        cb.append(g["cb_flat"][o:o + n])
        o += n
    for packing in ("dense", "ffd"):
        q = _Q(avg_bits=3.5, alpha=0.5, packing=packing)
        q.set_state(g["mu"], g["V"], g["var"], g["bits"], cb)
        assert q.code_size == int(g[f"code_size_{packing}"])
        codes = q.compress(g["X"])
        ref = g[f"codes_{packing}"]
        if not np.array_equal(codes, ref):
            # only coordinates within rounding distance of a codebook mid-point may differ (fp64 GEMM order)
            Y = opd.project(g["X"], g["mu"], g["V"])
            idx_ref = opd.indices(Y, g["bits"], cb)
            rows = np.nonzero((codes != ref).any(axis=1))[0]
            assert len(rows) <= 2
            for r in rows:
                near = False
                for d in range(24):
                    if g["bits"][d]:
                        mid = 0.5 * (cb[d][:-1] + cb[d][1:])
                        near |= np.min(np.abs(mid - Y[r, d])) <= 1e-12 * max(1.0, abs(Y[r, d]))
                assert near
        rec = q.decompress(ref)
        # the inverse projection is the fp32-accurate tensor-core GEMM (the reference: float64, then a cast):
        # reconstructions agree to fp32 rounding of the row, far inside north_star's 1e-4
        ref_rec = g[f"rec_{packing}"]
        assert rec.dtype == np.float32 and np.abs(rec - ref_rec).max() <= 2e-6 * np.abs(ref_rec).max()


def test_fit_and_reference_properties():
    X = _data()
    q = _Q(avg_bits=4, alpha=1.0)
    with pytest.raises(RuntimeError):
        q.compress(X)
    q.fit(X)
    codes = q.compress(X)
    xhat = q.decompress(codes)
    assert xhat.shape == X.shape and np.isfinite(xhat).all() and codes.shape == (3000, q.code_size)
    ids = np.array([0, 3, 50, 2999])
    np.testing.assert_allclose(q.decompress(codes[ids]), xhat[ids], rtol=1e-4, atol=1e-4)
    assert round(4 * 64) - 1 <= q.bits.sum() <= round(4 * 64)
    # own codes == oracle on own state
    assert (codes != opd.compress(X, q.mu, q.V, q.bits, q.cb)).any(axis=1).mean() < 2e-3
    m = {}
    for ab in (2, 4, 6):
        qq = _Q(avg_bits=ab, alpha=1.0)
        qq.fit(X)
        m[ab] = float(np.mean((X - qq.decompress(qq.compress(X))) ** 2))
    assert m[6] < m[4] < m[2]
    q0, q2 = _Q(avg_bits=3, alpha=0.0), _Q(avg_bits=3, alpha=2.0)
    q0.fit(X); q2.fit(X)
    assert q2.bits[:32].sum() > q0.bits[:32].sum()
    with pytest.raises(ValueError):
        q.compress(X[:, :10])
    with pytest.raises(NotImplementedError):
        _Q(avg_bits=4, codebook="exact")


def test_ffd_is_lossless_relayout():
    X = _data()
    qd, qf = _Q(avg_bits=4, alpha=1.0, packing="dense"), _Q(avg_bits=4, alpha=1.0, packing="ffd")
    qd.fit(X); qf.fit(X)
    np.testing.assert_array_equal(qf.bits, qd.bits)
    qf.set_state(qd.mu, qd.V, qd.var, qd.bits, qd.cb)       # identical state (eigh is not bit-reproducible across calls)
    xd, xf = qd.decompress(qd.compress(X)), qf.decompress(qf.compress(X))
    np.testing.assert_array_equal(xf, xd)
    assert qd.code_size <= qf.code_size <= 64
    assert qf.get_compression_ratio(X) == pytest.approx(64 * 4 / qf.code_size)


@pytest.mark.parametrize("packing", ["dense", "ffd"])
@pytest.mark.parametrize("metric", ["l2", "ip"])
def test_adc_scan_on_the_packed_layout_equals_search_on_reconstructions(packing, metric):
    """b200vq_perdim_scan_topk (the ADC loop of bench/ffd_speed.cpp:87-127 on the dense / FFD layout) against the
    distance to the reconstruction, the definition of FlatQuantizedIndex.search_with_scores (:57-68)."""
    from parity import assert_topk_parity
    X = _data(n=5000)
    Q = _data(seed=3, n=13)
    q = _Q(avg_bits=4, alpha=0.5, packing=packing, codebook="lloyd")
    q.fit(X)
    codes = q.compress(X)
    ids, vals = q.adc_search(codes, Q, 10, metric)
    Xhat = q.decompress(codes).astype(np.float64)
    if metric == "l2":
        true = ((Q[:, None, :].astype(np.float64) - Xhat[None]) ** 2).sum(-1)
    else:
        true = Q.astype(np.float64) @ Xhat.T
    scale = float(np.abs(true).max())
    assert_topk_parity(vals, ids, true.astype(np.float32), 10, 0 if metric == "l2" else 1, "perdim_scan", gap_rel=1e-5,
                       val_rel=1e-4, abs_tol=2e-6 * scale)
