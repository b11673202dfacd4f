"""GPU parity tests: extended RaBitQ (vs the reference's golden vectors) and 1-bit RaBitQ."""
import os

import numpy as np
import pytest

from oracle import rabitq as orq
from parity import assert_topk_parity, make_data

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def T():
    import torch
    import b200vq
    from b200vq import ops_sq
    b200vq.load()
    class NS:
        pass
    ns = NS()
    ns.torch, ns.ops_sq = torch, ops_sq
    ns.dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    return ns


def test_extended_rabitq_vs_reference_golden(T, golden_dir):
    from haag_vq.methods.extended_rabitq import ExtendedRaBitQuantizer
    g = np.load(os.path.join(golden_dir, "erabitq_D64.npz"))
    X = g["X"]
    for B in (1, 2, 3, 4, 8):
        q = ExtendedRaBitQuantizer(num_bits=B, seed=0)
        q.fit(X)
        assert np.array_equal(q.levels, g[f"levels{B}"]) and np.array_equal(q.P, g["P"])
        codes = q.compress(X[:96])
        ref = g[f"codes{B}"]
        ib = (64 * B + 7) // 8
        assert codes.shape == ref.shape and codes.dtype == np.uint8
        assert np.array_equal(codes[:, :ib], ref[:, :ib]), f"B={B}: packed level indices differ"
        f_got = codes[:, ib:].copy().view(np.float32)
        f_ref = ref[:, ib:].copy().view(np.float32)
        assert np.allclose(f_got, f_ref, rtol=1e-6), f"B={B}: (nrm, t) trailer"
        rec = q.decompress(ref)
        assert np.allclose(rec, g[f"recon{B}"], rtol=1e-4, atol=1e-6)
        # row-independent decode (tests/test_extended_rabitq.py:23-32 of the reference)
        ids = np.array([3, 77, 0])
        assert np.allclose(q.decompress(ref[ids]), rec[ids], atol=1e-6)


def _unpack_idx(codes, D, B):
    ib = (D * B + 7) // 8
    bits = np.unpackbits(codes[:, :ib], axis=1)[:, :D * B].reshape(codes.shape[0], D, B).astype(np.int64)
    return bits @ (1 << np.arange(B - 1, -1, -1))


def _assert_erabitq_codes_parity(X, st, codes, ref, D, B):
    """north_star gate: level indices bit-exact except where the oracle's fp64 s lies within 1e-6 relative of a
    level mid-point (the 'top-2 gap' of a 1-D nearest-level search); nrm exact, t within 1e-6."""
    gi, ri = _unpack_idx(codes, D, B), _unpack_idx(ref, D, B)
    rows, cols = np.nonzero(gi != ri)
    if rows.size:
        r = X[rows].astype(np.float64) - st["c"]
        o = r / np.maximum(np.linalg.norm(r, axis=1), 1e-12)[:, None]
        s = np.einsum("nd,dn->n", o, st["P"][:, cols]) * np.sqrt(D)
        mid = 0.5 * (st["levels"][:-1] + st["levels"][1:])
        near = np.abs(s[:, None] - mid[None, :]).min(1)
        assert np.all(near <= 1e-6 * np.maximum(1.0, np.abs(s))), \
            f"{int((near > 1e-6 * np.maximum(1.0, np.abs(s))).sum())} level indices differ away from a mid-point"
    ib = (D * B + 7) // 8
    fg, fr = codes[:, ib:].copy().view(np.float32), ref[:, ib:].copy().view(np.float32)
    assert (fg[:, 0] == fr[:, 0]).mean() > 0.999 and np.allclose(fg[:, 0], fr[:, 0], rtol=2e-7)
    assert np.allclose(fg[:, 1], fr[:, 1], rtol=1e-6)
    return int(rows.size)


@pytest.mark.parametrize("D,B,n", [(256, 4, 6000), (1536, 4, 12000), (1536, 1, 3000), (100, 3, 2000), (1536, 8, 2000)])
def test_extended_rabitq_codes_match_fp64_reference_at_size(T, D, B, n):
    """ExtendedRaBitQuantizer.compress with the rotation on the tensor cores + exact fp64 re-evaluation near the
    mid-points, against the oracle's fp64 numpy computation; D = 1536 is config C4's shape, D = 100 the padded-K path."""
    from haag_vq.methods.extended_rabitq import ExtendedRaBitQuantizer
    X = make_data(n, D, 5)
    st = orq.erabitq_fit(X, B, seed=0)
    ref = orq.erabitq_compress(X, st)
    q = ExtendedRaBitQuantizer(num_bits=B, seed=0)
    q.fit(X)
    codes = q.compress(X)
    _assert_erabitq_codes_parity(X, st, codes, ref, D, B)
    nfix = int(q._state()["nfix"].item())
    # 2^B - 1 mid-points, each flags a band of 2 eps ~ 4e-4 in s: the fp64 work stays a small fraction of the rotation
    assert 0 < nfix < 4e-4 * (1 << B) * n * D, f"{nfix} coordinates recomputed in fp64 for {n * D}"
    print(f"erabitq D={D} B={B}: {nfix / (n * D):.2e} of the coordinates re-evaluated in fp64")
    # float64 input takes the same path
    c64 = q.compress(X[:500].astype(np.float64))
    _assert_erabitq_codes_parity(X[:500], st, c64, ref[:500], D, B)
    Xhat = orq.erabitq_decompress(ref, st)
    rec = q.decompress(ref)
    assert np.abs(rec - Xhat).max() <= 1e-4 * np.abs(Xhat).max()
    assert abs(float(((rec - X) ** 2).mean()) / float(((Xhat - X) ** 2).mean()) - 1) < 1e-4


def test_extended_rabitq_at_size_and_scan(T):
    D, B, n = 256, 4, 6000
    X = make_data(n, D, 5)
    Q = make_data(12, D, 6)
    st = orq.erabitq_fit(X, B, seed=0)
    ref = orq.erabitq_compress(X, st)
    Xhat = orq.erabitq_decompress(ref, st)
    # code-domain scan == distance to the reconstruction
    for metric in (0, 1):
        if metric == 0:
            true = ((Q[:, None, :].astype(np.float64) - Xhat[None].astype(np.float64)) ** 2).sum(-1)
        else:
            true = Q.astype(np.float64) @ Xhat.T.astype(np.float64)
        true = true.astype(np.float32)
        qc = Q.astype(np.float64) - st["c"]
        qrot = ((qc if metric == 0 else Q.astype(np.float64)) @ st["P"]).astype(np.float32)
        qconst = ((qc * qc).sum(1) if metric == 0 else (Q.astype(np.float64) @ st["c"])).astype(np.float32)
        for impl in ("simt", "tc"):
            vals, ids = T.ops_sq.erabitq_scan_topk(T.dev(ref), D, B, T.dev(st["levels"].astype(np.float32)),
                                                   T.dev(qrot), T.dev(qconst), 10, metric, impl=impl)
            assert_topk_parity(vals.cpu().numpy(), ids.cpu().numpy().view(np.uint32), true, 10, metric,
                               "erabitq_scan", gap_rel=1e-4)


@pytest.mark.parametrize("D", [1536, 96, 100, 7])
@pytest.mark.parametrize("metric", [0, 1])
def test_rabitq1_encode_decode(T, D, metric):
    X = make_data(700, D, D, normalize=False)
    c = X.mean(0).astype(np.float32)
    for cent in (c, None):
        ref = orq.rabitq1_encode(X, cent, metric)
        got = T.ops_sq.rabitq1_encode(T.dev(X), None if cent is None else T.dev(cent), metric).cpu().numpy()
        nb = (D + 7) // 8
        assert got.shape == ref.shape == (700, nb + 8)
        assert np.array_equal(got[:, :nb], ref[:, :nb]), "sign bits"
        fg, fr = got[:, nb:].copy().view(np.float32), ref[:, nb:].copy().view(np.float32)
        # factor 0 of the IP metric is ||x-c||^2 - ||x||^2: a difference of two fp32 sums, so the
        # 1e-4 relative gate applies to the magnitude of the operands, not to the difference
        omc = X.astype(np.float64) - (0 if cent is None else cent.astype(np.float64))
        mag = (omc * omc).sum(1) + (X.astype(np.float64) ** 2).sum(1) if metric == 1 else (omc * omc).sum(1)
        assert np.all(np.abs(fg[:, 0] - fr[:, 0]) <= 1e-4 * mag + 1e-6), "factor 0"
        assert np.allclose(fg[:, 1], fr[:, 1], rtol=1e-4, atol=1e-6), "factor 1"
        rec = T.ops_sq.rabitq1_decode(T.dev(ref), D, None if cent is None else T.dev(cent)).cpu().numpy()
        assert np.allclose(rec, orq.rabitq1_decode(ref, D, cent), rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("D,qb", [(1536, 4), (96, 8), (100, 4), (64, 1)])
@pytest.mark.parametrize("metric", [0, 1])
def test_rabitq1_estimator_scan(T, D, qb, metric):
    n, nq, k = 9000, 9, 10
    X = make_data(n, D, D + 1)
    Q = make_data(nq, D, D + 2)
    c = X.mean(0).astype(np.float32)
    codes = orq.rabitq1_encode(X, c, metric)
    est = orq.rabitq1_estimate(codes, D, Q, c, qb, metric)
    planes, qfac = T.ops_sq.rabitq1_prepare_queries(T.dev(Q), T.dev(c), qb)
    vals, ids = T.ops_sq.rabitq1_scan_topk(T.dev(codes), D, planes, qfac, qb, k, metric)
    # estimated distances within 1e-4 relative (north_star); ids up to estimator ties
    assert_topk_parity(vals.cpu().numpy(), ids.cpu().numpy().view(np.uint32), est, k, metric, "rabitq1_scan",
                       gap_rel=1e-4)


def test_rabitq_index_contract(T, tmp_path):
    """tests/test_rabitq_index.py of the reference, on the drop-in."""
    from haag_vq.methods.search import RaBitQIndex
    X = make_data(3000, 64, 9)
    Q = make_data(11, 64, 10)
    idx = RaBitQIndex(qb=4)
    with pytest.raises(RuntimeError):
        idx.search(Q, 5)
    idx.fit(X, metric='l2')
    ids, d = idx.search_with_scores(Q, 10)
    assert ids.shape == (11, 10) and ids.dtype == np.uint32 and d.dtype == np.float32
    assert np.all(np.diff(d, axis=1) >= 0)
    assert idx.memory_footprint() == 3000 * (8 + 8)
    mse = idx.reconstruction_mse(X, np.arange(100))
    assert np.isfinite(mse) and mse < X.var()
    p = tmp_path / "r.idx"
    idx.save(p)
    idx2 = RaBitQIndex()
    idx2.load(p)
    assert idx2._qb == 4 and np.array_equal(idx2.search(Q, 10), ids)
    # recall against exact search is well above chance
    true = ((Q[:, None] - X[None]) ** 2).sum(-1)
    gt = np.argsort(true, axis=1)[:, :10]
    rec = np.mean([len(set(gt[i]) & set(ids[i])) / 10 for i in range(11)])
    assert rec > 0.3


@pytest.mark.parametrize("metric", [0, 1])
def test_rabitq1_scan_any_k_by_chained_pages(T, metric):
    """k > 128 (faiss IndexRaBitQ.search takes any k): pages chained by b200vq_rabitq1_scan_topk_after equal the
    oracle's full ranking; beyond ntotal the ids are 0xffffffff."""
    from haag_vq.methods.search import RaBitQIndex
    n, nq, D, qb, k = 700, 7, 96, 4, 300
    X = make_data(n, D, 31)
    X[100:140] = X[50]                      # 41 identical rows: ties are ordered by id across a page boundary
    Q = make_data(nq, D, 32)
    c = X.mean(0).astype(np.float32)
    codes = orq.rabitq1_encode(X, c, metric)
    est = orq.rabitq1_estimate(codes, D, Q, c, qb, metric)
    planes, qfac = T.ops_sq.rabitq1_prepare_queries(T.dev(Q), T.dev(c), qb)
    vals, ids = T.ops_sq.rabitq1_scan_topk(T.dev(codes), D, planes, qfac, qb, k, metric)
    v, i = vals.cpu().numpy(), ids.cpu().numpy().view(np.uint32)
    assert_topk_parity(v, i, est, k, metric, "rabitq1_scan_pages", gap_rel=1e-4)
    assert all(len(set(r)) == k for r in i)
    key = -v if metric == 1 else v
    for r in range(nq):
        order = np.lexsort((i[r], key[r]))
        assert np.array_equal(order, np.arange(k)), "pages must continue the (estimate, id) order"
    # the first page is the single-launch result, bit for bit
    v0, i0 = T.ops_sq.rabitq1_scan_topk(T.dev(codes), D, planes, qfac, qb, 128, metric)
    assert np.array_equal(i[:, :128], i0.cpu().numpy().view(np.uint32)) and np.array_equal(v[:, :128], v0.cpu().numpy())
    # k above ntotal through the index: all rows once, then sentinels
    idx = RaBitQIndex(qb=qb)
    idx.fit(X[:200], metric='l2' if metric == 0 else 'ip')
    ids2, d2 = idx.search_with_scores(Q, 260)
    assert ids2.shape == (nq, 260)
    for r in range(nq):
        assert sorted(ids2[r, :200].tolist()) == list(range(200))
        assert (ids2[r, 200:] == np.iinfo(np.uint32).max).all() and np.isinf(d2[r, 200:]).all()
    with pytest.raises(ValueError):
        idx.search(Q, 0)
