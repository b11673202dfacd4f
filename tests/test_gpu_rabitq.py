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


def test_extended_rabitq_at_size_and_scan(T):
    D, B, n = 256, 4, 6000
    X = make_data(n, D, 5)
    Q = make_data(12, D, 6)
    st = orq.erabitq_fit(X, B, seed=0)
    ref = orq.erabitq_compress(X, st)
    from haag_vq.methods.extended_rabitq import ExtendedRaBitQuantizer
    q = ExtendedRaBitQuantizer(num_bits=B, seed=0)
    q.fit(X)
    codes = q.compress(X)
    ib = (D * B + 7) // 8
    mism = (codes[:, :ib] != ref[:, :ib]).any(axis=1).mean()
    assert mism < 1e-3, f"{mism:.2e} of rows differ in level indices"
    Xhat = orq.erabitq_decompress(ref, st)
    assert np.allclose(q.decompress(ref), Xhat, rtol=1e-4, atol=1e-6)
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
        vals, ids = T.ops_sq.erabitq_scan_topk(T.dev(ref), D, B, T.dev(st["levels"].astype(np.float32)),
                                               T.dev(qrot), T.dev(qconst), 10, metric)
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
@pytest.mark.parametrize("B", [2, 4, 8])
def test_erabitq_scan_tensor_core_path_matches_simt(T, B, metric):
    """The estimator scan as an exact flat search over the decoded rows (b200vq_erabitq_decode_f32 +
    b200vq_flat_topk_tc) against the code-domain SIMT kernel: same top-k up to fp32 rounding of the
    differently associated sums (values within 1e-4 relative, ids equal outside near-ties)."""
    import torch
    n, D, nq, k = 70_000, 128, 33, 10
    g = torch.Generator(device="cuda").manual_seed(B * 10 + metric)
    levels = torch.sort(torch.randn(1 << B, generator=g, device="cuda"))[0].contiguous()
    ib = (D * B + 7) // 8
    codes = torch.randint(0, 256, (n, ib + 8), generator=g, device="cuda", dtype=torch.uint8)
    nrm = torch.rand(n, generator=g, device="cuda") + 0.5
    t = torch.rand(n, generator=g, device="cuda") * 0.2 + 0.9
    codes[:, ib:ib + 4] = nrm.view(torch.uint8).reshape(n, 4)
    codes[:, ib + 4:] = t.view(torch.uint8).reshape(n, 4)
    qrot = torch.randn(nq, D, generator=g, device="cuda")
    qconst = (qrot * qrot).sum(1) if metric == 0 else torch.randn(nq, generator=g, device="cuda")
    v0, i0 = T.ops_sq.erabitq_scan_topk(codes, D, B, levels, qrot, qconst, k, metric, impl="simt")
    v1, i1 = T.ops_sq.erabitq_scan_topk(codes, D, B, levels, qrot, qconst, k, metric, impl="tc")
    scale = v0.abs().max().item()
    assert torch.allclose(v0, v1, rtol=1e-4, atol=1e-5 * scale)
    same = (i0 == i1)
    assert same.float().mean().item() > 0.98
    # where the ids differ, the two candidates' scores are within rounding distance
    bad = torch.nonzero(~same)
    for a, b in bad.tolist():
        assert abs(v0[a, b].item() - v1[a, b].item()) <= 1e-4 * scale
