"""Code-domain estimator scan on the tensor cores (csrc/est_tc.cu, SURVEY.md §8 row a23 and the extended-RaBitQ
scan of config C4) through the C ABI: against a numpy restatement of the estimator formula
(external/saq/include/saq/caq_estimator.h:254-316, saq_estimator.h:183-190) and, bit for bit, against the
code-domain SIMT kernel b200vq_erabitq_scan_topk."""
import numpy as np
import pytest

from parity import assert_topk_parity

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def T():
    import torch
    import b200vq
    from b200vq import ops_est, ops_sq
    b200vq.load()

    class NS:
        pass
    ns = NS()
    ns.torch, ns.ops_est, ns.ops_sq = torch, ops_est, ops_sq
    ns.dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    return ns


def _pack_msb(codes: np.ndarray, bits: int) -> np.ndarray:
    """(n, len) integer codes -> (n, len * bits / 8) MSB-first bit string."""
    n, ln = codes.shape
    b = ((codes[:, :, None].astype(np.int64) >> np.arange(bits - 1, -1, -1)) & 1).astype(np.uint8)
    return np.packbits(b.reshape(n, ln * bits), axis=1)


def _segments(T, rng, n, plan, table_bits=None):
    """Random multi-segment rows. Returns the device CodeSegments and the decoded virtual rows (fp64)."""
    lens, bits, codes_d, scales_d, tabs_d, bases, cols = [], [], [], [], [], [], []
    for ln, b in plan:
        c = rng.integers(0, 1 << b, size=(n, ln))
        rs = (rng.random(n) * 0.5 + 0.25).astype(np.float32)
        lens.append(ln)
        bits.append(b)
        codes_d.append(T.dev(_pack_msb(c, b)))
        scales_d.append(T.dev(rs))
        if table_bits is not None and b == table_bits:
            tab = np.sort(rng.standard_normal(1 << b)).astype(np.float32)
            tabs_d.append(T.dev(tab))
            bases.append(0.0)
            val = tab[c].astype(np.float64)
        else:
            tabs_d.append(None)
            bases.append(-((1 << b) - 1) / 2.0)
            val = c.astype(np.float64) + bases[-1]
        cols.append(val * rs[:, None].astype(np.float64))
    segs = T.ops_est.CodeSegments(lens, bits, codes_d, scales_d, tabs_d, bases)
    return segs, np.concatenate(cols, axis=1)


@pytest.mark.parametrize("n", [3000, 50_000])
@pytest.mark.parametrize("metric", [0, 1])
def test_multi_segment_estimator_vs_numpy(T, n, metric):
    """SAQ-engine shaped rows: segments of 6 / 4 / 2 / 9 bits, value = code - (2^b - 1) / 2, per-(row, segment)
    scale. value(q, i) = C_q + B_i - 2 <x_i, q> (L2) or C_q + <x_i, q> (IP)."""
    rng = np.random.default_rng(n + metric)
    plan = [(128, 6), (64, 4), (64, 2), (32, 9)]
    segs, Xv = _segments(T, rng, n, plan)
    D = Xv.shape[1]
    nq, k = 37, 10
    Q = (rng.standard_normal((nq, D)) * 0.1).astype(np.float32)
    bias = (rng.random(n) * 3).astype(np.float32)
    cq = (rng.random(nq) * 2).astype(np.float32)
    norm2 = segs.row_norms()
    assert np.allclose(norm2.cpu().numpy(), (Xv * Xv).sum(1), rtol=1e-5)
    ip = Q.astype(np.float64) @ Xv.T
    true = (cq[:, None] + bias[None, :] - 2 * ip) if metric == 0 else (cq[:, None] + ip)
    vals, ids = segs.scan_topk(norm2, T.dev(bias), T.dev(Q), T.dev(cq), k, metric)
    mag = float(np.abs(ip).max() + np.abs(bias).max() + np.abs(cq).max())
    assert_topk_parity(vals.cpu().numpy(), ids.cpu().numpy().view(np.uint32), true.astype(np.float32), k, metric,
                       "est_scan", gap_rel=1e-5, abs_tol=2e-6 * mag)


@pytest.mark.parametrize("B,D,n", [(4, 128, 70_000), (2, 256, 40_000), (8, 64, 20_000), (1, 128, 5_000), (3, 96, 33_000),
                                   (4, 1536, 17_000)])
@pytest.mark.parametrize("metric", [0, 1])
def test_erabitq_scan_tc_equals_simt_bit_for_bit(T, B, D, n, metric):
    """One segment with a level table = extended RaBitQ: the tensor-core scan returns the SIMT kernel's ids and
    value BITS (same canonical arithmetic: one fp32 FMA chain over the dimensions)."""
    torch = T.torch
    nq, k = 33, 10
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
    assert torch.equal(i0, i1)
    assert torch.equal(v0.view(torch.int32), v1.view(torch.int32))


def test_survivor_list_overflow_is_redone_exactly(T):
    """Rows that all tie (identical codes and scale): every row passes the screen, the lists overflow, and the
    gated exact pass must return the (value, id)-ordered answer: ids 0 .. k-1."""
    torch = T.torch
    n, D, B, nq, k = 60_000, 64, 4, 5, 10
    g = torch.Generator(device="cuda").manual_seed(1)
    levels = torch.sort(torch.randn(16, generator=g, device="cuda"))[0].contiguous()
    ib = D * B // 8
    row = torch.randint(0, 256, (1, ib + 8), generator=g, device="cuda", dtype=torch.uint8)
    row[:, ib:ib + 4] = torch.tensor([1.25], device="cuda").view(torch.uint8)
    row[:, ib + 4:] = torch.tensor([0.75], device="cuda").view(torch.uint8)
    codes = row.repeat(n, 1).contiguous()
    qrot = torch.randn(nq, D, generator=g, device="cuda")
    qconst = (qrot * qrot).sum(1)
    v, i = T.ops_sq.erabitq_scan_topk(codes, D, B, levels, qrot, qconst, k, 0, impl="tc")
    assert torch.equal(i.cpu(), torch.arange(k, dtype=torch.int32).repeat(nq, 1))
    assert torch.equal(v[:, :1].expand(-1, k), v)


def test_est_scan_rejects_bad_layouts(T):
    torch = T.torch
    import b200vq
    codes = torch.zeros((100, 12), dtype=torch.uint8, device="cuda")
    with pytest.raises(b200vq.B200VQError):   # 24 dimensions: not a multiple of 32
        T.ops_est.CodeSegments([24], [4], [codes], [None], [None], [0.0]).row_norms()
    segs = T.ops_est.CodeSegments([32], [3], [codes], [None], [None], [-3.5])
    n2 = segs.row_norms()
    with pytest.raises(b200vq.B200VQError):
        segs.scan_topk(n2, n2, torch.zeros((2, 32), device="cuda"), None, 200, 0)     # k > 128
    v, i = segs.scan_topk(n2, n2, torch.zeros((3, 32), device="cuda"), None, 4, 0)
    assert v.shape == (3, 4) and torch.equal(i[0].cpu(), torch.arange(4, dtype=torch.int32))
