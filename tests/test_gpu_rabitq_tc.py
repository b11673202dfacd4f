"""Tensor-core 1-bit RaBitQ scan (csrc/rabitq_tc.cu) against the SIMT AND+popcount kernel it replaces
(faiss IndexRaBitQ.search at methods/search/rabitq_index.py:62-70): ids and value bits must be identical."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _setup(n, D, nq, qb, metric, seed=0, cluster=False):
    from b200vq import ops_sq
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = torch.randn((n, D), generator=g, device="cuda") * torch.linspace(2.0, 0.3, D, device="cuda")
    if cluster:   # many near-identical rows: long survivor lists
        x = x[:64].repeat(n // 64 + 1, 1)[:n] + 1e-3 * torch.randn((n, D), generator=g, device="cuda")
    q = torch.randn((nq, D), generator=g, device="cuda") * torch.linspace(2.0, 0.3, D, device="cuda")
    cen = x.mean(0).contiguous()
    codes = ops_sq.rabitq1_encode(x, cen, metric)
    planes, qfac = ops_sq.rabitq1_prepare_queries(q, cen, qb)
    return ops_sq, codes, planes, qfac


@pytest.mark.parametrize("n,D,nq,qb,k,metric", [
    (100_000, 1536, 300, 4, 10, "l2"),
    (100_000, 1536, 300, 4, 10, "ip"),
    (70_001, 192, 1000, 4, 100, "l2"),      # ragged last tile, D % 128 == 64, several query tiles
    (50_000, 1024, 7, 1, 1, "l2"),
    (40_000, 128, 257, 3, 33, "ip"),
])
def test_tc_scan_equals_simt_scan(n, D, nq, qb, k, metric):
    ops_sq, codes, planes, qfac = _setup(n, D, nq, qb, metric)
    v_tc, i_tc = ops_sq.rabitq1_scan_topk(codes, D, planes, qfac, qb, k, metric, impl="tc")
    v_s, i_s = ops_sq.rabitq1_scan_topk(codes, D, planes, qfac, qb, k, metric, impl="simt")
    assert torch.equal(i_tc, i_s)
    assert torch.equal(v_tc.view(torch.int32), v_s.view(torch.int32))


def test_tc_scan_overflowing_lists_fall_back_exactly():
    n, D, nq, qb, k = 80_000, 256, 64, 4, 10
    ops_sq, codes, planes, qfac = _setup(n, D, nq, qb, "l2", seed=3, cluster=True)
    v_tc, i_tc = ops_sq.rabitq1_scan_topk(codes, D, planes, qfac, qb, k, "l2", impl="tc")
    v_s, i_s = ops_sq.rabitq1_scan_topk(codes, D, planes, qfac, qb, k, "l2", impl="simt")
    assert torch.equal(v_tc.view(torch.int32), v_s.view(torch.int32))
    assert torch.equal(i_tc, i_s)


def test_tc_scan_uncovered_shapes_route_to_simt():
    # qb = 8 (values beyond E4M3's exact integers), D % 64 != 0, small n: same call, SIMT arithmetic
    for n, D, qb in ((40_000, 256, 8), (40_000, 200, 4), (500, 256, 4)):
        ops_sq, codes, planes, qfac = _setup(n, D, 20, qb, "l2", seed=5)
        a = ops_sq.rabitq1_scan_topk(codes, D, planes, qfac, qb, 5, "l2", impl="tc")
        b = ops_sq.rabitq1_scan_topk(codes, D, planes, qfac, qb, 5, "l2", impl="simt")
        assert torch.equal(a[1], b[1]) and torch.equal(a[0], b[0])
