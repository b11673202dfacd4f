"""Screened ADC scan (b200vq_adc_scan_topk_screen) == exact scan, through the C ABI.

The screen is a rigorous filter, so the ids must be IDENTICAL to the exact kernels' and the values
equal up to the fp32 summation order (the scan-layout kernel adds the table entries in a
lane-rotated order, the screen's exact finish in increasing sub-quantizer order)."""
import numpy as np
import pytest

from oracle import pq as opq

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def T():
    import torch
    import b200vq
    from b200vq import ops
    b200vq.load()

    class NS:
        pass
    ns = NS()
    ns.torch, ns.ops = torch, ops
    return ns


def _case(T, n, M, nq, seed, lut_kind="uniform"):
    g = T.torch.Generator(device="cuda").manual_seed(seed)
    codes = T.torch.randint(0, 256, (n, M), dtype=T.torch.uint8, device="cuda", generator=g)
    if lut_kind == "uniform":
        lut = T.torch.rand((nq, M, 256), device="cuda", generator=g)
    else:   # signed, wide dynamic range per sub-quantizer (inner-product-like tables)
        lut = T.torch.randn((nq, M, 256), device="cuda", generator=g) * T.torch.logspace(-2, 1, M, device="cuda")[None, :, None]
    return codes, lut.contiguous()


@pytest.mark.parametrize("n,M,nq,k,metric,kind", [
    (600_001, 32, 10, 10, "l2", "uniform"),
    (600_001, 32, 10, 10, "ip", "signed"),
    (540_000, 64, 5, 32, "l2", "signed"),
    (524_288, 96, 4, 1, "ip", "uniform"),
    (700_000, 32, 7, 64, "l2", "uniform"),
])
def test_screen_equals_exact_scan(T, n, M, nq, k, metric, kind):
    codes, lut = _case(T, n, M, nq, seed=n % 1000 + M + k, lut_kind=kind)
    rot = T.ops.pq_codes_to_scan_layout(codes)
    v0, i0 = T.ops.adc_scan_topk(rot, lut, k, metric, n=n, rot=True)
    v1, i1 = T.ops.adc_scan_topk(rot, lut, k, metric, n=n, rot=True, screen=True)
    assert T.torch.equal(i0, i1)
    assert T.torch.allclose(v0, v1, rtol=2e-6, atol=1e-6)
    # and against the oracle on a few queries
    sc = opq.adc_scores(codes.cpu().numpy(), lut[:2].cpu().numpy())
    ov, oi = opq.topk_sorted(sc, k, opq.L2 if metric == "l2" else opq.IP)
    assert np.array_equal(i1[:2].cpu().numpy().view(np.uint32), oi)


def test_screen_planted_neighbours_and_small_inputs_fall_back(T):
    n, M, nq, k = 650_000, 32, 8, 10
    codes, lut = _case(T, n, M, nq, seed=3)
    # plant rows that hit every table minimum (score far below everything else), inside and outside the sample
    amin = lut.argmin(dim=2).to(T.torch.uint8)          # (nq, M)
    for q in range(nq):
        codes[100 + q] = amin[q]
        codes[n - 1 - q] = amin[q]
    rot = T.ops.pq_codes_to_scan_layout(codes)
    v0, i0 = T.ops.adc_scan_topk(rot, lut, k, "l2", n=n, rot=True)
    v1, i1 = T.ops.adc_scan_topk(rot, lut, k, "l2", n=n, rot=True, screen=True)
    assert T.torch.equal(i0, i1)
    assert all(int(i1[q, 0]) == 100 + q and int(i1[q, 1]) == n - 1 - q for q in range(nq))
    small = rot[:32 * 1024 * M]
    a = T.ops.adc_scan_topk(small, lut, k, "l2", n=32 * 1024, rot=True)
    b = T.ops.adc_scan_topk(small, lut, k, "l2", n=32 * 1024, rot=True, screen=True)
    assert T.torch.equal(a[1], b[1]) and T.torch.equal(a[0], b[0])


def test_screen_overflow_goes_through_the_exact_fallback(T):
    """Every row identical: every row survives the screen, the per-query lists overflow and the gated
    exact kernels produce the result (ties broken by the lower id)."""
    n, M, nq, k = 600_000, 32, 6, 10
    codes = T.torch.full((n, M), 7, dtype=T.torch.uint8, device="cuda")
    codes[::3, 5] = 9                      # two distinct scores only
    _, lut = _case(T, 1, M, nq, seed=5)
    rot = T.ops.pq_codes_to_scan_layout(codes)
    v0, i0 = T.ops.adc_scan_topk(rot, lut, k, "l2", n=n, rot=True)
    v1, i1 = T.ops.adc_scan_topk(rot, lut, k, "l2", n=n, rot=True, screen=True)
    assert T.torch.equal(i0, i1) and T.torch.allclose(v0, v1, rtol=2e-6)


@pytest.mark.parametrize("M,n", [(16, 100_000), (80, 70_000), (48, 66_000), (16, 5_000), (16, 1_000_000)])
def test_padded_scan_layout_matches_row_layout(M, n):
    """M not a multiple of 32: M = 16 (BASELINE config 1) in its own half-group layout, other M padded to the scan
    layout's group width with all-zero tables, give the row-layout exact kernel's ids and value bits; n >= 65536 takes the
    screened path (threshold sample = n / 8 rows), smaller n the exact scan-layout kernel."""
    import torch
    from b200vq import ops, ops_sq, pq as dpq
    dev = torch.device("cuda")
    D = M * 8
    x = ops_sq.synth_rows(n, D, 0, 77, None, True, device=dev)
    q = ops_sq.synth_rows(37, D, 1 << 30, 78, None, True, device=dev)
    C = dpq.pq_train(x[:20000], M, 8, niter=3, seed=5)
    codec = dpq.PQCodec(C)
    codes = codec.encode(x)
    crot = ops.pq_codes_to_scan_layout(codec.pad_codes_for_scan(codes))
    mp = dpq.scan_layout_width(M)        # 16 -> the half-group layout (no padding), else the next multiple of 32
    assert mp == (16 if M == 16 else (M + 31) // 32 * 32) and crot.numel() == (n + 31) // 32 * 32 * mp
    for metric in ("l2", "ip"):
        v0, i0 = codec.search(codes, q, 10, metric)
        v1, i1 = codec.search(crot, q, 10, metric, n=n, rot=True, m_pad=mp)
        assert torch.equal(i0, i1)
        if n >= 65536:   # screened path: exact finish in the row-layout kernel's summation order
            assert torch.equal(v0.view(torch.int32), v1.view(torch.int32))
        else:            # exact scan-layout kernel: rotated summation order, last-bit differences
            assert torch.allclose(v0, v1, rtol=1e-6, atol=0)


def test_half_group_layout_m16_overflow_and_small_k(T):
    """M = 16 half-group kernels: the exact scan-layout kernel, the screen, its finish and the gated fallback."""
    torch = T.torch
    n, M, nq = 300_000, 16, 9
    codes, lut = _case(T, n, M, nq, seed=11)
    rot = T.ops.pq_codes_to_scan_layout(codes)
    assert rot.numel() == (n + 31) // 32 * 32 * 16
    for k in (1, 10, 64):
        v0, i0 = T.ops.adc_scan_topk(codes, lut, k, "l2")
        v1, i1 = T.ops.adc_scan_topk(rot, lut, k, "l2", n=n, rot=True)
        v2, i2 = T.ops.adc_scan_topk(rot, lut, k, "l2", n=n, rot=True, screen=True)
        assert torch.equal(i0, i1) and torch.allclose(v0, v1, rtol=1e-6, atol=0)
        assert torch.equal(i0, i2) and torch.equal(v0.view(torch.int32), v2.view(torch.int32))
    # every row identical but for one column: all rows survive, the lists overflow, the gated exact kernels answer
    same = torch.full((n, M), 3, dtype=torch.uint8, device="cuda")
    same[::5, 2] = 200
    rot = T.ops.pq_codes_to_scan_layout(same)
    # (reference = the exact scan-layout kernel: its lane-rotated summation order splits exact ties of identical
    # rows by an ulp, so the row-layout kernel's tie order is not comparable here; values are)
    v0, i0 = T.ops.adc_scan_topk(rot, lut, 10, "ip", n=n, rot=True)
    v1, i1 = T.ops.adc_scan_topk(rot, lut, 10, "ip", n=n, rot=True, screen=True)
    assert torch.equal(i0, i1) and torch.equal(v0, v1)
    vr, _ = T.ops.adc_scan_topk(same, lut, 10, "ip")
    assert torch.allclose(vr, v1, rtol=2e-6)


@pytest.mark.parametrize("n,M,k,metric,kind", [
    (100_000, 32, 10, "l2", "uniform"),       # small shard: radix-select threshold sample (adc_sample_kth_kernel<1, 32>)
    (262_144, 32, 64, "ip", "signed"),        # large k: k n / sample would crowd the lists - general sample path
    (262_144, 32, 10, "ip", "signed"),        # 320 expected survivors: general sample path
    (200_000, 32, 8, "ip", "signed"),         # the small-sample path near its limit
    (262_145, 32, 10, "l2", "uniform"),       # one row more: the general sample path
    (70_000, 16, 33, "ip", "signed"),         # half group, k > 32 (CAP = 256 instantiations)
    (65_536, 64, 5, "l2", "signed"),          # M = 64 keeps the general sample at every size
])
def test_small_shard_threshold_sample_paths(T, n, M, k, metric, kind):
    """Shards of at most 262 144 rows take their threshold from adc_sample_kth_kernel (M = 16 / 32); results must
    stay those of the exact kernels on both sides of the switch."""
    torch = T.torch
    codes, lut = _case(T, n, M, 11, seed=n % 97 + M, lut_kind=kind)
    rot = T.ops.pq_codes_to_scan_layout(codes)
    v0, i0 = T.ops.adc_scan_topk(codes, lut, k, metric)
    v1, i1 = T.ops.adc_scan_topk(rot, lut, k, metric, n=n, rot=True, screen=True)
    assert torch.equal(i0, i1)
    # values: the row-layout kernel and the scan-layout kernels add the M table entries in different orders (and a
    # query whose list overflows comes from the gated exact scan-layout kernel): equal to fp32 summation accuracy
    assert torch.allclose(v0, v1, rtol=2e-6, atol=1e-6)
    if kind == "uniform" and k <= 33:
        assert torch.equal(v0.view(torch.int32), v1.view(torch.int32))
    assert (v1[:, 1:] >= v1[:, :-1]).all() if metric == "l2" else (v1[:, 1:] <= v1[:, :-1]).all()


def test_adc_lut_unaligned_output_takes_the_scalar_kernel(T):
    """The 4 x 4-tile LUT kernel stores 128-bit words: an output buffer that is not 16-byte aligned goes through
    adc_lut_kernel; same bits either way."""
    torch = T.torch
    g = torch.Generator(device="cuda").manual_seed(5)
    M, dsub, nq = 8, 24, 37
    C = torch.randn((M, 256, dsub), device="cuda", generator=g)
    Q = torch.randn((nq, M * dsub), device="cuda", generator=g)
    ref = T.ops.adc_lut(Q, C, "l2")
    buf = torch.empty(nq * M * 256 + 1, device="cuda")
    out = buf[1:]                                   # 4-byte aligned only
    from b200vq import _lib
    rc = _lib.load().b200vq_adc_lut(Q.data_ptr(), nq, Q.stride(0), C.data_ptr(), M, 256, dsub, 0, out.data_ptr(), None)
    assert rc == 0
    assert torch.equal(out.view(nq, M, 256).view(torch.int32), ref.view(torch.int32))
