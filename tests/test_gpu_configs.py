"""BASELINE.json configurations at (or near) their full sizes, through size-independent properties:
idempotence (encode(decode(codes)) == codes), kernel-vs-kernel bit equality, shard/merge invariance,
round-trip bounds. Config 1 is also checked value by value against the CPU oracle (it is the
reference's own CPU-runnable case). Inputs are generated on the device (b200vq_synth_rows)."""
import numpy as np
import pytest

from parity import assert_codes_parity, assert_topk_parity

pytestmark = pytest.mark.gpu

from oracle import pq as opq  # noqa: E402


@pytest.fixture(scope="module")
def T():
    import torch
    import b200vq
    from b200vq import dist, ops, ops_sq, pq
    b200vq.load()

    class NS:
        pass
    ns = NS()
    ns.torch, ns.ops, ns.ops_sq, ns.pq, ns.dist = torch, ops, ops_sq, pq, dist
    ns.dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()

    def synth(n, D, row0=0, seed=1234, normalize=True):
        cs = torch.linspace(4.0, 0.2, D, device="cuda")
        return ops_sq.synth_rows(n, D, row0, seed, cs, normalize, device="cuda")
    ns.synth = synth
    return ns


def test_config1_pq_m16_100k_x_1536_against_oracle(T):
    """PQ M=16 nbits=8 on 100K x 1536: train on device, then codes, ADC top-10 and recall@10 vs the
    oracle with the SAME centroids (parity is conditional on shared trained state)."""
    n, D, M, nq, k = 100_000, 1536, 16, 64, 10
    X = T.synth(n, D)
    # queries = perturbed database rows, so that every query has a true near neighbour to recall
    Q = X[:: n // nq][:nq] + 0.02 * T.synth(nq, D, row0=1 << 41, seed=4321)
    Q = (Q / Q.norm(dim=1, keepdim=True)).contiguous()
    C = T.pq.pq_train(X, M, 8, niter=4, seed=1234)
    codec = T.pq.PQCodec(C)
    codes = codec.encode(X)                         # tcgen05 path (dsub = 96)
    assert T.torch.equal(codes, T.ops.pq_assign(X, C, impl="simt"))
    Xh, Ch = X.cpu().numpy(), C.cpu().numpy()
    ref, d1, d2 = opq.pq_assign(Xh, Ch, return_gap=True)
    assert_codes_parity(codes.cpu().numpy(), ref, d1, d2, "config 1 codes")
    vals, ids = codec.search(codes, Q, k, "l2")
    scores = opq.adc_scores(ref, opq.adc_lut(Q.cpu().numpy(), Ch, opq.L2))
    # scores of rows whose code differs inside the tie exemption are taken from the device codes
    diff_rows = np.nonzero((codes.cpu().numpy() != ref).any(1))[0]
    if diff_rows.size:
        scores[:, diff_rows] = opq.adc_scores(codes.cpu().numpy()[diff_rows], opq.adc_lut(Q.cpu().numpy(), Ch, opq.L2))
    ids_h = ids.cpu().numpy().view(np.uint32)
    n_mis = assert_topk_parity(vals.cpu().numpy(), ids_h, scores, k, 0, "config 1 ADC", gap_rel=1e-5)
    # exact ground truth on the device vs the oracle's flat search, then recall@10 of both arms
    gv, gi = T.ops.flat_topk(X, Q, k, "l2")
    gi_h = gi.cpu().numpy().view(np.uint32)
    flat = opq.flat_scores(Xh, Q.cpu().numpy(), opq.L2)
    g_mis = assert_topk_parity(gv.cpu().numpy(), gi_h, flat, k, 0, "config 1 ground truth", gap_rel=1e-5,
                               val_rel=1e-4, abs_tol=4e-6)   # |q|^2 + |x|^2 = 2: a few fp32 ulps of it
    gt = opq.topk_sorted(flat, k, opq.L2)[1]
    rec_dev = opq.recall_at_k(ids_h, gi_h, k)
    rec_ref = opq.recall_at_k(opq.topk_sorted(scores, k, opq.L2)[1], gt, k)
    assert abs(rec_dev - rec_ref) <= (n_mis + g_mis) / (nq * k) + 1e-12, (rec_dev, rec_ref, n_mis, g_mis)
    assert rec_ref >= 0.05, rec_ref   # at least the planted neighbour is found


def test_config5_shape_1m_rows_tc_equals_simt_and_is_idempotent(T):
    """PQ M=32 on 2^20 x 1024 (one streaming batch of config 5): tensor-core codes == exact SIMT codes,
    and re-encoding the reconstruction returns the same codes."""
    n, D, M = 1 << 20, 1024, 32
    X = T.synth(n, D)
    C = T.pq.pq_train(T.synth(65536, D, row0=1 << 40), M, 8, niter=4, seed=1234)
    codec = T.pq.PQCodec(C)
    stats = T.torch.zeros(8, dtype=T.torch.int64, device="cuda")
    tc = T.ops.pq_assign(X, C, codec.cnorm, impl="tc", stats=stats)
    simt = T.ops.pq_assign(X, C, codec.cnorm, impl="simt")
    assert T.torch.equal(tc, simt)
    assert int(stats[0]) < 0.2 * n * M
    rec = codec.decode(tc)
    again = T.ops.pq_assign(rec, C, codec.cnorm, impl="tc")
    # a centroid is its own nearest centroid unless two centroids coincide
    assert (again != tc).float().mean().item() < 1e-6
    # MSE through the device reduction equals the k-means objective / (n D)
    sse = T.ops.pq_sse(X, tc, C).item()
    err = T.ops_sq.sq_err_sum(X, rec).item()
    assert abs(sse - err) <= 1e-4 * abs(sse)


def test_config5_sharded_scan_equals_single_scan(T):
    """Row-sharding invariance of the ADC top-10: scanning 8 shards and merging (value, id) lists
    gives exactly the single-scan result (the NCCL path's arithmetic without the collective)."""
    n, M, nq, k, shards = 1_000_003, 32, 32, 10, 8
    g = T.torch.Generator(device="cuda").manual_seed(5)
    codes = T.torch.randint(0, 256, (n, M), dtype=T.torch.uint8, device="cuda", generator=g)
    lut = T.torch.rand((nq, M, 256), device="cuda", generator=g)
    v0, i0 = T.ops.adc_scan_topk(codes, lut, k, "l2")
    rot = T.ops.pq_codes_to_scan_layout(codes)
    v1, i1 = T.ops.adc_scan_topk(rot, lut, k, "l2", n=n, rot=True)
    # the scan layout adds the M table entries in a lane-dependent order: same ids, values to an ulp
    assert T.torch.equal(i0, i1) and T.torch.allclose(v0, v1, rtol=1e-6, atol=0)
    pv, pi = [], []
    for r in range(shards):
        s, e = T.dist.shard_bounds(n, shards, r)
        v, i = T.ops.adc_scan_topk(codes[s:e], lut, k, "l2")
        ids64 = (i.to(T.torch.int64) & 0xFFFFFFFF) + s
        pv.append(v)
        pi.append(ids64)
    mv, mi = T.ops.topk_merge(T.torch.stack(pv, 1).contiguous(), T.torch.stack(pi, 1).contiguous(), k, "l2")
    assert T.torch.equal(mi, i0.to(T.torch.int64) & 0xFFFFFFFF) and T.torch.equal(mv, v0)


def test_config3_sq8_1m_x_3072_round_trip(T):
    """SQ 8-bit on 1M x 3072: |x - decode(encode(x))| <= step/2 (+ fp32 slack) and the codes are a fixed
    point of encode(decode(.)). Processed in 4 batches of 262144 rows (3 GiB each)."""
    D, B, nb = 3072, 262_144, 4
    vmin = vmax = None
    for b in range(nb):
        X = T.synth(B, D, row0=b * B, normalize=False)
        vmin, vmax = T.ops_sq.sq_minmax(X, vmin, vmax)
        del X
    step = (vmax - vmin + 1e-8) / 255.0
    for b in range(nb):
        X = T.synth(B, D, row0=b * B, normalize=False)
        codes = T.ops_sq.sq_encode(X, vmin, vmax, 8)
        rec = T.ops_sq.sq_decode(codes, vmin, vmax, 8, D)
        err = (X - rec).abs()
        assert bool((err <= 0.5 * step * (1 + 1e-4) + 1e-6).all())
        assert T.torch.equal(T.ops_sq.sq_encode(rec, vmin, vmax, 8), codes)
        del X, codes, rec, err


def test_config4_rabitq1_1m_x_1536_shard_invariance_and_recall(T):
    """1-bit RaBitQ on 1M x 1536: estimator top-10 over the whole code array == merge of 4 shard scans;
    estimator top-100 contains the planted exact nearest neighbour of every query."""
    n, D, nq, qb = 1_000_000, 1536, 16, 4
    X = T.synth(n, D)
    Q = X[:: n // nq][:nq] + 0.05 * T.synth(nq, D, row0=1 << 41, seed=4321)   # planted neighbours
    Q = (Q / Q.norm(dim=1, keepdim=True)).contiguous()
    cent = X.mean(0).contiguous()
    codes = T.ops_sq.rabitq1_encode(X, cent, "l2")
    assert codes.shape == (n, D // 8 + 8)
    planes, qfac = T.ops_sq.rabitq1_prepare_queries(Q, cent, qb)
    v0, i0 = T.ops_sq.rabitq1_scan_topk(codes, D, planes, qfac, qb, 100, "l2")
    pv, pi = [], []
    for r in range(4):
        s, e = T.dist.shard_bounds(n, 4, r)
        v, i = T.ops_sq.rabitq1_scan_topk(codes[s:e], D, planes, qfac, qb, 100, "l2")
        pv.append(v)
        pi.append((i.to(T.torch.int64) & 0xFFFFFFFF) + s)
    mv, mi = T.ops.topk_merge(T.torch.stack(pv, 1).contiguous(), T.torch.stack(pi, 1).contiguous(), 100, "l2")
    assert T.torch.equal(mi, i0.to(T.torch.int64) & 0xFFFFFFFF) and T.torch.equal(mv, v0)
    gv, gi = T.ops.flat_topk(X, Q, 10, "l2")
    est = (i0.to(T.torch.int64) & 0xFFFFFFFF).cpu().numpy()
    gt = (gi.to(T.torch.int64) & 0xFFFFFFFF).cpu().numpy()
    assert np.array_equal(gt[:, 0], np.arange(nq) * (n // nq))           # exact search finds the plant
    assert all(gt[q, 0] in set(est[q]) for q in range(nq))                # and so does the estimator


def test_config2_opq_m32_1536_rotation_then_assign(T):
    """OPQ M=32 on 262144 x 1536 (dsub = 48): codes == assign(rotate(x)) on the exact kernel, decode
    inverts the rotation (orthonormal A), ADC on rotated queries == distance to the reconstruction."""
    n, D, M = 262_144, 1536, 32
    X = T.synth(n, D)
    A = T.dev(T.pq.random_rotation(D, 7))
    Y = T.pq.rotate(X[:65536], A)
    C = T.pq.pq_train(Y, M, 8, niter=3, seed=1234)
    codec = T.pq.PQCodec(C, rotation=A)
    codes = codec.encode(X)
    assert T.torch.equal(codes, T.ops.pq_assign(T.pq.rotate(X, A), C, impl="simt"))
    rec = codec.decode(codes[:4096])
    Q = T.synth(8, D, row0=1 << 41, seed=4321)
    vals, ids = codec.search(codes[:4096], Q, 10, "l2")
    true = ((Q[:, None, :].double() - rec[None].double()) ** 2).sum(-1).float().cpu().numpy()
    assert_topk_parity(vals.cpu().numpy(), ids.cpu().numpy().view(np.uint32), true, 10, 0, "config 2 ADC",
                       gap_rel=1e-4, val_rel=2e-4)
