"""GPU parity tests (through the C ABI) for PQ assign / decode / k-means / ADC / flat search."""
import os

import numpy as np
import pytest

from parity import assert_codes_parity, assert_topk_parity, make_data

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def T():
    import torch
    import b200vq
    from b200vq import ops, ops_sq, pq
    b200vq.load()
    class NS:
        pass
    ns = NS()
    ns.torch, ns.ops, ns.ops_sq, ns.pq = torch, ops, ops_sq, pq
    ns.dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    return ns


from oracle import pq as opq  # noqa: E402


@pytest.mark.parametrize("n,D,M,ksub,impl", [
    (1000, 64, 8, 256, "simt"), (1000, 64, 8, 256, "auto"),
    (777, 48, 8, 256, "simt"),          # dsub = 6 (the reference tests' 256x48 shape family)
    (513, 30, 6, 16, "simt"),           # odd dsub = 5, small codebook
    (2049, 96, 1, 256, "simt"),         # one sub-quantizer, dsub = 96
    (4096, 1536, 16, 256, "auto"),      # config-1 shape (M=16, dsub=96)
    (4096, 1024, 32, 256, "auto"),      # config-5 shape (M=32, dsub=32)
    (1500, 1536, 32, 256, "auto"),      # config-2 shape (dsub=48), ragged n
    (3, 16, 2, 4, "simt"),
])
def test_assign_matches_oracle(T, n, D, M, ksub, impl):
    X = make_data(n, D, seed=n + D, aniso=D <= 128)
    Xt = make_data(max(ksub, 300), D, seed=7, aniso=D <= 128)
    C = opq.pq_init_centroids(Xt, M, ksub, seed=11)
    ref, d1, d2 = opq.pq_assign(X, C, return_gap=True)
    got = T.ops.pq_assign(T.dev(X), T.dev(C), impl=impl).cpu().numpy()
    assert_codes_parity(got, ref, d1, d2, f"pq_assign[{impl}] n={n} D={D} M={M}")


def test_assign_golden_fixture_and_distances(T, golden_dir):
    g = np.load(os.path.join(golden_dir, "pq_oracle_D64.npz"))
    codes, dist = T.ops.pq_assign(T.dev(g["X"]), T.dev(g["C"]), want_dist=True, impl="simt")
    assert_codes_parity(codes.cpu().numpy(), g["codes"], g["d1"], g["d2"], "golden codes")
    # (||x||^2 + ||c||^2) - 2 x.c cancels: fp32 rounding is relative to ||x||^2 + ||c||^2, not to d
    X, C, codes = g["X"], g["C"], g["codes"]
    dsub = C.shape[2]
    mag = np.stack([(X[:, m * dsub:(m + 1) * dsub] ** 2).sum(1) + (C[m][codes[:, m]] ** 2).sum(1)
                    for m in range(C.shape[0])], axis=1)
    assert np.all(np.abs(dist.cpu().numpy() - g["d1"]) <= 1e-4 * g["d1"] + 2e-6 * mag)


def test_assign_edge_cases(T):
    C = opq.pq_init_centroids(make_data(300, 32, 1), 4, 256, seed=0)
    empty = T.torch.empty((0, 32), dtype=T.torch.float32, device="cuda")
    assert T.ops.pq_assign(empty, T.dev(C)).shape == (0, 4)
    # strided rows (ldx > D): a column slice of a wider matrix
    wide = T.dev(make_data(100, 64, 2))
    view = wide[:, :32]
    ref = opq.pq_assign(view.cpu().numpy(), C)
    assert np.array_equal(T.ops.pq_assign(view, T.dev(C), impl="simt").cpu().numpy(), ref)
    # exact duplicate centroid: lowest index wins
    C2 = C.copy()
    C2[2, 200] = C2[2, 10]
    got = T.ops.pq_assign(wide[:, 32:], T.dev(C2), impl="simt").cpu().numpy()
    assert not np.any(got[:, 2] == 200)
    from b200vq import B200VQError
    with pytest.raises(B200VQError):
        T.ops.pq_assign(wide, T.dev(C))  # D mismatch


def test_decode_is_exact_and_by_id(T):
    X = make_data(900, 96, 3)
    C = opq.pq_init_centroids(X, 12, 256, seed=2)
    codes = opq.pq_assign(X, C)
    got = T.ops.pq_decode(T.dev(codes), T.dev(C)).cpu().numpy()
    assert np.array_equal(got, opq.pq_decode(codes, C))
    ids = np.array([5, 0, 899, 5, 17], dtype=np.int64)
    got = T.ops.pq_decode(T.dev(codes), T.dev(C), ids=T.dev(ids)).cpu().numpy()
    assert np.array_equal(got, opq.pq_decode(codes[ids], C))
    C5 = opq.pq_init_centroids(make_data(300, 30, 4), 6, 16, seed=1)  # dsub = 5: scalar path
    c5 = opq.pq_assign(make_data(50, 30, 5), C5)
    assert np.array_equal(T.ops.pq_decode(T.dev(c5), T.dev(C5)).cpu().numpy(), opq.pq_decode(c5, C5))


def test_kmeans_step_and_training_objective(T):
    X = make_data(3000, 32, 6)
    C0 = opq.pq_init_centroids(X, 4, 256, seed=9)
    # one Lloyd step from shared state
    codes = opq.pq_assign(X, C0)
    Cref, _ = opq.pq_lloyd_update(X, codes, C0)
    torch = T.torch
    Cd = T.dev(C0.copy())
    sums = torch.empty_like(Cd)
    counts = torch.empty((4, 256), dtype=torch.int32, device="cuda")
    cd = torch.empty((3000, 4), dtype=torch.uint8, device="cuda")
    ns = torch.zeros(1, dtype=torch.int32, device="cuda")
    T.ops.pq_kmeans_step(T.dev(X), Cd, sums, counts, cd, ns)
    assert np.allclose(Cd.cpu().numpy(), Cref, rtol=1e-5, atol=1e-6)
    # full training: same init, objective within 1e-3 relative of the oracle's
    Cg = T.pq.pq_train(T.dev(X), 4, 8, niter=10, seed=9).cpu().numpy()
    Co = opq.pq_train(X, 4, 8, niter=10, seed=9)
    sg = opq.pq_sse(X, opq.pq_assign(X, Cg), Cg)
    so = opq.pq_sse(X, opq.pq_assign(X, Co), Co)
    assert abs(sg - so) <= 1e-3 * so
    dev_sse = T.ops.pq_sse(T.dev(X), T.dev(opq.pq_assign(X, Cg)), T.dev(Cg)).item()
    assert abs(dev_sse - sg) <= 1e-4 * sg
    with pytest.raises(RuntimeError):
        T.pq.pq_train(T.dev(X[:100]), 4, 8)


def test_kmeans_empty_cluster_split(T):
    X = make_data(2000, 16, 8)
    C0 = opq.pq_init_centroids(X, 2, 256, seed=1)
    C0[1, 33] = 1e6
    codes = opq.pq_assign(X, C0)
    Cref, nsr = opq.pq_lloyd_update(X, codes, C0)
    torch = T.torch
    Cd = T.dev(C0.copy())
    sums = torch.empty_like(Cd)
    counts = torch.empty((2, 256), dtype=torch.int32, device="cuda")
    cd = torch.empty((2000, 2), dtype=torch.uint8, device="cuda")
    ns = torch.zeros(1, dtype=torch.int32, device="cuda")
    T.ops.pq_kmeans_step(T.dev(X), Cd, sums, counts, cd, ns)
    assert int(ns.item()) == nsr >= 1
    assert np.allclose(Cd.cpu().numpy(), Cref, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("metric", [0, 1])
@pytest.mark.parametrize("n,D,M,k,nq", [(5000, 64, 8, 10, 7), (70000, 128, 32, 10, 33), (300, 48, 12, 100, 3),
                                        (20000, 192, 64, 10, 5)])
def test_adc_scan_topk_matches_oracle(T, metric, n, D, M, k, nq):
    X = make_data(n, D, seed=n)
    Q = make_data(nq, D, seed=n + 1)
    C = opq.pq_train(X[:3000], M, 8, niter=2, seed=3)
    codes = opq.pq_assign(X, C)
    lut_ref = opq.adc_lut(Q, C, metric)
    lut = T.ops.adc_lut(T.dev(Q), T.dev(C), metric)
    assert np.allclose(lut.cpu().numpy(), lut_ref, rtol=1e-5, atol=1e-7)
    scores = opq.adc_scores(codes, lut_ref)
    vals, ids = T.ops.adc_scan_topk(T.dev(codes), lut, k, metric)
    assert_topk_parity(vals.cpu().numpy(), ids.cpu().numpy().view(np.uint32), scores, k, metric, "adc_scan")
    if M % 32 == 0 and M <= 96:
        rot = T.ops.pq_codes_to_scan_layout(T.dev(codes))
        vals, ids = T.ops.adc_scan_topk(rot, lut, k, metric, n=n, rot=True)
        assert_topk_parity(vals.cpu().numpy(), ids.cpu().numpy().view(np.uint32), scores, k, metric, "adc_scan_rot")


def test_adc_equals_reference_decompress_search(T):
    """Parity anchor of SURVEY.md F6: ADC top-k == FlatQuantizedIndex's decompress-then-cdist top-k."""
    X = make_data(20000, 64, 21)
    Q = make_data(20, 64, 22)
    C = opq.pq_train(X[:4000], 8, 8, niter=3, seed=1)
    codes = opq.pq_assign(X, C)
    R = opq.pq_decode(codes, C)
    q64, x64 = Q.astype(np.float64), R.astype(np.float64)
    d = ((q64 * q64).sum(1)[:, None] + (x64 * x64).sum(1)[None] - 2 * q64 @ x64.T).astype(np.float32)
    codec = T.pq.PQCodec(T.dev(C))
    vals, ids = codec.search(T.dev(codes), T.dev(Q), 10, "l2")
    assert_topk_parity(vals.cpu().numpy(), ids.cpu().numpy().view(np.uint32), d, 10, 0, "adc vs decompress")


@pytest.mark.parametrize("metric", [0, 1])
@pytest.mark.parametrize("n,D,k,nq", [(3000, 48, 10, 5), (40000, 96, 100, 70), (129, 33, 10, 64), (50, 8, 50, 2)])
def test_flat_topk_matches_oracle(T, metric, n, D, k, nq):
    X = make_data(n, D, seed=n + 5)
    Q = make_data(nq, D, seed=n + 6)
    scores = opq.flat_scores(X, Q, metric)
    vals, ids = T.ops.flat_topk(T.dev(X), T.dev(Q), k, metric)
    assert_topk_parity(vals.cpu().numpy(), ids.cpu().numpy().view(np.uint32), scores, k, metric, "flat_topk",
                       gap_rel=2e-6)


def test_flat_topk_row_scale_is_normalized_ip(T):
    X = make_data(5000, 40, 31, normalize=False)
    Q = make_data(9, 40, 32)
    norms = np.maximum(np.linalg.norm(X, axis=1), 1e-12).astype(np.float32)
    scale = (1.0 / norms).astype(np.float32)
    scores = opq.flat_scores(X, Q, 1, row_scale=scale)
    vals, ids = T.ops.flat_topk(T.dev(X), T.dev(Q), 10, 1, row_scale=T.dev(scale))
    assert_topk_parity(vals.cpu().numpy(), ids.cpu().numpy().view(np.uint32), scores, 10, 1, "scaled ip",
                       gap_rel=2e-6)
    sq = T.ops.row_sqnorms(T.dev(X)).cpu().numpy()
    assert np.allclose(np.sqrt(sq), norms, rtol=1e-5)


def test_topk_merge_is_stable_by_value_then_id(T):
    torch = T.torch
    rng = np.random.default_rng(0)
    nq, P, k = 6, 8, 10
    vals = np.sort(rng.integers(0, 20, size=(nq, P, k)).astype(np.float32), axis=2)  # many ties
    ids = rng.permutation(nq * P * k).reshape(nq, P, k).astype(np.int64) + (1 << 33)
    ids[0, 1, 5:] = -1
    for metric in (0, 1):
        v = vals if metric == 0 else -vals
        ov, oi = T.ops.topk_merge(T.dev(v), T.dev(ids), k, metric)
        ov, oi = ov.cpu().numpy(), oi.cpu().numpy()
        for q in range(nq):
            fv, fi = vals[q].ravel(), ids[q].ravel()
            keep = fi >= 0
            order = np.lexsort((fi[keep], fv[keep]))[:k]
            assert np.array_equal(oi[q], fi[keep][order])
            assert np.array_equal(np.abs(ov[q]), fv[keep][order])
