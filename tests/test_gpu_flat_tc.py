"""Tensor-core exact search (b200vq_flat_topk_tc) == the exact SIMT kernel (b200vq_flat_topk), through the
C ABI: the screen is rigorous, so ids and values must be IDENTICAL (the final scores are computed with the
SIMT kernel's arithmetic)."""
import numpy as np
import pytest

from oracle import pq as opq
from parity import assert_topk_parity

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def T():
    import torch
    import b200vq
    from b200vq import ops, ops_sq
    b200vq.load()

    class NS:
        pass
    ns = NS()
    ns.torch, ns.ops, ns.ops_sq = torch, ops, ops_sq

    def synth(n, D, row0=0, seed=1234, normalize=True):
        cs = torch.linspace(4.0, 0.2, D, device="cuda")
        return ops_sq.synth_rows(n, D, row0, seed, cs, normalize, device="cuda")
    ns.synth = synth
    return ns


@pytest.mark.parametrize("n,D,nq,k,metric,normalize", [
    (100_000, 1536, 37, 10, "l2", True),       # config-1 ground truth shape
    (100_000, 1536, 300, 100, "ip", True),     # GT k = 100, more than one query tile
    (70_001, 96, 5, 1, "l2", False),           # ragged rows, un-normalised, tiny D
    (300_000, 256, 64, 32, "ip", False),
    (150_000, 1024, 257, 128, "l2", False),
])
def test_tc_equals_simt(T, n, D, nq, k, metric, normalize):
    X = T.synth(n, D, normalize=normalize)
    Q = T.synth(nq, D, row0=1 << 41, seed=4321, normalize=normalize)
    v0, i0 = T.ops.flat_topk(X, Q, k, metric, impl="simt")
    v1, i1 = T.ops.flat_topk(X, Q, k, metric, impl="tc")
    assert T.torch.equal(i0, i1), f"{int((i0 != i1).sum())} ids differ"
    assert T.torch.equal(v0, v1)


def test_tc_planted_duplicates_row_scale_and_oracle(T):
    n, D, nq, k = 80_000, 128, 16, 10
    X = T.synth(n, D, normalize=False)
    Q = X[:: n // nq][:nq] * 1.001            # queries almost equal to database rows
    X[n - 5:] = X[:5]                         # exact duplicates far apart: ties broken by the lower id
    nrm = T.ops.row_sqnorms(X)
    rs = (1.0 / nrm.sqrt()).contiguous()      # build_scaled_ip_index: rows scaled to unit norm
    a = T.ops.flat_topk(X, Q, k, "ip", xnorm=nrm, row_scale=rs, impl="simt")
    b = T.ops.flat_topk(X, Q, k, "ip", xnorm=nrm, row_scale=rs, impl="tc")
    assert T.torch.equal(a[1], b[1]) and T.torch.equal(a[0], b[0])
    c = T.ops.flat_topk(X, Q, k, "l2", impl="simt")
    d = T.ops.flat_topk(X, Q, k, "l2", impl="tc")
    assert T.torch.equal(c[1], d[1]) and T.torch.equal(c[0], d[0])
    sc = opq.flat_scores(X.cpu().numpy(), Q.cpu().numpy(), opq.L2)
    assert_topk_parity(d[0].cpu().numpy(), d[1].cpu().numpy().view(np.uint32), sc, k, 0, "flat tc vs oracle",
                       gap_rel=1e-5, val_rel=1e-4, abs_tol=1e-5 * float(nrm.max()))


def test_tc_overflow_falls_back_to_simt(T):
    """All rows identical: every row survives every screen, the lists overflow, the gated SIMT launch answers."""
    n, D, nq, k = 70_000, 64, 3, 10
    X = T.torch.ones((n, D), device="cuda") * 0.5
    X[::7, 3] = 0.25
    Q = T.synth(nq, D, seed=9, normalize=False)
    a = T.ops.flat_topk(X, Q, k, "l2", impl="simt")
    b = T.ops.flat_topk(X, Q, k, "l2", impl="tc")
    assert T.torch.equal(a[1], b[1]) and T.torch.equal(a[0], b[0])


def test_flat_topk_beyond_128_is_the_exact_ranking_and_index_accepts_large_k():
    """k > 128: chained continuations (b200vq_flat_topk_after) concatenate to the exact (value, id) ranking."""
    import numpy as np
    import torch
    from b200vq import ops
    from haag_vq.methods.scalar_quantization import ScalarQuantizer
    from haag_vq.methods.search.flat_quantized_index import FlatQuantizedIndex
    rng = np.random.default_rng(5)
    X = rng.standard_normal((3000, 40)).astype(np.float32)
    X[100:140] = X[50]                      # 41 identical rows: ties across a continuation boundary
    Q = rng.standard_normal((9, 40)).astype(np.float32)
    Q[0] = X[50]
    xd, qd = torch.from_numpy(X).cuda(), torch.from_numpy(Q).cuda()
    for metric in ("l2", "ip"):
        for k in (129, 300, 3000):
            v, i = ops.flat_topk(xd, qd, k, metric, impl="simt")
            v1, i1 = ops.flat_topk(xd, qd, 128, metric, impl="simt")
            assert torch.equal(v[:, :128], v1) and torch.equal(i[:, :128], i1)
            vh, ih = v.cpu().numpy(), i.cpu().numpy().view(np.uint32)
            # a permutation prefix, sorted by (value, id) with the kernel's own values
            sgn = -1.0 if metric == "ip" else 1.0
            for q in range(9):
                assert len(set(ih[q])) == k
                key = list(zip((sgn * vh[q]).tolist(), ih[q].tolist()))
                assert key == sorted(key)
            if k == 3000:
                assert (np.sort(ih, axis=1) == np.arange(3000, dtype=np.uint32)).all()
    idx = FlatQuantizedIndex(ScalarQuantizer(num_bits=8))
    idx.fit(X, "l2")
    ids, d = idx.search_with_scores(Q, 200)
    assert ids.shape == (9, 200) and (np.diff(d, axis=1) >= 0).all()
    ids128, _ = idx.search_with_scores(Q, 128)
    assert np.array_equal(ids[:, :128], ids128)
