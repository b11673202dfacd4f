"""Data-fit per-dimension codebooks on the device (csrc/codebook.cu) against oracle/codebook.py, the restatement of
the SAQ engine's build_codebook_lloyd that is pinned on the reference's own compiled builder, and the
RankAwareQuantizer(codebook="lloyd") / registry paths that consume them (rank_aware_quantization.py:189-214)."""
import numpy as np
import pytest

from oracle import codebook as ocb

pytestmark = pytest.mark.gpu


def _columns(n, seed):
    rng = np.random.default_rng(seed)
    cols = [
        rng.standard_normal(n) * 0.3,
        rng.standard_normal(n) ** 2,
        np.concatenate([rng.normal(-2, 0.2, n // 3), rng.normal(1, 0.5, n - n // 3)]),
        np.round(rng.standard_normal(n) * 3),                  # few distinct values: the degenerate branch at 5+ bits
        rng.standard_normal(n) * 1e-3 + 5.0,
        rng.uniform(-1, 1, n),
        rng.standard_normal(n) * 40,
        np.zeros(n) + 0.25,                                     # constant column
        rng.laplace(0, 1, n),
    ]
    return np.stack(cols).astype(np.float32)


@pytest.mark.parametrize("n,seed", [(3000, 0), (777, 5), (20000, 2)])
def test_codebook_lloyd_kernel_matches_pinned_oracle(n, seed):
    import torch
    from b200vq import ops_sq
    cols = _columns(n, seed)
    D = cols.shape[0]
    bits = np.array([4, 3, 2, 6, 5, 1, 6, 3, 0], dtype=np.int32)
    cen, sizes, costs = ops_sq.codebook_lloyd(torch.from_numpy(cols).cuda(), torch.from_numpy(bits).cuda(), seed=seed)
    cen, sizes, costs = cen.cpu().numpy(), sizes.cpu().numpy(), costs.cpu().numpy()
    exact = 0
    for d in range(D):
        b = int(bits[d])
        oc, ob = ocb.build_codebook_lloyd(cols[d], max(b, 1), seed=seed, levels=[b])
        ref = ob[b]
        assert sizes[d] == len(ref), (d, sizes[d], len(ref))
        got = cen[d, :sizes[d]]
        # same algorithm, same draws; the device's prefix sums are a parallel scan (fp64 rounding of the order)
        assert np.allclose(got, ref, rtol=3e-7, atol=1e-9), (d, b, np.abs(got - ref).max())
        assert np.isclose(costs[d], oc[b], rtol=1e-5, atol=1e-12), (d, costs[d], oc[b])
        exact += int(np.array_equal(got, ref))
    assert exact >= D - 1, f"only {exact} of {D} codebooks are bit-identical to the oracle"


def test_uniform_bit_rate_and_costs_fall_with_bits():
    import torch
    from b200vq import ops_sq
    cols = torch.from_numpy(_columns(5000, 1)[[0, 1, 2, 5, 8]]).cuda()
    prev = None
    for b in range(0, 7):
        cen, sizes, costs = ops_sq.codebook_lloyd(cols, b, seed=0)
        assert (sizes == (1 << b)).all()
        c = cen[:, :1 << b]
        assert bool((c[:, 1:] >= c[:, :-1]).all())
        if prev is not None:
            assert bool((costs <= prev * 1.0001).all())
        prev = costs


def test_rank_aware_lloyd_codebooks_and_registry():
    from haag_vq.benchmarks.method_registry import build_quantizer
    from haag_vq.benchmarks.quantizer_adapters import FaissQuantizerAdapter
    from haag_vq.methods.rank_aware_quantization import RankAwareQuantizer
    rng = np.random.default_rng(0)
    d = 64
    A = np.diag(np.linspace(4.0, 0.2, d)) @ rng.standard_normal((d, d))
    X = ((rng.standard_normal((6000, d)) ** 3) @ A).astype(np.float32)          # heavy tails: data-fit beats Gaussian
    mse = {}
    for cbk in ("gaussian", "lloyd"):
        q = RankAwareQuantizer(avg_bits=4, alpha=0.5, packing="ffd", codebook=cbk)
        q.fit(X)
        codes = q.compress(X)
        assert codes.shape == (6000, q.code_size)
        mse[cbk] = float(np.mean((X - q.decompress(codes)) ** 2))
        if cbk == "lloyd":
            # the engine's builder on the float64 projection of the same sample, at every dimension's bit-rate
            Y = (X.astype(np.float64) - q.mu) @ q.V
            for dd in (0, 1, 7, 20, 40):
                b = int(q.bits[dd])
                if b == 0:
                    continue
                _, ob = ocb.build_codebook_lloyd(Y[:, dd].astype(np.float32), b, levels=[b])
                scale = float(np.abs(ob[b]).max())
                assert np.allclose(q.cb[dd], np.sort(ob[b].astype(np.float64)), rtol=0, atol=2e-5 * scale), dd
    assert mse["lloyd"] < mse["gaussian"]
    for name, alpha in (("rankaware", 0.5), ("perdim_mse", 0.0)):
        a = build_quantizer(name, 3.0, d)
        assert isinstance(a, FaissQuantizerAdapter)
        a.fit(X)
        R = a.reconstruct(np.arange(100, dtype=np.uint32))
        assert R.shape == (100, d) and float(np.mean((R - X[:100]) ** 2)) < float(np.mean(X[:100] ** 2))
    with pytest.raises(NotImplementedError):
        build_quantizer("rankaware_exact", 3.0, d)


def test_engine_greedy_allocator_and_ours_registry_name():
    """Method ``ours``: the engine with the greedy allocator on the per-dimension Lloyd MSE table (bit-rates 0..13).
    The device table against the pinned oracle builder on a few dimensions, the plan against the oracle's
    restatement of BitAllocatorGreedy on the SAME table, and the adapter contract."""
    import torch
    from b200vq import saq_engine as E
    from haag_vq.benchmarks.method_registry import build_quantizer
    rng = np.random.default_rng(3)
    n, D = 3000, 128
    X = (rng.standard_normal((n, D)) * np.linspace(3.0, 0.05, D)).astype(np.float32)
    eng = E.SaqEngine(3.0, allocator="greedy").fit(torch.from_numpy(X).cuda())
    tab = eng.mse_table
    assert tab.shape == (128, 14) and np.all(np.diff(tab, axis=1) <= 1e-6 * tab[:, :1] + 1e-12)
    Y = ((X - eng.mean.cpu().numpy()) @ eng.rot.cpu().numpy()).astype(np.float32)
    for d in (0, 5, 60):
        oc, _ = ocb.build_codebook_lloyd(Y[:, d], 13)
        # the device projects with the 3xTF32 GEMM: sample values differ from numpy's in the last bit
        assert np.allclose(tab[d, :9], oc[:9], rtol=2e-3, atol=1e-9), d
        # at hundreds of centroids on 3000 points a last-bit change of a sample value moves the local optimum
        assert np.allclose(tab[d, 9:12], oc[9:12], rtol=0.25, atol=1e-9), d
    plan = ocb.allocate_greedy(tab, 64, 13, int(3.0 * 128) + 64, 64)
    assert eng.plan == [(int(a), int(b)) for a, b in plan]
    assert sum(a * b + (64 if b else 0) for a, b in eng.plan) <= int(3.0 * 128) + 64
    R = eng.reconstruct(torch.arange(n, device="cuda")).cpu().numpy()
    dp = E.SaqEngine(3.0).fit(torch.from_numpy(X).cuda())
    Rd = dp.reconstruct(torch.arange(n, device="cuda")).cpu().numpy()
    assert float(((R - X) ** 2).mean()) < 1.5 * float(((Rd - X) ** 2).mean())
    for name in ("ours", "ours_exact"):
        q = build_quantizer(name, 3.0, D)
        q.fit(X)
        assert q.reconstruct(np.arange(10)).shape == (10, D)
