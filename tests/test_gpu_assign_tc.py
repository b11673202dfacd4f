"""tcgen05 PQ assign (b200vq_pq_assign_tc): must return the SAME codes as the exact SIMT kernel
(bit-identical, no exemption) and match the oracle under the 1e-6 top-2 gate."""
import numpy as np
import pytest

from parity import assert_codes_parity, make_data

pytestmark = pytest.mark.gpu

from oracle import pq as opq  # noqa: E402


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
    ns.dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    return ns


def _centroids(D, M, seed, trained_rows=2000):
    Xt = make_data(trained_rows, D, seed=seed, aniso=D <= 256)
    C = opq.pq_init_centroids(Xt, M, 256, seed=seed + 1)
    return C


@pytest.mark.parametrize("n,D,M", [
    (128, 32, 1),          # one tile, one sub-quantizer, dsub = 32 (128B swizzle)
    (1000, 64, 8),         # dsub = 8  (32B swizzle), KS = 8
    (777, 128, 8),         # dsub = 16 (64B swizzle), ragged n
    (4096, 1024, 32),      # config-5 shape: dsub = 32, KS = 2
    (1500, 1536, 32),      # config-2 shape: dsub = 48 (64B swizzle, 3 chunks)
    (4096, 1536, 16),      # config-1 shape: dsub = 96 (128B swizzle, 3 chunks)
    (3000, 512, 8),        # dsub = 64
    (2500, 1024, 8),       # dsub = 128 (single A stage)
    (20000, 256, 8),       # many tiles per CTA, several work items
    (1, 64, 2),            # single row
])
def test_tc_equals_simt_and_oracle(T, n, D, M):
    X = make_data(n, D, seed=n + D + M, aniso=D <= 256)
    C = _centroids(D, M, seed=5)
    xd, cd = T.dev(X), T.dev(C)
    simt = T.ops.pq_assign(xd, cd, impl="simt").cpu().numpy()
    stats = T.torch.zeros(3, dtype=T.torch.int64, device="cuda")
    tc = T.ops.pq_assign(xd, cd, impl="tc", stats=stats).cpu().numpy()
    ndiff = int((tc != simt).sum())
    assert ndiff == 0, f"tc vs simt: {ndiff} of {tc.size} codes differ (first rows {np.nonzero((tc != simt).any(1))[0][:5]})"
    ref, d1, d2 = opq.pq_assign(X, C, return_gap=True)
    assert_codes_parity(tc, ref, d1, d2, f"pq_assign[tc] n={n} D={D} M={M}")
    st = stats.cpu().numpy()
    # the screen must leave most pairs with a single candidate
    assert st[0] <= 0.5 * n * M + 64, f"re-evaluated pairs: {st[0]} of {n * M}"


def test_tc_near_duplicate_centroids_and_strided_rows(T):
    """Duplicate / near-duplicate centroids force 2- and >=3-candidate pairs through the exact
    paths (pairwise re-evaluation and the fix-up list); rows are a column slice (ldx > D)."""
    n, D, M = 3000, 128, 4
    wide = make_data(n, 2 * D, seed=3)
    C = _centroids(D, M, seed=9)
    C[1, 200] = C[1, 10]                       # exact duplicate: lowest index must win
    C[2, 17] = C[2, 16] * np.float32(1 + 1e-7)  # near duplicates inside the TF32 margin
    C[2, 18] = C[2, 16] * np.float32(1 - 1e-7)
    C[3, 100:140] = C[3, 100] + np.float32(1e-6) * np.arange(40, dtype=np.float32)[:, None]
    wd = T.dev(wide)
    view = wd[:, D:]
    cd = T.dev(C)
    simt = T.ops.pq_assign(view, cd, impl="simt").cpu().numpy()
    stats = T.torch.zeros(3, dtype=T.torch.int64, device="cuda")
    tc = T.ops.pq_assign(view, cd, impl="tc", stats=stats).cpu().numpy()
    assert np.array_equal(tc, simt)
    assert not np.any(tc[:, 1] == 200)
    assert int(stats[0]) > 0


def test_tc_large_magnitudes_and_zero_rows(T):
    n, D, M = 2048, 256, 8
    X = make_data(n, D, seed=1, normalize=False) * np.float32(1000.0)
    X[::7] = 0
    C = _centroids(D, M, seed=2) * np.float32(1000.0)
    xd, cd = T.dev(X), T.dev(C)
    assert np.array_equal(T.ops.pq_assign(xd, cd, impl="tc").cpu().numpy(),
                          T.ops.pq_assign(xd, cd, impl="simt").cpu().numpy())
