"""Tensor-core rotation GEMM (csrc/rotate_tc.cu) against fp64 and against the fp32 library sgemm it
replaces (faiss OPQMatrix.apply at optimized_product_quantization.py:30-31)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _strict_mm(x, w):
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        return x @ w.t()
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev


def _err(y, ref64):
    """max over rows of |y - ref|_inf / |ref|_2-row-scale: error in units of the row norm."""
    scale = ref64.norm(dim=1, keepdim=True).clamp_min(1e-30)
    return ((y.double() - ref64).abs() / scale).max().item()


@pytest.mark.parametrize("n,K,N", [(1000, 64, 64), (4096, 1536, 1536), (333, 96, 40), (129, 1024, 1000), (5000, 384, 384)])
def test_rotate_tc_accuracy(n, K, N):
    from b200vq import ops
    g = torch.Generator(device="cuda").manual_seed(7)
    x = torch.randn((n, K), generator=g, device="cuda") * torch.linspace(4.0, 0.2, K, device="cuda")
    w = torch.linalg.qr(torch.randn((max(N, K), max(N, K)), generator=g, device="cuda"))[0][:N, :K].contiguous()
    rot = ops.RotationTC(w)
    assert rot.applies(x)
    y = rot(x)
    assert y.shape == (n, N)
    ref = x.double() @ w.double().t()
    e_tc, e_lib = _err(y, ref), _err(_strict_mm(x, w), ref)
    # fp32 class: a few 2^-22 per product, random signs over K; the library sgemm is the yardstick
    assert e_tc < 2e-6, (e_tc, e_lib)
    assert e_tc < 16 * max(e_lib, 3e-8), (e_tc, e_lib)


def test_rotate_tc_strided_rows_and_tail():
    from b200vq import ops
    g = torch.Generator(device="cuda").manual_seed(3)
    big = torch.randn((700, 256), generator=g, device="cuda")
    x = big[:, 64:192]                      # row stride 256, 128 columns, 16-byte aligned
    w = torch.randn((50, 128), generator=g, device="cuda")
    rot = ops.RotationTC(w)
    assert rot.applies(x)
    out = torch.full((700, 52), 7.0, device="cuda")
    y = rot(x, out=out[:, :50])
    ref = x.double() @ w.double().t()
    assert _err(y, ref) < 2e-6
    assert torch.all(out[:, 50:] == 7.0)    # columns beyond N untouched


def test_rotate_tc_special_values():
    from b200vq import ops
    x = torch.zeros((128, 32), device="cuda")
    x[0, 0] = 1.0
    x[1, 1] = -3.0000002
    x[2, :] = 1e-30
    x[3, 5] = 3.0e38
    w = torch.eye(32, device="cuda")
    y = ops.RotationTC(w)(x)
    assert torch.equal(y[:3], x[:3])         # exact for 1-term products: hi + lo recombine to the fp32 value
    assert torch.isfinite(y[3, 5]) and abs(y[3, 5].item() / 3.0e38 - 1) < 1e-6


def test_opq_encode_uses_tensor_core_rotation_and_matches_library_path():
    from b200vq import ops, pq as dpq
    g = torch.Generator(device="cuda").manual_seed(11)
    D, M = 256, 16
    x = torch.randn((20000, D), generator=g, device="cuda")
    A = torch.from_numpy(dpq.random_rotation(D, 5)).cuda()
    C = dpq.pq_train(dpq.rotate(x[:8192], A, impl="lib"), M, 8, niter=4, seed=1)
    y_tc, y_lib = dpq.rotate(x, A, impl="tc"), dpq.rotate(x, A, impl="lib")
    assert (y_tc - y_lib).abs().max().item() < 1e-5
    c_tc, c_lib = ops.pq_assign(y_tc, C), ops.pq_assign(y_lib, C)
    assert (c_tc != c_lib).float().mean().item() < 1e-4   # only near-ties of the assignment can move


def test_rotation_split_is_owned_by_its_matrix():
    """A freed rotation's address is reused by the next (D, D) tensor: the prepared split must not be
    (round-1 advisor finding: a cache keyed on data_ptr applied the stale rotation)."""
    from b200vq import pq as dpq
    g = torch.Generator(device="cuda").manual_seed(3)
    D = 128
    x = torch.randn((2048, D), generator=g, device="cuda")
    seen = set()
    for seed in range(6):
        A = torch.from_numpy(dpq.random_rotation(D, seed)).cuda()
        seen.add(A.data_ptr())
        for transpose in (True, False):
            y = dpq.rotate(x, A, transpose=transpose, impl="tc")
            ref = x.double() @ (A.double().t() if transpose else A.double())
            assert (y.double() - ref).abs().max().item() < 1e-5, (seed, transpose)
        del A
    assert len(seen) < 6, "the allocator never reused an address: the test did not exercise the hazard"
    # in-place update of a live matrix: the version counter invalidates the split
    A = torch.from_numpy(dpq.random_rotation(D, 100)).cuda()
    dpq.rotate(x, A, impl="tc")
    A.copy_(torch.from_numpy(dpq.random_rotation(D, 101)).cuda())
    assert (dpq.rotate(x, A, impl="tc").double() - x.double() @ A.double().t()).abs().max().item() < 1e-5


def test_opq_train_tensor_core_rotation_matches_library_rotation_and_oracle():
    """opq_train rebinds A every iteration; with the tensor-core rotation (>= 1024 training rows) the
    iterates must follow the library-sgemm run and the CPU oracle (same seeds, shared recipe)."""
    from b200vq import ops, pq as dpq
    from oracle import pq as opq
    rng = np.random.default_rng(5)
    n, D, M, nbits = 3000, 64, 8, 4
    Xh = (rng.standard_normal((n, D)) * np.linspace(3.0, 0.3, D) @ rng.standard_normal((D, D)) / np.sqrt(D)).astype(np.float32)
    x = torch.from_numpy(Xh).cuda()
    kw = dict(niter=5, niter_pq=2, niter_pq_0=6, seed=9)
    A_tc = dpq.opq_train(x, M, nbits, impl="tc", **kw)
    A_lib = dpq.opq_train(x, M, nbits, impl="lib", **kw)
    A_or = opq.opq_train(Xh, M, nbits, **kw)

    def objective(A):
        A = np.asarray(A, dtype=np.float32)
        assert np.abs(A @ A.T - np.eye(D)).max() < 1e-4
        Y = opq.opq_apply(Xh, A)
        C = opq.pq_train(Y, M, nbits, niter=8, seed=1)
        return opq.pq_sse(Y, opq.pq_assign(Y, C), C)
    o_tc, o_lib, o_or, o_0 = objective(A_tc.cpu().numpy()), objective(A_lib.cpu().numpy()), objective(A_or), \
        objective(opq.random_rotation(D, 9))
    assert abs(o_tc - o_lib) <= 0.02 * o_lib and abs(o_tc - o_or) <= 0.02 * o_or
    assert o_tc < 0.97 * o_0, "training did not improve on the start rotation"
    # two quantizers of the same D fitted back to back keep their own rotations
    from haag_vq.methods.optimized_product_quantization import OptimizedProductQuantizer
    q1, q2 = OptimizedProductQuantizer(M=8, B=8), OptimizedProductQuantizer(M=8, B=8)
    q1.fit(Xh)
    c1 = q1.compress(Xh)
    q2.fit(Xh[::-1].copy() * 1.5)
    assert np.array_equal(q1.compress(Xh), c1)
    Y1 = q1.opq.apply(Xh)
    A1 = q1.opq.A.reshape(D, D)
    assert np.abs(Y1 - Xh @ A1.T).max() < 1e-4
    assert np.mean((Xh - q1.decompress(c1)) ** 2) < np.mean(Xh ** 2)


@pytest.mark.parametrize("n,K,N,tr", [(700, 64, 96, False), (700, 64, 96, True), (4096, 1536, 1536, True), (130, 96, 40, False),
                                      (1, 128, 128, False), (257, 160, 300, True)])
def test_rotate_tc_ex_fused_centering_scaling_bias_and_transposed_store(n, K, N, tr):
    """b200vq_rotate_tc_ex: y = row_scale * ((x - x_sub) W^T) + col_bias, optionally stored transposed
    (the SAQ engine's projection + rotation, ExtendedRaBitQuantizer's rotation and inverse)."""
    from b200vq import ops
    g = torch.Generator(device="cuda").manual_seed(n + K + N)
    x = torch.randn((n, K), generator=g, device="cuda") * torch.linspace(3.0, 0.3, K, device="cuda")
    w = torch.randn((N, K), generator=g, device="cuda") / K ** 0.5
    sub = torch.randn(K, generator=g, device="cuda")
    rs = torch.rand(n, generator=g, device="cuda") + 0.5
    bias = torch.randn(N, generator=g, device="cuda")
    r = ops.RotationTC(w)
    ref = ((x.double() - sub.double()) @ w.double().t()) * rs.double()[:, None] + bias.double()
    y = r(x, x_sub=sub, row_scale=rs, col_bias=bias, transpose_out=tr)
    if tr:
        assert y.shape == (N, n)
        y = y.t()
    assert _err(y, ref) < 2e-6
    # each option alone
    assert _err(r(x, x_sub=sub), (x.double() - sub.double()) @ w.double().t()) < 2e-6
    y2 = r(x, col_bias=bias, transpose_out=tr)
    assert _err(y2.t() if tr else y2, x.double() @ w.double().t() + bias.double()) < 2e-6
    assert torch.equal(r(x), r(x, transpose_out=True).t())      # same sums, other store
