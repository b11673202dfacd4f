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
