"""FFD pack / unpack kernels (csrc/ffd.cu) against the reference's golden bytes and the pinned oracle;
the reference's own round-trip tests (tests/test_ffd_packing.py:37-60) re-stated. Bit-exact."""
import os

import numpy as np
import pytest

from oracle import ffd as offd

pytestmark = pytest.mark.gpu


def test_ffd_matches_reference_golden(golden_dir):
    from haag_vq.methods.ffd_packing import ffd_decode, ffd_encode, ffd_layout
    g = np.load(os.path.join(golden_dir, "ffd_kat.npz"))
    for i in range(int(g["ncases"])):
        b = g[f"bits{i}"]
        bi, bo, nb = ffd_layout(b)
        packed = ffd_encode(g[f"codes{i}"].astype(np.int64), b, bi, bo, nb)
        assert packed.dtype == np.uint8 and np.array_equal(packed, g[f"packed{i}"])
        out = ffd_decode(packed, b, bi, bo, len(b))
        assert out.dtype == np.int64 and np.array_equal(out, g[f"codes{i}"])


@pytest.mark.parametrize("D,N,cap", [(256, 1000, 8), (1536, 20000, 8), (40, 50, 8), (77, 301, 6), (5, 3, 8)])
def test_ffd_round_trip_and_oracle(D, N, cap):
    from haag_vq.methods.ffd_packing import ffd_decode, ffd_encode, ffd_layout
    rng = np.random.default_rng(D + N)
    b = rng.integers(0, cap + 1, size=D)
    codes = np.zeros((N, D), dtype=np.int64)
    for d in range(D):
        if b[d] > 0:
            codes[:, d] = rng.integers(0, 1 << int(b[d]), size=N)
    bi, bo, nb = ffd_layout(b, cap)
    packed = ffd_encode(codes, b, bi, bo, nb, cap)
    assert packed.shape == (N, nb)
    assert np.array_equal(packed, offd.ffd_encode(codes, b, bi, bo, nb, cap))
    assert np.array_equal(ffd_decode(packed, b, bi, bo, D, cap), codes)


def test_ffd_all_widths_and_empty():
    from haag_vq.methods.ffd_packing import ffd_decode, ffd_encode, ffd_layout
    for width in range(1, 9):
        b = np.full(40, width)
        codes = np.random.default_rng(width).integers(0, 1 << width, size=(50, 40)).astype(np.int64)
        bi, bo, nb = ffd_layout(b)
        assert np.array_equal(ffd_decode(ffd_encode(codes, b, bi, bo, nb), b, bi, bo, 40), codes)
    b = np.array([3, 0, 5])
    bi, bo, nb = ffd_layout(b)
    assert ffd_encode(np.zeros((0, 3), np.int64), b, bi, bo, nb).shape == (0, nb)
    z = np.zeros(4, np.int64)
    bi, bo, nb = ffd_layout(z)
    assert nb == 0 and ffd_encode(np.zeros((7, 4), np.int64), z, bi, bo, nb).shape == (7, 0)
    assert np.array_equal(ffd_decode(np.zeros((7, 0), np.uint8), z, bi, bo, 4), np.zeros((7, 4), np.int64))
