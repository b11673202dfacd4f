"""FFD layout (host logic of haag_vq/methods/ffd_packing.py) and the oracle against the reference's golden
layouts and known answers (reference tests/test_ffd_packing.py:5-72). CPU only."""
import os

import numpy as np
import pytest

from haag_vq.methods.ffd_packing import byte_cap_minus, ffd_layout
from oracle import ffd as offd


def test_layout_matches_reference_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "ffd_kat.npz"))
    for i in range(int(g["ncases"])):
        b = g[f"bits{i}"]
        for fn in (ffd_layout, offd.ffd_layout):
            bi, bo, nb = fn(b)
            assert np.array_equal(bi, g[f"byte_idx{i}"]) and np.array_equal(bo, g[f"bit_off{i}"])
            assert nb == int(g[f"n_bytes{i}"])
        assert np.array_equal(offd.ffd_encode(g[f"codes{i}"], b, bi, bo, nb), g[f"packed{i}"])
        assert np.array_equal(offd.ffd_decode(g[f"packed{i}"], b, bi, bo, len(b)), g[f"codes{i}"])
    assert int(g["n_bytes0"]) == 2 and int(g["n_bytes1"]) == 2      # the reference's orphaned-4 known answers


def test_layout_properties_and_other_caps():
    rng = np.random.default_rng(0)
    for cap in (8, 8, 8, 6, 5, 3):
        b = rng.integers(0, cap + 1, size=300)
        bi, bo, nb = ffd_layout(b, cap)
        o = offd.ffd_layout(b, cap)
        assert np.array_equal(bi, o[0]) and np.array_equal(bo, o[1]) and nb == o[2]
        assert np.all(bi[b == 0] == -1)
        assert np.all((bo[b > 0] >= 0) & (bo[b > 0] + b[b > 0] <= cap))
        assert int(np.ceil(b.sum() / cap)) <= nb <= int((b > 0).sum())
        for byte in range(nb):
            iv = sorted((bo[d], bo[d] + b[d]) for d in np.nonzero(bi == byte)[0])
            assert all(e1 <= s2 for (_, e1), (s2, _) in zip(iv, iv[1:]))
    assert byte_cap_minus(3, 2) == 3
    with pytest.raises(ValueError):
        ffd_layout(np.array([9]))
