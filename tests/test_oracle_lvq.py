"""oracle/lvq.py against the golden vectors produced by the reference's own LVQQuantizer
(oracle/make_golden_lvq.py; reference src/haag_vq/methods/lvq_quantization.py:44-141). CPU only."""
import os

import numpy as np
import pytest

from oracle import lvq as olvq


@pytest.mark.parametrize("D", [40, 61])
def test_oracle_reproduces_reference_bytes(golden_dir, D):
    g = np.load(os.path.join(golden_dir, f"lvq_D{D}.npz"))
    X, Xq = g["X"], g["Xq"]
    mu = olvq.lvq_fit(X)
    assert np.array_equal(mu.view(np.uint32), g["mu"].view(np.uint32))
    for B in (1, 2, 3, 4, 5, 8):
        codes = olvq.lvq_compress(Xq, mu, B)
        assert codes.shape == (Xq.shape[0], olvq.lvq_code_size(D, B))
        assert np.array_equal(codes, g[f"codes{B}"])
        assert np.array_equal(olvq.lvq_decompress(codes, mu, B).view(np.uint32), g[f"rec{B}"].view(np.uint32))
        # row independence (reference tests/test_lvq_quantization.py: decompress(codes[ids]) == decompress(codes)[ids])
        ids = np.array([7, 0, 33])
        assert np.array_equal(olvq.lvq_decompress(codes[ids], mu, B), olvq.lvq_decompress(codes, mu, B)[ids])
