"""LVQ on the GPU (csrc/lvq.cu) against the reference's golden bytes and the pinned oracle:
bit-exact codes, reconstructions and mean (integer / byte work and fp32 arithmetic evaluated in
the reference's order). Reference: src/haag_vq/methods/lvq_quantization.py:44-141."""
import os
import pickle

import numpy as np
import pytest

from oracle import lvq as olvq

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("D", [40, 61])
def test_lvq_matches_reference_golden(golden_dir, D):
    from haag_vq.methods.lvq_quantization import LVQQuantizer
    g = np.load(os.path.join(golden_dir, f"lvq_D{D}.npz"))
    X, Xq = g["X"], g["Xq"]
    for B in (1, 2, 3, 4, 5, 8):
        q = LVQQuantizer(num_bits=B)
        with pytest.raises(RuntimeError):
            q.compress(Xq)
        q.fit(X)
        assert np.array_equal(q.mu.view(np.uint32), g["mu"].view(np.uint32))      # numpy's summation order
        codes = q.compress(Xq)
        assert codes.dtype == np.uint8 and np.array_equal(codes, g[f"codes{B}"])
        rec = q.decompress(codes)
        assert rec.dtype == np.float32 and np.array_equal(rec.view(np.uint32), g[f"rec{B}"].view(np.uint32))
        assert q.get_compression_ratio(X) == pytest.approx(float(g[f"ratio{B}"]))
        ids = np.array([7, 0, 33])
        assert np.array_equal(q.decompress(codes[ids]), rec[ids])
    with pytest.raises(ValueError):
        LVQQuantizer(num_bits=9)


@pytest.mark.parametrize("D,B", [(1536, 8), (1536, 4), (1024, 2), (387, 5), (7, 3), (3072, 1), (3072, 8), (2048, 2), (1540, 6),
                                 (512, 6)])
def test_lvq_matches_oracle_larger(D, B):
    from haag_vq.methods.lvq_quantization import LVQQuantizer
    rng = np.random.default_rng(D + B)
    X = (rng.standard_normal((3000, D)) * np.linspace(3.0, 0.1, D)).astype(np.float32)
    X /= np.linalg.norm(X, axis=1, keepdims=True)
    q = LVQQuantizer(num_bits=B)
    q.fit(X)
    mu = olvq.lvq_fit(X)
    assert np.array_equal(q.mu.view(np.uint32), mu.view(np.uint32))
    codes = q.compress(X)
    assert np.array_equal(codes, olvq.lvq_compress(X, mu, B))
    rec = q.decompress(codes)
    assert np.array_equal(rec.view(np.uint32), olvq.lvq_decompress(codes, mu, B).view(np.uint32))
    # quantisation error is bounded by half a step per dimension
    delta = codes[:, -4:].copy().view(np.float32).reshape(-1)
    assert np.all(np.abs(rec - X).max(axis=1) <= 0.5001 * delta + 1e-6)
    # idempotence: re-encoding the reconstruction reproduces the indices' reconstruction
    rec2 = q.decompress(q.compress(rec))
    assert np.abs(rec2 - rec).max() <= 1e-5
    # pickle round trip keeps the trained state
    q2 = pickle.loads(pickle.dumps(q))
    assert np.array_equal(q2.compress(X[:10]), codes[:10])


def test_lvq_registry_and_empty():
    from haag_vq.benchmarks.method_registry import build_quantizer
    X = np.random.default_rng(0).standard_normal((256, 48)).astype(np.float32)
    q = build_quantizer("lvq", 4.0, 48)
    q.fit(X)
    R = q.reconstruct(np.arange(256, dtype=np.uint32))
    assert R.shape == X.shape and R.dtype == np.float32
    assert q.code_bytes() == 256 * (48 * 4 // 8 + 8) + 4 * 256
    assert np.mean((R - X) ** 2) < np.mean(X ** 2) * 0.05
    from haag_vq.methods.lvq_quantization import LVQQuantizer
    l = LVQQuantizer(4)
    l.fit(X)
    assert l.compress(X[:0]).shape == (0, 32) and l.decompress(np.zeros((0, 32), np.uint8)).shape == (0, 48)
