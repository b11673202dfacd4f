"""Golden fixtures for LVQ from the REFERENCE's own class (authoring container only).

    python oracle/make_golden_lvq.py

Imports src/haag_vq/methods/lvq_quantization.py (numpy only) from /root/reference, runs
fit / compress / decompress on seeded inputs for several (D, num_bits), asserts that oracle/lvq.py
reproduces every byte (this pins the oracle), and writes tests/golden/lvq_D{D}.npz.
"""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SRC = "/root/reference/src"
OUT = os.path.join(ROOT, "tests", "golden")


def main() -> None:
    assert os.path.isdir(REF_SRC), "reference tree not present"
    sys.path.insert(0, REF_SRC)
    sys.path.insert(1, ROOT)
    from haag_vq.methods.lvq_quantization import LVQQuantizer as RefLVQ
    from oracle import lvq as olvq

    rng = np.random.default_rng(20260511)
    for D in (40, 61):                       # 61: D * B is not a multiple of 8 for odd B (ragged last byte)
        X = (rng.standard_normal((257, D)) * np.linspace(2.5, 0.05, D) + 0.3).astype(np.float32)
        X[5] = X[5, 0]                       # constant row: span == 0 -> delta = float32 tiny
        X[6] = 0.0
        Xq = X[:96].copy()
        out = {"X": X, "Xq": Xq, "source": "reference LVQQuantizer (lvq_quantization.py)"}
        for B in (1, 2, 3, 4, 5, 8):
            ref = RefLVQ(num_bits=B)
            ref.fit(X)
            codes = ref.compress(Xq)
            rec = ref.decompress(codes)
            mu = olvq.lvq_fit(X)
            assert np.array_equal(mu.view(np.uint32), ref.mu.view(np.uint32)), "LVQ mu"
            assert np.array_equal(codes, olvq.lvq_compress(Xq, mu, B)), f"LVQ compress B={B}"
            assert np.array_equal(rec.view(np.uint32), olvq.lvq_decompress(codes, mu, B).view(np.uint32)), f"LVQ dec B={B}"
            out["mu"] = ref.mu
            out[f"codes{B}"] = codes
            out[f"rec{B}"] = rec.astype(np.float32)
            out[f"ratio{B}"] = np.float64(ref.get_compression_ratio(X))
        np.savez_compressed(os.path.join(OUT, f"lvq_D{D}.npz"), **out)
        print("wrote", f"lvq_D{D}.npz")


if __name__ == "__main__":
    main()
