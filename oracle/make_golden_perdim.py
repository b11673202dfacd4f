"""Golden vectors for RankAwareQuantizer (codebook="gaussian") from the REFERENCE class (authoring container
only; numpy suffices for this configuration):   python oracle/make_golden_perdim.py
Writes tests/golden/perdim_D24.npz: inputs, the reference's trained state (mu, V, var, bits, codebooks,
Lloyd-Max tables), its codes for both packings and its reconstructions; asserts oracle/perdim.py agrees."""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SRC = "/root/reference/src"


def main() -> None:
    assert os.path.isdir(REF_SRC), "reference tree not present"
    sys.path.insert(0, REF_SRC)
    sys.path.insert(1, ROOT)
    from haag_vq.methods.rank_aware_quantization import RankAwareQuantizer as Ref
    from oracle import perdim as opd

    rng = np.random.default_rng(20260513)
    D = 24
    A = np.diag(np.linspace(4.0, 0.2, D)) @ rng.standard_normal((D, D))
    X = (rng.standard_normal((600, D)) @ A).astype(np.float32)
    out = {"X": X, "source": "reference RankAwareQuantizer(codebook='gaussian')"}
    for packing in ("dense", "ffd"):
        q = Ref(avg_bits=3.5, alpha=0.5, packing=packing)
        q.fit(X)
        codes = q.compress(X)
        rec = q.decompress(codes)
        assert np.array_equal(codes, opd.compress(X, q.mu, q.V, q.bits, q.cb, packing)), packing
        assert np.array_equal(rec, opd.decompress(codes, q.mu, q.V, q.bits, q.cb, packing)), packing
        out.update({"mu": q.mu, "V": q.V, "var": q.var, "bits": q.bits, "Dg": q._Dg,
                    "cb_flat": np.concatenate(q.cb), "cb_len": np.array([len(c) for c in q.cb]),
                    f"codes_{packing}": codes, f"rec_{packing}": rec, f"code_size_{packing}": np.int64(q.code_size)})
        for b in range(q.max_bits + 1):
            out[f"levels{b}"] = q._levels[b]
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "perdim_D24.npz"), **out)
    print("wrote perdim_D24.npz; bits:", out["bits"].tolist())


if __name__ == "__main__":
    main()
