"""Generate the golden fixtures under tests/golden/ from the REFERENCE's own code.

Run in the authoring container only (needs /root/reference):

    python oracle/make_golden.py

What it does
  1. imports the reference classes that run here without faiss
     (ScalarQuantizer, ExtendedRaBitQuantizer, exact_search helpers,
     method_registry helpers) from /root/reference/src;
  2. runs them on small seeded inputs and checks that the oracle restatements
     (oracle/sq.py, oracle/rabitq.py) reproduce them bit for bit — this is what
     PINS those oracles;
  3. writes inputs + reference outputs as .npz fixtures (a few hundred KB) that
     travel to the GPU box, where /root/reference does not exist.

The faiss-backed paths (PQ / OPQ / 1-bit RaBitQ / IndexFlat) cannot be run here
(faiss is absent), so their fixtures are produced by the oracle restatement and
are labelled ``source='oracle (parity unpinned)'`` inside the file.
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
    from haag_vq.methods.scalar_quantization import ScalarQuantizer as RefSQ
    from haag_vq.methods.extended_rabitq import ExtendedRaBitQuantizer as RefERQ
    from haag_vq.benchmarks import exact_search as ref_es
    from haag_vq.benchmarks import method_registry as ref_reg
    from oracle import pq as opq
    from oracle import rabitq as orq
    from oracle import sq as osq

    os.makedirs(OUT, exist_ok=True)
    rng = np.random.default_rng(20240611)

    # ---------------------------------------------------------------- SQ
    for D in (48, 33):
        X = (rng.standard_normal((300, D)) * np.linspace(3.0, 0.1, D)).astype(np.float32)
        Xq = X[:64].copy()  # rows inside the fit range
        out = {"X": X, "Xq": Xq, "source": "reference ScalarQuantizer (scalar_quantization.py)"}
        for bits in (4, 8, 16):
            ref = RefSQ(num_bits=bits)
            ref.fit(X)
            codes = ref.compress(Xq)
            rec = ref.decompress(codes)
            assert np.array_equal(codes, osq.sq_compress(Xq, ref.min, ref.max, bits)), f"SQ{bits} compress"
            assert np.array_equal(rec, osq.sq_decompress(codes, ref.min, ref.max, bits)), f"SQ{bits} decompress"
            out[f"codes{bits}"] = codes
            out[f"recon{bits}"] = np.asarray(rec, dtype=np.float32)
            out["vmin"], out["vmax"] = ref.min, ref.max
        np.savez_compressed(os.path.join(OUT, f"sq_D{D}.npz"), **out)
    print("SQ oracle pinned against the reference (4/8/16 bit, D=48 and odd D=33)")

    # ---------------------------------------------------------------- Extended RaBitQ
    D = 64
    X = rng.standard_normal((256, D)).astype(np.float32)
    X /= np.linalg.norm(X, axis=1, keepdims=True)
    out = {"X": X, "source": "reference ExtendedRaBitQuantizer (extended_rabitq.py)"}
    for B in (1, 2, 3, 4, 8):
        ref = RefERQ(num_bits=B, seed=0)
        ref.fit(X)
        codes = ref.compress(X[:96])
        rec = ref.decompress(codes)
        st = orq.erabitq_fit(X, B, seed=0)
        assert np.array_equal(st["levels"], ref.levels) and np.array_equal(st["P"], ref.P)
        assert np.array_equal(codes, orq.erabitq_compress(X[:96], st)), f"ERQ B={B} compress"
        assert np.array_equal(rec, orq.erabitq_decompress(codes, st)), f"ERQ B={B} decompress"
        out[f"codes{B}"] = codes
        out[f"recon{B}"] = rec
        out[f"levels{B}"] = ref.levels
        out["P"], out["c"] = ref.P, ref.c
    np.savez_compressed(os.path.join(OUT, "erabitq_D64.npz"), **out)
    print("Extended-RaBitQ oracle pinned against the reference (B=1,2,3,4,8)")

    # ---------------------------------------------------------------- exact_search helpers / registry KATs
    Xn = rng.standard_normal((200, 24)).astype(np.float32)
    norms = ref_es.compute_exact_norms(Xn)
    kat = {
        "X": Xn, "norms": norms,
        "recall_ret": np.array([[0, 1, 2]]), "recall_gt": np.array([[0, 9, 2]]),
        "recall_1": ref_es.recall_at_ks(np.array([[0, 1, 2]]), np.array([[0, 9, 2]]), ks=(1, 3))[1],
        "recall_3": ref_es.recall_at_ks(np.array([[0, 1, 2]]), np.array([[0, 9, 2]]), ks=(1, 3))[3],
        "divisor_cases": np.array([[1536, 600, ref_reg.largest_divisor_leq(1536, 600)],
                                   [30, 7, ref_reg.largest_divisor_leq(30, 7)],
                                   [1024, 128, ref_reg.largest_divisor_leq(1024, 128)],
                                   [7, 100, ref_reg.largest_divisor_leq(7, 100)],
                                   [12, 0, ref_reg.largest_divisor_leq(12, 0)]]),
        "pq_M_cases": np.array([[bpd10, Dd, ref_reg._pq_subquantizers(bpd10 / 10.0, Dd)]
                                for bpd10 in (5, 10, 20, 40, 80) for Dd in (48, 1024, 1536, 3072)]),
        "mse_zero": ref_es.reconstruction_mse(Xn, lambda ids: Xn[ids], np.arange(50)),
        "mse_one": ref_es.reconstruction_mse(Xn, lambda ids: Xn[ids] + 1.0, np.arange(50)),
        "source": "reference exact_search.py / method_registry.py",
    }
    np.savez_compressed(os.path.join(OUT, "helpers_kat.npz"), **kat)
    print("exact_search / registry known answers recorded")

    # ---------------------------------------------------------------- faiss-backed paths: oracle only
    D, M = 64, 8
    X = rng.standard_normal((2048, D)).astype(np.float32)
    Q = rng.standard_normal((16, D)).astype(np.float32)
    C = opq.pq_train(X, M, 8, niter=4, seed=7)
    codes, d1, d2 = opq.pq_assign(X, C, return_gap=True)
    rv, ri = opq.adc_search(codes, C, Q, 10, opq.L2)
    fv, fi = opq.flat_search(X, Q, 10, opq.L2)
    r1 = orq.rabitq1_encode(X[:256], X.mean(0), orq.L2)
    np.savez_compressed(os.path.join(OUT, "pq_oracle_D64.npz"), X=X, Q=Q, C=C, codes=codes, d1=d1, d2=d2,
                        adc_vals=rv, adc_ids=ri, flat_vals=fv, flat_ids=fi, rabitq1_codes=r1,
                        rabitq1_est=orq.rabitq1_estimate(r1, D, Q, X.mean(0), 4, orq.L2),
                        source="oracle (parity unpinned: faiss absent)")
    print("PQ / ADC / flat / RaBitQ-1 regression fixture written (oracle-generated, unpinned)")


if __name__ == "__main__":
    main()
