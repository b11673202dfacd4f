"""Golden vectors for FFD packing from the REFERENCE module (authoring container only):
    python oracle/make_golden_ffd.py
Loads src/haag_vq/methods/ffd_packing.py from /root/reference by path (numpy only), checks oracle/ffd.py
against it and writes tests/golden/ffd_kat.npz."""
from __future__ import annotations

import importlib.util
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference/src/haag_vq/methods/ffd_packing.py"


def main() -> None:
    assert os.path.isfile(REF), "reference tree not present"
    spec = importlib.util.spec_from_file_location("ref_ffd_packing", REF)
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    sys.path.insert(0, ROOT)
    from oracle import ffd as offd

    rng = np.random.default_rng(20260512)
    out = {"source": "reference ffd_packing.py"}
    cases = [np.array([4, 3, 3, 2, 2, 2]), np.array([4, 4, 3, 3, 2]), rng.integers(0, 9, size=300),
             rng.integers(1, 9, size=97), np.full(40, 5), rng.integers(0, 5, size=64)]
    for i, b in enumerate(cases):
        b = b.astype(np.int64)
        bi, bo, nb = ref.ffd_layout(b)
        o = offd.ffd_layout(b)
        assert np.array_equal(bi, o[0]) and np.array_equal(bo, o[1]) and nb == o[2], i
        codes = np.zeros((33, len(b)), dtype=np.int64)
        for d in range(len(b)):
            if b[d]:
                codes[:, d] = rng.integers(0, 1 << int(b[d]), size=33)
        packed = ref.ffd_encode(codes, b, bi, bo, nb)
        assert np.array_equal(packed, offd.ffd_encode(codes, b, bi, bo, nb))
        assert np.array_equal(ref.ffd_decode(packed, b, bi, bo, len(b)), offd.ffd_decode(packed, b, bi, bo, len(b)))
        assert np.array_equal(ref.ffd_decode(packed, b, bi, bo, len(b)), codes)
        out.update({f"bits{i}": b, f"byte_idx{i}": bi, f"bit_off{i}": bo, f"n_bytes{i}": np.int64(nb),
                    f"codes{i}": codes.astype(np.uint8), f"packed{i}": packed})
    out["ncases"] = np.int64(len(cases))
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "ffd_kat.npz"), **out)
    print("wrote ffd_kat.npz; n_bytes:", [int(out[f"n_bytes{i}"]) for i in range(len(cases))])


if __name__ == "__main__":
    main()
