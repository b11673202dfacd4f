"""Golden vectors for the 1-D Lloyd codebook builder, produced by the REFERENCE's own code
(oracle/_ref/libsaq_ref_codebook.so = external/saq/src/preprocessing/codebook_builder.cpp compiled in place).
Run in the authoring container (needs /root/reference to build the .so); writes tests/golden/codebook_lloyd.npz."""
import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def ref_lib():
    lib = ctypes.CDLL(os.path.join(ROOT, "oracle", "_ref", "libsaq_ref_codebook.so"))
    P = ctypes.c_void_p
    lib.ref_codebook_lloyd.argtypes = [P, ctypes.c_long, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                       ctypes.c_float, ctypes.c_ulonglong, ctypes.c_long, P, P, P]
    lib.ref_codebook_exact.argtypes = [P, ctypes.c_long, ctypes.c_int, P, P, P]
    return lib


def ref_lloyd(lib, values, max_bits, init=2, restarts=1, max_iters=50, tol=1e-6, seed=0, sample_size=0):
    v = np.ascontiguousarray(values, dtype=np.float32)
    cen = np.zeros((max_bits + 1) << max_bits, dtype=np.float32)
    sizes = np.zeros(max_bits + 1, dtype=np.int32)
    costs = np.zeros(max_bits + 1, dtype=np.float32)
    lib.ref_codebook_lloyd(v.ctypes.data, len(v), max_bits, init, restarts, max_iters, tol, seed, sample_size,
                           cen.ctypes.data, sizes.ctypes.data, costs.ctypes.data)
    return costs, [cen[(b << max_bits):(b << max_bits) + sizes[b]].copy() for b in range(max_bits + 1)]


def columns():
    rng = np.random.default_rng(7)
    return {
        "gauss": rng.standard_normal(5000).astype(np.float32) * 0.3,
        "skew": (rng.standard_normal(3000) ** 2).astype(np.float32),
        "bimodal": np.concatenate([rng.normal(-2, 0.2, 1500), rng.normal(1, 0.5, 2500)]).astype(np.float32),
        "ties": np.round(rng.standard_normal(4000) * 3).astype(np.float32),            # few distinct values
        "tiny": rng.standard_normal(40).astype(np.float32),
    }


if __name__ == "__main__":
    lib = ref_lib()
    out = {}
    for name, col in columns().items():
        out[f"{name}_x"] = col
        mb = 6 if name != "tiny" else 5
        costs, books = ref_lloyd(lib, col, mb)
        out[f"{name}_costs"] = costs
        for b, cb in enumerate(books):
            out[f"{name}_cb{b}"] = cb
    col = columns()["gauss"]
    costs, books = ref_lloyd(lib, col, 4, seed=3, sample_size=1000)
    out["sampled_costs"] = costs
    for b, cb in enumerate(books):
        out[f"sampled_cb{b}"] = cb
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "codebook_lloyd.npz"), **out)
    print("wrote", len(out), "arrays")
