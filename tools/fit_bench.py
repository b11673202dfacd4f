"""Diagnostic: wall-clock of the drop-in fit / compress / decompress calls (host arrays in, numpy out)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "vector-quantization_b200"), ROOT]
import numpy as np
import torch
from haag_vq.methods.product_quantization import ProductQuantizer
from haag_vq.methods.optimized_product_quantization import OptimizedProductQuantizer
from haag_vq.methods.scalar_quantization import ScalarQuantizer
from haag_vq.methods.lvq_quantization import LVQQuantizer
from haag_vq.methods.extended_rabitq import ExtendedRaBitQuantizer

n, D = 100_000, 1536
rng = np.random.default_rng(0)
X = (rng.standard_normal((n, D), dtype=np.float32) * np.linspace(4.0, 0.2, D, dtype=np.float32))
X /= np.linalg.norm(X, axis=1, keepdims=True)


def t(fn):
    torch.cuda.synchronize(); a = time.perf_counter(); r = fn(); torch.cuda.synchronize()
    return time.perf_counter() - a, r

for name, q in (("PQ M=16", ProductQuantizer(M=16, B=8)), ("OPQ M=32", OptimizedProductQuantizer(M=32, B=8)),
                ("SQ 8", ScalarQuantizer(8)), ("LVQ 8", LVQQuantizer(8)), ("ExtRaBitQ B=4", ExtendedRaBitQuantizer(4))):
    q.fit(X[:4096]) if name.startswith("Ext") else None
    tf, _ = t(lambda: q.fit(X))
    tc, codes = t(lambda: q.compress(X))
    tc2, codes = t(lambda: q.compress(X))
    td, _ = t(lambda: q.decompress(codes))
    print(f"{name:14s} fit {tf*1e3:9.1f} ms   compress {tc2*1e3:8.1f} ms   decompress {td*1e3:8.1f} ms   ({n}x{D}, host arrays)")
