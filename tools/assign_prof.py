"""Profiling driver: a few plain tcgen05 assign calls on bench-like data (for ncu -k regex:assign_tc)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "vector-quantization_b200"), ROOT]
import torch
from b200vq import ops, ops_sq, pq as dpq

D, M = int(sys.argv[1]) if len(sys.argv) > 1 else 1024, int(sys.argv[2]) if len(sys.argv) > 2 else 32
n = int(sys.argv[3]) if len(sys.argv) > 3 else 1 << 20
dev = torch.device("cuda")
cs = torch.linspace(4.0, 0.2, D, device=dev)
xt = ops_sq.synth_rows(65536, D, 1 << 40, 1234, cs, True, device=dev)
C = dpq.pq_train(xt, M, 8, niter=4, seed=1234, impl="simt")
x = ops_sq.synth_rows(n, D, 0, 1234, cs, True, device=dev)
codes = ops.pq_assign(x, C, impl="tc")
for _ in range(3):
    ops.pq_assign(x, C, impl="tc", out=codes)
torch.cuda.synchronize()
print("ok", int(codes.sum()))
