"""One tensor-core flat search call (for ncu launch lists)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "vector-quantization_b200"), ROOT]
import torch
from b200vq import ops, ops_sq
n, D, nq, k = (int(a) for a in (sys.argv[1:5] if len(sys.argv) > 4 else (1000000, 1536, 1000, 100)))
dev = torch.device("cuda")
cs = torch.linspace(4.0, 0.2, D, device=dev)
X = ops_sq.synth_rows(n, D, 0, 1234, cs, True, device=dev)
Q = ops_sq.synth_rows(nq, D, 1 << 41, 4321, cs, True, device=dev)
xn = ops.row_sqnorms(X)
for _ in range(2):
    v, i = ops.flat_topk(X, Q, k, "l2", xnorm=xn, impl="tc")
torch.cuda.synchronize()
print("done", v[0, :3].tolist())
