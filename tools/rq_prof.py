"""One tensor-core RaBitQ scan call at config C4 (for launch lists under ncu). Diagnostic."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "vector-quantization_b200"), ROOT]
import torch
from b200vq import ops_sq
n, D, nq = 1_000_000, 1536, 1000
dev = torch.device("cuda")
cs = torch.linspace(4.0, 0.2, D, device=dev)
x = ops_sq.synth_rows(n, D, 0, 1234, cs, True, device=dev)
q = ops_sq.synth_rows(nq, D, 0, 1235, cs, True, device=dev)
cen = x[:65536].mean(0).contiguous()
codes = ops_sq.rabitq1_encode(x, cen, "l2")
planes, qfac = ops_sq.rabitq1_prepare_queries(q, cen, 4)
for _ in range(2):
    v, i = ops_sq.rabitq1_scan_topk(codes, D, planes, qfac, 4, 10, "l2")
torch.cuda.synchronize()
print("ok", v[0, :3].tolist())
