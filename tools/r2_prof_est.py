"""One ext-RaBitQ encode batch + one code-domain estimator scan at C4's shape, for ncu launch lists."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "vector-quantization_b200"), ROOT]
import torch

from b200vq import ops_sq
from haag_vq.methods.extended_rabitq import ExtendedRaBitQuantizer

dev = torch.device("cuda")
D, B, nq = 1536, 4, 1000
n = int(os.environ.get("ERQ_ROWS", 1_000_000))
x = ops_sq.synth_rows(n, D, 0, 1234, torch.linspace(4.0, 0.2, D, device=dev), True, device=dev)
q = ExtendedRaBitQuantizer(B, seed=0)
q.fit(x[:4096].cpu().numpy())
st = q._state()
codes = torch.empty((n, q.code_size), dtype=torch.uint8, device=dev)
nb = 131_072
for rep in range(2):
    for s in range(0, n, nb):
        xs = x[s:s + nb]
        nrm = ops_sq.resid_norms_f64(xs, st["c"])
        y = st["fwd"](xs, x_sub=st["c32"])
        ops_sq.erabitq_pack_tc(y, xs, st["c"], st["PT"], nrm, B, st["levels"], st["bounds"], st["err_rel"],
                               2 * st["err_abs"], out=codes[s:s + nb])
        if rep == 1:
            break
Q = ops_sq.synth_rows(nq, D, 0, 77, torch.linspace(4.0, 0.2, D, device=dev), True, device=dev)
sc = ops_sq.ERaBitQScanner(codes, D, B, st["levels32"])
qrot, qconst = q.rotate_queries(Q, "l2")
for _ in range(2):
    sc.search(qrot, qconst, 10, "l2")
torch.cuda.synchronize()
print("done")
