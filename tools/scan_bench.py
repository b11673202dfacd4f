"""Diagnostic: exact scan-layout kernel vs the screened scan at the bench shape."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "vector-quantization_b200"), ROOT]
import torch
from b200vq import ops, ops_sq, pq as dpq

n = int(sys.argv[1]) if len(sys.argv) > 1 else 6_625_000
nq = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
D, M, k = 1024, 32, 10
dev = torch.device("cuda")
cs = torch.linspace(4.0, 0.2, D, device=dev)
C = dpq.pq_train(ops_sq.synth_rows(65536, D, 1 << 40, 1234, cs, True, device=dev), M, 8, niter=4, seed=1234)
codec = dpq.PQCodec(C)
codes = torch.empty((n, M), dtype=torch.uint8, device=dev)
for s in range(0, n, 1 << 20):
    e = min(n, s + (1 << 20))
    codec.encode(ops_sq.synth_rows(e - s, D, s, 1234, cs, True, device=dev), out=codes[s:e])
rot = ops.pq_codes_to_scan_layout(codes)
Q = ops_sq.synth_rows(nq, D, 1 << 41, 4321, cs, True, device=dev)
lut = codec.luts(Q, "l2")
res = {}
for name, kw in (("exact", {}), ("screen", dict(screen=True))):
    v, i = ops.adc_scan_topk(rot, lut, k, "l2", n=n, rot=True, **kw)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(3):
        v, i = ops.adc_scan_topk(rot, lut, k, "l2", n=n, rot=True, **kw)
    b.record(); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 3
    res[name] = (v, i)
    print(f"{name}: {ms:.2f} ms  {n * nq / ms / 1e6:.1f} G codes/s")
print("ids equal:", torch.equal(res["exact"][1], res["screen"][1]),
      "max |dv|:", (res["exact"][0] - res["screen"][0]).abs().max().item())
d = res["exact"][1] != res["screen"][1]
ve, vs = res["exact"][0], res["screen"][0]
print("positions differing:", int(d.sum()), "of", d.numel(), "| max rel value gap there:",
      ((ve - vs).abs() / vs.abs().clamp_min(1e-30))[d].max().item() if d.any() else 0.0)
vr, ir = ops.adc_scan_topk(codes, lut, k, "l2")          # row layout: sub-quantizers summed in increasing order
print("screen vs row-layout exact kernel: ids equal", torch.equal(ir, res["screen"][1]), "values equal",
      torch.equal(vr, res["screen"][0]))
