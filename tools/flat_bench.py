"""Diagnostic: exact flat top-k (ground truth) and row-layout ADC scan timings."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "vector-quantization_b200"), ROOT]
import torch
from b200vq import ops, ops_sq

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
D = int(sys.argv[2]) if len(sys.argv) > 2 else 1536
nq = int(sys.argv[3]) if len(sys.argv) > 3 else 1000
k = int(sys.argv[4]) if len(sys.argv) > 4 else 100
dev = torch.device("cuda")
cs = torch.linspace(4.0, 0.2, D, device=dev)
X = ops_sq.synth_rows(n, D, 0, 1234, cs, True, device=dev)
Q = ops_sq.synth_rows(nq, D, 1 << 41, 4321, cs, True, device=dev)


def timed(fn, reps=2):
    fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


xn = ops.row_sqnorms(X)
res = {}
for impl in ("simt", "tc"):
    for metric in ("l2", "ip"):
        ms = timed(lambda: ops.flat_topk(X, Q, k, metric, xnorm=xn, impl=impl))
        res[(impl, metric)] = ops.flat_topk(X, Q, k, metric, xnorm=xn, impl=impl)
        print(f"flat_topk[{impl}] {metric} {n}x{D} nq={nq} k={k}: {ms:.1f} ms = {2.0 * n * D * nq / ms / 1e9:.1f} TFLOP/s")
for metric in ("l2", "ip"):
    print(metric, "tc == simt:", torch.equal(res[("simt", metric)][1], res[("tc", metric)][1]),
          torch.equal(res[("simt", metric)][0], res[("tc", metric)][0]))
if os.environ.get("FLAT_ONLY"):
    sys.exit(0)
g = torch.Generator(device="cuda").manual_seed(1)
for M in (16, 32):
    codes = torch.randint(0, 256, (n, M), dtype=torch.uint8, device=dev, generator=g)
    lut = torch.rand((nq, M, 256), device=dev, generator=g)
    ms = timed(lambda: ops.adc_scan_topk(codes, lut, 10, "l2"))
    print(f"row-layout adc scan M={M}: {ms:.1f} ms = {n * nq / ms / 1e6:.1f} G codes/s")
