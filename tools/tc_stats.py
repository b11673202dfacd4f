"""Diagnostic: candidate statistics and timing of the tcgen05 assign kernel on bench-like data."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "vector-quantization_b200"), ROOT]
import torch
from b200vq import ops, ops_sq, pq as dpq

D, M = int(sys.argv[1]) if len(sys.argv) > 1 else 1024, int(sys.argv[2]) if len(sys.argv) > 2 else 32
n = int(sys.argv[3]) if len(sys.argv) > 3 else 1 << 20
dev = torch.device("cuda")
cs = torch.linspace(4.0, 0.2, D, device=dev)
xt = ops_sq.synth_rows(65536, D, 1 << 40, 1234, cs, True, device=dev)
C = dpq.pq_train(xt, M, 8, niter=8, seed=1234, impl=os.environ.get("TRAIN_IMPL", "auto"))
x = ops_sq.synth_rows(n, D, 0, 1234, cs, True, device=dev)
stats = torch.zeros(8, dtype=torch.int64, device=dev)
codes = ops.pq_assign(x, C, impl="tc", stats=stats)
torch.cuda.synchronize()
st = stats.tolist()
print("mma-thread cycles (sum over CTAs): a_full wait, t_empty wait, b_full wait, total:", st[3:7])
print(f"pairs={n*M} re-evaluated={st[0]} ({st[0]/(n*M):.4f}) candidates={st[1]} fixup={st[2]} ({st[2]/(n*M):.5f})")
for impl in ("tc", "simt"):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(3):
        ops.pq_assign(x, C, impl=impl, out=codes)
    b.record(); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 3
    print(f"{impl}: {ms:.3f} ms  {n/ms/1e3:.1f} M vec/s  {2*D*256*n/ms/1e9:.1f} TFLOP/s")
