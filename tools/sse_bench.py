"""Diagnostic: streaming-distortion kernels (b200vq_pq_sse, b200vq_sq_err_sum) at a resident batch."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "vector-quantization_b200"), ROOT]
import torch
from b200vq import ops, ops_sq, pq as dpq

n, D, M = 1 << 20, 1024, 32
dev = torch.device("cuda")
x = ops_sq.synth_rows(n, D, 0, 1, None, True, device=dev)
C = dpq.pq_train(x[:65536], M, 8, niter=3, seed=1)
codec = dpq.PQCodec(C)
codes = codec.encode(x)
y = codec.decode(codes)


def timed(fn, reps=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        r = fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps, r

t, r = timed(lambda: ops.pq_sse(x, codes, C))
print(f"pq_sse {n}x{D}: {t:.3f} ms  {n*(4*D+M)/t/1e6:.0f} GB/s  value {float(r):.6f}")
t2, r2 = timed(lambda: ops_sq.sq_err_sum(x, y))
print(f"sq_err_sum {n}x{D}: {t2:.3f} ms  {n*8*D/t2/1e6:.0f} GB/s  value {float(r2):.6f}  rel diff {abs(float(r)-float(r2))/float(r):.2e}")
