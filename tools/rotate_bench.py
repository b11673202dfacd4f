"""Timing + accuracy of the tensor-core rotation GEMM against the strict-fp32 library sgemm
(config C2 shape: 1M x 1536 rotated by a 1536 x 1536 matrix, in batches). Diagnostic, not bench.py."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "vector-quantization_b200"), ROOT]
import torch
from b200vq import ops, pq as dpq

n = int(sys.argv[1]) if len(sys.argv) > 1 else 262144
D = int(sys.argv[2]) if len(sys.argv) > 2 else 1536
dev = torch.device("cuda")
g = torch.Generator(device=dev).manual_seed(1)
x = torch.randn((n, D), generator=g, device=dev)
A = torch.from_numpy(dpq.random_rotation(D, 5)).to(dev)


def timed(fn, reps=3):
    fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


rot = ops.RotationTC(A)
y = torch.empty((n, D), device=dev)
t_tc = timed(lambda: rot(x, out=y))
t_lib = timed(lambda: dpq.rotate(x, A, impl="lib"))
fl = 2.0 * n * D * D
print(f"rotate {n}x{D}: tc {t_tc:.2f} ms = {fl/t_tc/1e9:.1f} TFLOP/s useful ({3*fl/t_tc/1e9:.1f} issued) | "
      f"library sgemm {t_lib:.2f} ms = {fl/t_lib/1e9:.1f} TFLOP/s")
m = min(n, 8192)
ref = x[:m].double() @ A.double().t()
sc = ref.norm(dim=1, keepdim=True)
print("max |err| / |row|: tc %.3g, library %.3g" % (((y[:m].double() - ref).abs() / sc).max().item(),
                                                   ((dpq.rotate(x[:m], A, impl="lib").double() - ref).abs() / sc).max().item()))
print("rms err / rms y: tc %.3g, library %.3g" % (((y[:m].double() - ref).pow(2).mean().sqrt() / ref.pow(2).mean().sqrt()).item(),
      ((dpq.rotate(x[:m], A, impl="lib").double() - ref).pow(2).mean().sqrt() / ref.pow(2).mean().sqrt()).item()))
