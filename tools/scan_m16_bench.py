"""ADC search at M = 16 (BASELINE config 1's sub-quantizer count): row-layout exact kernel vs codes padded to
the scan layout (screened scan), same ids and value bits. usage: python tools/scan_m16_bench.py [n] [nq]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "vector-quantization_b200"), ROOT]
import torch
from b200vq import ops, ops_sq, pq as dpq

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
nq = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
D, M = 1536, 16
dev = torch.device("cuda")
cs = torch.linspace(4.0, 0.2, D, device=dev)
C = dpq.pq_train(ops_sq.synth_rows(65536, D, 1 << 40, 1234, cs, True, device=dev), M, 8, niter=4, seed=1234)
codec = dpq.PQCodec(C)
codes = torch.empty((n, M), dtype=torch.uint8, device=dev)
for s in range(0, n, 1 << 18):
    e = min(n, s + (1 << 18))
    codec.encode(ops_sq.synth_rows(e - s, D, s, 1234, cs, True, device=dev), out=codes[s:e])
q = ops_sq.synth_rows(nq, D, 1 << 41, 99, cs, True, device=dev)
crot = ops.pq_codes_to_scan_layout(codec.pad_codes_for_scan(codes))


def timed(fn, reps=5):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        r = fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps, r

t0, (v0, i0) = timed(lambda: codec.search(codes, q, 10, "l2"))
t1, (v1, i1) = timed(lambda: codec.search(crot, q, 10, "l2", n=n, rot=True))
print(f"n={n} nq={nq} M=16: row layout {t0:.3f} ms ({n*nq/t0/1e6:.1f} G codes/s)  padded scan layout {t1:.3f} ms "
      f"({n*nq/t1/1e6:.1f} G codes/s)  identical={bool(torch.equal(i0, i1) and torch.equal(v0.view(torch.int32), v1.view(torch.int32)))}")
if not torch.equal(i0, i1) or not torch.equal(v0.view(torch.int32), v1.view(torch.int32)):
    di = (i0 != i1); dv = (v0.view(torch.int32) != v1.view(torch.int32))
    print("id mismatches", int(di.sum()), "value-bit mismatches", int(dv.sum()), "max |dv|", float((v0 - v1).abs().max()))
    r = torch.nonzero(di | dv)[:5]
    for a, b in r.tolist():
        print(a, b, int(i0[a, b]), int(i1[a, b]), float(v0[a, b]), float(v1[a, b]))
    # which one agrees with the per-row exact score?
    lut = codec.luts(q, "l2")
    for a, b in r.tolist()[:3]:
        for ii in (int(i0[a, b]), int(i1[a, b])):
            sc = lut[a, torch.arange(M, device=dev), codes[ii].long()]
            acc = torch.zeros((), device=dev)
            for m in range(M):
                acc = acc + sc[m]
            print("   row", ii, "sequential fp32 sum", float(acc))
