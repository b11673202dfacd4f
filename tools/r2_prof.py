"""Profiling drivers for ncu (one short workload per mode, warmed): scan = the bench's screened ADC scan over one
C5 shard; rotate = rotate_tc 262144 x 1536; caq = all-segment CAQ encode at 99 000 x 1536. Diagnostic."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "vector-quantization_b200"), ROOT]
import torch

from b200vq import ops, ops_sq, pq as dpq, saq_engine as E

mode = sys.argv[1]
dev = torch.device("cuda")


def data(n, D, seed=1234, row0=0):
    return ops_sq.synth_rows(n, D, row0, seed, torch.linspace(4.0, 0.2, D, device=dev), True, device=dev)


if mode == "scan":
    n, D, M, nq = 6_625_000, 1024, 32, 1000
    C = dpq.pq_train(data(65536, D, row0=1 << 40), M, 8, niter=4, seed=1234)
    codec = dpq.PQCodec(C)
    codes = torch.empty((n, M), dtype=torch.uint8, device=dev)
    for s in range(0, n, 1 << 20):
        e = min(n, s + (1 << 20))
        codec.encode(data(e - s, D, row0=s), out=codes[s:e])
    rot = ops.pq_codes_to_scan_layout(codes)
    Q = data(nq, D, seed=4321, row0=1 << 41)
    lut = codec.luts(Q, "l2")
    for _ in range(2):
        v, i = ops.adc_scan_topk(rot, lut, 10, "l2", n=n, rot=True, screen=True)
elif mode == "c1adc":
    # C1 search: 100 000 codes, M = 16 padded to the scan layout, 1000 queries
    n, D, M, nq = 100_000, 1536, 16, 1000
    x = data(n, D)
    C = dpq.pq_train(x[:65536], M, 8, niter=4, seed=1234)
    codec = dpq.PQCodec(C)
    crot = ops.pq_codes_to_scan_layout(codec.pad_codes_for_scan(codec.encode(x)))
    Q = data(nq, D, seed=1235)
    mp = dpq.scan_layout_width(M)
    for _ in range(3):
        v, i = codec.search(crot, Q, 10, "l2", n=n, rot=True, m_pad=mp)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20):
        v, i = codec.search(crot, Q, 10, "l2", n=n, rot=True, m_pad=mp)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 20
    # survivors per query: the screen's counters inside the scratch block (layout of plan_screen, csrc/scan.cu)
    up = lambda v: (v + 255) & ~255
    k_ = 10
    o = up(nq * k_ * 4); o = up(o + nq * k_ * 4); o = up(o + nq * 1 * k_ * 8)
    ng = (nq + 3) // 4
    o = up(o + ng * mp * 256 * 4); o = up(o + ng * 16)
    sc = ops.scratch(0, crot.device, "scan_screen")
    cnt = sc[o:o + nq * 4].view(torch.int32)
    print("survivors per query: mean %.0f  median %.0f  max %d" % (cnt.float().mean().item(), cnt.float().median().item(), cnt.max().item()))
    print(f"c1adc search {ms:.3f} ms  {n * nq / ms / 1e6:.1f} G codes/s  tune={os.environ.get('B200VQ_TUNE_SAMPLE_DIV')},"
          f"{os.environ.get('B200VQ_TUNE_SAMPLE_MIN')}")
elif mode == "c1enc":
    # C1 encode: 100 000 x 1536, M = 16 (dsub = 96)
    n, D, M = 100_000, 1536, 16
    x = data(n, D)
    C = dpq.pq_train(x[:65536], M, 8, niter=4, seed=1234)
    codec = dpq.PQCodec(C)
    codes = codec.encode(x)
    for _ in range(3):
        codec.encode(x, out=codes)
    torch.cuda.synchronize()
elif mode == "rotate":
    n, D = 262_144, 1536
    x = data(n, D)
    A = torch.from_numpy(dpq.random_rotation(D, 5)).to(dev)
    for _ in range(2):
        v = dpq.rotate(x, A, impl="tc")
elif mode == "caq":
    n, D = 99_000, 1536
    x = data(n, D)
    eng = E.SaqEngine(4.0).fit(x)
    oT = eng._project(x, transpose_out=True)
    for _ in range(2):
        v, _ = E.caq_encode_t_segments(oT, eng.plan)
elif mode == "assign":
    # C5 shape: 1 Mi x 1024, M = 32; prints the screen's statistics and the event time of the whole call
    n, D, M = 1 << 20, 1024, int(sys.argv[2]) if len(sys.argv) > 2 else 32
    x = data(n, D)
    C = dpq.pq_train(data(65536, D, row0=1 << 40), M, 8, niter=4, seed=1234)
    cn = ops.pq_centroid_norms(C)
    st = torch.zeros(8, dtype=torch.int64, device=dev)
    out = ops.pq_assign(x, C, cn, impl="tc", stats=st)
    torch.cuda.synchronize()
    print("stats pairs / candidates / fix-up entries:", st[:3].tolist(), "of", n * M)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(3):
        ops.pq_assign(x, C, cn, out=out, impl="tc")
    a.record()
    for _ in range(10):
        ops.pq_assign(x, C, cn, out=out, impl="tc")
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 10
    print(f"assign_tc {ms:.3f} ms  {2.0 * n * D * 256 / ms / 1e9:.1f} TFLOP/s")
else:
    raise SystemExit("mode: scan | rotate | caq | assign")
torch.cuda.synchronize()
print("ok", mode)
