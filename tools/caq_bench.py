"""Timing of the SAQ-engine encode path on one B200 at the reference's own recorded size
(99 000 x 1536, run_remaining_log.txt:13-15 / gpu-encode-analysis.md:45-49) and at config C3
(SAQ rows of 3072 dims), next to the REFERENCE's CUDA encoder (oracle/_ref/libsaq_ref_gpu.so,
sequential and block-Jacobi modes) on the same rotated segments. Diagnostic, not bench.py."""
import ctypes
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "vector-quantization_b200"), ROOT]
import numpy as np
import torch

from b200vq import ops_sq, saq_engine as E

n = int(sys.argv[1]) if len(sys.argv) > 1 else 99_000
D = int(sys.argv[2]) if len(sys.argv) > 2 else 1536
avg_bits = float(sys.argv[3]) if len(sys.argv) > 3 else 4.0
dev = torch.device("cuda")
cs = torch.linspace(4.0, 0.2, D, device=dev)
X = ops_sq.synth_rows(n, D, 0, 1234, cs, True, device=dev)


def timed(fn, reps=3):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


t0 = time.time()
eng = E.SaqEngine(avg_bits).fit(X)
torch.cuda.synchronize()
print(f"fit {n}x{D} avg_bits={avg_bits}: {time.time() - t0:.2f} s wall (PCA + plan + rotate + encode + pack)")
print("plan", eng.plan, "recon_mse/dim", eng.recon_mse())

ref = None
path = os.path.join(ROOT, "oracle", "_ref", "libsaq_ref_gpu.so")
if os.path.exists(path):
    ref = ctypes.CDLL(path)
    P = ctypes.c_void_p
    ref.ref_saq_gpu_caq_encode.argtypes = [P] * 9 + [ctypes.c_size_t] * 3 + [ctypes.c_int, ctypes.c_float, ctypes.c_int]

Xp = eng._project(X)
tot_mine = tot_seq = tot_jac = 0.0
for seg, Pm in zip(eng.segs, eng.P):
    ln, bits, off = seg["len"], seg["bits"], seg["off"]
    if bits == 0:
        continue
    oT = E._mm(Pm.t(), Xp[:, off:off + ln].t())
    t_enc = timed(lambda: E.caq_encode_t(oT, bits, seg["c_rot"]))
    codesT, _ = E.caq_encode_t(oT, bits, seg["c_rot"])
    t_pack = timed(lambda: E.caq_pack_t(codesT, bits))
    tot_mine += t_enc + t_pack
    line = f"seg len={ln} bits={bits}: encode {t_enc:.2f} ms + pack {t_pack:.3f} ms"
    if ref is not None and ln <= 1024:
        o = oT.t().contiguous()
        cen = seg["c_rot"]
        cid = torch.zeros(n, dtype=torch.int32, device=dev)
        f = [torch.zeros(n, device=dev) for _ in range(4)]
        short = torch.zeros((n, ln // 8), dtype=torch.uint8, device=dev)
        long_ = torch.zeros((n, max(4, ln * (bits - 1) // 8)), dtype=torch.uint8, device=dev)
        for mode, name in ((1, "sequential"), (0, "jacobi")):
            def run():
                short.zero_(); long_.zero_()
                ref.ref_saq_gpu_caq_encode(o.data_ptr(), cen.data_ptr(), cid.data_ptr(), *[t.data_ptr() for t in f],
                                           short.data_ptr(), long_.data_ptr(), ln, n, bits, 6, ctypes.c_float(1e-8), mode)
            t = timed(run, reps=2)
            line += f" | reference {name} {t:.2f} ms"
            if mode:
                tot_seq += t
            else:
                tot_jac += t
    print(line)
print(f"TOTAL encode+pack: mine {tot_mine:.2f} ms = {n / tot_mine / 1e3:.2f} M vec/s | reference CUDA sequential "
      f"{tot_seq:.2f} ms, block-Jacobi {tot_jac:.2f} ms (same B200)")
ids = torch.arange(0, n, max(1, n // 100000), device=dev)
t_dec = timed(lambda: eng.reconstruct(ids))
print(f"reconstruct {ids.numel()} rows: {t_dec:.2f} ms")
