"""Per-config kernel table: every hot-path kernel of SURVEY 8(a) at its BASELINE config shape, timed on
one B200 with CUDA events (inputs resident, > L2, 3 warm-ups), with the algorithmic bytes / FLOP of
SURVEY 8(d) against MEASURED_PEAKS.json. Writes one JSON document (default profiles/r1_configs.json).
Diagnostic companion of bench.py (which carries the C5 headline); no oracle, no CPU path."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "vector-quantization_b200"), ROOT]
import numpy as np
import torch

from b200vq import ops, ops_sq, pq as dpq

dev = torch.device("cuda")
peaks = {"hbm_gbps": 6551.0, "tf32_tflops": 695.0}
try:
    mp = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    peaks["hbm_gbps"] = float(mp.get("hbm_gbs", peaks["hbm_gbps"]))
    peaks["tf32_tflops"] = float(mp.get("bf16_tflops_sustained", 2 * peaks["tf32_tflops"])) / 2
except Exception:
    pass
rows = []


def timed(fn, reps=5, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def add(config, kernel, ms, units, unit_name, bytes_=None, flops=None, note=""):
    r = {"config": config, "kernel": kernel, "ms": round(ms, 4), "units_per_s": units / ms * 1e3, "unit": unit_name}
    if bytes_ is not None:
        r["GBps"] = bytes_ / ms / 1e6
        r["frac_hbm"] = r["GBps"] / peaks["hbm_gbps"]
    if flops is not None:
        r["TFLOPs"] = flops / ms / 1e9
        r["frac_tf32"] = r["TFLOPs"] / peaks["tf32_tflops"]
    if note:
        r["note"] = note
    rows.append(r)
    tail = (f" {r['GBps']:8.0f} GB/s ({r['frac_hbm']:.2f} of HBM)" if bytes_ is not None else "") + \
           (f" {r['TFLOPs']:7.1f} TFLOP/s ({r['frac_tf32']:.2f} of TF32)" if flops is not None else "")
    print(f"{config:3s} {kernel:34s} {ms:9.3f} ms {r['units_per_s']:.3e} {unit_name}/s{tail}  {note}", flush=True)


def data(n, D, seed=1234, normalize=True):
    cs = torch.linspace(4.0, 0.2, D, device=dev)
    return ops_sq.synth_rows(n, D, 0, seed, cs, normalize, device=dev)


def c1():
    n, D, M, nq = 100_000, 1536, 16, 1000
    x = data(n, D)
    q = data(nq, D, seed=1235)
    C = dpq.pq_train(x[:65536], M, 8, niter=6, seed=1234)
    codec = dpq.PQCodec(C)
    codes = codec.encode(x)
    add("C1", "pq_assign_tc dsub=96", timed(lambda: codec.encode(x, out=codes)), n, "vectors",
        bytes_=n * (4 * D + M), flops=2.0 * n * D * 256)
    out = torch.empty((n, D), device=dev)
    add("C1", "pq_decode", timed(lambda: ops.pq_decode(codes, C, out=out)), n, "vectors", bytes_=n * (M + 4 * D))
    lut = codec.luts(q, "l2")
    add("C1", "adc_lut (1000 queries)", timed(lambda: codec.luts(q, "l2")), nq, "queries", flops=2.0 * nq * D * 256)
    add("C1", "adc_scan_topk row layout M=16 k=10", timed(lambda: ops.adc_scan_topk(codes, lut, 10, "l2")), n * nq,
        "codes", bytes_=float(n) * M * nq, note="query tile 1")
    crot = ops.pq_codes_to_scan_layout(codec.pad_codes_for_scan(codes))
    add("C1", "ADC search M=16 padded to the scan layout (LUT + screened scan, k=10)",
        timed(lambda: codec.search(crot, q, 10, "l2", n=n, rot=True, m_pad=32)), n * nq, "codes",
        bytes_=float(n) * M * nq / 4, note="query tile 4; default FlatQuantizedIndex path")
    xn = ops.row_sqnorms(x)
    add("C1", "flat_topk_tc k=100 (ground truth)", timed(lambda: ops.flat_topk(x, q, 100, "l2", xnorm=xn)), n * nq,
        "pairs", flops=2.0 * n * nq * D)
    add("C1", "flat_topk simt k=100", timed(lambda: ops.flat_topk(x, q, 100, "l2", xnorm=xn, impl="simt"), reps=2, warm=1),
        n * nq, "pairs", flops=2.0 * n * nq * D)


def c2():
    n, D, M = 262_144, 1536, 32         # one resident batch of the 1M-row config
    x = data(n, D)
    A = torch.from_numpy(dpq.random_rotation(D, 5)).to(dev)
    y = dpq.rotate(x, A)
    C = dpq.pq_train(y[:65536], M, 8, niter=6, seed=1234)
    codec = dpq.PQCodec(C, rotation=A)
    codes = codec.encode(x)
    add("C2", "rotate_tc 1536x1536 (3xTF32)", timed(lambda: dpq.rotate(x, A, impl="tc")), n, "vectors",
        bytes_=n * 8 * D, flops=2.0 * n * D * D, note="issued tensor work is 3x")
    add("C2", "library sgemm rotation (replaced)", timed(lambda: dpq.rotate(x, A, impl="lib")), n, "vectors",
        flops=2.0 * n * D * D)
    add("C2", "pq_assign_tc dsub=48", timed(lambda: ops.pq_assign(y, C, codec.cnorm, out=codes)), n, "vectors",
        bytes_=n * (4 * D + M), flops=2.0 * n * D * 256)
    add("C2", "OPQ encode = rotate + assign", timed(lambda: codec.encode(x, out=codes)), n, "vectors",
        flops=2.0 * n * D * (D + 256))
    add("C2", "OPQ decode = gather + reverse rotate", timed(lambda: codec.decode(codes)), n, "vectors",
        flops=2.0 * n * D * D)


def c3():
    n, D = 262_144, 3072
    x = data(n, D, normalize=False)
    vmin, vmax = ops_sq.sq_minmax(x)
    add("C3", "sq_minmax", timed(lambda: ops_sq.sq_minmax(x)), n, "vectors", bytes_=n * 4 * D)
    for bits in (8, 4, 16):
        codes = ops_sq.sq_encode(x, vmin, vmax, bits)
        cb = codes.numel() * codes.element_size() / n
        add("C3", f"sq_encode {bits}-bit", timed(lambda: ops_sq.sq_encode(x, vmin, vmax, bits, out=codes)), n, "vectors",
            bytes_=n * (4 * D + cb))
        out = torch.empty((n, D), device=dev)
        add("C3", f"sq_decode {bits}-bit", timed(lambda: ops_sq.sq_decode(codes, vmin, vmax, bits, D, out=out)), n,
            "vectors", bytes_=n * (4 * D + cb))
    # SAQ (CLI contract): 8 segments, widths from {0,2,4,6,8} at an average of 4 bits
    seg_bits = np.array([8, 8, 6, 4, 4, 2, 0, 0], dtype=np.int32)
    dim_bits = torch.from_numpy(np.repeat(seg_bits, D // 8)).to(dev)
    off = torch.from_numpy(np.concatenate([[0], np.cumsum(np.repeat(seg_bits, D // 8))[:-1]]).astype(np.int32)).to(dev)
    code_bytes = int(seg_bits.sum() * (D // 8) + 7) // 8
    codes = ops_sq.saq_encode(x, vmin, vmax, dim_bits, off, code_bytes)
    add("C3", "saq_encode avg 4 bits", timed(lambda: ops_sq.saq_encode(x, vmin, vmax, dim_bits, off, code_bytes)), n,
        "vectors", bytes_=n * (4 * D + code_bytes))
    add("C3", "saq_decode avg 4 bits", timed(lambda: ops_sq.saq_decode(codes, vmin, vmax, dim_bits, off, D)), n,
        "vectors", bytes_=n * (4 * D + code_bytes))
    # LVQ (registry method "lvq"): per-row lo / delta, B-bit MSB-first packing
    mu = ops_sq.col_mean_seq(x[:65536])
    for bits in (8, 4):
        cs = ops_sq.lvq_code_size(D, bits)
        lc = ops_sq.lvq_encode(x, mu, bits)
        add("C3", f"lvq_encode {bits}-bit", timed(lambda: ops_sq.lvq_encode(x, mu, bits, out=lc)), n, "vectors",
            bytes_=n * (4 * D + cs))
        add("C3", f"lvq_decode {bits}-bit", timed(lambda: ops_sq.lvq_decode(lc, mu, bits)), n, "vectors",
            bytes_=n * (4 * D + cs))


def c4():
    n, D, nq = 1_000_000, 1536, 1000
    x = data(n, D)
    q = data(nq, D, seed=1235)
    cen = x[:65536].mean(0).contiguous()
    codes = ops_sq.rabitq1_encode(x, cen, "l2")
    add("C4", "rabitq1_encode", timed(lambda: ops_sq.rabitq1_encode(x, cen, "l2")), n, "vectors",
        bytes_=n * (4 * D + D // 8 + 8))
    add("C4", "rabitq1_decode", timed(lambda: ops_sq.rabitq1_decode(codes, D, cen)), n, "vectors",
        bytes_=n * (4 * D + D // 8 + 8))
    planes, qfac = ops_sq.rabitq1_prepare_queries(q, cen, 4)
    add("C4", "rabitq1_scan_topk_tc qb=4 k=10", timed(lambda: ops_sq.rabitq1_scan_topk(codes, D, planes, qfac, 4, 10, "l2")),
        n * nq, "codes", bytes_=float(n) * (D // 8 + 8) * ((nq + 255) // 256), flops=2.0 * n * nq * D,
        note="query tile 256; FLOP = FP8 tensor work, fraction quoted against TF32 peak/1 (FP8 peak is 4x)")
    add("C4", "rabitq1_scan_topk simt qb=4 k=10",
        timed(lambda: ops_sq.rabitq1_scan_topk(codes, D, planes, qfac, 4, 10, "l2", impl="simt"), reps=2, warm=1),
        n * nq, "codes", bytes_=float(n) * (D // 8 + 8) * nq, note="query tile 1")
    del x
    # extended RaBitQ, B = 4: pack / unpack on fp64 rotated rows (the rotation itself is an fp64 library GEMM)
    nb, B = 131_072, 4
    g = torch.Generator(device=dev).manual_seed(5)
    srot = torch.randn((nb, D), generator=g, device=dev, dtype=torch.float64)
    nrm = torch.rand((nb,), generator=g, device=dev, dtype=torch.float64) + 0.5
    from haag_vq.methods.extended_rabitq import _lloyd_1d_normal
    levels = torch.from_numpy(np.asarray(_lloyd_1d_normal(1 << B, 0), dtype=np.float64)).to(dev)
    bounds = (0.5 * (levels[:-1] + levels[1:])).contiguous()
    ecodes = ops_sq.erabitq_pack(srot, nrm, B, levels, bounds)
    add("C4", "erabitq_pack B=4 (fp64 in)", timed(lambda: ops_sq.erabitq_pack(srot, nrm, B, levels, bounds)), nb,
        "vectors", bytes_=nb * (8 * D + D * B // 8 + 8))
    add("C4", "erabitq_unpack B=4 (fp64 out)", timed(lambda: ops_sq.erabitq_unpack(ecodes, D, B, levels)), nb, "vectors",
        bytes_=nb * (8 * D + D * B // 8 + 8))
    # extended RaBitQ estimator scan in the code domain (BASELINE config 4): <levels[idx], q P> per (code, query)
    qrot = torch.randn((nq, D), generator=g, device=dev, dtype=torch.float32)
    qconst = torch.rand((nq,), generator=g, device=dev, dtype=torch.float32)
    lv32 = levels.to(torch.float32).contiguous()
    add("C4", f"erabitq_scan_topk simt B=4 k=10 ({nb} codes)",
        timed(lambda: ops_sq.erabitq_scan_topk(ecodes, D, B, lv32, qrot, qconst, 10, "l2", impl="simt"), reps=1, warm=1),
        nb * nq, "codes", bytes_=float(nb) * (D * B // 8 + 8) * nq, note="query tile 1")
    add("C4", f"erabitq_scan_topk tc B=4 k=10 ({nb} codes; decode + flat search)",
        timed(lambda: ops_sq.erabitq_scan_topk(ecodes, D, B, lv32, qrot, qconst, 10, "l2", impl="tc"), reps=3, warm=2),
        nb * nq, "codes", bytes_=float(nb) * (D * B // 8 + 8) * ((nq + 255) // 256), flops=2.0 * nb * nq * D,
        note="query tile 256")
    big = ecodes.repeat(8, 1)[:1_000_000].contiguous()
    add("C4", "erabitq_scan_topk tc B=4 k=10 (1M codes; decode + flat search)",
        timed(lambda: ops_sq.erabitq_scan_topk(big, D, B, lv32, qrot, qconst, 10, "l2", impl="tc"), reps=3, warm=2),
        1_000_000 * nq, "codes", bytes_=1e6 * (D * B // 8 + 8) * ((nq + 255) // 256), flops=2.0 * 1e6 * nq * D,
        note="query tile 256")
    del big
    P = torch.randn((D, D), generator=g, device=dev, dtype=torch.float64)
    add("C4", "fp64 rotation GEMM (library)", timed(lambda: torch.matmul(srot, P), reps=2, warm=1), nb, "vectors",
        flops=2.0 * nb * D * D)


if __name__ == "__main__":
    which = sys.argv[1].split(",") if len(sys.argv) > 1 else ["c1", "c2", "c3", "c4"]
    out_path = sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "gpurun_out", "r1_configs.json")
    for w in which:
        {"c1": c1, "c2": c2, "c3": c3, "c4": c4}[w]()
        torch.cuda.empty_cache()
    os.makedirs(os.path.dirname(out_path), exist_ok=True)
    json.dump({"peaks": peaks, "rows": rows}, open(out_path, "w"), indent=1)
