"""Per-config kernel table: every hot-path kernel of SURVEY 8(a) at its BASELINE config shape (C1-C4), timed on
one B200 with CUDA events (inputs resident, > L2, warm-ups first), with the algorithmic bytes / FLOP of
SURVEY 8(d) against MEASURED_PEAKS.json. bench.py imports run() and prints the rows as "configs" in its JSON line
(N = 1, rank 0, outside the timed region); run as a script it writes one JSON document. No oracle, no CPU path."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "vector-quantization_b200"), ROOT]
import numpy as np
import torch

from b200vq import ops, ops_sq, pq as dpq

dev = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cuda")
peaks = {"hbm_gbps": 6551.0, "tf32_tflops": 695.0}
try:
    mp = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    peaks["hbm_gbps"] = float(mp.get("hbm_gbs", peaks["hbm_gbps"]))
    peaks["tf32_tflops"] = float(mp.get("bf16_tflops_sustained", 2 * peaks["tf32_tflops"])) / 2
except Exception:
    pass
rows = []
QUICK = False   # bench.py: skip the rows that only time a replaced (SIMT / library) path


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
    print(f"{config:3s} {kernel:34s} {ms:9.3f} ms {r['units_per_s']:.3e} {unit_name}/s{tail}  {note}", flush=True,
          file=sys.stderr)


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
    add("C1", "ADC search M=16 half-group scan layout (LUT + screened scan, k=10)",
        timed(lambda: codec.search(crot, q, 10, "l2", n=n, rot=True, m_pad=dpq.scan_layout_width(M))), n * nq, "codes",
        bytes_=float(n) * M * nq / 4, note="query tile 4; default FlatQuantizedIndex path")
    xn = ops.row_sqnorms(x)
    add("C1", "flat_topk_tc k=100 (ground truth)", timed(lambda: ops.flat_topk(x, q, 100, "l2", xnorm=xn)), n * nq,
        "pairs", flops=2.0 * n * nq * D)
    if not QUICK:
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
    if not QUICK:
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
    xn = ops.row_sqnorms(x)
    codes = ops_sq.rabitq1_encode(x, cen, "l2")
    add("C4", "rabitq1_encode", timed(lambda: ops_sq.rabitq1_encode(x, cen, "l2")), n, "vectors",
        bytes_=n * (4 * D + D // 8 + 8))
    add("C4", "rabitq1_decode", timed(lambda: ops_sq.rabitq1_decode(codes, D, cen)), n, "vectors",
        bytes_=n * (4 * D + D // 8 + 8))
    planes, qfac = ops_sq.rabitq1_prepare_queries(q, cen, 4)
    add("C4", "rabitq1_scan_topk_tc qb=4 k=10", timed(lambda: ops_sq.rabitq1_scan_topk(codes, D, planes, qfac, 4, 10, "l2")),
        n * nq, "codes", bytes_=float(n) * (D // 8 + 8) * ((nq + 255) // 256), flops=2.0 * n * nq * D,
        note="query tile 256; FLOP = FP8 tensor work, fraction quoted against TF32 peak/1 (FP8 peak is 4x)")
    if not QUICK:
      add("C4", "rabitq1_scan_topk simt qb=4 k=10",
        timed(lambda: ops_sq.rabitq1_scan_topk(codes, D, planes, qfac, 4, 10, "l2", impl="simt"), reps=2, warm=1),
        n * nq, "codes", bytes_=float(n) * (D // 8 + 8) * nq, note="query tile 1")
    # extended RaBitQ, B = 4 (registry method "rabitq"): encode = residual norms + rotation GEMM on the tensor cores
    # (centroid subtraction fused) + level search with exact fp64 re-evaluation near the mid-points
    from haag_vq.methods.extended_rabitq import ExtendedRaBitQuantizer
    nb, B = 131_072, 4
    eq = ExtendedRaBitQuantizer(B, seed=0)
    eq.fit(x[:4096].cpu().numpy())
    st = eq._state()
    ecodes = torch.empty((n, eq.code_size), dtype=torch.uint8, device=dev)

    def enc(s, e):
        xs = x[s:e]
        nrm = ops_sq.resid_norms_f64(xs, st["c"])
        y = st["fwd"](xs, x_sub=st["c32"])
        ops_sq.erabitq_pack_tc(y, xs, st["c"], st["PT"], nrm, B, st["levels"], st["bounds"], st["err_rel"],
                               2 * st["err_abs"], out=ecodes[s:e], stats=st["nfix"])
    add("C4", "extended RaBitQ encode B=4 (norms + rotate_tc + pack with fp64 fix-up)", timed(lambda: enc(0, nb)), nb,
        "vectors", bytes_=nb * (4 * D + eq.code_size), flops=2.0 * nb * D * D, note="fp64 indices; issued tensor work is 3x")
    xb = x[:nb]
    nrm = ops_sq.resid_norms_f64(xb, st["c"])
    yb = st["fwd"](xb, x_sub=st["c32"])
    add("C4", "  rotate_tc_ex 1536x1536 (x - c fused)", timed(lambda: st["fwd"](xb, x_sub=st["c32"])), nb, "vectors",
        flops=2.0 * nb * D * D)
    add("C4", "  erabitq_pack_tc B=4", timed(lambda: ops_sq.erabitq_pack_tc(yb, xb, st["c"], st["PT"], nrm, B, st["levels"],
                                                                         st["bounds"], st["err_rel"], 2 * st["err_abs"],
                                                                         out=ecodes[:nb])), nb, "vectors",
        bytes_=nb * (4 * D + eq.code_size))
    for s0 in range(0, n, nb):
        enc(s0, min(n, s0 + nb))
    add("C4", "extended RaBitQ decode B=4 (decode_f32 + rotate_tc_ex)",
        timed(lambda: st["rev"](ops_sq.erabitq_decode_f32(ecodes[:nb], D, B, st["levels32"])[0], col_bias=st["c32d"])), nb,
        "vectors", flops=2.0 * nb * D * D)
    sc = ops_sq.ERaBitQScanner(ecodes, D, B, st["levels32"])
    qrot, qconst = eq.rotate_queries(q, "l2")
    add("C4", "extended RaBitQ estimator scan B=4 k=10 (1M codes; code-domain, tcgen05)",
        timed(lambda: sc.search(qrot, qconst, 10, "l2"), reps=3, warm=2), n * nq, "codes",
        bytes_=float(n) * eq.code_size * ((nq + 255) // 256), flops=2.0 * n * nq * D, note="query tile 256; codes only")
    add("C4", "exact ground truth flat_topk_tc k=10 (1M x 1536)", timed(lambda: ops.flat_topk(x, q, 10, "l2", xnorm=xn), reps=3, warm=2),
        n * nq, "pairs", flops=2.0 * n * nq * D)
    ev, ei = sc.search(qrot, qconst, 10, "l2")
    gv, gi = ops.flat_topk(x, q, 10, "l2", xnorm=xn)
    rec = float(np.mean([len(set(a) & set(b)) / 10 for a, b in zip(ei.cpu().numpy().tolist(), gi.cpu().numpy().tolist())]))
    rows.append({"config": "C4", "kernel": "recall@10 of the extended RaBitQ B=4 scan vs exact search", "value": rec})
    rv, ri = ops_sq.rabitq1_scan_topk(codes, D, planes, qfac, 4, 10, "l2")
    rec1 = float(np.mean([len(set(a) & set(b)) / 10 for a, b in zip(ri.cpu().numpy().tolist(), gi.cpu().numpy().tolist())]))
    rows.append({"config": "C4", "kernel": "recall@10 of the 1-bit RaBitQ scan (qb=4) vs exact search", "value": rec1})
    del x, ecodes, sc


def saq_engine():
    """SAQ engine (registry "saq_paper") at the reference's own recorded size 99 000 x 1536, avg 4 bits."""
    from b200vq import saq_engine as E
    n, D, nq = 99_000, 1536, 1000
    x = data(n, D)
    q = data(nq, D, seed=1235)
    eng = E.SaqEngine(4.0).fit(x)
    add("SAQ", "engine projection + segment rotations (one GEMM, transposed store)",
        timed(lambda: eng._project(x, transpose_out=True)), n, "vectors", flops=2.0 * n * D * D)
    oT = eng._project(x, transpose_out=True)

    def encode_all():
        codesT, _ = E.caq_encode_t_segments(oT, eng.plan)
        for seg in eng.segs:
            if seg["bits"]:
                E.caq_pack_t(codesT[seg["off"]:seg["off"] + seg["len"]], seg["bits"], out=seg["packed"])
    bits_total = sum(ln * b for ln, b in eng.plan)
    add("SAQ", f"CAQ encode + pack, plan {eng.plan}", timed(encode_all), n, "vectors",
        bytes_=n * (4 * D + bits_total / 8 + 16 * len(eng.plan)), note="sequential code adjustment = CPU-identical codes")
    eng.search(q, 10)
    add("SAQ", "engine estimator search k=10 (code-domain, tcgen05)", timed(lambda: eng.search(q, 10)), n * nq, "codes",
        flops=2.0 * n * nq * (D - 192))
    ids = torch.arange(n, device=dev)
    add("SAQ", "engine reconstruct (decode + one GEMM, bias fused)", timed(lambda: eng.reconstruct(ids)), n, "vectors",
        flops=2.0 * n * D * D)


def ivf_group():
    """IVF shell at the reference's ratios: 256 rows per cluster, ~5 % of the clusters probed."""
    import time
    from haag_vq.methods.scalar_quantization import ScalarQuantizer
    from haag_vq.methods.search.ivf_quantized_index import IvfQuantizedIndex
    from b200vq import host, ivf as divf
    n, D, K, nprobe, nq, k = 262144, 1536, 1024, 50, 1000, 10
    xd = data(n, D)
    qd = data(nq, D, seed=99)
    X, Q = xd.cpu().numpy(), qd.cpu().numpy()
    t0 = time.perf_counter()
    C = divf.kmeans(xd, K, seed=0)
    torch.cuda.synchronize()
    t_km = time.perf_counter() - t0
    rows.append({"config": "IVF", "kernel": f"coarse k-means K={K}, 20 iterations, {n} x {D} (tensor-core assignment)",
                 "ms": round(t_km * 1e3, 1), "units_per_s": n * 20 / t_km, "unit": "point-iterations",
                 "TFLOPs": 2.0 * n * K * D * 20 / t_km / 1e12})
    print(f"IVF kmeans {t_km:.2f} s", file=sys.stderr, flush=True)
    idx = IvfQuantizedIndex(lambda: ScalarQuantizer(num_bits=8), K=K, nprobe=nprobe)
    t0 = time.perf_counter()
    idx.fit(X, "l2", centroids=C.cpu().numpy())
    t_fit = time.perf_counter() - t0
    rows.append({"config": "IVF", "kernel": f"assign + {K} per-cluster SQ8 fit/compress/decompress (host loop)",
                 "ms": round(t_fit * 1e3, 1), "units_per_s": n / t_fit, "unit": "vectors"})
    d = idx._dev
    probe = idx._probe(qd)
    frac = nprobe / K
    if not QUICK:
        ms = timed(lambda: divf.ivf_search(d["recon"], d["rnorm"], d["gid"], d["row_off"], K, qd, probe, k, "l2",
                                           impl="simt"))
        add("IVF", f"ivf_search simt nprobe={nprobe}/{K} k={k} (cluster-major exact scan of reconstructions)", ms, nq,
            "queries", flops=2.0 * nq * n * frac * D, note=f"{frac:.3f} of the rows per query")
    ms = timed(lambda: divf.ivf_search(d["recon"], d["rnorm"], d["gid"], d["row_off"], K, qd, probe, k, "l2",
                                       rowcluster=d["rowcluster"], impl="tc"))
    add("IVF", f"ivf_search tc nprobe={nprobe}/{K} k={k} (tcgen05 screen of all rows, probe lists as a mask; default)", ms,
        nq, "queries", flops=2.0 * nq * n * D, note="FLOP = issued (all rows); same results as the simt kernel")
    ms_p = timed(lambda: idx._probe(qd))
    add("IVF", "probe lists: exact flat search over the centroids", ms_p, nq, "queries", flops=2.0 * nq * K * D)
    ids, _ = idx.search_with_scores(Q, k)
    _, gt = ops.flat_topk(xd, qd, k, "l2")
    gt = gt.cpu().numpy().view(np.uint32)
    rec = float(np.mean([len(set(ids[i]) & set(gt[i])) / k for i in range(nq)]))
    rows.append({"config": "IVF", "kernel": f"recall@{k} of the IVF-SQ8 search vs exact search", "value": rec})
    print(f"IVF recall@{k} {rec:.4f}", file=sys.stderr, flush=True)


def run(which=("c1", "c2", "c3", "c4", "saq", "ivf"), device=None, quick=False):
    """Run the named config groups; returns the list of rows (and leaves it in `rows`)."""
    global dev, QUICK
    QUICK = bool(quick)
    if device is not None:
        dev = device
    rows.clear()
    fns = {"c1": c1, "c2": c2, "c3": c3, "c4": c4, "saq": saq_engine, "ivf": ivf_group}
    for w in which:
        fns[w]()
        torch.cuda.empty_cache()
    return list(rows)


if __name__ == "__main__":
    which = sys.argv[1].split(",") if len(sys.argv) > 1 else ["c1", "c2", "c3", "c4", "saq", "ivf"]
    out_path = sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "gpurun_out", "r2_configs.json")
    run(which)
    os.makedirs(os.path.dirname(out_path), exist_ok=True)
    json.dump({"peaks": peaks, "rows": rows}, open(out_path, "w"), indent=1)
