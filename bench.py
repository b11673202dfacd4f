#!/usr/bin/env python
"""bench.py — PQ streaming encode + ADC top-10 scan on the MS MARCO shape (BASELINE config 5).

Workload (weak scaling): every GPU owns a shard of 6 625 000 x 1024 fp32 rows (8 shards =
53 M rows), PQ M=32 nbits=8. One step = encode one batch of 2^20 synthetic rows that are
already resident in HBM (b200vq_pq_assign*) + scan the shard's (n, 32) uint8 codes with
1000 per-query LUTs for the top-10 (b200vq_adc_lut, b200vq_adc_scan_topk*), followed, when
N > 1, by the NCCL all-gather + k-way merge of the per-shard lists.

  value            vectors/s encoded, whole job (all ranks), device-timed, inputs in HBM
  adc.value        ADC codes/s scanned = total rows x queries / scan time
  e2e              same two numbers through the reference-facing haag_vq objects with HOST
                   (pinned) buffers: ProductQuantizer.compress(X) and
                   FlatQuantizedIndex.search_with_scores(Q, 10), copies inside the timed region
  roofline         ADC scan kernel vs the measured HBM copy bandwidth (algorithmic bytes =
                   rows x M per query pass)
  roofline_assign  PQ assign kernel vs the tensor pipe (TF32 = measured bf16 peak / 2)
  cpu_baseline     the CPU oracle timed on a bounded sample on this box's host cores

`--impl reference` times the reference's CPU path for the same step (faiss-semantics
numpy/BLAS nearest-centroid assign + the ADC loop of bench/ffd_speed.cpp restated in C with
OpenMP, all host threads) on a bounded sample and prints the same line.
"""
from __future__ import annotations

import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "vector-quantization_b200")
for _p in (PKG, ROOT):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

D, M, KSUB, K = 1024, 32, 256, 10
SHARD_ROWS = 6_625_000          # 53 M / 8
BATCH_ROWS = 1 << 20
NQ = 1000
E2E_ROWS = 1 << 18              # host batch of the e2e encode leg (1 GiB pinned)
TRAIN_ROWS = 65_536             # 256 * ksub, faiss' training subset size
print_line = None
METRIC = "vectors/s encoded + ADC codes/s scanned at 1/2/4/8 B200, % of HBM roofline"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--rows-per-gpu", type=int, default=SHARD_ROWS)
    ap.add_argument("--batch-rows", type=int, default=BATCH_ROWS)
    ap.add_argument("--nq", type=int, default=NQ)
    ap.add_argument("--assign-impl", default="auto", choices=["auto", "simt", "tc"])
    ap.add_argument("--scan-layout", default="auto", choices=["auto", "rows", "rot"])
    ap.add_argument("--scan-impl", default="auto", choices=["auto", "exact", "screen"],
                    help="screen = four-queries-per-lookup screened scan (scan layout only), same results")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the C1-C4 kernel table (N = 1 only)")
    ap.add_argument("--no-parity", action="store_true", help="skip the untimed result check of the scan / NCCL merge")
    return ap.parse_args()


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return {"hbm_gbs": float(d["hbm_gbs"]), "bf16_tflops": float(d["bf16_tflops"]),
                "bf16_tflops_sustained": float(d.get("bf16_tflops_sustained", d["bf16_tflops"])),
                "source": "measured"}
    # fallback stated in the profiling recipe
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []
        self.thread = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                 "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return
        self.thread = threading.Thread(target=self._pump, daemon=True)
        self.thread.start()

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1]))
                mx = float(f[2])
            except ValueError:
                continue
            for nm, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": (sm[len(sm) // 2] if sm else None), "sm_max_mhz": mx,
                "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------- CPU legs ---
def synth_host(n: int, seed: int) -> np.ndarray:
    """Host twin of the device generator's distribution (anisotropic Gaussian, unit rows)."""
    rng = np.random.default_rng(seed)
    X = rng.standard_normal((n, D), dtype=np.float32) * np.linspace(4.0, 0.2, D, dtype=np.float32)
    X /= np.maximum(np.linalg.norm(X, axis=1, keepdims=True), 1e-12)
    return np.ascontiguousarray(X, dtype=np.float32)


def _load_oracle_adc():
    path = os.path.join(ROOT, "oracle", "_ref", "liboracle_adc.so")
    if not os.path.exists(path):
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "_ref/liboracle_adc.so"], check=True)
    lib = ctypes.CDLL(path)
    lib.oracle_adc_scores.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_int,
                                      ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
    lib.oracle_adc_scores.restype = None
    lib.oracle_adc_topk.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
                                    ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
    lib.oracle_adc_topk.restype = ctypes.c_int
    return lib


def _load_oracle_pq():
    path = os.path.join(ROOT, "oracle", "_ref", "liboracle_pq.so")
    if not os.path.exists(path):
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "_ref/liboracle_pq.so"], check=True)
    lib = ctypes.CDLL(path)
    lib.oracle_pq_assign_dec.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_void_p, ctypes.c_int,
                                     ctypes.c_int, ctypes.c_void_p]
    lib.oracle_pq_assign_dec.restype = ctypes.c_int
    # all host cores, whatever OMP_NUM_THREADS says (torchrun sets it to 1 for its workers); liboracle_adc.so shares
    # the same OpenMP runtime
    lib.oracle_pq_set_threads(ctypes.c_int(os.cpu_count() or 1))
    return lib


_PQ_LIB = None


def cpu_assign(X, C):
    """Nearest-centroid codes on the host cores: oracle/pq_assign.c (C / OpenMP, all threads), the compiled form of
    oracle.pq.pq_assign (same arithmetic, same codes; tests/test_oracle.py) - an order of magnitude faster than the
    numpy form, so that the CPU arm is not an under-estimate of what faiss' compute_codes does on the same cores."""
    global _PQ_LIB
    if _PQ_LIB is None:
        _PQ_LIB = _load_oracle_pq()
    codes = np.empty((X.shape[0], M), dtype=np.uint8)
    Cc = np.ascontiguousarray(C, dtype=np.float32)
    rc = _PQ_LIB.oracle_pq_assign_dec(X.ctypes.data, X.shape[0], D, Cc.ctypes.data, M, KSUB, codes.ctypes.data)
    if rc != 0:
        raise RuntimeError(f"oracle_pq_assign_dec failed: {rc}")
    return codes


def cpu_step(X, C, codes_db, Q, adc_lib, scores_buf):
    """One pass of the reference CPU path on a bounded sample: nearest-centroid assign (C / OpenMP restatement of
    faiss compute_codes) + ADC scores (C loop, OpenMP) + argpartition top-k."""
    from oracle import pq as opq
    t0 = time.perf_counter()
    cpu_assign(X, C)
    t1 = time.perf_counter()
    lut = opq.adc_lut(Q, C, opq.L2)
    n = codes_db.shape[0]
    # top-k inside the OpenMP scan (queries over threads, no n-sized score matrix): round 1 wrote all scores and
    # selected with a single-threaded argpartition, which was half of the leg
    nq = Q.shape[0]
    vals = np.empty((nq, K), dtype=np.float32)
    ids = np.empty((nq, K), dtype=np.int64)
    rc = adc_lib.oracle_adc_topk(codes_db.ctypes.data, n, M, KSUB, lut.ctypes.data, nq, K, vals.ctypes.data,
                                 ids.ctypes.data)
    if rc != 0:
        raise RuntimeError(f"oracle_adc_topk failed: {rc}")
    t2 = time.perf_counter()
    return t1 - t0, t2 - t1


def cpu_setup(enc_rows: int, db_rows: int, nq: int):
    from oracle import pq as opq
    X = synth_host(enc_rows, 11)
    C = opq.pq_init_centroids(synth_host(4096, 12), M, KSUB, seed=1)
    rng = np.random.default_rng(13)
    codes_db = rng.integers(0, 256, size=(db_rows, M), dtype=np.uint8)
    Q = synth_host(nq, 14)
    scores = None      # (round 1 kept an (nq, n) score matrix here)
    return X, C, codes_db, Q, _load_oracle_adc(), scores


CPU_SAMPLE = dict(enc_rows=524_288, db_rows=1_000_000, nq=64)


def cpu_baseline_leg(reps: int = 2):
    cores = os.cpu_count() or 1
    X, C, codes_db, Q, lib, scores = cpu_setup(**CPU_SAMPLE)
    cpu_step(X[:2048], C, codes_db[:65536], Q[:2], lib, scores)  # warm BLAS / OpenMP pools
    te = ts = 0.0
    for _ in range(reps):
        a, b = cpu_step(X, C, codes_db, Q, lib, scores)
        te += a
        ts += b
    return {"value": CPU_SAMPLE["enc_rows"] * reps / te, "unit": "vectors/s", "cores": cores, "kind": "port",
            "adc": {"value": CPU_SAMPLE["db_rows"] * CPU_SAMPLE["nq"] * reps / ts, "unit": "codes/s"},
            "sample": (f"oracle port (faiss is not vendored in the reference): C/OpenMP pq_assign of "
                       f"{CPU_SAMPLE['enc_rows']}x{D} rows + C/OpenMP ADC scan of {CPU_SAMPLE['db_rows']} codes x "
                       f"{CPU_SAMPLE['nq']} queries with top-{K} kept inside the scan, {reps} reps")}


def run_reference(args):
    """--impl reference: the CPU path, rank 0 only, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    X, C, codes_db, Q, lib, scores = cpu_setup(**CPU_SAMPLE)
    for _ in range(max(1, min(args.warmup, 3))):
        cpu_step(X[:4096], C, codes_db[:131072], Q[:2], lib, scores)
    te = ts = 0.0
    for _ in range(args.steps):
        a, b = cpu_step(X, C, codes_db, Q, lib, scores)
        te += a
        ts += b
    enc = CPU_SAMPLE["enc_rows"] * args.steps / te
    adc = CPU_SAMPLE["db_rows"] * CPU_SAMPLE["nq"] * args.steps / ts
    sample = (f"per step: pq_assign of {CPU_SAMPLE['enc_rows']}x{D} rows (C/OpenMP, faiss semantics) + ADC scan of "
              f"{CPU_SAMPLE['db_rows']} codes x {CPU_SAMPLE['nq']} queries with top-{K} (C/OpenMP restatement of bench/ffd_speed.cpp)")
    line = {"impl": "reference", "metric": METRIC, "value": enc, "unit": "vectors/s",
            "adc": {"value": adc, "unit": "codes/s"}, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": (te + ts) / args.steps * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args),
            "cpu_baseline": {"value": enc, "unit": "vectors/s", "cores": cores, "kind": "port",
                             "adc": {"value": adc, "unit": "codes/s"}, "sample": sample},
            "e2e": {"value": enc, "unit": "vectors/s", "adc": {"value": adc, "unit": "codes/s"},
                    "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print_line(line)


def workload_config(args):
    return {"workload": f"PQ M={M} nbits=8 streaming encode + ADC top-{K}, synthetic {args.rows_per_gpu * args.gpus}x{D} "
                        f"(MS MARCO shape, BASELINE config 5), {args.rows_per_gpu} rows per GPU",
            "rows_per_gpu": args.rows_per_gpu, "batch_rows": args.batch_rows, "D": D, "M": M, "ksub": KSUB,
            "queries": args.nq, "k": K, "parallelism": f"row-shard x{args.gpus}",
            "l2": "inputs larger than L2 (4 GiB row batches rotate; 212 MB of codes per scan)"}


# ----------------------------------------------------------------- GPU arm ---
def run_b200(args):
    import torch
    import torch.distributed as tdist

    import b200vq
    from b200vq import _lib, dist as vdist, ops, ops_sq
    from b200vq import pq as dpq

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the b200 arm has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        tdist.init_process_group("nccl", device_id=dev)
    if world != args.gpus and rank == 0:
        print(f"bench.py: WORLD_SIZE={world} but --gpus {args.gpus}; using WORLD_SIZE", file=sys.stderr)
    b200vq.load()
    peaks = measured_peaks()

    n, B, nq = args.rows_per_gpu, min(args.batch_rows, args.rows_per_gpu), args.nq
    row_offset = rank * n
    col_scale = torch.linspace(4.0, 0.2, D, device=dev)

    # trained state: trained on every rank from the same generator rows, then rank 0's copy is broadcast
    xt = ops_sq.synth_rows(TRAIN_ROWS, D, 1 << 40, 1234, col_scale, True, device=dev)
    C = dpq.pq_train(xt, M, 8, niter=8, seed=1234)
    del xt
    if world > 1:
        # ONE trained state for the whole job: the Lloyd sums are accumulated with float atomics, so two ranks
        # training on the same rows end up a few ulps apart; rank 0's centroids are the index's
        tdist.broadcast(C, src=0)
    codec = dpq.PQCodec(C)
    impl = args.assign_impl
    if impl == "auto":
        impl = "tc" if ops.tc_available() else "simt"

    # streaming encode of the whole shard (untimed): generator batch -> codes
    codes = torch.empty((n, M), dtype=torch.uint8, device=dev)
    xbuf = [torch.empty((B, D), dtype=torch.float32, device=dev) for _ in range(2)]
    starts = list(range(0, n, B))
    for i, s in enumerate(starts):
        e = min(n, s + B)
        ops_sq.synth_rows(e - s, D, row_offset + s, 1234, col_scale, True, out=xbuf[i & 1][:e - s])
        codec.encode(xbuf[i & 1][:e - s], out=codes[s:e], impl=impl)
    # the two timed batches are the first two full batches of the shard
    full = [s for s in starts if s + B <= n][:2] or [0]
    for j, s in enumerate(full):
        ops_sq.synth_rows(B, D, row_offset + s, 1234, col_scale, True, out=xbuf[j])
    Qd = ops_sq.synth_rows(nq, D, 1 << 41, 4321, col_scale, True, device=dev)
    torch.cuda.synchronize()

    def time_ms(fn, reps=2):
        fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / reps

    lut0 = codec.luts(Qd[:min(nq, 64)], "l2")
    layout = args.scan_layout
    codes_rot = None
    if layout in ("auto", "rot") and M % 32 == 0:
        codes_rot = ops.pq_codes_to_scan_layout(codes)
        if layout == "auto":
            t_rows = time_ms(lambda: ops.adc_scan_topk(codes, lut0, K, "l2"))
            t_rot = time_ms(lambda: ops.adc_scan_topk(codes_rot, lut0, K, "l2", n=n, rot=True))
            layout = "rot" if t_rot < t_rows else "rows"
            if layout == "rows":
                codes_rot = None
    else:
        layout = "rows"
    del lut0

    screen = layout == "rot" and args.scan_impl in ("auto", "screen")
    if args.scan_impl == "screen" and not screen:
        raise SystemExit("--scan-impl screen needs the scan layout (M % 32 == 0)")
    q_tile = 4 if screen else 1

    def scan(lut):
        if layout == "rot":
            return ops.adc_scan_topk(codes_rot, lut, K, "l2", n=n, rot=True, screen=screen)
        return ops.adc_scan_topk(codes, lut, K, "l2")

    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
    marks = []

    def step(i, record):
        j = i % len(full)
        s = full[j]
        e0, e1, e2, e3, e4 = ev(), ev(), ev(), ev(), ev()
        e0.record()
        codec.encode(xbuf[j], out=codes[s:s + B], impl=impl)
        e1.record()
        lut = codec.luts(Qd, "l2")
        e2.record()
        vals, ids = scan(lut)
        e3.record()
        if world > 1:
            vals, ids = vdist.sharded_topk(vals, ids, row_offset, K, "l2")
        e4.record()
        if record:
            marks.append((e0, e1, e2, e3, e4))
        return vals, ids

    def barrier():
        if world > 1:
            tdist.barrier()
        torch.cuda.synchronize()

    for i in range(max(3, args.warmup)):
        step(i, False)
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    calls0 = _lib.abi_calls
    t0, t1 = ev(), ev()
    barrier()
    t0.record()
    for i in range(args.steps):
        step(i, True)
    t1.record()
    barrier()
    n_calls = _lib.abi_calls - calls0
    clocks = sampler.stop() if sampler else None
    tot = t0.elapsed_time(t1)
    enc = sum(a.elapsed_time(b) for a, b, _, _, _ in marks)
    scn = sum(b.elapsed_time(e) for _, b, _, _, e in marks)
    knl = sum(c.elapsed_time(d) for _, _, c, d, _ in marks)
    t = torch.tensor([tot, enc, scn, knl], dtype=torch.float64, device=dev)
    if world > 1:
        tdist.all_reduce(t, op=tdist.ReduceOp.MAX)
    tot, enc, scn, knl = (float(v) for v in t.tolist())
    steps = args.steps
    enc_vps = B * world * steps / (enc * 1e-3)
    adc_cps = float(n) * world * nq * steps / (scn * 1e-3)

    # roofline: algorithmic bytes per launch / average launch duration (per GPU)
    # rows x M code bytes per query-tile pass (N * M * ceil(Q / Q_tile)); the launch duration is
    # the whole scan call (for the screened scan: sample + prepare + screen + finish + gated fallback)
    scan_bytes = float(n) * M * ((nq + q_tile - 1) // q_tile)
    scan_kernel = "adc_screen_kernel" if screen else ("adc_scan_rot_kernel" if layout == "rot" else "adc_scan_kernel")
    scan_gbs = scan_bytes / (knl / steps * 1e-3) / 1e9
    assign_tflops = 2.0 * D * KSUB * B / (enc / steps * 1e-3) / 1e12
    tf32_peak = peaks["bf16_tflops_sustained"] / 2.0
    # DRAM bytes of one scan launch from the committed ncu --set full capture, valid for that launch shape only
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            rec = json.load(f).get(scan_kernel)
        if rec and (rec["rows"], rec["queries"], rec["M"]) == (n, nq, M):
            traffic = rec["dram_bytes_per_launch"]
    except (OSError, ValueError, KeyError):
        traffic = None
    sm_mhz = (clocks or {}).get("sm_mhz") or 1965.0
    lds_floor_ms = float(n) * M * nq / q_tile / 32.0 / (148 * sm_mhz * 1e6) * 1e3
    line = {
        "metric": METRIC, "value": enc_vps, "unit": "vectors/s",
        "adc": {"value": adc_cps, "unit": "codes/s"},
        "n_gpus": world, "steps": steps, "warmup": max(3, args.warmup), "ms_per_step": tot / steps,
        "ms_encode_per_step": enc / steps, "ms_scan_per_step": scn / steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": dict(workload_config(args), assign_impl=impl, scan_layout=layout,
                                             scan_impl="screen" if screen else "exact", scan_query_tile=q_tile),
        "roofline": {"kernel": scan_kernel, "bound": "hbm",
                     "achieved": scan_gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                     "frac": scan_gbs / peaks["hbm_gbs"], "traffic": traffic, "peak_source": peaks["source"],
                     "algorithmic_bytes_per_launch": scan_bytes, "query_tile": q_tile,
                     "table_lookups_per_s": float(n) * M * nq / (knl / steps * 1e-3),
                     # what the number is and is not: the query passes are served by L2, so DRAM sees `traffic`
                     # bytes per call (dram_gbs); the kernel's own ceiling is the shared-memory look-up rate:
                     # lookups / (4 queries per LDS) / 32 lanes per warp-LDS / (SMs x clock) = lds_floor_ms
                     "dram_gbs": (traffic / (knl / steps * 1e-3) / 1e9) if traffic else None,
                     "dram_frac": (traffic / (knl / steps * 1e-3) / 1e9 / peaks["hbm_gbs"]) if traffic else None,
                     "lds_floor_ms": lds_floor_ms, "lds_frac": lds_floor_ms / (knl / steps),
                     "bound_note": "hbm per SURVEY 8(d) accounting (algorithmic code bytes x query-tile passes); "
                                   "actual limiter = shared-memory LUT look-ups (lds_frac), DRAM at dram_frac"},
        "roofline_assign": {"kernel": "pq_assign_tc_kernel" if impl == "tc" else "pq_assign_simt_kernel",
                            "bound": "tensor", "achieved": assign_tflops, "peak": tf32_peak, "unit": "TFLOP/s",
                            "frac": assign_tflops / tf32_peak,
                            "peak_source": peaks["source"] + " (bf16 sustained / 2 = dense TF32)",
                            "hbm_gbs": (4.0 * D + M) * B / (enc / steps * 1e-3) / 1e9},
        "gpu_launches": int(n_calls + steps),  # every ABI call is >= 1 kernel; the scan call adds its merge kernel
        "clocks": clocks,
    }
    ctx = dict(torch=torch, tdist=tdist, codec=codec, codes=codes, C=C, Qd=Qd, world=world, rank=rank,
               dev=dev, n=n, nq=nq, row_offset=row_offset, col_scale=col_scale, xbuf=xbuf, step=step, scan=scan,
               codes_rot=codes_rot, layout=layout)
    if not args.no_parity:
        line["parity_check"] = parity_check(ctx)
    return line, ctx


def parity_check(ctx):
    """Untimed result check on hardware. N = 1: the screened scan against the exact row-layout kernel (ids and value
    bits). N > 1: every rank's codes are all-gathered (53 M x 32 B = 1.7 GB at N = 8), rank 0 scans them as ONE shard
    and the NCCL-merged (value, global id) lists must equal that single-GPU answer bit for bit."""
    torch, tdist = ctx["torch"], ctx["tdist"]
    from b200vq import ops
    codec, codes, Qd, world, rank, n = ctx["codec"], ctx["codes"], ctx["Qd"], ctx["world"], ctx["rank"], ctx["n"]
    vals, ids = ctx["step"](0, False)
    if world == 1:
        lut = codec.luts(Qd, "l2")
        v0, i0 = ops.adc_scan_topk(codes, lut, K, "l2")
        ok = bool(torch.equal(ids.view(torch.int32), i0.view(torch.int32))) and \
            bool(torch.equal(vals.view(torch.int32), v0.view(torch.int32)))
        return "ok (N=1: scan path == exact row-layout scan, ids and value bits)" if ok else "MISMATCH (N=1)"
    allc = torch.empty((world * n, M), dtype=torch.uint8, device=ctx["dev"])
    tdist.all_gather_into_tensor(allc, codes)
    res = "ok"
    if rank == 0:
        v1, i1 = codec.search(allc, Qd, K, "l2")                      # exact row-layout scan of the whole corpus
        g1 = i1.view(torch.int32).to(torch.int64) & 0xFFFFFFFF
        bad_i = int((ids.to(torch.int64) != g1).sum().item())
        bad_v = int((vals.view(torch.int32) != v1.view(torch.int32)).sum().item())
        res = (f"ok (N={world}: NCCL all-gather + merge == single-GPU scan of all {world * n} codes, ids and value bits)"
               if bad_i == 0 and bad_v == 0 else
               f"MISMATCH (N={world}: {bad_i} ids, {bad_v} values of {ids.numel()} differ; first rows "
               f"{ids[0, :4].tolist()} vs {g1[0, :4].tolist()}, {vals[0, :4].tolist()} vs {v1[0, :4].tolist()})")
    del allc
    tdist.barrier()
    return res


def run_e2e(args, ctx):
    """Same step through the reference-facing objects with host buffers (pinned numpy in,
    numpy out): ProductQuantizer.compress + FlatQuantizedIndex.search_with_scores."""
    torch, tdist = ctx["torch"], ctx["tdist"]
    from b200vq import dist as vdist, host, ops_sq
    from haag_vq.methods.product_quantization import ProductQuantizer
    from haag_vq.methods.search.flat_quantized_index import FlatQuantizedIndex

    world, dev, nq = ctx["world"], ctx["dev"], ctx["nq"]
    rows = min(E2E_ROWS, args.rows_per_gpu)
    pq = ProductQuantizer._from_state({"type": "ProductQuantizer", "M": M, "B": 8, "D": D,
                                       "flat_centroids": ctx["C"].cpu().numpy().reshape(-1)})
    index = FlatQuantizedIndex(pq)
    index.attach_device_codes(ctx["codes"], D, "l2")
    xh = torch.empty((rows, D), dtype=torch.float32, pin_memory=True)
    xh.copy_(ctx["xbuf"][0][:rows])
    qh = torch.empty((nq, D), dtype=torch.float32, pin_memory=True)
    qh.copy_(ctx["Qd"])
    torch.cuda.synchronize()
    Xh, Qh = xh.numpy(), qh.numpy()

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            tdist.barrier()           # the two legs are timed separately: no rank's H2D skew leaks into the search leg
            torch.cuda.synchronize()

    def one():
        sync_all()
        a = time.perf_counter()
        codes_h = pq.compress(Xh)
        torch.cuda.synchronize()
        b = time.perf_counter()
        sync_all()
        c = time.perf_counter()
        ids, sc = index.search_with_scores(Qh, K)
        if world > 1:
            v, i = vdist.sharded_topk(host.to_device(sc), host.to_device(ids.view(np.int32)),
                                      ctx["row_offset"], K, "l2")
            sc, ids = host.to_host(v), host.to_host(i)
        torch.cuda.synchronize()
        d = time.perf_counter()
        assert codes_h.shape == (rows, M) and ids.shape == (nq, K)
        return b - a, d - c

    one()
    reps = max(1, min(args.steps, 3))
    te = ts = 0.0
    for _ in range(reps):
        a, b = one()
        te += a
        ts += b
    mine = te
    t = torch.tensor([te, ts], dtype=torch.float64, device=dev)
    if world > 1:
        tdist.all_reduce(t, op=tdist.ReduceOp.MAX)
    te, ts = (float(v) for v in t.tolist())
    h2d = rows * D * 4
    per_rank = torch.tensor([h2d * reps / mine / 1e9], dtype=torch.float64, device=dev)
    gathered = [torch.zeros_like(per_rank) for _ in range(world)]
    if world > 1:
        tdist.all_gather(gathered, per_rank)
    else:
        gathered = [per_rank]
    return {"value": rows * world * reps / te, "unit": "vectors/s",
            "adc": {"value": float(ctx["n"]) * world * nq * reps / ts, "unit": "codes/s"},
            "h2d_bytes_per_step": rows * D * 4 + nq * D * 4, "d2h_bytes_per_step": rows * M + nq * K * 8,
            "h2d_gbs_per_rank": [round(float(g.item()), 2) for g in gathered],
            "api": "haag_vq ProductQuantizer.compress(X) + FlatQuantizedIndex.search_with_scores(Q, 10); "
                   f"pinned host X ({rows}x{D}) and Q, numpy results; codes of the shard resident in HBM",
            "timing": "host perf_counter around each call, device synchronised and ranks aligned by a barrier before "
                      "each of the two legs, max over ranks; the encode leg is bound by the host link (h2d_gbs_per_rank: "
                      "input bytes / encode time of every rank; all GPUs of the box share the host's PCIe / DRAM)"}


def run_configs(ctx):
    """C1-C4 + SAQ-engine + IVF kernel table (tools/config_bench.py), CUDA-event timed on this box, outside the timed
    region of the headline: one row per kernel with its algorithmic bytes / FLOP and fraction of the measured peak."""
    torch = ctx["torch"]
    for key in ("codes", "codes_rot", "xbuf"):
        ctx[key] = None                      # free the C5 shard before the 1M x 1536 / 3072 configs
    ctx["step"] = ctx["scan"] = None
    import gc
    gc.collect()
    torch.cuda.empty_cache()
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    try:
        import config_bench
        rows = config_bench.run(("c1", "c2", "c3", "c4", "saq", "ivf"), device=ctx["dev"], quick=True)
        for r in rows:
            for k, v in list(r.items()):
                if isinstance(v, float):
                    r[k] = float(f"{v:.5g}")
        return {"peaks": config_bench.peaks, "timing": "CUDA events, 3 warm-ups, inputs resident in HBM and larger than L2",
                "rows": rows}
    except Exception as e:  # the headline line must still print
        return {"error": f"{type(e).__name__}: {e}"}


def main():
    args = parse_args()
    # libraries (NCCL's version banner) write to fd 1: keep stdout clean for the ONE JSON line
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)

    def emit(obj):
        os.write(real_stdout, (json.dumps(obj) + "\n").encode())
    global print_line
    print_line = emit
    if args.impl == "reference":
        run_reference(args)
        return
    line, ctx = run_b200(args)
    if not args.no_e2e:
        line["e2e"] = run_e2e(args, ctx)
    if ctx["rank"] == 0:
        if ctx["world"] == 1 and not args.no_configs:
            line["configs"] = run_configs(ctx)
        if ctx["world"] == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline_leg()
        print_line(line)
    if ctx["world"] > 1:
        ctx["tdist"].barrier()
        ctx["tdist"].destroy_process_group()


if __name__ == "__main__":
    main()
