"""Round-2 kernels at their BASELINE shapes on one B200 (CUDA events, inputs resident): the code-domain
estimator scan (ext-RaBitQ C4, SAQ engine), ext-RaBitQ encode with the tensor-core rotation, SAQ engine fit."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "vector-quantization_b200"), ROOT]
import numpy as np
import torch

from b200vq import ops, ops_sq, saq_engine as E

dev = torch.device("cuda")
out = {}


def timed(fn, reps=3, warm=2):
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


def data(n, D, seed=1234):
    return ops_sq.synth_rows(n, D, 0, seed, torch.linspace(4.0, 0.2, D, device=dev), True, device=dev)


which = sys.argv[1].split(",") if len(sys.argv) > 1 else ["erq", "saq"]
if "erq" in which:
    from haag_vq.methods.extended_rabitq import ExtendedRaBitQuantizer
    D, B, nq = 1536, 4, 1000
    n = int(os.environ.get("ERQ_ROWS", 1_000_000))
    x = data(n, D)
    q = ExtendedRaBitQuantizer(B, seed=0)
    q.fit(x[:4096].cpu().numpy())
    st = q._state()
    nb = 131_072
    xb = x[:nb]
    codes = torch.empty((n, q.code_size), dtype=torch.uint8, device=dev)

    def enc(s, e):
        xs = x[s:e]
        nrm = ops_sq.resid_norms_f64(xs, st["c"])
        y = st["fwd"](xs, x_sub=st["c32"])
        ops_sq.erabitq_pack_tc(y, xs, st["c"], st["PT"], nrm, B, st["levels"], st["bounds"], st["err_rel"],
                               2 * st["err_abs"], out=codes[s:e], stats=st["nfix"])
    out["erq_encode_131072_ms"] = timed(lambda: enc(0, nb))
    nrm = ops_sq.resid_norms_f64(xb, st["c"])
    y = st["fwd"](xb, x_sub=st["c32"])
    out["erq_resid_norms_ms"] = timed(lambda: ops_sq.resid_norms_f64(xb, st["c"]))
    out["erq_rotate_ms"] = timed(lambda: st["fwd"](xb, x_sub=st["c32"]))
    out["erq_pack_tc_ms"] = timed(lambda: ops_sq.erabitq_pack_tc(y, xb, st["c"], st["PT"], nrm, B, st["levels"], st["bounds"],
                                                                 st["err_rel"], 2 * st["err_abs"], out=codes[:nb]))
    for s in range(0, n, nb):
        enc(s, min(n, s + nb))
    torch.cuda.synchronize()
    Q = data(nq, D, seed=77)
    sc = ops_sq.ERaBitQScanner(codes, D, B, st["levels32"])
    qrot, qconst = q.rotate_queries(Q, "l2")
    out["erq_scan_prep_ms"] = timed(lambda: ops_sq.ERaBitQScanner(codes, D, B, st["levels32"]))
    for k in (10, 100):
        out[f"erq_est_scan_{n}_k{k}_ms"] = timed(lambda: sc.search(qrot, qconst, k, "l2"))
    v, i = sc.search(qrot, qconst, 10, "l2")
    # recall@10 against the exact search
    xn = ops.row_sqnorms(x)
    gv, gi = ops.flat_topk(x, Q, 10, "l2", xnorm=xn)
    rec = np.mean([len(set(a) & set(b)) / 10 for a, b in zip(i.cpu().numpy().tolist(), gi.cpu().numpy().tolist())])
    out["erq_recall10_B4"] = float(rec)
    out["erq_est_codes_per_s"] = n * nq / out[f"erq_est_scan_{n}_k10_ms"] * 1e3
    out["erq_encode_vec_per_s"] = nb / out["erq_encode_131072_ms"] * 1e3
    del x, codes, sc
    torch.cuda.empty_cache()

if "saq" in which:
    n, D = 99_000, 1536
    X = data(n, D)
    for ab in (4.0,):
        t0 = time.time()
        eng = E.SaqEngine(ab).fit(X)
        torch.cuda.synchronize()
        out[f"saq_fit_{n}_ab{ab}_wall_s"] = time.time() - t0
        out["saq_plan"] = eng.plan
        out["saq_project_T_ms"] = timed(lambda: eng._project(X, transpose_out=True))
        oT = eng._project(X, transpose_out=True)
        tot = 0.0
        for seg in eng.segs:
            t = timed(lambda: E.caq_encode_t(oT[seg["off"]:seg["off"] + seg["len"]], seg["bits"], None))
            out[f"saq_caq_encode_len{seg['len']}_b{seg['bits']}_ms"] = t
            tot += t
        out["saq_caq_encode_total_ms"] = tot
        Q = data(1000, D, seed=5)
        eng.search(Q, 10)
        out["saq_search_99000x1000_k10_ms"] = timed(lambda: eng.search(Q, 10))
        ids = torch.arange(0, n, device=dev)
        out["saq_reconstruct_99000_ms"] = timed(lambda: eng.reconstruct(ids))

print(json.dumps(out, indent=1))
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "r2_new_paths.json"), "w"), indent=1)
