"""CPU oracle for the SAQ engine path.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py). Restates the reference's C++ engine
``external/saq`` (commit 73e730e, vendored source; it cannot be built here: Eigen 3.4 / glog /
fmt / gflags are fetched from the network by its CMake, external/saq/CMakeLists.txt:22-76).

  pca_fit            src/preprocessing/pca_eigen.cpp:17-47
  kmeans1_row        src/preprocessing/kmeans_eigen.cpp:17-26,40-56  (K = 1: Lloyd stops before its
                     first update, the centroid stays the k-means++ seed row)
  plan_dp            src/quantization_plan.cpp:144-255 (C: oracle/caq.c; pure Python twin below)
  equal_segmentation src/quantization_plan.cpp:125-138
  caq_encode         include/saq/caq_encoder.h:67-140,170-261 (C: oracle/caq.c; pure Python twin below)
  fit                src/ivf_index.cpp:186-243 + include/saq/saq_quantizer.h:49-83 + quantizer.h:56-104
  decompress         src/ivf_index.cpp:245-374
  accurate_dist      include/saq/caq_estimator.h:254-283,341-375 + saq_estimator.h:183-190

PARITY UNPINNED against the CPU engine. Three details cannot be recovered from the source and are
fixed here (and shared with the device path as trained state): Eigen's float reduction order
(sequential float sums are used), the per-segment rotators (Eigen ``Random`` = unseeded
``std::rand``; a seeded numpy QR is used) and the standard library behind
``std::uniform_int_distribution`` (libstdc++ >= 11 multiply-shift is used).
The encoder is pinned against the reference's own CUDA twin instead (tests/test_gpu_caq.py).
"""
from __future__ import annotations

import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
BLK = 64          # kDimPaddingSize (defines.h:29)
MAX_BITS = 13     # KMaxQuantizeBits (defines.h:27)
NBF = 64          # kNumShortFactors * sizeof(float) * 8 (quantization_plan.h:166, .cpp:149)


def _lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "_ref", "liboracle_caq.so")
        if not os.path.exists(path):
            raise RuntimeError(f"{path} missing: run `make -C oracle`")
        lib = ctypes.CDLL(path)
        P = ctypes.c_void_p
        lib.oracle_caq_encode.argtypes = [P, ctypes.c_int64, ctypes.c_int, ctypes.c_int64, ctypes.c_int,
                                          ctypes.c_int, ctypes.c_float, P, P, P, P, P, P]
        lib.oracle_caq_encode.restype = None
        lib.oracle_saq_plan_dp.argtypes = [P, ctypes.c_int, ctypes.c_float, P, P]
        lib.oracle_saq_plan_dp.restype = ctypes.c_int
        _LIB = lib
    return _LIB


def pad64(d: int) -> int:
    return (d + BLK - 1) // BLK * BLK


# ------------------------------------------------------------------ preprocessing
def pca_fit(X: np.ndarray):
    """mean (D,), rotation (D, D) with columns by DESCENDING eigenvalue, variances (D,) fp32."""
    X = np.asarray(X, dtype=np.float32)
    n = X.shape[0]
    mean = X.mean(axis=0, dtype=np.float32)
    Xc = X - mean
    cov = (Xc.T @ Xc).astype(np.float64) / float(n - 1)
    evals, evecs = np.linalg.eigh(cov)
    rot = np.ascontiguousarray(evecs[:, ::-1].astype(np.float32))
    var = np.maximum(evals[::-1], 0.0).astype(np.float32)
    return mean, rot, var


def mt19937_first(seed: int) -> int:
    """First output of std::mt19937(seed)."""
    mt = [0] * 624
    mt[0] = seed & 0xFFFFFFFF
    for i in range(1, 624):
        mt[i] = (1812433253 * (mt[i - 1] ^ (mt[i - 1] >> 30)) + i) & 0xFFFFFFFF
    y = (mt[0] & 0x80000000) | (mt[1] & 0x7FFFFFFF)
    v = mt[397] ^ (y >> 1) ^ (0x9908B0DF if y & 1 else 0)
    v ^= v >> 11
    v ^= (v << 7) & 0x9D2C5680
    v ^= (v << 15) & 0xEFC60000
    v ^= v >> 18
    return v & 0xFFFFFFFF


def kmeans1_row(n: int, seed: int) -> int:
    """Row index std::uniform_int_distribution<int>(0, n-1) draws first from mt19937(seed)
    (libstdc++ >= 11: 64-bit multiply-shift; the rejection branch is not reachable on the first
    draw for n < 2^31 unless low < n % ..., which is then handled by the caller re-drawing)."""
    r = mt19937_first(seed)
    prod = r * n
    low = prod & 0xFFFFFFFF
    if low < n:
        thr = ((1 << 32) - n) % n
        if low < thr:
            raise NotImplementedError("rejection on the first draw: pick another seed")
    return prod >> 32


# ------------------------------------------------------------------ plan
def equal_segmentation(d_pad: int, avg_bits: float, num_segs: int):
    b = int(avg_bits)
    plan, d = [], 0
    for i in range(num_segs):
        if d >= d_pad:
            break
        t = pad64((d_pad - d) // (num_segs - i))
        plan.append((t, b))
        d += t
    return plan


def plan_dp(variance: np.ndarray, avg_bits: float):
    var = np.ascontiguousarray(variance, dtype=np.float32)
    d_pad = var.shape[0]
    assert d_pad % BLK == 0
    nmax = d_pad // BLK + 1
    sl = np.zeros(nmax, dtype=np.int32)
    sb = np.zeros(nmax, dtype=np.int32)
    ns = _lib().oracle_saq_plan_dp(var.ctypes.data, d_pad, ctypes.c_float(avg_bits), sl.ctypes.data, sb.ctypes.data)
    if ns < 0:
        raise MemoryError("plan_dp")
    return [(int(sl[i]), int(sb[i])) for i in range(ns)]


def plan_dp_py(variance: np.ndarray, avg_bits: float):
    """Pure-Python twin of the DP with dict-sparse tables (small d_pad only)."""
    var = np.asarray(variance, dtype=np.float32)
    d_pad = var.shape[0]
    i_end = d_pad // BLK
    tot_bits = int(np.float32(avg_bits) * np.float32(d_pad) + np.float32(NBF))
    max_segs = max(1, i_end if avg_bits < 2 else i_end // 2)
    blk = [float(np.float32(var[i * BLK:(i + 1) * BLK].astype(np.float32).cumsum(dtype=np.float32)[-1]))
           for i in range(i_end)]
    INF = float("inf")
    f = {(0, 0, 0): (0.0, 0)}
    ans = None
    ans_val = INF
    for ns in range(max_segs + 1):
        for i in range(i_end + 1):
            for ub in range(tot_bits + 1):
                cur = f.get((ns, i, ub))
                if cur is None:
                    continue
                cv = cur[0]
                if i == i_end:
                    if cv * 1.01 < ans_val:
                        ans, ans_val = (ns, i, ub), cv
                    continue
                if ns == max_segs:
                    continue
                vs = 0.0
                for j in range(1, i_end - i + 1):
                    vs += blk[i + j - 1]
                    for b in range(1, MAX_BITS + 1):
                        bn = ub + b * j * BLK + NBF
                        if bn > tot_bits:
                            break
                        v = vs / (1 << b)
                        key = (ns + 1, i + j, bn)
                        old = f.get(key)
                        if old is None or old[0] > cv + v:
                            f[key] = (cv + v, (i << 4) + b)
                key = (ns + 1, i_end, ub)
                old = f.get(key)
                if old is None or cv + vs < old[0]:
                    f[key] = (cv + vs, (i << 4) + 0)
    # the reference compares against f[0][i_end][tot_bits] = DBL_MAX first, so any finite terminal wins
    plan = []
    if ans is None:
        return plan
    ns, i, B = ans
    while i > 0:
        s = f[(ns, i, B)][1]
        prev_i, bits = s >> 4, s & 0xF
        ln = (i - prev_i) * BLK
        plan.append((ln, bits))
        ns -= 1
        i = prev_i
        if bits:
            B -= bits * ln + NBF
    return plan[::-1]


def make_rotators(plan, seed: int):
    """Per-segment orthogonal P (len, len): P = Q^T of a QR of a random matrix (rotator.h:43-48).
    TRAINED STATE: the reference draws from unseeded std::rand; only orthogonality is specified."""
    rng = np.random.default_rng(seed)
    out = []
    for ln, _ in plan:
        Q, _r = np.linalg.qr(rng.standard_normal((ln, ln)))
        out.append(np.ascontiguousarray(Q.T.astype(np.float32)))
    return out


# ------------------------------------------------------------------ encoder
def caq_encode(o: np.ndarray, bits: int, rd_lmt: int = 6, adj_eps: float = 1e-8):
    """o (n, len) fp32 rotated residuals of one segment -> dict(codes u16 (n,len), l2norm, rescale,
    error, mse, cos: (n,) fp32)."""
    o = np.asarray(o, dtype=np.float32)
    if o.strides[1] != 4:
        o = np.ascontiguousarray(o)
    n, ln = o.shape
    codes = np.zeros((n, ln), dtype=np.uint16)
    outs = [np.zeros(n, dtype=np.float32) for _ in range(5)]
    if n:
        _lib().oracle_caq_encode(o.ctypes.data, n, ln, o.strides[0] // 4, bits, rd_lmt, ctypes.c_float(adj_eps),
                                 codes.ctypes.data, *[a.ctypes.data for a in outs])
    return dict(codes=codes, l2norm=outs[0], rescale=outs[1], error=outs[2], mse=outs[3], cos=outs[4])


def caq_encode_row_py(o: np.ndarray, bits: int, rd_lmt: int = 6, adj_eps: float = 1e-8):
    """Pure-Python twin of one row (bits >= 1): returns (codes list, ip_o_oa, oa_l2sqr)."""
    f32 = np.float32
    o = [f32(v) for v in o]
    ln = len(o)
    code_max = (1 << bits) - 1
    v_mx = float(max(max(o), -min(o)))
    v_mi = -v_mx
    delta = (v_mx - v_mi) / (code_max + 1)
    fs = f32(0)
    for v in o:
        fs = f32(fs + v)
    vec_sum = float(fs)
    code, ip_o_code, l2, cs = [], 0.0, 0, 0
    for v in o:
        cf = f32(0)
        if delta:
            cf = f32(np.floor(f32(f32(v - f32(v_mi)) / f32(delta))))
            cf = min(cf, f32(code_max))
        ip_o_code += float(f32(cf * v))
        c = int(cf)
        code.append(c)
        l2 += c * c
        cs += c
    ip = ip_o_code * delta + (v_mi + 0.5 * delta) * vec_sum
    oa2 = delta * delta * l2 + (delta * delta + 2 * delta * v_mi) * cs
    oa2 += (0.25 * delta * delta + delta * v_mi + v_mi * v_mi) * ln
    if rd_lmt and oa2:
        re_eps = float(f32(adj_eps)) * oa2
        adj, rnd = 1, 1
        while adj:
            adj = 0
            for j in range(ln):
                cur = 0
                of = float(o[j])
                oa = (code[j] + 0.5) * delta + v_mi
                c = code[j]
                tmp = oa2 - oa * oa
                ipd = delta * of
                while c < code_max:
                    nq = oa + delta
                    nl = tmp + nq * nq
                    nip = ip + ipd
                    if (ip * ip + re_eps) * nl >= nip * nip * oa2:
                        break
                    c += 1; ip = nip; oa = nq; oa2 = nl; cur += 1
                while c > 0:
                    nq = oa - delta
                    nl = tmp + nq * nq
                    nip = ip - ipd
                    if (ip * ip + re_eps) * nl >= nip * nip * oa2:
                        break
                    c -= 1; ip = nip; oa = nq; oa2 = nl; cur += 1
                if code[j] != c:
                    code[j] = c
                    adj += cur
            if rd_lmt and rnd >= rd_lmt:
                break
            rnd += 1
            cl2 = cip = 0.0
            for j in range(ln):
                q = (code[j] + 0.5) * delta + v_mi
                cip += q * float(o[j])
                cl2 += q * q
            oa2, ip = cl2, cip
    return code, ip, oa2


# ------------------------------------------------------------------ engine pipeline
def _seq_norm_f32(a: np.ndarray) -> np.ndarray:
    """Row-wise sqrt of a sequential float32 sum of squares (Eigen .norm(), order fixed here)."""
    acc = np.zeros(a.shape[0], dtype=np.float32)
    for j in range(a.shape[1]):
        acc = acc + a[:, j] * a[:, j]
    return np.sqrt(acc)


def fit(X: np.ndarray, avg_bits: float, apply_pca: bool = True, seed: int = 0, rot_seed: int = 20240607,
        seg_eqseg: int = 0, state_in: dict | None = None) -> dict:
    """IVF(n, D, K=1, cfg).fit(X, apply_pca, 1, seed): PCA -> K=1 centroid -> DP plan on the PCA
    variances -> per segment: rotate, subtract the rotated centroid, CAQ-encode every row.
    `state_in` (mean, rot, var, cent_row, plan, P) replaces the trained state when given, so the
    device path and the oracle can be compared conditional on shared state."""
    X = np.ascontiguousarray(X, dtype=np.float32)
    n, D = X.shape
    d_pad = pad64(D)
    st = dict(n=n, D=D, d_pad=d_pad, avg_bits=float(avg_bits), apply_pca=bool(apply_pca))
    if state_in is not None:
        for k in ("mean", "rot", "var", "cent_row", "plan", "P"):
            st[k] = state_in[k]
    else:
        if apply_pca:
            st["mean"], st["rot"], st["var"] = pca_fit(X)
        else:
            st["mean"] = np.zeros(D, np.float32)
            st["rot"] = np.eye(D, dtype=np.float32)
            st["var"] = np.ones(D, np.float32)
        st["cent_row"] = kmeans1_row(n, seed)
        var_pad = np.zeros(d_pad, np.float32)
        var_pad[:D] = st["var"]
        st["plan"] = equal_segmentation(d_pad, avg_bits, seg_eqseg) if seg_eqseg > 0 else plan_dp(var_pad, avg_bits)
        st["P"] = make_rotators(st["plan"], rot_seed)
    Xp = ((X - st["mean"]) @ st["rot"]).astype(np.float32) if apply_pca else X
    cen = Xp[st["cent_row"]].copy()
    st["centroid"] = cen
    segs, off = [], 0
    for (ln, bits), P in zip(st["plan"], st["P"]):
        cp = max(0, min(ln, D - off))
        v = np.zeros((n, ln), np.float32)
        v[:, :cp] = Xp[:, off:off + cp]
        c = np.zeros(ln, np.float32)
        c[:cp] = cen[off:off + cp]
        c_rot = (c[None, :] @ P).astype(np.float32)[0]
        o = ((v @ P).astype(np.float32) - c_rot).astype(np.float32)
        enc = caq_encode(o, bits)
        enc.update(len=ln, bits=bits, off=off, c_rot=c_rot)
        segs.append(enc)
        off += ln
    st["segs"] = segs
    return st


def decompress(st: dict, ids: np.ndarray) -> np.ndarray:
    ids = np.asarray(ids, dtype=np.int64)
    D = st["D"]
    o_rot = np.zeros((ids.shape[0], st["d_pad"]), np.float32)
    for seg, P in zip(st["segs"], st["P"]):
        ln, bits, off = seg["len"], seg["bits"], seg["off"]
        cen_unrot = (seg["c_rot"][None, :] @ P.T).astype(np.float32)[0]
        if bits == 0:
            sv = np.broadcast_to(cen_unrot, (ids.shape[0], ln)).copy()
        else:
            delta = np.float32(2.0) / np.float32(1 << bits)
            oa = (seg["codes"][ids].astype(np.float32) + np.float32(0.5)) * delta + np.float32(-1.0)
            nrm = _seq_norm_f32(oa)
            ok = nrm > np.float32(1e-9)
            scale = np.where(ok, seg["l2norm"][ids] / np.where(ok, nrm, np.float32(1)), np.float32(1)).astype(np.float32)
            oa = oa * scale[:, None]
            sv = (oa @ P.T).astype(np.float32) + cen_unrot
        o_rot[:, off:off + ln] = sv
    o_rot = o_rot[:, :D]
    if st["apply_pca"]:
        return ((o_rot @ st["rot"].T).astype(np.float32) + st["mean"]).astype(np.float32)
    return o_rot


def accurate_dist(st: dict, Q: np.ndarray, ids: np.ndarray | None = None) -> np.ndarray:
    """Engine estimator, L2: sum over segments of |o|^2 + |q|^2 - 2 rescale <q, oa_unit> with
    q = (query_pca_seg P) - centroid_rot; 0-bit segments contribute |o|^2 + |q|^2.
    Q is given in the ORIGINAL space (the PCA transform is applied here). Returns (nq, m) fp32."""
    Q = np.ascontiguousarray(Q, dtype=np.float32)
    Qp = ((Q - st["mean"]) @ st["rot"]).astype(np.float32) if st["apply_pca"] else Q
    n = st["n"]
    ids = np.arange(n) if ids is None else np.asarray(ids, dtype=np.int64)
    out = np.zeros((Q.shape[0], ids.shape[0]), np.float32)
    D = st["D"]
    for seg, P in zip(st["segs"], st["P"]):
        ln, bits, off = seg["len"], seg["bits"], seg["off"]
        cp = max(0, min(ln, D - off))
        q = np.zeros((Q.shape[0], ln), np.float32)
        q[:, :cp] = Qp[:, off:off + cp]
        q = ((q @ P).astype(np.float32) - seg["c_rot"]).astype(np.float32)
        q2 = np.einsum("qd,qd->q", q, q, dtype=np.float32)
        o2 = (seg["l2norm"][ids] * seg["l2norm"][ids]).astype(np.float32)
        if bits == 0:
            out += o2[None, :] + q2[:, None]
            continue
        delta = np.float32(2.0 / (1 << bits))
        codes = seg["codes"][ids].astype(np.float32)
        ip = (q @ codes.T).astype(np.float32) * delta + (np.float32(-1.0) + delta / np.float32(2)) * q.sum(1, dtype=np.float32)[:, None]
        ip_o_q = seg["rescale"][ids][None, :] * ip
        out += (o2[None, :] + q2[:, None] - np.float32(2) * ip_o_q).astype(np.float32)
    return out
