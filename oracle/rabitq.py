"""CPU oracle: Extended (multi-bit) RaBitQ and 1-bit RaBitQ (faiss semantics).

TEST INFRASTRUCTURE ONLY — see oracle/__init__.py.

Extended RaBitQ restates src/haag_vq/methods/extended_rabitq.py:6-44,90-198
(PINNED against the reference class by oracle/make_golden.py).

1-bit RaBitQ restates faiss RaBitQuantizer / IndexRaBitQ as driven by
src/haag_vq/methods/rabit_quantization.py:20-29 and
src/haag_vq/methods/search/rabitq_index.py:42-70. PARITY UNPINNED: faiss
(``faiss-cpu`` >= 1.11, unpinned) is not in the reference tree nor in this image;
restated from faiss/impl/RaBitQuantizer.cpp (compute_codes_core, decode_core,
RaBitDistanceComputerQ) and Gao & Long 2024.
"""
from __future__ import annotations

import numpy as np

L2, IP = 0, 1


# ------------------------------------------------------------------ extended
def lloyd_1d_normal(num_levels: int, seed: int, n_samples: int = 200_000, max_iter: int = 100,
                    tol: float = 1e-7) -> np.ndarray:
    """extended_rabitq.py:6-44 — sampled Lloyd-Max levels for N(0,1)."""
    rng = np.random.default_rng(seed)
    samples = rng.standard_normal(n_samples)
    qs = (np.arange(num_levels) + 0.5) / num_levels
    levels = np.quantile(samples, qs)
    for _ in range(max_iter):
        boundaries = 0.5 * (levels[:-1] + levels[1:])
        idx = np.searchsorted(boundaries, samples)
        new_levels = levels.copy()
        for k in range(num_levels):
            mask = idx == k
            if np.any(mask):
                new_levels[k] = samples[mask].mean()
        new_levels.sort()
        shift = float(np.max(np.abs(new_levels - levels)))
        levels = new_levels
        if shift < tol:
            break
    return levels.astype(np.float64)


def erabitq_fit(X: np.ndarray, num_bits: int, seed: int = 0) -> dict:
    """extended_rabitq.py:90-106."""
    X = np.asarray(X)
    D = X.shape[1]
    c = X.mean(axis=0).astype(np.float64)
    rng = np.random.default_rng(seed)
    Qm, _ = np.linalg.qr(rng.standard_normal((D, D)))
    return dict(c=c, P=Qm.astype(np.float64), levels=lloyd_1d_normal(2 ** num_bits, seed=seed), D=D,
                B=int(num_bits))


def erabitq_compress(X: np.ndarray, st: dict) -> np.ndarray:
    """extended_rabitq.py:125-170."""
    X = np.asarray(X, dtype=np.float64)
    N, D = X.shape
    B, levels = st["B"], st["levels"]
    r = X - st["c"]
    nrm = np.linalg.norm(r, axis=1)
    o = r / np.maximum(nrm, 1e-12)[:, None]
    s = (o @ st["P"]) * np.sqrt(D)
    idx = np.searchsorted(0.5 * (levels[:-1] + levels[1:]), s).astype(np.int64)
    s_hat = levels[idx]
    num = np.einsum("nd,nd->n", s, s_hat)
    den = np.einsum("nd,nd->n", s_hat, s_hat)
    t = np.where(den > 1e-12, num / den, 1.0)
    ib = (D * B + 7) // 8
    out = np.zeros((N, ib + 8), dtype=np.uint8)
    pos = np.arange(B - 1, -1, -1)
    bits = ((idx[:, :, None] >> pos) & 1).astype(np.uint8).reshape(N, D * B)
    out[:, :ib] = np.packbits(bits, axis=1)
    out[:, ib:ib + 4] = nrm.astype(np.float32).view(np.uint8).reshape(N, 4)
    out[:, ib + 4:ib + 8] = t.astype(np.float32).view(np.uint8).reshape(N, 4)
    return out


def erabitq_unpack(codes: np.ndarray, st: dict):
    D, B = st["D"], st["B"]
    ib = (D * B + 7) // 8
    codes = np.ascontiguousarray(codes, dtype=np.uint8)
    N = codes.shape[0]
    bits = np.unpackbits(codes[:, :ib], axis=1)[:, :D * B].reshape(N, D, B).astype(np.int64)
    idx = bits @ (1 << np.arange(B - 1, -1, -1)).astype(np.int64)
    nrm = codes[:, ib:ib + 4].copy().view(np.float32).reshape(N).astype(np.float64)
    t = codes[:, ib + 4:ib + 8].copy().view(np.float32).reshape(N).astype(np.float64)
    return idx, nrm, t


def erabitq_decompress(codes: np.ndarray, st: dict) -> np.ndarray:
    """extended_rabitq.py:173-198."""
    idx, nrm, t = erabitq_unpack(codes, st)
    o_hat = (st["levels"][idx] / np.sqrt(st["D"])) * t[:, None]
    return ((o_hat @ st["P"].T) * nrm[:, None] + st["c"]).astype(np.float32)


# ------------------------------------------------------------------ 1-bit (faiss)
FLT_EPS = np.float32(np.finfo(np.float32).eps)


def rabitq1_code_size(D: int) -> int:
    return (D + 7) // 8 + 8


def rabitq1_encode(X: np.ndarray, centroid: np.ndarray | None, metric: int) -> np.ndarray:
    """RaBitQuantizer::compute_codes_core (fp32)."""
    X = np.ascontiguousarray(X, dtype=np.float32)
    n, D = X.shape
    c = np.zeros(D, dtype=np.float32) if centroid is None else centroid.astype(np.float32)
    omc = X - c
    norm = np.einsum("nd,nd->n", omc, omc, dtype=np.float32)
    orl2 = np.einsum("nd,nd->n", X, X, dtype=np.float32)
    bits = (omc > 0)
    dp = np.abs(omc).sum(axis=1, dtype=np.float32)
    inv_d_sqrt = np.float32(1.0) / np.sqrt(np.float32(D))
    sq = np.sqrt(norm)
    inv_norm = np.where(np.abs(norm) < FLT_EPS, np.float32(1), np.float32(1) / sq).astype(np.float32)
    dpn = (dp * inv_norm * inv_d_sqrt).astype(np.float32)
    inv_dp = np.where(np.abs(dpn) < FLT_EPS, np.float32(1), np.float32(1) / dpn).astype(np.float32)
    f0 = norm - orl2 if metric == IP else norm
    f1 = (inv_dp * sq).astype(np.float32)
    nb = (D + 7) // 8
    out = np.zeros((n, nb + 8), dtype=np.uint8)
    out[:, :nb] = np.packbits(bits, axis=1, bitorder="little")
    out[:, nb:nb + 4] = f0.astype(np.float32).view(np.uint8).reshape(n, 4)
    out[:, nb + 4:nb + 8] = f1.view(np.uint8).reshape(n, 4)
    return out


def rabitq1_decode(codes: np.ndarray, D: int, centroid: np.ndarray | None) -> np.ndarray:
    """RaBitQuantizer::decode_core."""
    nb = (D + 7) // 8
    n = codes.shape[0]
    c = np.zeros(D, dtype=np.float32) if centroid is None else centroid.astype(np.float32)
    bits = np.unpackbits(np.ascontiguousarray(codes[:, :nb]), axis=1, bitorder="little")[:, :D]
    dpm = codes[:, nb + 4:nb + 8].copy().view(np.float32).reshape(n)
    inv_d_sqrt = np.float32(1.0) / np.sqrt(np.float32(D))
    v = (bits.astype(np.float32) - np.float32(0.5)) * dpm[:, None] * np.float32(2) * inv_d_sqrt
    return (v + c).astype(np.float32)


def rabitq1_prepare_query(q: np.ndarray, centroid: np.ndarray | None, qb: int):
    """RaBitDistanceComputerQ::set_query."""
    q = q.astype(np.float32)
    D = q.shape[0]
    c = np.zeros(D, dtype=np.float32) if centroid is None else centroid.astype(np.float32)
    r = q - c
    l2 = np.float32(np.dot(r, r))
    nrm = np.float32(np.dot(q, q))
    vmin, vmax = r.min(), r.max()
    delta = np.float32((vmax - vmin) / np.float32((1 << qb) - 1))
    inv_delta = np.float32(1) / delta
    vq = (r - vmin) * inv_delta
    vqq = np.floor(np.abs(vq) + 0.5) * np.sign(vq)  # std::round
    vqq = vqq.astype(np.int64)
    sum_q = int(vqq.sum())
    qq = np.clip(vqq, 0, 255).astype(np.uint8)
    inv_d = np.float32(1.0) / np.sqrt(np.float32(D))
    c1 = np.float32(2) * delta * inv_d
    c2 = np.float32(2) * vmin * inv_d
    c34 = inv_d * (delta * np.float32(sum_q) + np.float32(D) * vmin)
    return qq, (c1, c2, np.float32(c34), l2, nrm)


def rabitq1_estimate(codes: np.ndarray, D: int, Q: np.ndarray, centroid, qb: int, metric: int):
    """Estimated distance of every (query, code): RaBitDistanceComputerQ::distance_to_code.
    Returns (nq, n) fp32: squared L2 estimate, or inner-product estimate for IP."""
    nb = (D + 7) // 8
    n = codes.shape[0]
    bits = np.unpackbits(np.ascontiguousarray(codes[:, :nb]), axis=1, bitorder="little")[:, :D]
    bits = bits.astype(np.int64)
    f0 = codes[:, nb:nb + 4].copy().view(np.float32).reshape(n)
    f1 = codes[:, nb + 4:nb + 8].copy().view(np.float32).reshape(n)
    sumb = bits.sum(1).astype(np.float32)
    out = np.empty((Q.shape[0], n), dtype=np.float32)
    for i in range(Q.shape[0]):
        qq, (c1, c2, c34, l2, nrm) = rabitq1_prepare_query(Q[i], centroid, qb)
        dot = (bits @ qq.astype(np.int64)).astype(np.float32)  # == sum_j popc(plane_j & code) << j
        final = c1 * dot
        final = final + c2 * sumb
        final = final - c34
        pre = f0 + l2 - np.float32(2) * f1 * final
        out[i] = pre if metric == L2 else np.float32(-0.5) * (pre - nrm)
    return out
