"""CPU oracle: product quantizer, OPQ, ADC scan and exact flat search (numpy).

TEST INFRASTRUCTURE ONLY — see oracle/__init__.py.

PARITY UNPINNED: the arithmetic restated here lives in faiss (``faiss-cpu``,
unpinned, API level >= 1.11; reference pyproject.toml:24), which is neither
vendored in the reference tree nor installed in this image, and the reference's
tests hold no golden codes/centroids/distances for it. The
restatement follows the reference's call sites and faiss' published behaviour:

  ProductQuantizer.train/compute_codes/decode
      <- src/haag_vq/methods/product_quantization.py:58-86
  OPQMatrix.train/apply/reverse_transform
      <- src/haag_vq/methods/optimized_product_quantization.py:21-34
  IndexFlatL2 / IndexFlatIP .search
      <- src/haag_vq/benchmarks/exact_search.py:32-51,72-77,
         src/haag_vq/benchmarks/search_bench.py:25-56
  decompress-then-distance search (the parity anchor of the ADC scan)
      <- src/haag_vq/methods/search/flat_quantized_index.py:45-76

Every source of randomness (training subset, centroid seeds, OPQ initial
rotation) is an explicit input so the CUDA path and the oracle share state.
"""
from __future__ import annotations

import numpy as np

L2, IP = 0, 1
MAX_POINTS_PER_CENTROID = 256  # faiss ClusteringParameters default
KMEANS_NITER = 25              # faiss ClusteringParameters default
SPLIT_EPS = np.float32(1.0 / 1024.0)


# ------------------------------------------------------------------ helpers
def subsample_ids(n: int, nmax: int, seed: int) -> np.ndarray:
    """Training subset (faiss subsamples to 256*ksub points with a seeded permutation)."""
    if n <= nmax:
        return np.arange(n, dtype=np.int64)
    return np.random.default_rng(seed).permutation(n)[:nmax].astype(np.int64)


def pq_init_centroids(X: np.ndarray, M: int, ksub: int, seed: int) -> np.ndarray:
    """Initial centroids: ksub distinct training rows per sub-quantizer (seed + m)."""
    X = np.ascontiguousarray(X, dtype=np.float32)
    n, D = X.shape
    dsub = D // M
    if n < ksub:
        raise RuntimeError(
            f"Number of training points ({n}) should be at least as large as number of clusters ({ksub})")
    C = np.empty((M, ksub, dsub), dtype=np.float32)
    for m in range(M):
        ids = np.random.default_rng(seed + m).permutation(n)[:ksub]
        C[m] = X[ids, m * dsub:(m + 1) * dsub]
    return C


def centroid_norms(C: np.ndarray) -> np.ndarray:
    return np.einsum("mkd,mkd->mk", C, C, dtype=np.float32).astype(np.float32)


def pq_sub_distances(Xm: np.ndarray, Cm: np.ndarray) -> np.ndarray:
    """faiss exhaustive_L2sqr_blas: (||x||^2 + ||c||^2) - 2 x.c in fp32, clamped at 0."""
    xn = np.einsum("nd,nd->n", Xm, Xm, dtype=np.float32)
    cn = np.einsum("kd,kd->k", Cm, Cm, dtype=np.float32)
    ip = Xm @ Cm.T  # sgemm
    d = (xn[:, None] + cn[None, :]) - np.float32(2.0) * ip
    np.maximum(d, np.float32(0), out=d)
    return d


def pq_assign(X: np.ndarray, C: np.ndarray, return_gap: bool = False, chunk: int = 16384):
    """compute_codes for nbits <= 8: per sub-quantizer argmin, first index wins ties.

    return_gap=True also returns (d1, d2): best and second-best distance per
    (row, sub-quantizer), used for the 1e-6 relative top-2 exemption.
    """
    X = np.ascontiguousarray(X, dtype=np.float32)
    M, ksub, dsub = C.shape
    n = X.shape[0]
    codes = np.empty((n, M), dtype=np.uint8)
    d1 = np.empty((n, M), dtype=np.float32) if return_gap else None
    d2 = np.empty((n, M), dtype=np.float32) if return_gap else None
    for s in range(0, n, chunk):
        e = min(n, s + chunk)
        for m in range(M):
            d = pq_sub_distances(X[s:e, m * dsub:(m + 1) * dsub], C[m])
            k = np.argmin(d, axis=1)
            codes[s:e, m] = k
            if return_gap:
                if ksub > 1:
                    part = np.partition(d, 1, axis=1)
                    d1[s:e, m] = part[:, 0]
                    d2[s:e, m] = part[:, 1]
                else:
                    d1[s:e, m] = d[:, 0]
                    d2[s:e, m] = np.inf
    if return_gap:
        return codes, d1, d2
    return codes


def pq_assign_exact64(X: np.ndarray, C: np.ndarray, chunk: int = 8192) -> np.ndarray:
    """Float64 ground truth of the same argmin (to study fp32 tie noise)."""
    X = np.asarray(X, dtype=np.float64)
    C = np.asarray(C, dtype=np.float64)
    M, ksub, dsub = C.shape
    n = X.shape[0]
    codes = np.empty((n, M), dtype=np.uint8)
    for s in range(0, n, chunk):
        e = min(n, s + chunk)
        for m in range(M):
            xm = X[s:e, m * dsub:(m + 1) * dsub]
            d = (xm * xm).sum(1)[:, None] + (C[m] * C[m]).sum(1)[None, :] - 2.0 * xm @ C[m].T
            codes[s:e, m] = np.argmin(d, axis=1)
    return codes


def pq_decode(codes: np.ndarray, C: np.ndarray) -> np.ndarray:
    M, ksub, dsub = C.shape
    codes = np.asarray(codes)
    n = codes.shape[0]
    out = np.empty((n, M * dsub), dtype=np.float32)
    for m in range(M):
        out[:, m * dsub:(m + 1) * dsub] = C[m][codes[:, m]]
    return out


def pq_lloyd_update(X: np.ndarray, codes: np.ndarray, C: np.ndarray):
    """One centroid update: mean of the assigned rows (fp32 sums x 1/count);
    empty clusters split the largest cluster with the +-1/1024 perturbation of
    faiss' split_clusters (donor choice made deterministic: largest count,
    lowest index)."""
    M, ksub, dsub = C.shape
    newC = C.copy()
    n_split = 0
    for m in range(M):
        xm = X[:, m * dsub:(m + 1) * dsub]
        cnt = np.bincount(codes[:, m], minlength=ksub).astype(np.int64)
        sums = np.zeros((ksub, dsub), dtype=np.float32)
        np.add.at(sums, codes[:, m], xm)
        nz = cnt > 0
        newC[m][nz] = sums[nz] * (np.float32(1.0) / cnt[nz].astype(np.float32))[:, None]
        for ci in range(ksub):
            if cnt[ci] != 0:
                continue
            cj = int(np.argmax(cnt))
            v = newC[m][cj].copy()
            even = (np.arange(dsub) % 2) == 0
            plus, minus = v * (np.float32(1) + SPLIT_EPS), v * (np.float32(1) - SPLIT_EPS)
            newC[m][ci] = np.where(even, plus, minus)
            newC[m][cj] = np.where(even, minus, plus)
            h = cnt[cj]
            cnt[ci] = h // 2
            cnt[cj] = h - h // 2
            n_split += 1
    return newC, n_split


def pq_train(X: np.ndarray, M: int, nbits: int = 8, niter: int = KMEANS_NITER, seed: int = 1234,
             centroids0: np.ndarray | None = None) -> np.ndarray:
    """faiss ProductQuantizer.train: M independent Lloyd k-means (ksub = 2**nbits)."""
    X = np.ascontiguousarray(X, dtype=np.float32)
    n, D = X.shape
    assert D % M == 0
    ksub = 1 << nbits
    ids = subsample_ids(n, MAX_POINTS_PER_CENTROID * ksub, seed)
    Xt = X[ids] if ids.size != n else X
    C = pq_init_centroids(Xt, M, ksub, seed) if centroids0 is None else centroids0.astype(np.float32).copy()
    if Xt.shape[0] == ksub and centroids0 is None:
        return C  # faiss: as many points as centroids -> copy
    for _ in range(niter):
        codes = pq_assign(Xt, C)
        C, _ = pq_lloyd_update(Xt, codes, C)
    return C


def pq_sse(X: np.ndarray, codes: np.ndarray, C: np.ndarray) -> float:
    R = pq_decode(codes, C)
    diff = X.astype(np.float64) - R.astype(np.float64)
    return float((diff * diff).sum())


# ------------------------------------------------------------------ OPQ
def random_rotation(D: int, seed: int) -> np.ndarray:
    """Seeded random orthonormal matrix (faiss initialises OPQ with one)."""
    A = np.random.default_rng(seed).standard_normal((D, D))
    Qm, _ = np.linalg.qr(A)
    return np.ascontiguousarray(Qm.astype(np.float32))


def opq_apply(X: np.ndarray, A: np.ndarray) -> np.ndarray:
    """OPQMatrix.apply: y = x A^T (A is d_out x d_in, no bias)."""
    return np.ascontiguousarray(np.asarray(X, dtype=np.float32) @ A.T)


def opq_reverse(Y: np.ndarray, A: np.ndarray) -> np.ndarray:
    """reverse_transform for an orthonormal A: x = y A."""
    return np.ascontiguousarray(np.asarray(Y, dtype=np.float32) @ A)


def opq_train(X: np.ndarray, M: int, nbits: int = 8, niter: int = 50, niter_pq: int = 4,
              niter_pq_0: int = 40, seed: int = 123, A0: np.ndarray | None = None) -> np.ndarray:
    """OPQMatrix.train (Ge et al. 2013, non-parametric): alternate PQ training on
    the rotated data with the orthogonal-Procrustes update of the rotation."""
    X = np.ascontiguousarray(X, dtype=np.float32)
    n, D = X.shape
    ksub = 1 << nbits
    ids = subsample_ids(n, MAX_POINTS_PER_CENTROID * 256, seed)
    Xt = X[ids] if ids.size != n else X
    A = random_rotation(D, seed) if A0 is None else A0.astype(np.float32).copy()
    C = None
    for it in range(niter):
        Y = opq_apply(Xt, A)
        if it == 0:
            C = pq_train(Y, M, nbits, niter=niter_pq_0, seed=seed)
        else:
            C = pq_train(Y, M, nbits, niter=niter_pq, seed=seed, centroids0=C)
        R = pq_decode(pq_assign(Y, C), C)
        # maximise tr(Rot^T X^T R): X^T R = U S V^T  ->  Rot = U V^T, A = Rot^T
        U, _, Vt = np.linalg.svd(Xt.T.astype(np.float64) @ R.astype(np.float64))
        A = np.ascontiguousarray((U @ Vt).T.astype(np.float32))
    return A


# ------------------------------------------------------------------ search
def topk_sorted(scores: np.ndarray, k: int, metric: int):
    """Exact top-k ordered by (value, id): ascending for L2, descending for IP."""
    nq, n = scores.shape
    k = min(k, n)
    key = scores if metric == L2 else -scores
    ids = np.empty((nq, k), dtype=np.uint32)
    vals = np.empty((nq, k), dtype=np.float32)
    for i in range(nq):
        if k < n:
            cand = np.argpartition(key[i], k - 1)[:k]
            kth = key[i][cand].max()
            cand = np.nonzero(key[i] <= kth)[0]  # keep every tie of the k-th value
        else:
            cand = np.arange(n)
        order = np.lexsort((cand, key[i][cand]))[:k]
        sel = cand[order]
        ids[i] = sel
        vals[i] = scores[i][sel]
    return vals, ids


def adc_lut(Q: np.ndarray, C: np.ndarray, metric: int) -> np.ndarray:
    """lut[q, m, k] = ||q_m - c_mk||^2 (L2) or q_m . c_mk (IP)."""
    Q = np.asarray(Q, dtype=np.float32)
    M, ksub, dsub = C.shape
    nq = Q.shape[0]
    lut = np.empty((nq, M, ksub), dtype=np.float32)
    for m in range(M):
        qm = Q[:, m * dsub:(m + 1) * dsub]
        if metric == L2:
            diff = qm[:, None, :] - C[m][None, :, :]
            lut[:, m, :] = np.einsum("qkd,qkd->qk", diff, diff, dtype=np.float32)
        else:
            lut[:, m, :] = qm @ C[m].T
    return lut


def adc_scores(codes: np.ndarray, lut: np.ndarray) -> np.ndarray:
    """score[q, i] = sum_m lut[q, m, codes[i, m]] (bench/ffd_speed.cpp:120-126)."""
    nq, M, ksub = lut.shape
    n = codes.shape[0]
    out = np.zeros((nq, n), dtype=np.float32)
    for m in range(M):
        out += lut[:, m, :][:, codes[:, m]]
    return out


def adc_search(codes: np.ndarray, C: np.ndarray, Q: np.ndarray, k: int, metric: int):
    return topk_sorted(adc_scores(codes, adc_lut(Q, C, metric)), k, metric)


def decompress_search(Xhat: np.ndarray, Q: np.ndarray, k: int, metric: int):
    """FlatQuantizedIndex.search_with_scores on the reconstructions
    (flat_quantized_index.py:57-76): float64 squared L2 cast to fp32, or fp32 IP."""
    Q = np.asarray(Q, dtype=np.float32)
    if metric == L2:
        q64, x64 = Q.astype(np.float64), Xhat.astype(np.float64)
        d = ((q64 * q64).sum(1)[:, None] + (x64 * x64).sum(1)[None, :] - 2.0 * q64 @ x64.T)
        d = np.maximum(d, 0).astype(np.float32)
    else:
        d = (Q @ Xhat.T).astype(np.float32)
    return topk_sorted(d, k, metric)


def flat_scores(X: np.ndarray, Q: np.ndarray, metric: int, row_scale: np.ndarray | None = None):
    """faiss IndexFlatL2 / IndexFlatIP distances in fp32."""
    X = np.asarray(X, dtype=np.float32)
    Q = np.asarray(Q, dtype=np.float32)
    if row_scale is not None:
        X = X * row_scale.astype(np.float32)[:, None]
    ip = Q @ X.T
    if metric == IP:
        return ip
    qn = np.einsum("qd,qd->q", Q, Q, dtype=np.float32)
    xn = np.einsum("nd,nd->n", X, X, dtype=np.float32)
    d = (qn[:, None] + xn[None, :]) - np.float32(2.0) * ip
    return np.maximum(d, np.float32(0))


def flat_search(X: np.ndarray, Q: np.ndarray, k: int, metric: int,
                row_scale: np.ndarray | None = None, chunk: int = 1000):
    nq = Q.shape[0]
    k = min(k, X.shape[0])
    vals = np.empty((nq, k), dtype=np.float32)
    ids = np.empty((nq, k), dtype=np.uint32)
    for s in range(0, nq, chunk):
        e = min(nq, s + chunk)
        v, i = topk_sorted(flat_scores(X, Q[s:e], metric, row_scale), k, metric)
        vals[s:e], ids[s:e] = v, i
    return vals, ids


def recall_at_k(ret: np.ndarray, gt: np.ndarray, k: int) -> float:
    """exact_search.recall_at_ks for one k (set semantics, exact_search.py:80-101)."""
    kk_r, kk_g = min(k, ret.shape[1]), min(k, gt.shape[1])
    denom = min(kk_r, kk_g)
    if denom == 0 or ret.shape[0] == 0:
        return 0.0
    tot = 0.0
    for i in range(ret.shape[0]):
        tot += len(set(gt[i, :kk_g].tolist()) & set(ret[i, :kk_r].tolist())) / denom
    return tot / ret.shape[0]
