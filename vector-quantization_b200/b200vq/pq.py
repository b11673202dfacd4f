"""Device-resident product quantizer / OPQ (train, encode, decode, ADC search).

Host orchestration only; the arithmetic is in libb200vq.so. Mirrors the faiss
objects the reference drives (ProductQuantizer, OPQMatrix): reference call
sites src/haag_vq/methods/product_quantization.py:58-86 and
src/haag_vq/methods/optimized_product_quantization.py:21-34. All randomness
(training subset, seeds, OPQ start rotation) follows the same seeded numpy
recipe as the CPU oracle so both share trained state.
"""
from __future__ import annotations

import numpy as np
import torch

from . import ops

MAX_POINTS_PER_CENTROID = 256
KMEANS_NITER = 25


def subsample_ids(n: int, nmax: int, seed: int) -> np.ndarray:
    if n <= nmax:
        return np.arange(n, dtype=np.int64)
    return np.random.default_rng(seed).permutation(n)[:nmax].astype(np.int64)


def _take_rows(x: torch.Tensor, ids: np.ndarray) -> torch.Tensor:
    if ids.size == x.shape[0]:
        return x
    return x.index_select(0, torch.from_numpy(ids).to(x.device))


def pq_init_centroids(xt: torch.Tensor, M: int, ksub: int, seed: int) -> torch.Tensor:
    n, D = xt.shape
    dsub = D // M
    if n < ksub:
        raise RuntimeError(
            f"Number of training points ({n}) should be at least as large as number of clusters ({ksub})")
    C = torch.empty((M, ksub, dsub), dtype=torch.float32, device=xt.device)
    for m in range(M):
        ids = np.random.default_rng(seed + m).permutation(n)[:ksub]
        C[m] = xt.index_select(0, torch.from_numpy(ids).to(xt.device))[:, m * dsub:(m + 1) * dsub]
    return C


def pq_train(x: torch.Tensor, M: int, nbits: int = 8, niter: int = KMEANS_NITER, seed: int = 1234,
             centroids0: torch.Tensor | None = None, impl: str = "auto") -> torch.Tensor:
    """M independent Lloyd k-means on the device. Returns centroids (M, ksub, dsub)."""
    n, D = x.shape
    if D % M != 0:
        raise AssertionError("D must be divisible by M (number of subquantizers)")
    if nbits < 1 or nbits > 8:
        raise ValueError("b200vq supports 1 <= B <= 8 bits per sub-quantizer index")
    ksub = 1 << nbits
    xt = _take_rows(x, subsample_ids(n, MAX_POINTS_PER_CENTROID * ksub, seed))
    if centroids0 is None:
        C = pq_init_centroids(xt, M, ksub, seed)
        if xt.shape[0] == ksub:
            return C
    else:
        C = centroids0.to(torch.float32).clone().contiguous()
    dsub = D // M
    nt = xt.shape[0]
    sums = torch.empty((M, ksub, dsub), dtype=torch.float32, device=x.device)
    counts = torch.empty((M, ksub), dtype=torch.int32, device=x.device)
    codes = torch.empty((nt, M), dtype=torch.uint8, device=x.device)
    nsplit = torch.zeros(1, dtype=torch.int32, device=x.device)
    for _ in range(niter):
        ops.pq_kmeans_step(xt, C, sums, counts, codes, nsplit, impl=impl)
    return C


def random_rotation(D: int, seed: int) -> np.ndarray:
    A = np.random.default_rng(seed).standard_normal((D, D))
    Qm, _ = np.linalg.qr(A)
    return np.ascontiguousarray(Qm.astype(np.float32))


TC_ROTATE_MIN_ROWS = 1024


def _rotation_tc(A: torch.Tensor, transpose: bool) -> ops.RotationTC:
    """RotationTC for y = x A^T (transpose) or y = x A. The prepared operand split is owned by the
    tensor OBJECT it was made from (an attribute on A, tagged with A's in-place version counter), so it
    lives and dies with that matrix: a new matrix never finds another one's split, whatever address
    the allocator hands it."""
    slot = A.__dict__.setdefault("_b200vq_rotation_tc", {})
    ent = slot.get(transpose)
    if ent is None or ent[0] != A._version:
        ent = slot[transpose] = (A._version, ops.RotationTC(A if transpose else A.t().contiguous()))
    return ent[1]


def rotate(x: torch.Tensor, A: torch.Tensor, transpose: bool = True, impl: str = "auto") -> torch.Tensor:
    """y = x A^T (OPQMatrix.apply) or y = x A (reverse_transform), fp32-accurate.
    impl "tc": the three-pass TF32 tensor-core GEMM (csrc/rotate_tc.cu); "lib": the library sgemm
    in strict fp32 (TF32 off); "auto": tc for batches of >= 1024 rows it covers, else lib."""
    if impl not in ("auto", "tc", "lib"):
        raise ValueError(f"rotate: unknown impl {impl!r}")
    if impl != "lib" and x.is_cuda and (impl == "tc" or x.shape[0] >= TC_ROTATE_MIN_ROWS):
        r = _rotation_tc(A, transpose)
        if r.applies(x):
            return r(x)
        if impl == "tc":
            raise ops.B200VQError("rotate: shape not covered by the tensor-core rotation")
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        return torch.matmul(x, A.t() if transpose else A)
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev


def opq_train(x: torch.Tensor, M: int, nbits: int = 8, niter: int = 50, niter_pq: int = 4,
              niter_pq_0: int = 40, seed: int = 123, A0: np.ndarray | None = None,
              impl: str = "auto") -> torch.Tensor:
    """OPQMatrix.train: alternate PQ training on rotated data with the Procrustes update.
    impl selects the rotation GEMM (see rotate); every iterate A is a new tensor with its own split."""
    n, D = x.shape
    xt = _take_rows(x, subsample_ids(n, MAX_POINTS_PER_CENTROID * 256, seed))
    A = torch.from_numpy(random_rotation(D, seed) if A0 is None else np.asarray(A0, dtype=np.float32)).to(x.device)
    C = None
    xt64 = None
    for it in range(niter):
        y = rotate(xt, A, impl=impl)
        if it == 0:
            C = pq_train(y, M, nbits, niter=niter_pq_0, seed=seed)
        else:
            C = pq_train(y, M, nbits, niter=niter_pq, seed=seed, centroids0=C)
        r = ops.pq_decode(ops.pq_assign(y, C), C)
        if xt64 is None:
            xt64 = xt.to(torch.float64)
        U, _, Vt = torch.linalg.svd(xt64.t() @ r.to(torch.float64))
        A = (U @ Vt).t().to(torch.float32).contiguous()
    return A


def scan_layout_width(M: int) -> int:
    """Sub-quantizers per row in the scan layout: 16 (half group), else the next multiple of 32."""
    return 16 if M == 16 else (M + 31) // 32 * 32


class PQCodec:
    """Trained PQ state on one device + encode/decode/search helpers."""

    def __init__(self, centroids: torch.Tensor, rotation: torch.Tensor | None = None):
        self.centroids = centroids.contiguous()
        self.M, self.ksub, self.dsub = centroids.shape
        self.D = self.M * self.dsub
        self.cnorm = ops.pq_centroid_norms(self.centroids)
        self.rotation = rotation  # (D, D) A of OPQ, or None

    def encode(self, x: torch.Tensor, out: torch.Tensor | None = None, impl: str = "auto") -> torch.Tensor:
        if self.rotation is not None:
            x = rotate(x, self.rotation)
        return ops.pq_assign(x, self.centroids, self.cnorm, out=out, impl=impl)

    def decode(self, codes: torch.Tensor, ids: torch.Tensor | None = None) -> torch.Tensor:
        r = ops.pq_decode(codes, self.centroids, ids=ids)
        if self.rotation is not None:
            r = rotate(r, self.rotation, transpose=False)
        return r

    def luts(self, queries: torch.Tensor, metric) -> torch.Tensor:
        if self.rotation is not None:
            queries = rotate(queries, self.rotation)  # orthonormal: distances are preserved
        return ops.adc_lut(queries, self.centroids, metric)

    def pad_codes_for_scan(self, codes: torch.Tensor) -> torch.Tensor:
        """(n, M) codes -> (n, M_pad) with M_pad the next multiple of 32 (the scan layout's group width; M = 16 has
        its own half-group layout and is not padded); the extra sub-quantizers hold code 0 and their tables are all
        zero (search(m_pad=...)), so every score keeps its bits: x + 0.0 is exact and the real sub-quantizers are
        still summed first, in order."""
        n, M = codes.shape
        mp = scan_layout_width(M)
        if mp == M:
            return codes
        out = torch.zeros((n, mp), dtype=torch.uint8, device=codes.device)
        out[:, :M] = codes
        return out

    def search(self, codes: torch.Tensor, queries: torch.Tensor, k: int, metric, n: int | None = None,
               rot: bool = False, q_batch: int = 4096, screen: bool | None = None, m_pad: int | None = None):
        """ADC top-k over (n, M) codes. Returns (vals (nq,k) fp32, ids (nq,k) int32-as-uint32).
        screen (scan layout only; default on): four-queries-per-lookup screened scan, same results.
        m_pad: the codes were padded to m_pad sub-quantizers (pad_codes_for_scan); the tables are padded with zeros."""
        if screen is None:
            screen = rot
        if rot and m_pad is None:
            m_pad = scan_layout_width(self.M)
        nq = queries.shape[0]
        vals = torch.empty((nq, k), dtype=torch.float32, device=queries.device)
        ids = torch.empty((nq, k), dtype=torch.int32, device=queries.device)
        for s in range(0, nq, q_batch):
            e = min(nq, s + q_batch)
            lut = self.luts(queries[s:e], metric)
            if m_pad is not None and m_pad != self.M:
                lp = torch.zeros((e - s, m_pad, self.ksub), dtype=torch.float32, device=lut.device)
                lp[:, :self.M] = lut
                lut = lp
            v, i = ops.adc_scan_topk(codes, lut, k, metric, n=n, rot=rot, screen=bool(screen and rot))
            vals[s:e], ids[s:e] = v, i
        return vals, ids
