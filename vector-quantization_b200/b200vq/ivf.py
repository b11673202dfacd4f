"""Device side of the IVF shell (haag_vq.methods.search.ivf_quantized_index): coarse k-means, cluster assignment and
the cluster-major exact search over resident reconstructions (csrc/ivf.cu).

Reference: src/haag_vq/methods/search/ivf_quantized_index.py:45-146. Its coarse quantiser is ``faiss.Kmeans(D, K,
seed=0, spherical=(metric == 'ip'))`` (20 Lloyd iterations, at most 256 training points per centroid, random initial
centroids, empty clusters re-seeded by splitting a big one). faiss is a third-party dependency that is not vendored
in the reference tree, so the centroids cannot be reproduced bit for bit: this module restates the published
algorithm (faiss/Clustering.cpp) and parity is anchored one step later, on everything computed FROM the centroids
(oracle/ivf.py takes the fitted centroids as an input).

Assignment scores |c|^2 - 2 x.c come from the fp32-accurate tensor-core GEMM (RotationTC with -2 C as the matrix and
|c|^2 as the column bias) followed by a row arg-min, in blocks of rows so that the score matrix stays below 1 GiB.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib, ops
from .ops import B200VQError, RotationTC, check, metric_id, scratch, _p, _rows, _stream

MAX_POINTS_PER_CENTROID = 256    # faiss ClusteringParameters defaults
NITER = 20
SCORE_BYTES = 1 << 30


def row_argmin(y: torch.Tensor) -> torch.Tensor:
    ptr, n, ld, K = _rows(y, "y")
    idx = torch.empty((n,), dtype=torch.int32, device=y.device)
    check(_lib.load().b200vq_row_argmin(ptr, n, ld, K, idx.data_ptr(), None, _stream()), "row_argmin")
    return idx


class CoarseAssigner:
    """Nearest centroid (L2) or largest inner product (spherical) of row batches against fixed centroids."""

    def __init__(self, centroids: torch.Tensor, spherical: bool = False):
        self.K, self.D = centroids.shape
        self.spherical = spherical
        Dp = (self.D + 31) // 32 * 32          # the GEMM wants K % 32 == 0: zero columns change nothing
        w = torch.zeros((self.K, Dp), dtype=torch.float32, device=centroids.device)
        w[:, :self.D] = centroids
        self.Dp = Dp
        if spherical:
            self.gemm = RotationTC(-w)         # arg-min of -x.c
            self.bias = None
        else:
            self.gemm = RotationTC(-2.0 * w)
            self.bias = ops.row_sqnorms(w).contiguous()
        self.block = max(1024, int(SCORE_BYTES // (4 * ((self.K + 3) // 4 * 4))))

    def __call__(self, x: torch.Tensor) -> torch.Tensor:
        n = x.shape[0]
        out = torch.empty((n,), dtype=torch.int32, device=x.device)
        if self.Dp != self.D:
            xp = torch.zeros((n, self.Dp), dtype=torch.float32, device=x.device)
            xp[:, :self.D] = x
            x = xp
        for s in range(0, n, self.block):
            e = min(n, s + self.block)
            y = self.gemm(x[s:e], col_bias=self.bias)
            out[s:e] = row_argmin(y)
        return out


def kmeans(x: torch.Tensor, K: int, niter: int = NITER, seed: int = 0, spherical: bool = False) -> torch.Tensor:
    """faiss.Kmeans-style Lloyd on the device. x (n, D) fp32 on the GPU; returns (K, D) fp32 centroids."""
    n, D = x.shape
    if n < K:
        raise B200VQError(f"kmeans: {n} training points for {K} centroids")
    g = torch.Generator(device="cpu")
    g.manual_seed(int(seed) + 1234)
    if n > K * MAX_POINTS_PER_CENTROID:      # subsample like faiss (random permutation, first K * 256)
        perm = torch.randperm(n, generator=g)[:K * MAX_POINTS_PER_CENTROID].to(x.device)
        x = x[perm].contiguous()
        n = x.shape[0]
    init = torch.randperm(n, generator=g)[:K].to(x.device)
    C = x[init].clone()
    if spherical:
        C = C / C.norm(dim=1, keepdim=True).clamp_min(1e-20)
    for _ in range(max(1, niter)):
        a = CoarseAssigner(C, spherical)(x).long()
        sums = torch.zeros((K, D), dtype=torch.float64, device=x.device)
        sums.index_add_(0, a, x.double())
        cnt = torch.bincount(a, minlength=K)
        C = (sums / cnt.clamp_min(1).unsqueeze(1).double()).float()
        cnt_h = cnt.cpu().numpy().astype(np.int64)
        empty = np.flatnonzero(cnt_h == 0)
        if empty.size:
            # faiss split_clusters: an empty cluster takes a copy of a big cluster's centroid, both are nudged apart
            # by +-1/1024 on alternating coordinates and the points are shared out
            eps = 1.0 / 1024.0
            sign = torch.ones(D, device=x.device)
            sign[1::2] = -1.0
            for ci in empty:
                cj = int(np.argmax(cnt_h))
                C[ci] = C[cj] * (1.0 + eps * sign)
                C[cj] = C[cj] * (1.0 - eps * sign)
                cnt_h[ci] = cnt_h[cj] // 2
                cnt_h[cj] -= cnt_h[ci]
        if spherical:
            C = C / C.norm(dim=1, keepdim=True).clamp_min(1e-20)
    return C.contiguous()


def ivf_search(recon: torch.Tensor, rnorm: torch.Tensor | None, gid: torch.Tensor, row_off: torch.Tensor, K: int,
               queries: torch.Tensor, probe: torch.Tensor, k: int, metric, rowcluster: torch.Tensor | None = None,
               impl: str = "auto"):
    """Exact top-k of every query over the rows of the clusters it probes. recon (N, D) fp32 sorted by cluster,
    gid (N) int32 global ids, row_off (K + 1) int64, probe (nq, nprobe) int32, rowcluster (N) int32 (tensor-core path).
    impl: "simt" = cluster-major exact kernel (b200vq_ivf_search); "tc" / "auto" = tensor-core screen with the probe
    lists as a mask (b200vq_ivf_search_tc, same results), queries it flags are redone by the SIMT kernel."""
    rptr, N, ldr, D = _rows(recon, "recon")
    qptr, nq, ldq, Dq = _rows(queries, "queries")
    if D != Dq:
        raise B200VQError(f"ivf_search: dimension mismatch {D} vs {Dq}")
    if row_off.dtype != torch.int64 or row_off.numel() != K + 1:
        raise B200VQError("ivf_search: row_off must be int64 with K + 1 entries")
    if impl not in ("auto", "tc", "simt"):
        raise B200VQError(f"ivf_search: unknown impl {impl!r}")
    probe = probe.contiguous()
    nprobe = probe.shape[1]
    mid = metric_id(metric)
    lib = _lib.load()
    if k > 128:
        # pages of 128: each continues after the last (distance, id) of the previous one (b200vq_ivf_search_after)
        vals = torch.empty((nq, k), dtype=torch.float32, device=recon.device)
        ids = torch.empty((nq, k), dtype=torch.int32, device=recon.device)
        vals[:, :128], ids[:, :128] = ivf_search(recon, rnorm, gid, row_off, K, queries, probe, 128, metric, rowcluster,
                                                 impl)
        if mid == ops.METRIC_L2 and rnorm is None:
            rnorm = ops.row_sqnorms(recon)
        sc = scratch(lib.b200vq_ivf_search_scratch_bytes(nq, nprobe, K, 128), recon.device, "ivf")
        done = 128
        while done < k:
            kk = min(128, k - done)
            av, ai = vals[:, done - 1].contiguous(), ids[:, done - 1].contiguous()
            v1 = torch.empty((nq, kk), dtype=torch.float32, device=recon.device)
            i1 = torch.empty((nq, kk), dtype=torch.int32, device=recon.device)
            check(lib.b200vq_ivf_search_after(rptr, ldr, D, _p(rnorm), gid.data_ptr(), row_off.data_ptr(), K, qptr, nq,
                                              ldq, probe.data_ptr(), nprobe, mid, kk, av.data_ptr(), ai.data_ptr(),
                                              v1.data_ptr(), i1.data_ptr(), sc.data_ptr(), sc.numel(), _stream()),
                  "ivf_search_after")
            vals[:, done:done + kk], ids[:, done:done + kk] = v1, i1
            done += kk
        return vals, ids
    vals = torch.empty((nq, k), dtype=torch.float32, device=recon.device)
    ids = torch.empty((nq, k), dtype=torch.int32, device=recon.device)
    tc_ok = (rowcluster is not None and N >= 1024 and D % 32 == 0 and ldr % 4 == 0 and ldq % 4 == 0 and
             rptr % 16 == 0 and qptr % 16 == 0 and ops.tc_available() and nq > 0)
    if impl == "tc" and not tc_ok:
        raise B200VQError("ivf_search(impl='tc'): needs N >= 1024, D % 32 == 0, aligned rows and the row -> cluster map")
    if impl != "simt" and tc_ok:
        if rnorm is None:
            rnorm = ops.row_sqnorms(recon)
        flags = torch.empty((nq,), dtype=torch.int32, device=recon.device)
        sc = scratch(lib.b200vq_ivf_search_tc_scratch_bytes(N, nq, K, k), recon.device, "ivf_tc")
        check(lib.b200vq_ivf_search_tc(rptr, N, ldr, D, rnorm.data_ptr(), gid.data_ptr(), rowcluster.data_ptr(), K, qptr,
                                       nq, ldq, probe.data_ptr(), nprobe, mid, k, vals.data_ptr(), ids.data_ptr(),
                                       flags.data_ptr(), None, sc.data_ptr(), sc.numel(), _stream()),
              "ivf_search_tc")
        redo = torch.nonzero(flags, as_tuple=False).flatten()      # host sync; the results go to the host anyway
        if redo.numel():
            v2, i2 = ivf_search(recon, rnorm, gid, row_off, K, queries[redo].contiguous(), probe[redo].contiguous(), k,
                                metric, impl="simt")
            vals[redo], ids[redo] = v2, i2
        return vals, ids
    sc = scratch(lib.b200vq_ivf_search_scratch_bytes(nq, nprobe, K, k), recon.device, "ivf")
    check(lib.b200vq_ivf_search(rptr, ldr, D, _p(rnorm), gid.data_ptr(), row_off.data_ptr(), K, qptr, nq, ldq,
                                probe.data_ptr(), nprobe, mid, k, vals.data_ptr(), ids.data_ptr(), sc.data_ptr(),
                                sc.numel(), _stream()), "ivf_search")
    return vals, ids
