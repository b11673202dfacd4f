"""Tensor-level wrappers over the C ABI: torch CUDA tensors in, torch CUDA tensors out.

torch is used for device memory, streams and (elsewhere) torch.distributed only;
all arithmetic below runs in the hand-written sm_100a kernels of libb200vq.so.
"""
from __future__ import annotations

import torch

from . import _lib
from ._lib import METRIC_IP, METRIC_L2, B200VQError, check

_scratch_cache: dict = {}


def require_cuda() -> None:
    if not torch.cuda.is_available():
        raise B200VQError("b200vq needs a CUDA device (sm_100a); there is no CPU fallback")


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _chk(t: torch.Tensor, dtype, name: str) -> torch.Tensor:
    if not t.is_cuda:
        raise B200VQError(f"{name}: expected a CUDA tensor")
    if t.dtype != dtype:
        raise B200VQError(f"{name}: expected dtype {dtype}, got {t.dtype}")
    return t


def _rows(t: torch.Tensor, name: str):
    """(n, D) tensor with unit column stride -> (ptr, n, ld, D)."""
    if t.dim() != 2 or (t.shape[1] > 1 and t.stride(1) != 1):
        raise B200VQError(f"{name}: expected a row-major 2-D tensor")
    ld = t.stride(0) if t.shape[0] > 1 else max(t.shape[1], t.stride(0))
    return t.data_ptr(), t.shape[0], ld, t.shape[1]


def _p(t):
    return None if t is None else t.data_ptr()


def scratch(nbytes: int, device, key: str = "default") -> torch.Tensor:
    """Grow-only per-device scratch buffers (caller-owned from the ABI's point of view)."""
    k = (key, torch.device(device).index)
    buf = _scratch_cache.get(k)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, device=device)
        _scratch_cache[k] = buf
    return buf


def metric_id(metric) -> int:
    if metric in ("l2", "L2", METRIC_L2):
        return METRIC_L2
    if metric in ("ip", "IP", METRIC_IP):
        return METRIC_IP
    raise ValueError(f"unknown metric {metric!r}")


# ------------------------------------------------------------------ PQ
def pq_centroid_norms(centroids: torch.Tensor) -> torch.Tensor:
    _chk(centroids, torch.float32, "centroids")
    M, ksub, dsub = centroids.shape
    out = torch.empty((M, ksub), dtype=torch.float32, device=centroids.device)
    check(_lib.load().b200vq_pq_centroid_norms(centroids.data_ptr(), M, ksub, dsub, out.data_ptr(), _stream()),
          "pq_centroid_norms")
    return out


def pq_assign(x: torch.Tensor, centroids: torch.Tensor, cnorm: torch.Tensor | None = None,
              out: torch.Tensor | None = None, want_dist: bool = False, impl: str = "auto",
              stats: torch.Tensor | None = None):
    """codes (n, M) uint8 = nearest centroid per sub-quantizer.

    impl: "simt" exact fp32 kernel, "tc" tcgen05 screen + exact re-rank, "auto" picks
    "tc" when the shape qualifies (ksub == 256, dsub % 8 == 0, 16-byte aligned rows).
    """
    _chk(x, torch.float32, "x")
    _chk(centroids, torch.float32, "centroids")
    M, ksub, dsub = centroids.shape
    ptr, n, ld, D = _rows(x, "x")
    if D != M * dsub:
        raise B200VQError(f"pq_assign: D={D} != M*dsub={M * dsub}")
    centroids = centroids.contiguous()
    if cnorm is None:
        cnorm = pq_centroid_norms(centroids)
    if out is None:
        out = torch.empty((n, M), dtype=torch.uint8, device=x.device)
    lib = _lib.load()
    tc_ok = (ksub == 256 and dsub % 8 == 0 and 8 <= dsub <= 128 and ld % 4 == 0 and ptr % 16 == 0
             and not want_dist and tc_available())
    if impl == "tc" and not tc_ok:
        raise B200VQError("pq_assign(impl='tc'): shape/alignment not supported by the tcgen05 path")
    if impl == "tc" or (impl == "auto" and tc_ok):
        nbytes = lib.b200vq_pq_assign_tc_scratch_bytes2(n, M, dsub)
        sc = scratch(nbytes, x.device, "assign_tc")
        check(lib.b200vq_pq_assign_tc(ptr, n, ld, centroids.data_ptr(), cnorm.data_ptr(), M, ksub, dsub,
                                      out.data_ptr(), sc.data_ptr(), sc.numel(), _p(stats), _stream()),
              "pq_assign_tc")
        return out
    dist = torch.empty((n, M), dtype=torch.float32, device=x.device) if want_dist else None
    check(lib.b200vq_pq_assign_f32(ptr, n, ld, centroids.data_ptr(), cnorm.data_ptr(), M, ksub, dsub,
                                   out.data_ptr(), _p(dist), _stream()), "pq_assign_f32")
    return (out, dist) if want_dist else out


_tc_state = {"checked": False, "ok": False}


def tc_available() -> bool:
    """True when the tcgen05 assign kernel is compiled in (probe with n = 0)."""
    if not _tc_state["checked"]:
        lib = _lib.load()
        rc = lib.b200vq_pq_assign_tc(None, 0, 0, None, None, 1, 256, 8, None, None, 0, None, None)
        _tc_state["ok"] = rc != -3  # B200VQ_ERR_UNSUPPORTED
        _tc_state["checked"] = True
    return _tc_state["ok"]


def pq_decode(codes: torch.Tensor, centroids: torch.Tensor, ids: torch.Tensor | None = None,
              out: torch.Tensor | None = None) -> torch.Tensor:
    _chk(codes, torch.uint8, "codes")
    _chk(centroids, torch.float32, "centroids")
    M, ksub, dsub = centroids.shape
    codes = codes.contiguous()
    if codes.dim() != 2 or codes.shape[1] != M:
        raise B200VQError(f"pq_decode: codes must be (n, {M})")
    n = codes.shape[0] if ids is None else ids.shape[0]
    if out is None:
        out = torch.empty((n, M * dsub), dtype=torch.float32, device=codes.device)
    lib = _lib.load()
    if ids is None:
        check(lib.b200vq_pq_decode(codes.data_ptr(), n, centroids.data_ptr(), M, ksub, dsub, out.data_ptr(),
                                   out.stride(0), _stream()), "pq_decode")
    else:
        _chk(ids, torch.int64, "ids")
        check(lib.b200vq_pq_decode_ids(codes.data_ptr(), ids.contiguous().data_ptr(), n, centroids.data_ptr(),
                                       M, ksub, dsub, out.data_ptr(), out.stride(0), _stream()),
              "pq_decode_ids")
    return out


def pq_kmeans_step(x: torch.Tensor, centroids: torch.Tensor, sums: torch.Tensor, counts: torch.Tensor,
                   codes: torch.Tensor, nsplit: torch.Tensor, impl: str = "simt") -> None:
    """One Lloyd iteration in place: assign, accumulate, finalize."""
    M, ksub, dsub = centroids.shape
    ptr, n, ld, D = _rows(x, "x")
    lib = _lib.load()
    pq_assign(x, centroids, out=codes, impl=impl)
    sums.zero_()
    counts.zero_()
    check(lib.b200vq_pq_kmeans_accumulate(ptr, n, ld, codes.data_ptr(), M, ksub, dsub, sums.data_ptr(),
                                          counts.data_ptr(), _stream()), "pq_kmeans_accumulate")
    check(lib.b200vq_pq_kmeans_finalize(sums.data_ptr(), counts.data_ptr(), M, ksub, dsub,
                                        centroids.data_ptr(), nsplit.data_ptr(), _stream()),
          "pq_kmeans_finalize")


def pq_sse(x: torch.Tensor, codes: torch.Tensor, centroids: torch.Tensor) -> torch.Tensor:
    M, ksub, dsub = centroids.shape
    ptr, n, ld, D = _rows(x, "x")
    out = torch.zeros(1, dtype=torch.float64, device=x.device)
    check(_lib.load().b200vq_pq_sse(ptr, n, ld, codes.data_ptr(), centroids.data_ptr(), M, ksub, dsub,
                                    out.data_ptr(), _stream()), "pq_sse")
    return out


# ------------------------------------------------------------------ ADC / search
def adc_lut(queries: torch.Tensor, centroids: torch.Tensor, metric) -> torch.Tensor:
    _chk(queries, torch.float32, "queries")
    M, ksub, dsub = centroids.shape
    ptr, nq, ld, D = _rows(queries, "queries")
    if D != M * dsub:
        raise B200VQError(f"adc_lut: D={D} != M*dsub={M * dsub}")
    lut = torch.empty((nq, M, ksub), dtype=torch.float32, device=queries.device)
    check(_lib.load().b200vq_adc_lut(ptr, nq, ld, centroids.data_ptr(), M, ksub, dsub, metric_id(metric),
                                     lut.data_ptr(), _stream()), "adc_lut")
    return lut


def _topk_out(nq, k, device):
    return (torch.empty((nq, k), dtype=torch.float32, device=device),
            torch.empty((nq, k), dtype=torch.int32, device=device))


def adc_scan_topk(codes: torch.Tensor, lut: torch.Tensor, k: int, metric, n: int | None = None,
                  rot: bool = False, screen: bool = False):
    """Top-k of sum_m lut[q, m, codes[i, m]]. codes: (n, M) uint8 row-major, or the
    scan layout (rot=True, n must then be given). screen=True (scan layout only) uses the
    four-queries-per-lookup screened scan: same results, faster for large n.
    Returns (vals fp32, ids int32 view of uint32)."""
    _chk(codes, torch.uint8, "codes")
    _chk(lut, torch.float32, "lut")
    nq, M, ksub = lut.shape
    if n is None:
        n = codes.shape[0]
    lib = _lib.load()
    vals, ids = _topk_out(nq, k, codes.device)
    if screen:
        if not rot:
            raise B200VQError("adc_scan_topk(screen=True) needs the scan layout (rot=True)")
        nbytes = lib.b200vq_adc_scan_screen_scratch_bytes(n, M, nq, k)
        sc = scratch(nbytes, codes.device, "scan_screen")
        check(lib.b200vq_adc_scan_topk_screen(codes.data_ptr(), n, M, ksub, lut.data_ptr(), nq, metric_id(metric),
                                              k, vals.data_ptr(), ids.data_ptr(), sc.data_ptr(), sc.numel(),
                                              _stream()), "adc_scan_topk_screen")
        return vals, ids
    nbytes = lib.b200vq_adc_scan_scratch_bytes(n, nq, k)
    sc = scratch(nbytes, codes.device, "scan")
    fn = lib.b200vq_adc_scan_topk_rot if rot else lib.b200vq_adc_scan_topk
    check(fn(codes.data_ptr(), n, M, ksub, lut.data_ptr(), nq, metric_id(metric), k, vals.data_ptr(),
             ids.data_ptr(), sc.data_ptr(), sc.numel(), _stream()),
          "adc_scan_topk_rot" if rot else "adc_scan_topk")
    return vals, ids


def pq_codes_to_scan_layout(codes: torch.Tensor) -> torch.Tensor:
    _chk(codes, torch.uint8, "codes")
    n, M = codes.shape
    out = torch.empty((((n + 31) // 32) * 32 * M,), dtype=torch.uint8, device=codes.device)
    check(_lib.load().b200vq_pq_codes_to_scan_layout(codes.contiguous().data_ptr(), n, M, out.data_ptr(),
                                                     _stream()), "pq_codes_to_scan_layout")
    return out


def pq_pack_bits(codes: torch.Tensor, nbits: int) -> torch.Tensor:
    """(n, M) uint8 indices -> faiss' packed layout for nbits < 8: (n, ceil(M nbits / 8)) little-endian bit stream."""
    _chk(codes, torch.uint8, "codes")
    codes = codes.contiguous()
    n, M = codes.shape
    out = torch.empty((n, (M * nbits + 7) // 8), dtype=torch.uint8, device=codes.device)
    check(_lib.load().b200vq_pq_pack_bits(codes.data_ptr(), n, M, nbits, out.data_ptr(), _stream()), "pq_pack_bits")
    return out


def pq_unpack_bits(packed: torch.Tensor, M: int, nbits: int) -> torch.Tensor:
    _chk(packed, torch.uint8, "packed")
    packed = packed.contiguous()
    n = packed.shape[0]
    if packed.shape[1] != (M * nbits + 7) // 8:
        raise B200VQError(f"pq_unpack_bits: rows have {packed.shape[1]} bytes, expected {(M * nbits + 7) // 8}")
    out = torch.empty((n, M), dtype=torch.uint8, device=packed.device)
    check(_lib.load().b200vq_pq_unpack_bits(packed.data_ptr(), n, M, nbits, out.data_ptr(), _stream()), "pq_unpack_bits")
    return out


class RotationTC:
    """W (N, K) prepared for y = x W^T on the tensor cores (three TF32 passes over hi/lo operand
    splits, fp32-accurate; csrc/rotate_tc.cu). The split of W is made once and reused."""

    def __init__(self, w: torch.Tensor):
        require_cuda()
        w = _chk(w, torch.float32, "w")
        wptr, self.N, ldw, self.K = _rows(w, "w")
        lib = _lib.load()
        self.wsplit = torch.empty(lib.b200vq_rotate_tc_scratch_bytes(self.N, self.K), dtype=torch.uint8,
                                  device=w.device)
        check(lib.b200vq_rotate_tc_prepare(wptr, ldw, self.N, self.K, self.wsplit.data_ptr(), self.wsplit.numel(),
                                           _stream()), "rotate_tc_prepare")

    def applies(self, x: torch.Tensor) -> bool:
        if x.dim() != 2 or x.shape[1] != self.K or x.dtype != torch.float32 or x.shape[0] == 0:
            return False
        _, n, ldx, _ = _rows(x, "x")
        return bool(_lib.load().b200vq_rotate_tc_applies(n, ldx, self.N, self.K, (self.N + 3) // 4 * 4)) and \
            x.data_ptr() % 16 == 0

    def __call__(self, x: torch.Tensor, out: torch.Tensor | None = None, *, x_sub: torch.Tensor | None = None,
                 row_scale: torch.Tensor | None = None, col_bias: torch.Tensor | None = None,
                 transpose_out: bool = False) -> torch.Tensor:
        """y = row_scale[:, None] * ((x - x_sub) W^T) + col_bias (each optional, fused into the GEMM).
        transpose_out: the result is returned as a (N, n) tensor (row stride >= n) holding y^T."""
        x = _chk(x, torch.float32, "x")
        xptr, n, ldx, K = _rows(x, "x")
        if K != self.K:
            raise B200VQError(f"rotate_tc: x has {K} columns, W has {self.K}")
        for t, m, name in ((x_sub, self.K, "x_sub"), (row_scale, n, "row_scale"), (col_bias, self.N, "col_bias")):
            if t is not None and (_chk(t, torch.float32, name).numel() != m or not t.is_contiguous()):
                raise B200VQError(f"rotate_tc: {name} must be a contiguous fp32 vector of {m} entries")
        if transpose_out:
            if out is None:
                out = torch.empty((self.N, (n + 3) // 4 * 4), dtype=torch.float32, device=x.device)[:, :n]
            if out.shape != (self.N, n) or (n > 1 and out.stride(1) != 1):
                raise B200VQError("rotate_tc: transposed output must be (N, n) with unit column stride")
            yptr, ldy = out.data_ptr(), (out.stride(0) if self.N > 1 else max(n, out.stride(0)))
        else:
            if out is None:
                ldy = (self.N + 3) // 4 * 4
                buf = torch.empty((n, ldy), dtype=torch.float32, device=x.device)
                out = buf[:, :self.N]
            yptr, _, ldy, _ = _rows(out, "out")
        check(_lib.load().b200vq_rotate_tc_ex(xptr, n, ldx, self.wsplit.data_ptr(), self.N, self.K, _p(x_sub),
                                              _p(row_scale), _p(col_bias), int(transpose_out), yptr, ldy, _stream()),
              "rotate_tc")
        return out


def row_sqnorms(x: torch.Tensor) -> torch.Tensor:
    _chk(x, torch.float32, "x")
    ptr, n, ld, D = _rows(x, "x")
    out = torch.empty((n,), dtype=torch.float32, device=x.device)
    check(_lib.load().b200vq_row_sqnorms(ptr, n, ld, D, out.data_ptr(), _stream()), "row_sqnorms")
    return out


def flat_topk(x: torch.Tensor, queries: torch.Tensor, k: int, metric, xnorm: torch.Tensor | None = None,
              row_scale: torch.Tensor | None = None, impl: str = "auto"):
    """Exact top-k of queries against raw rows. impl: "simt" = exact fp32 kernel; "tc" / "auto" = tensor-core
    screen + exact re-scoring (same results; the library falls back to the SIMT kernel for shapes the screen
    does not cover). xnorm = squared norms of the raw rows (computed here when absent)."""
    _chk(x, torch.float32, "x")
    _chk(queries, torch.float32, "queries")
    ptr, n, ld, D = _rows(x, "x")
    qptr, nq, ldq, Dq = _rows(queries, "queries")
    if D != Dq:
        raise B200VQError(f"flat_topk: dimension mismatch {D} vs {Dq}")
    mid = metric_id(metric)
    if mid == METRIC_L2 and xnorm is None:
        xnorm = row_sqnorms(x)
    lib = _lib.load()
    if impl not in ("auto", "tc", "simt"):
        raise B200VQError(f"flat_topk: unknown impl {impl!r}")
    if k > 128:
        # one launch ranks at most 128 rows per query: chain continuations, each starting after the last entry of
        # the previous one in the same (key, id) order -- the concatenation is the exact top-k
        vals, ids = _topk_out(nq, k, x.device)
        v0, i0 = flat_topk(x, queries, 128, metric, xnorm, row_scale, impl)
        vals[:, :128], ids[:, :128] = v0, i0
        nbytes = lib.b200vq_flat_topk_scratch_bytes(n, nq, 128)
        sc = scratch(nbytes, x.device, "scan")
        done = 128
        while done < k:
            kk = min(128, k - done)
            av, ai = vals[:, done - 1].contiguous(), ids[:, done - 1].contiguous()
            v1, i1 = _topk_out(nq, kk, x.device)
            check(lib.b200vq_flat_topk_after(ptr, n, ld, D, _p(xnorm), _p(row_scale), qptr, nq, ldq, mid, kk,
                                             av.data_ptr(), ai.data_ptr(), v1.data_ptr(), i1.data_ptr(), sc.data_ptr(),
                                             sc.numel(), _stream()), "flat_topk_after")
            vals[:, done:done + kk], ids[:, done:done + kk] = v1, i1
            done += kk
        return vals, ids
    vals, ids = _topk_out(nq, k, x.device)
    if impl != "simt" and tc_available() and k <= 128:
        if xnorm is None:
            xnorm = row_sqnorms(x)
        nbytes = lib.b200vq_flat_topk_tc_scratch_bytes(n, nq, k)
        sc = scratch(nbytes, x.device, "flat_tc")
        check(lib.b200vq_flat_topk_tc(ptr, n, ld, D, _p(xnorm), _p(row_scale), qptr, nq, ldq, mid, k,
                                      vals.data_ptr(), ids.data_ptr(), sc.data_ptr(), sc.numel(), _stream()),
              "flat_topk_tc")
        return vals, ids
    nbytes = lib.b200vq_flat_topk_scratch_bytes(n, nq, k)
    sc = scratch(nbytes, x.device, "scan")
    check(lib.b200vq_flat_topk(ptr, n, ld, D, _p(xnorm), _p(row_scale), qptr, nq, ldq, mid, k,
                               vals.data_ptr(), ids.data_ptr(), sc.data_ptr(), sc.numel(), _stream()),
          "flat_topk")
    return vals, ids


def topk_merge(vals: torch.Tensor, ids: torch.Tensor, k: int, metric):
    """vals (nq, P, k) fp32, ids (nq, P, k) int64 -> merged (nq, k)."""
    _chk(vals, torch.float32, "vals")
    _chk(ids, torch.int64, "ids")
    nq, P, kk = vals.shape
    out_v = torch.empty((nq, k), dtype=torch.float32, device=vals.device)
    out_i = torch.empty((nq, k), dtype=torch.int64, device=vals.device)
    if kk != k:
        raise B200VQError("topk_merge: partial lists must have k entries")
    check(_lib.load().b200vq_topk_merge(vals.contiguous().data_ptr(), ids.contiguous().data_ptr(), nq, P, k,
                                        metric_id(metric), out_v.data_ptr(), out_i.data_ptr(), _stream()),
          "topk_merge")
    return out_v, out_i
