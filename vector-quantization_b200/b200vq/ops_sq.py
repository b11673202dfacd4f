"""Tensor-level wrappers: scalar quantizers, RaBitQ family, utilities."""
from __future__ import annotations

import torch

from . import _lib
from ._lib import B200VQError, check
from .ops import _chk, _p, _rows, _stream, _topk_out, metric_id, scratch


# ------------------------------------------------------------------ SQ
def sq_minmax(x: torch.Tensor, vmin: torch.Tensor | None = None, vmax: torch.Tensor | None = None):
    """Column-wise min/max; pass the previous (vmin, vmax) to accumulate over batches."""
    _chk(x, torch.float32, "x")
    ptr, n, ld, D = _rows(x, "x")
    if vmin is None:
        vmin = torch.full((D,), float("inf"), dtype=torch.float32, device=x.device)
        vmax = torch.full((D,), float("-inf"), dtype=torch.float32, device=x.device)
    check(_lib.load().b200vq_sq_minmax(ptr, n, ld, D, vmin.data_ptr(), vmax.data_ptr(), _stream()),
          "sq_minmax")
    return vmin, vmax


def sq_code_shape(n: int, D: int, bits: int):
    if bits == 4:
        return (n, (D + 1) // 2), torch.uint8
    if bits == 8:
        return (n, D), torch.uint8
    if bits == 16:
        return (n, D), torch.int16  # bit pattern of uint16
    raise ValueError(f"num_bits must be 4, 8, or 16, got {bits}")


def sq_encode(x: torch.Tensor, vmin: torch.Tensor, vmax: torch.Tensor, bits: int,
              out: torch.Tensor | None = None) -> torch.Tensor:
    _chk(x, torch.float32, "x")
    ptr, n, ld, D = _rows(x, "x")
    shape, dt = sq_code_shape(n, D, bits)
    if out is None:
        out = torch.empty(shape, dtype=dt, device=x.device)
    check(_lib.load().b200vq_sq_encode(ptr, n, ld, D, vmin.data_ptr(), vmax.data_ptr(), bits, out.data_ptr(),
                                       _stream()), "sq_encode")
    return out


def sq_decode(codes: torch.Tensor, vmin: torch.Tensor, vmax: torch.Tensor, bits: int, D: int,
              out: torch.Tensor | None = None) -> torch.Tensor:
    codes = codes.contiguous()
    n = codes.shape[0]
    if out is None:
        out = torch.empty((n, D), dtype=torch.float32, device=codes.device)
    check(_lib.load().b200vq_sq_decode(codes.data_ptr(), n, D, vmin.data_ptr(), vmax.data_ptr(), bits,
                                       out.data_ptr(), out.stride(0), _stream()), "sq_decode")
    return out


def saq_encode(x: torch.Tensor, vmin, vmax, dim_bits: torch.Tensor, bit_offset: torch.Tensor,
               code_bytes: int) -> torch.Tensor:
    _chk(x, torch.float32, "x")
    _chk(dim_bits, torch.int32, "dim_bits")
    _chk(bit_offset, torch.int32, "bit_offset")
    ptr, n, ld, D = _rows(x, "x")
    out = torch.empty((n, code_bytes), dtype=torch.uint8, device=x.device)
    check(_lib.load().b200vq_saq_encode(ptr, n, ld, D, vmin.data_ptr(), vmax.data_ptr(), dim_bits.data_ptr(),
                                        bit_offset.data_ptr(), code_bytes, out.data_ptr(), _stream()),
          "saq_encode")
    return out


def saq_decode(codes: torch.Tensor, vmin_fill, vmax, dim_bits, bit_offset, D: int) -> torch.Tensor:
    _chk(codes, torch.uint8, "codes")
    codes = codes.contiguous()
    n, code_bytes = codes.shape
    out = torch.empty((n, D), dtype=torch.float32, device=codes.device)
    check(_lib.load().b200vq_saq_decode(codes.data_ptr(), n, D, vmin_fill.data_ptr(), vmax.data_ptr(),
                                        dim_bits.data_ptr(), bit_offset.data_ptr(), code_bytes,
                                        out.data_ptr(), out.stride(0), _stream()), "saq_decode")
    return out


# ------------------------------------------------------------------ extended RaBitQ
def erabitq_pack(srot: torch.Tensor, nrm: torch.Tensor, B: int, levels: torch.Tensor,
                 bounds: torch.Tensor) -> torch.Tensor:
    _chk(srot, torch.float64, "srot")
    _chk(nrm, torch.float64, "nrm")
    srot = srot.contiguous()
    n, D = srot.shape
    out = torch.empty((n, (D * B + 7) // 8 + 8), dtype=torch.uint8, device=srot.device)
    check(_lib.load().b200vq_erabitq_pack(srot.data_ptr(), nrm.data_ptr(), n, D, B, levels.data_ptr(),
                                          bounds.data_ptr(), out.data_ptr(), _stream()), "erabitq_pack")
    return out


def erabitq_unpack(codes: torch.Tensor, D: int, B: int, levels: torch.Tensor):
    _chk(codes, torch.uint8, "codes")
    codes = codes.contiguous()
    n = codes.shape[0]
    shat = torch.empty((n, D), dtype=torch.float64, device=codes.device)
    nrm = torch.empty((n,), dtype=torch.float64, device=codes.device)
    check(_lib.load().b200vq_erabitq_unpack(codes.data_ptr(), n, D, B, levels.data_ptr(), shat.data_ptr(),
                                            nrm.data_ptr(), _stream()), "erabitq_unpack")
    return shat, nrm


def erabitq_decode_f32(codes: torch.Tensor, D: int, B: int, levels_f32: torch.Tensor):
    """x (n, D) fp32 = nrm t / sqrt(D) * levels[idx] and its squared row norms (b200vq_erabitq_decode_f32)."""
    _chk(codes, torch.uint8, "codes")
    codes = codes.contiguous()
    n = codes.shape[0]
    x = torch.empty((n, D), dtype=torch.float32, device=codes.device)
    xn = torch.empty((n,), dtype=torch.float32, device=codes.device)
    check(_lib.load().b200vq_erabitq_decode_f32(codes.data_ptr(), n, D, B, levels_f32.data_ptr(), x.data_ptr(),
                                                x.stride(0), xn.data_ptr(), _stream()), "erabitq_decode_f32")
    return x, xn


def resid_norms_f64(x: torch.Tensor, c: torch.Tensor) -> torch.Tensor:
    """|x_i - c| in fp64 (x fp32 or fp64 rows, c fp64)."""
    _chk(c, torch.float64, "c")
    if x.dtype not in (torch.float32, torch.float64) or not x.is_cuda:
        raise B200VQError("resid_norms_f64: x must be a CUDA fp32 / fp64 matrix")
    ptr, n, ld, D = _rows(x, "x")
    out = torch.empty((n,), dtype=torch.float64, device=x.device)
    check(_lib.load().b200vq_resid_norms_f64(ptr, int(x.dtype == torch.float64), n, ld, D, c.data_ptr(), out.data_ptr(),
                                             _stream()), "resid_norms_f64")
    return out


def erabitq_rotation_error_bound(D: int, c_norm: float) -> tuple:
    """(err_rel, err_abs): |y - (x - c) P| <= err_rel |x - c| + err_abs for y = rotate_tc_ex(x, P32^T, x_sub = c32).
    Operand splits 3 * 2^-22, twelve TMEM accumulation steps per chunk (measured 2^-25.6 each, taken as 2^-25),
    one round-to-nearest register add per 32-column chunk, fp32 rounding of x - c32, of P and of c; Cauchy-Schwarz
    over the row (|P[:, d]| = 1), times 1.25."""
    u = 2.0 ** -24
    rel = 1.25 * (3 * 2.0 ** -22 + 12 * 2.0 ** -25 + (D // 32 + 2) * u)
    return rel, 1.25 * u * float(c_norm)


def erabitq_pack_tc(y: torch.Tensor | None, x: torch.Tensor, c: torch.Tensor, PT: torch.Tensor, nrm: torch.Tensor,
                    B: int, levels: torch.Tensor, bounds: torch.Tensor, err_rel: float, err_abs: float,
                    out: torch.Tensor | None = None, stats: torch.Tensor | None = None,
                    deferred: bool = True) -> torch.Tensor:
    """Extended-RaBitQ codes from the fp32 rotated residuals y (or, y None, all-exact fp64) with exact fp64
    re-evaluation of the coordinates inside the rotation's error bound of a level mid-point."""
    for t, name in ((c, "c"), (PT, "PT"), (nrm, "nrm"), (levels, "levels"), (bounds, "bounds")):
        _chk(t, torch.float64, name)
    ptr, n, ldx, D = _rows(x, "x")
    if PT.shape != (D, D) or not PT.is_contiguous():
        raise B200VQError("erabitq_pack_tc: PT must be a contiguous (D, D) fp64 matrix")
    yptr, ldy = None, 0
    if y is not None:
        _chk(y, torch.float32, "y")
        yptr, ny, ldy, Dy = _rows(y, "y")
        if (ny, Dy) != (n, D):
            raise B200VQError("erabitq_pack_tc: y and x disagree")
    if out is None:
        out = torch.empty((n, (D * B + 7) // 8 + 8), dtype=torch.uint8, device=x.device)
    lib = _lib.load()
    if yptr is not None and n > 0 and deferred:
        # fix-ups visited coordinate by coordinate in a second kernel (same codes; see b200vq_erabitq_pack_tc2)
        sc = scratch(lib.b200vq_erabitq_pack_tc2_scratch_bytes(n, D), x.device, "erq_pack")
        check(lib.b200vq_erabitq_pack_tc2(yptr, ldy, ptr, int(x.dtype == torch.float64), ldx, c.data_ptr(), PT.data_ptr(),
                                          nrm.data_ptr(), n, D, B, levels.data_ptr(), bounds.data_ptr(), err_rel, err_abs,
                                          out.data_ptr(), _p(stats), sc.data_ptr(), sc.numel(), _stream()),
              "erabitq_pack_tc2")
        return out
    check(lib.b200vq_erabitq_pack_tc(yptr, ldy, ptr, int(x.dtype == torch.float64), ldx, c.data_ptr(),
                                     PT.data_ptr(), nrm.data_ptr(), n, D, B, levels.data_ptr(), bounds.data_ptr(),
                                     err_rel, err_abs, out.data_ptr(), _p(stats), _stream()), "erabitq_pack_tc")
    return out


def erabitq_rowscale(codes: torch.Tensor, D: int, B: int) -> torch.Tensor:
    _chk(codes, torch.uint8, "codes")
    codes = codes.contiguous()
    a = torch.empty((codes.shape[0],), dtype=torch.float32, device=codes.device)
    check(_lib.load().b200vq_erabitq_rowscale(codes.data_ptr(), codes.shape[0], D, B, a.data_ptr(), _stream()),
          "erabitq_rowscale")
    return a


class ERaBitQScanner:
    """Extended-RaBitQ codes prepared for the code-domain estimator scan on the tensor cores (csrc/est_tc.cu):
    row scales a_i and |x_i|^2 are extracted once, every search then reads only the code bytes."""

    def __init__(self, codes: torch.Tensor, D: int, B: int, levels_f32: torch.Tensor):
        from .ops_est import CodeSegments
        _chk(codes, torch.uint8, "codes")
        if D % 32 or codes.stride(0) % 4 or codes.data_ptr() % 4:
            raise B200VQError("ERaBitQScanner: needs D % 32 == 0 and 4-byte aligned code rows")
        self.codes, self.D, self.B = codes, D, B
        self.a = erabitq_rowscale(codes, D, B)
        self.segs = CodeSegments([D], [B], [codes], [self.a], [levels_f32.contiguous()], [0.0])
        self.norm2 = self.segs.row_norms()

    def search(self, qrot: torch.Tensor, qconst: torch.Tensor, k: int, metric):
        return self.segs.scan_topk(self.norm2, self.norm2, qrot, qconst, k, metric)


def erabitq_scan_topk(codes: torch.Tensor, D: int, B: int, levels_f32: torch.Tensor, qrot: torch.Tensor,
                      qconst: torch.Tensor, k: int, metric, impl: str = "auto"):
    """Extended-RaBitQ estimator top-k: qc + a^2 sum l^2 - 2 a <l, q> (L2) / qc + a <l, q> (IP), a = nrm t / sqrt(D).
    impl "simt": one level look-up per (dimension, code, query); "tc" / "auto": the code-domain scan on the tensor
    cores (TF32 screen straight from the code bytes + exact scoring of the survivors). Same arithmetic of the returned
    values either way (one fp32 FMA chain in dimension order), hence the same ids and value bits."""
    from .ops import tc_available
    _chk(codes, torch.uint8, "codes")
    codes = codes.contiguous()
    n = codes.shape[0]
    nq = qrot.shape[0]
    if impl not in ("auto", "tc", "simt"):
        raise B200VQError(f"erabitq_scan_topk: unknown impl {impl!r}")
    if impl != "simt" and tc_available() and D % 32 == 0 and k <= 128 and (impl == "tc" or n >= 16384):
        return ERaBitQScanner(codes, D, B, levels_f32).search(qrot.contiguous(), qconst, k, metric)
    lib = _lib.load()
    vals, ids = _topk_out(nq, k, codes.device)
    sc = scratch(lib.b200vq_erabitq_scan_scratch_bytes(n, nq, k), codes.device, "scan")
    check(lib.b200vq_erabitq_scan_topk(codes.data_ptr(), n, D, B, levels_f32.data_ptr(),
                                       qrot.contiguous().data_ptr(), qconst.data_ptr(), nq, metric_id(metric),
                                       k, vals.data_ptr(), ids.data_ptr(), sc.data_ptr(), sc.numel(),
                                       _stream()), "erabitq_scan_topk")
    return vals, ids


# ------------------------------------------------------------------ 1-bit RaBitQ
def lvq_code_size(D: int, bits: int) -> int:
    return (D * bits + 7) // 8 + 8


def lvq_encode(x: torch.Tensor, mu: torch.Tensor, bits: int, out: torch.Tensor | None = None) -> torch.Tensor:
    _chk(x, torch.float32, "x")
    _chk(mu, torch.float32, "mu")
    ptr, n, ld, D = _rows(x, "x")
    if mu.numel() != D:
        raise B200VQError(f"lvq_encode: mu has {mu.numel()} entries, rows have {D}")
    if out is None:
        out = torch.empty((n, lvq_code_size(D, bits)), dtype=torch.uint8, device=x.device)
    check(_lib.load().b200vq_lvq_encode(ptr, n, ld, D, mu.data_ptr(), bits, out.data_ptr(), _stream()), "lvq_encode")
    return out


def lvq_decode(codes: torch.Tensor, mu: torch.Tensor, bits: int) -> torch.Tensor:
    _chk(codes, torch.uint8, "codes")
    codes = codes.contiguous()
    D, n = mu.numel(), codes.shape[0]
    if codes.shape[1] != lvq_code_size(D, bits):
        raise B200VQError(f"lvq_decode: code rows have {codes.shape[1]} bytes, expected {lvq_code_size(D, bits)}")
    out = torch.empty((n, D), dtype=torch.float32, device=codes.device)
    check(_lib.load().b200vq_lvq_decode(codes.data_ptr(), n, D, mu.data_ptr(), bits, out.data_ptr(), out.stride(0),
                                        _stream()), "lvq_decode")
    return out


def col_mean_seq(x: torch.Tensor) -> torch.Tensor:
    """Column means in numpy's X.mean(axis=0) order (sequential fp32 row adds, then / n)."""
    _chk(x, torch.float32, "x")
    ptr, n, ld, D = _rows(x, "x")
    out = torch.empty((D,), dtype=torch.float32, device=x.device)
    check(_lib.load().b200vq_col_mean_seq(ptr, n, ld, D, out.data_ptr(), _stream()), "col_mean_seq")
    return out


def rabitq1_encode(x: torch.Tensor, centroid: torch.Tensor | None, metric) -> torch.Tensor:
    _chk(x, torch.float32, "x")
    ptr, n, ld, D = _rows(x, "x")
    out = torch.empty((n, (D + 7) // 8 + 8), dtype=torch.uint8, device=x.device)
    check(_lib.load().b200vq_rabitq1_encode(ptr, n, ld, D, _p(centroid), metric_id(metric), out.data_ptr(),
                                            _stream()), "rabitq1_encode")
    return out


def rabitq1_decode(codes: torch.Tensor, D: int, centroid: torch.Tensor | None) -> torch.Tensor:
    _chk(codes, torch.uint8, "codes")
    codes = codes.contiguous()
    n = codes.shape[0]
    out = torch.empty((n, D), dtype=torch.float32, device=codes.device)
    check(_lib.load().b200vq_rabitq1_decode(codes.data_ptr(), n, D, _p(centroid), out.data_ptr(),
                                            out.stride(0), _stream()), "rabitq1_decode")
    return out


def rabitq1_prepare_queries(queries: torch.Tensor, centroid: torch.Tensor | None, qb: int):
    _chk(queries, torch.float32, "queries")
    ptr, nq, ld, D = _rows(queries, "queries")
    W = (D + 31) // 32
    planes = torch.empty((nq, qb, W), dtype=torch.int32, device=queries.device)
    qfac = torch.empty((nq, 8), dtype=torch.float32, device=queries.device)
    check(_lib.load().b200vq_rabitq1_prepare_queries(ptr, nq, ld, D, _p(centroid), qb, planes.data_ptr(),
                                                     qfac.data_ptr(), _stream()), "rabitq1_prepare_queries")
    return planes, qfac


def rabitq1_scan_topk(codes: torch.Tensor, D: int, planes: torch.Tensor, qfac: torch.Tensor, qb: int,
                      k: int, metric, impl: str = "auto"):
    """impl "auto"/"tc": tensor-core scan (csrc/rabitq_tc.cu; it falls back by itself for shapes it does
    not cover); "simt": the AND+popcount kernel. Same ids and values either way."""
    _chk(codes, torch.uint8, "codes")
    codes = codes.contiguous()
    n = codes.shape[0]
    nq = planes.shape[0]
    lib = _lib.load()
    vals, ids = _topk_out(nq, k, codes.device)
    if impl not in ("auto", "tc", "simt"):
        raise ValueError(f"rabitq1_scan_topk: unknown impl {impl!r}")
    if k > 128:
        # one launch ranks at most 128 rows per query: pages, each starting after the last entry of the previous one
        # in the same (estimate, id) order -- the concatenation is the exact top-k
        v0, i0 = rabitq1_scan_topk(codes, D, planes, qfac, qb, 128, metric, impl)
        vals[:, :128], ids[:, :128] = v0, i0
        sc = scratch(lib.b200vq_rabitq1_scan_scratch_bytes(n, nq, 128), codes.device, "scan")
        done = 128
        while done < k:
            kk = min(128, k - done)
            av, ai = vals[:, done - 1].contiguous(), ids[:, done - 1].contiguous()
            v1, i1 = _topk_out(nq, kk, codes.device)
            check(lib.b200vq_rabitq1_scan_topk_after(codes.data_ptr(), n, D, planes.data_ptr(), qfac.data_ptr(), nq, qb,
                                                     metric_id(metric), kk, av.data_ptr(), ai.data_ptr(),
                                                     v1.data_ptr(), i1.data_ptr(), sc.data_ptr(), sc.numel(),
                                                     _stream()), "rabitq1_scan_topk_after")
            vals[:, done:done + kk], ids[:, done:done + kk] = v1, i1
            done += kk
        return vals, ids
    if impl != "simt":
        sc = scratch(lib.b200vq_rabitq1_scan_tc_scratch_bytes(n, D, nq, k), codes.device, "scan")
        check(lib.b200vq_rabitq1_scan_topk_tc(codes.data_ptr(), n, D, planes.data_ptr(), qfac.data_ptr(), nq, qb,
                                              metric_id(metric), k, vals.data_ptr(), ids.data_ptr(), sc.data_ptr(),
                                              sc.numel(), _stream()), "rabitq1_scan_topk_tc")
        return vals, ids
    sc = scratch(lib.b200vq_rabitq1_scan_scratch_bytes(n, nq, k), codes.device, "scan")
    check(lib.b200vq_rabitq1_scan_topk(codes.data_ptr(), n, D, planes.data_ptr(), qfac.data_ptr(), nq, qb,
                                       metric_id(metric), k, vals.data_ptr(), ids.data_ptr(), sc.data_ptr(),
                                       sc.numel(), _stream()), "rabitq1_scan_topk")
    return vals, ids


# ------------------------------------------------------------------ utilities
def synth_rows(n: int, D: int, row0: int, seed: int, col_scale: torch.Tensor | None = None,
               normalize: bool = False, out: torch.Tensor | None = None, device="cuda") -> torch.Tensor:
    if out is None:
        out = torch.empty((n, D), dtype=torch.float32, device=device)
    check(_lib.load().b200vq_synth_rows(out.data_ptr(), n, out.stride(0), D, row0, seed, _p(col_scale),
                                        int(normalize), _stream()), "synth_rows")
    return out


def sq_err_sum(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    _chk(a, torch.float32, "a")
    _chk(b, torch.float32, "b")
    pa, n, lda, D = _rows(a, "a")
    pb, nb, ldb, Db = _rows(b, "b")
    if (n, D) != (nb, Db):
        raise B200VQError("sq_err_sum: shape mismatch")
    out = torch.zeros(1, dtype=torch.float64, device=a.device)
    check(_lib.load().b200vq_sq_err_sum(pa, lda, pb, ldb, n, D, out.data_ptr(), _stream()), "sq_err_sum")
    return out


# ------------------------------------------------------------------ data-fit per-dimension codebooks
def codebook_prefix(cols: torch.Tensor):
    """(sorted columns, prefix sums, prefix sums of squares) of the (D, n) sample columns (library sort / scan)."""
    D, n = cols.shape
    srt = torch.sort(cols, dim=1)[0].contiguous()
    v = srt.to(torch.float64)
    ps = torch.zeros((D, n + 1), dtype=torch.float64, device=cols.device)
    psq = torch.zeros((D, n + 1), dtype=torch.float64, device=cols.device)
    torch.cumsum(v, dim=1, out=ps[:, 1:])
    torch.cumsum(v * v, dim=1, out=psq[:, 1:])
    return srt, ps, psq


def codebook_lloyd(cols: torch.Tensor, bits, seed: int = 0, max_iters: int = 50, tol: float = 1e-6, max_bits: int = 8,
                   presorted=None):
    """1-D k-means codebooks (the SAQ engine's build_codebook_lloyd) for every row of ``cols`` (D, n) fp32 = the sample
    of one dimension each. bits: an int (same bit-rate for all) or a (D,) int32 device tensor of per-dimension
    bit-rates in [0, max_bits] (max_bits <= 13). Returns (centroids (D, 2^max_bits) fp32, sizes (D,) int32, costs (D,)
    fp32 = MSE on the sample). presorted: the (sorted, ps, psq) triple of an earlier call on the same columns.
    The sort and the fp64 prefix sums are library calls (train time); the seeding and the Lloyd iterations are
    b200vq_codebook_lloyd."""
    _chk(cols, torch.float32, "cols")
    D, n = cols.shape
    srt, ps, psq = presorted if presorted is not None else codebook_prefix(cols)
    cen = torch.zeros((D, 1 << max_bits), dtype=torch.float32, device=cols.device)
    sizes = torch.zeros((D,), dtype=torch.int32, device=cols.device)
    costs = torch.zeros((D,), dtype=torch.float32, device=cols.device)
    lib = _lib.load()
    sc = scratch(lib.b200vq_codebook_lloyd_scratch_bytes(n, D), cols.device, "codebook")
    if isinstance(bits, torch.Tensor):
        _chk(bits, torch.int32, "bits")
        bptr, fixed = bits.data_ptr(), 0
    else:
        bptr, fixed = None, int(bits)
    check(lib.b200vq_codebook_lloyd(srt.data_ptr(), srt.stride(0), ps.data_ptr(), psq.data_ptr(), n, D, bptr, fixed,
                                    max_bits, seed, max_iters, tol, cen.data_ptr(), sizes.data_ptr(), costs.data_ptr(), sc.data_ptr(),
                                    sc.numel(), _stream()), "codebook_lloyd")
    return cen, sizes, costs
