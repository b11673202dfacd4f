"""RankAwareQuantizer on the B200 kernels.

Mirrors src/haag_vq/methods/rank_aware_quantization.py:57-329: centre -> PCA rotate -> per-dimension
scalar codebook scaled by sqrt(var_d), with per-dimension bit widths from a greedy whose marginal gain is
var_d^(1+alpha) * (Dg[b] - Dg[b+1]); codes are per-row bit strings (``packing="dense"``, MSB-first) or FFD
byte layouts (``packing="ffd"``). Same constructor, public state (mu, V, var, bits, cb, D, code_size) and
errors as the reference.

Device work: mean / covariance (fp64 accumulation) and the eigendecomposition in ``fit`` (train time); the
projection GEMMs on the tensor cores (``b200vq_rotate_tc_ex``, fp32-accurate) with the level search finished in
exact fp64 wherever the GEMM's error bound reaches a codebook mid-point (``b200vq_perdim_encode_tc``), and the
inverse gather + GEMM (``b200vq_perdim_decode_f32``). The 1-D Lloyd-Max tables for N(0,1) and the greedy allocation are host
training logic (a few thousand scalar steps). ``codebook="gaussian"`` only: the data-fit codebooks
('lloyd', 'exact') come from the SAQ engine's C++ builders (external/saq), which are not built here.
Given the same trained state, codes are identical to the reference's wherever a projected coordinate is
not within rounding distance (1e-12 relative) of a codebook mid-point: the fp64 GEMM's summation order is
the library's, not numpy's.
"""
from __future__ import annotations

import functools

import numpy as np
import torch

from b200vq import _lib, host
from b200vq._lib import check
from b200vq.ops import _stream

from .base_quantizer import BaseQuantizer
from .ffd_packing import ffd_layout


@functools.lru_cache(maxsize=64)   # deterministic in its arguments; a fit needs nine tables
def _lloyd_1d_normal(num_levels: int, seed: int, n_samples: int = 200_000, max_iter: int = 100, tol: float = 1e-7):
    """(levels, normalised MSE) of the Lloyd-Max quantizer of N(0,1) on a seeded sample
    (rank_aware_quantization.py:7-54): quantile start, nearest-level assignment by mid-points,
    centroid update, stop when no level moves by more than ``tol``."""
    x = np.random.default_rng(seed).standard_normal(n_samples)
    power = float(np.mean(x ** 2))
    if num_levels == 1:
        return np.array([0.0]), float(np.mean(x ** 2) / power)
    levels = np.quantile(x, (np.arange(num_levels) + 0.5) / num_levels)
    for _ in range(max_iter):
        cell = np.searchsorted(0.5 * (levels[:-1] + levels[1:]), x)
        moved = levels.copy()
        for k in range(num_levels):
            members = cell == k
            if np.any(members):
                moved[k] = x[members].mean()
        moved.sort()
        step = float(np.max(np.abs(moved - levels)))
        levels = moved
        if step < tol:
            break
    cell = np.searchsorted(0.5 * (levels[:-1] + levels[1:]), x)
    return levels.astype(np.float64), float(np.mean((x - levels[cell]) ** 2) / power)


class RankAwareQuantizer(BaseQuantizer):
    def __init__(self, avg_bits: float, alpha: float = 1.0, max_bits: int = 8, seed: int = 0,
                 packing: str = "dense", codebook: str = "gaussian"):
        if max_bits < 1 or max_bits > 8:
            raise ValueError("max_bits must be in [1, 8]")
        if packing not in ("dense", "ffd"):
            raise ValueError("packing must be 'dense' or 'ffd'")
        if codebook not in ("gaussian", "lloyd", "exact"):
            raise ValueError("codebook must be 'gaussian', 'lloyd', or 'exact'")
        if codebook == "exact":
            raise NotImplementedError(
                "codebook='exact' (the engine's optimal-DP 1-D k-means, build_codebook_exact) is not built on the "
                "B200 path; 'lloyd' (the method's production codebook) and 'gaussian' are. No fallback is provided")
        self.codebook = codebook
        self.avg_bits = float(avg_bits)
        self.alpha = float(alpha)
        self.max_bits = int(max_bits)
        self.seed = int(seed)
        self.packing = packing
        self.mu = self.V = self.var = self.bits = self.cb = None
        self.D = None
        self._offsets = None
        self.code_size = None
        self._ffd_byte_idx = self._ffd_bit_off = self._ffd_n_bytes = None
        self._dev = None

    def __getstate__(self):
        st = dict(self.__dict__)
        st["_dev"] = None
        return st

    # ------------------------------------------------------------------ fit
    def fit(self, X: np.ndarray) -> None:
        X = np.asarray(X)
        N, D = X.shape
        self.D = int(D)
        dev = host.device()
        S = torch.zeros(D, dtype=torch.float64, device=dev)
        G = torch.zeros((D, D), dtype=torch.float64, device=dev)
        rows = max(1, min(N, (1 << 28) // max(1, 8 * D)))
        for s in range(0, N, rows):
            blk = host.to_device(np.ascontiguousarray(X[s:s + rows])).to(torch.float64)
            S += blk.sum(dim=0)
            G += blk.t() @ blk
        mu = S / N
        C = G / N - torch.outer(mu, mu)
        w, Vt = torch.linalg.eigh(C)                       # ascending
        order = torch.argsort(w, descending=True)
        self.mu = host.to_host(mu).copy()
        self.var = np.clip(host.to_host(w[order]), 1e-12, None)
        self.V = host.to_host(Vt[:, order]).copy()

        self._levels = [None] * (self.max_bits + 1)
        Dg = np.empty(self.max_bits + 1, dtype=np.float64)
        for b in range(self.max_bits + 1):
            lv, Dg[b] = _lloyd_1d_normal(2 ** b, seed=self.seed + b)
            self._levels[b] = lv.copy()
        Dg[0] = 1.0
        self._Dg = Dg
        self.bits = self._allocate(self.var, Dg, int(round(self.avg_bits * D)))
        scale = np.sqrt(self.var)
        if self.codebook == "gaussian":
            self.cb = [self._levels[int(b)] * scale[d] for d, b in enumerate(self.bits)]
        else:
            self.cb = self._fit_lloyd_codebooks(X)
        self._finish_layout()

    CB_SAMPLE = 200_000

    def _fit_lloyd_codebooks(self, X: np.ndarray):
        """Per-dimension data-fit codebooks (rank_aware_quantization.py:189-214): the engine's build_codebook_lloyd on
        the projected column of a fixed ~200k-row sample at the dimension's own bit-rate. Projection on the tensor
        cores, then one CTA per dimension (b200vq_codebook_lloyd: k-means++ seeding with mt19937_64(0), Lloyd over
        the sorted column). The codebooks are trained state (the fp32-accurate projection differs from the
        reference's float64-then-cast in the last bit of some sample values)."""
        from b200vq import ops, ops_sq
        N, D = X.shape
        if N > self.CB_SAMPLE:
            idx = np.sort(np.random.default_rng(0).choice(N, self.CB_SAMPLE, replace=False))
            Xs = np.ascontiguousarray(X[idx], dtype=np.float32)
        else:
            Xs = np.ascontiguousarray(X, dtype=np.float32)
        Dp = (D + 31) // 32 * 32
        Wf = np.zeros((D, Dp), np.float32)
        Wf[:, :D] = self.V.astype(np.float32).T
        mu32 = np.zeros(Dp, np.float32)
        mu32[:D] = self.mu.astype(np.float32)
        fwd = ops.RotationTC(host.to_device(Wf))
        xd = self._pad_cols(host.to_device(Xs), Dp)
        yT = fwd(xd, x_sub=host.to_device(mu32), transpose_out=True).contiguous()      # (D, n): one sample column per row
        bits = host.to_device(self.bits.astype(np.int32))
        cen, sizes, _ = ops_sq.codebook_lloyd(yT, bits, seed=0)
        cen_h, sizes_h = host.to_host(cen), host.to_host(sizes)
        cb = []
        for d, b in enumerate(self.bits):
            if int(b) == 0:
                cb.append(np.array([0.0], dtype=np.float64))
                continue
            c = np.sort(cen_h[d, :sizes_h[d]].astype(np.float64))
            if c.size < (1 << int(b)):        # fewer distinct sample values than levels: pad like the reference's encoders need
                c = np.concatenate([c, np.full((1 << int(b)) - c.size, c[-1])])
            cb.append(c)
        return cb

    def _allocate(self, var: np.ndarray, Dg: np.ndarray, total: int) -> np.ndarray:
        """Greedy of rank_aware_quantization.py:148-181: one bit at a time to the dimension with the largest
        var^(1+alpha) * (Dg[b] - Dg[b+1]); first index wins ties (np.argmax); stops when every dimension is capped."""
        D = var.shape[0]
        weight = var ** (1.0 + self.alpha)
        drop = Dg[:-1] - Dg[1:]
        bits = np.zeros(D, dtype=np.int64)
        gain = weight * drop[0]
        for _ in range(total):
            d = int(np.argmax(gain))
            if not np.isfinite(gain[d]):
                break
            bits[d] += 1
            gain[d] = weight[d] * drop[bits[d]] if bits[d] < self.max_bits else -np.inf
        return bits

    def _finish_layout(self) -> None:
        self.bits = np.asarray(self.bits, dtype=np.int64)
        self._offsets = np.concatenate(([0], np.cumsum(self.bits))).astype(np.int64)
        if self.packing == "ffd":
            self._ffd_byte_idx, self._ffd_bit_off, n_bytes = ffd_layout(self.bits)
            self._ffd_n_bytes = int(n_bytes)
            self.code_size = int(n_bytes)
        else:
            self.code_size = (int(self.bits.sum()) + 7) // 8
        self._dev = None

    def set_state(self, mu, V, var, bits, cb) -> None:
        """Adopt trained state (e.g. exported by the reference class): parity tests and load paths."""
        self.mu = np.asarray(mu, dtype=np.float64)
        self.V = np.asarray(V, dtype=np.float64)
        self.var = np.asarray(var, dtype=np.float64)
        self.bits = np.asarray(bits, dtype=np.int64)
        self.cb = [np.asarray(c, dtype=np.float64) for c in cb]
        self.D = int(self.mu.shape[0])
        self._finish_layout()

    # -------------------------------------------------------------- device tables
    def _tables(self):
        if self._dev is None:
            from b200vq import ops, ops_sq
            D = self.D
            bits = self.bits
            if self.packing == "ffd":
                pos = np.where(bits > 0, self._ffd_byte_idx * 8 + self._ffd_bit_off, 0)
            else:
                pos = self._offsets[:-1]
            bounds = [0.5 * (c[:-1] + c[1:]) if b > 0 else np.zeros(0) for c, b in zip(self.cb, bits)]
            boff = np.concatenate(([0], np.cumsum([len(x) for x in bounds])))[:-1]
            cboff = np.concatenate(([0], np.cumsum([len(c) for c in self.cb])))[:-1]
            t = lambda a, dt: host.to_device(np.ascontiguousarray(a, dtype=dt))
            Dp = (D + 31) // 32 * 32              # the GEMM wants K % 32 == 0: zero columns are exact
            V32 = self.V.astype(np.float32)
            Wf = np.zeros((D, Dp), np.float32)    # y = (x - mu) V: W[d][k] = V[k][d]
            Wf[:, :D] = V32.T
            Wr = np.zeros((D, Dp), np.float32)    # x^ = y^ V^T: W[j][d] = V[j][d]
            Wr[:, :D] = V32
            mu32 = np.zeros(Dp, np.float32)
            mu32[:D] = self.mu.astype(np.float32)
            err_rel, err_abs = ops_sq.erabitq_rotation_error_bound(Dp, float(np.linalg.norm(self.mu)))
            self._dev = {
                "mu": t(self.mu, np.float64), "VT": t(self.V.T, np.float64),
                "bits": t(bits, np.uint8), "pos": t(pos, np.int32),
                "bounds": t(np.concatenate(bounds + [np.zeros(1)]), np.float64), "boff": t(boff, np.int32),
                "cb32": t(np.concatenate(list(self.cb)), np.float32), "cboff": t(cboff, np.int32),
                "fwd": ops.RotationTC(host.to_device(Wf)), "rev": ops.RotationTC(host.to_device(Wr)),
                "mu32": host.to_device(mu32), "mu32d": host.to_device(mu32[:D].copy()), "Dp": Dp,
                "err_rel": err_rel, "err_abs": err_abs,
                "nfix": torch.zeros(1, dtype=torch.int64, device=host.device()),
            }
        return self._dev

    @staticmethod
    def _pad_cols(xb: torch.Tensor, Dp: int) -> torch.Tensor:
        if xb.shape[1] == Dp and xb.stride(0) % 4 == 0 and xb.data_ptr() % 16 == 0:
            return xb
        xp = torch.zeros((xb.shape[0], Dp), dtype=torch.float32, device=xb.device)
        xp[:, :xb.shape[1]] = xb
        return xp

    # ------------------------------------------------------------- compress
    def compress(self, X: np.ndarray) -> np.ndarray:
        """Projection on the tensor cores (fp32-accurate 3xTF32 GEMM, mean subtraction fused) + nearest-level search
        with an exact fp64 re-evaluation of every coordinate inside the GEMM's error bound of a codebook mid-point
        (b200vq_perdim_encode_tc): the level indices are those of the reference's float64 projection."""
        if self.V is None:
            raise RuntimeError("Quantizer must be fit before compress().")
        from b200vq import ops_sq
        X = np.ascontiguousarray(X)
        N, D = X.shape
        if D != self.D:
            raise ValueError(f"compress() got D={D}, but fit() saw D={self.D}")
        if X.dtype not in (np.float32, np.float64):
            X = X.astype(np.float64)
        T = self._tables()
        out = torch.empty((N, self.code_size), dtype=torch.uint8, device=host.device())
        lib = _lib.load()

        def f(xb, s, e):
            nrm = ops_sq.resid_norms_f64(xb, T["mu"])
            x32 = xb if xb.dtype == torch.float32 else xb.to(torch.float32)
            y = T["fwd"](self._pad_cols(x32, T["Dp"]), x_sub=T["mu32"])
            rel = T["err_rel"] + (2.0 ** -24 if xb.dtype == torch.float64 else 0.0)
            check(lib.b200vq_perdim_encode_tc(y.data_ptr(), y.stride(0), xb.data_ptr(), int(xb.dtype == torch.float64),
                                              xb.stride(0), T["mu"].data_ptr(), T["VT"].data_ptr(), nrm.data_ptr(), e - s, D,
                                              T["bounds"].data_ptr(), T["boff"].data_ptr(), T["bits"].data_ptr(),
                                              T["pos"].data_ptr(), self.code_size, rel, 2.0 * T["err_abs"],
                                              out[s:e].data_ptr(), T["nfix"].data_ptr(), _stream()), "perdim_encode_tc")
        if N:
            host.stream_rows(X, f, batch_rows=host.batch_rows_for(D * 4 * 4, N))
        return host.to_host(out)

    # ----------------------------------------------------------- decompress
    def decompress(self, codes: np.ndarray) -> np.ndarray:
        if self.V is None:
            raise RuntimeError("Quantizer must be fit before decompress().")
        codes = np.ascontiguousarray(codes, dtype=np.uint8)
        N, D = codes.shape[0], self.D
        T = self._tables()
        out = np.empty((N, D), dtype=np.float32)
        rows = max(1, min(max(N, 1), (1 << 28) // max(1, 8 * D)))
        lib = _lib.load()
        for s in range(0, N, rows):
            e = min(N, s + rows)
            c = host.to_device(codes[s:e])
            Yh = torch.zeros((e - s, T["Dp"]), dtype=torch.float32, device=c.device)
            check(lib.b200vq_perdim_decode_f32(c.data_ptr(), e - s, self.code_size, D, T["bits"].data_ptr(),
                                               T["pos"].data_ptr(), T["cb32"].data_ptr(), T["cboff"].data_ptr(),
                                               Yh.data_ptr(), Yh.stride(0), _stream()), "perdim_decode_f32")
            out[s:e] = host.to_host(T["rev"](Yh, col_bias=T["mu32d"]))       # y^ V^T + mu on the tensor cores
        return out

    # ---- code-domain search (not part of the reference surface; the ADC loop of bench/ffd_speed.cpp:87-127)
    def adc_search(self, codes, Q: np.ndarray, k: int, metric: str = "l2"):
        """Top-k of |q - x^_i|^2 (l2) or q . x^_i (ip) straight from the packed codes (dense or FFD layout):
        with y^ the decoded PCA coordinates and V orthonormal, |q - x^|^2 = |(q - mu) V - y^|^2 and
        q . x^ = (q V) . y^ + q . mu. codes: (n, code_size) uint8 numpy array or device tensor.
        Returns (ids uint32 (nq, k), values fp32 (nq, k))."""
        from b200vq import ops
        from b200vq.ops import _topk_out, metric_id, scratch
        T = self._tables()
        cd = codes if isinstance(codes, torch.Tensor) else host.to_device(np.ascontiguousarray(codes, dtype=np.uint8))
        n, D = int(cd.shape[0]), self.D
        qd = host.to_device(np.ascontiguousarray(Q, dtype=np.float32))
        nq = int(qd.shape[0])
        lib = _lib.load()
        if metric == "l2":
            qp = T["fwd"](self._pad_cols(qd, T["Dp"]), x_sub=T["mu32"]).contiguous()
            qc = ops.row_sqnorms(qp)
            bias = self._row_sqnorms(cd)
        else:
            qp = T["fwd"](self._pad_cols(qd, T["Dp"])).contiguous()
            qc = (qd * T["mu32d"]).sum(1).contiguous()
            bias = None
        vals, ids = _topk_out(nq, k, cd.device)
        sc = scratch(lib.b200vq_perdim_scan_scratch_bytes(n, nq, k), cd.device, "scan")
        check(lib.b200vq_perdim_scan_topk(cd.data_ptr(), n, self.code_size, D, T["bits"].data_ptr(), T["pos"].data_ptr(),
                                          T["cb32"].data_ptr(), T["cboff"].data_ptr(),
                                          bias.data_ptr() if bias is not None else None, qp.data_ptr(), nq, qp.stride(0),
                                          qc.data_ptr(), metric_id(metric), k, vals.data_ptr(), ids.data_ptr(),
                                          sc.data_ptr(), sc.numel(), _stream()), "perdim_scan_topk")
        return host.to_host(ids).view(np.uint32), host.to_host(vals)

    def _row_sqnorms(self, cd: torch.Tensor) -> torch.Tensor:
        """|y^_i|^2 of every stored row (decode in row batches + b200vq_row_sqnorms)."""
        from b200vq import ops
        T = self._tables()
        n = int(cd.shape[0])
        out = torch.empty((n,), dtype=torch.float32, device=cd.device)
        step = max(1, (1 << 28) // max(1, 4 * self.D))
        lib = _lib.load()
        for s in range(0, n, step):
            e = min(n, s + step)
            yh = torch.empty((e - s, self.D), dtype=torch.float32, device=cd.device)
            check(lib.b200vq_perdim_decode_f32(cd[s:e].data_ptr(), e - s, self.code_size, self.D, T["bits"].data_ptr(),
                                               T["pos"].data_ptr(), T["cb32"].data_ptr(), T["cboff"].data_ptr(),
                                               yh.data_ptr(), yh.stride(0), _stream()), "perdim_decode_f32")
            out[s:e] = ops.row_sqnorms(yh)
        return out

    def get_compression_ratio(self, X: np.ndarray) -> float:
        return float(int(X.shape[1]) * 4 / self.code_size)
