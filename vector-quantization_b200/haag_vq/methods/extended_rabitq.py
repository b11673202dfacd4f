"""ExtendedRaBitQuantizer (multi-bit RaBitQ) on the B200.

Mirrors src/haag_vq/methods/extended_rabitq.py:47-204. Model state (centroid,
seeded QR rotation P, sampled Lloyd-Max levels) is data-independent apart from
the mean and is derived with exactly the reference's seeded numpy recipe on the
host (tiny, fit-time only). The per-vector work runs in libb200vq.so: residual
norms in fp64, the O(N D^2) rotation as the fp32-accurate tensor-core GEMM
(b200vq_rotate_tc_ex, centroid subtraction fused), level indices from it with an
exact fp64 re-evaluation of every coordinate that lies within the GEMM's error
bound of a level mid-point (b200vq_erabitq_pack_tc) - so the indices are those of
the reference's fp64 computation - and the inverse path as decode + the same GEMM.
"""
import numpy as np
import torch

from b200vq import host, ops_sq

from .base_quantizer import BaseQuantizer


def _lloyd_1d_normal(num_levels: int, seed: int, n_samples: int = 200_000, max_iter: int = 100,
                     tol: float = 1e-7) -> np.ndarray:
    """Sampled Lloyd-Max levels for N(0,1) (extended_rabitq.py:6-44). Kept on the host:
    it is seeded model state, not per-vector arithmetic."""
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


class ExtendedRaBitQuantizer(BaseQuantizer):
    def __init__(self, num_bits: int = 4, seed: int = 0):
        if num_bits < 1 or num_bits > 8:
            raise ValueError("num_bits must be in [1, 8]")
        self.num_bits = int(num_bits)
        self.seed = int(seed)
        self.c = None
        self.P = None
        self.levels = None
        self.D = None
        self._eps = 1e-12
        self._dev = None

    def __getstate__(self):
        st = dict(self.__dict__)
        st["_dev"] = None
        return st

    def fit(self, X: np.ndarray) -> None:
        X = np.asarray(X)
        N, D = X.shape
        self.D = int(D)
        self.c = X.mean(axis=0).astype(np.float64)
        rng = np.random.default_rng(self.seed)
        Q, _ = np.linalg.qr(rng.standard_normal((D, D)))
        self.P = Q.astype(np.float64)
        self.levels = _lloyd_1d_normal(2 ** self.num_bits, seed=self.seed)
        self._dev = None

    @property
    def _index_bytes(self) -> int:
        return (self.D * self.num_bits + 7) // 8

    @property
    def code_size(self) -> int:
        return self._index_bytes + 8

    def _state(self):
        if self._dev is None:
            import torch as _t
            from b200vq import ops
            bounds = 0.5 * (self.levels[:-1] + self.levels[1:])
            if bounds.size == 0:
                bounds = np.zeros(1)
            D = self.D
            Dp = (D + 31) // 32 * 32            # the GEMM wants K % 32 == 0: zero columns are exact
            P32 = self.P.astype(np.float32)
            Wf = np.zeros((D, Dp), np.float32)  # s = o P: W[d][k] = P[k][d]
            Wf[:, :D] = P32.T
            Wr = np.zeros((D, Dp), np.float32)  # x^ = o^ P^T: W[j][k] = P[j][k]
            Wr[:, :D] = P32
            c32 = np.zeros(Dp, np.float32)
            c32[:D] = self.c.astype(np.float32)
            err_rel, err_abs = ops_sq.erabitq_rotation_error_bound(Dp, float(np.linalg.norm(self.c)))
            self._dev = dict(c=host.to_device(self.c), PT=host.to_device(np.ascontiguousarray(self.P.T)),
                             levels=host.to_device(self.levels), bounds=host.to_device(bounds),
                             levels32=host.to_device(self.levels.astype(np.float32)),
                             fwd=ops.RotationTC(host.to_device(Wf)), rev=ops.RotationTC(host.to_device(Wr)),
                             c32=host.to_device(c32), c32d=host.to_device(c32[:D].copy()), Dp=Dp,
                             err_rel=err_rel, err_abs=err_abs,
                             nfix=_t.zeros(1, dtype=_t.int64, device=host.device()))
        return self._dev

    def _pad_cols(self, xb: torch.Tensor, Dp: int) -> torch.Tensor:
        if xb.shape[1] == Dp and xb.stride(0) % 4 == 0 and xb.data_ptr() % 16 == 0:
            return xb
        xp = torch.zeros((xb.shape[0], Dp), dtype=torch.float32, device=xb.device)
        xp[:, :xb.shape[1]] = xb
        return xp

    def compress(self, X: np.ndarray) -> np.ndarray:
        if self.P is None:
            raise RuntimeError("Quantizer must be fit before compress().")
        X = np.ascontiguousarray(X)
        if X.ndim != 2 or X.shape[1] != self.D:
            raise ValueError(f"compress() got D={X.shape[-1]}, but fit() saw D={self.D}")
        if X.dtype not in (np.float32, np.float64):
            X = X.astype(np.float64)
        st = self._state()
        N, D = X.shape
        out = torch.empty((N, self.code_size), dtype=torch.uint8, device=host.device())
        def f(xb, s, e):
            nrm = ops_sq.resid_norms_f64(xb, st["c"])
            x32 = xb if xb.dtype == torch.float32 else xb.to(torch.float32)
            y = st["fwd"](self._pad_cols(x32, st["Dp"]), x_sub=st["c32"])       # (x - c32) P32, fp32-accurate
            # an fp64 input is screened through its fp32 cast: 2^-24 (|x - c| + |c|) more rotation error
            rel = st["err_rel"] + (2.0 ** -24 if xb.dtype == torch.float64 else 0.0)
            ops_sq.erabitq_pack_tc(y, xb, st["c"], st["PT"], nrm, self.num_bits, st["levels"], st["bounds"],
                                   rel, 2.0 * st["err_abs"], out=out[s:e], stats=st["nfix"])
        host.stream_rows(X, f, batch_rows=host.batch_rows_for(D * 4 * 4, N))
        return host.to_host(out)

    def decompress(self, codes: np.ndarray) -> np.ndarray:
        if self.P is None:
            raise RuntimeError("Quantizer must be fit before decompress().")
        codes = np.ascontiguousarray(codes, dtype=np.uint8)
        st = self._state()
        N = codes.shape[0]
        out = np.empty((N, self.D), dtype=np.float32)
        step = host.batch_rows_for(self.D * 4 * 4, max(N, 1))
        for s in range(0, N, step):
            e = min(N, s + step)
            # o^ nrm = levels[idx] t nrm / sqrt(D) (b200vq_erabitq_decode_f32), then x^ = (o^ nrm) P^T + c
            xs, _ = ops_sq.erabitq_decode_f32(host.to_device(codes[s:e]), self.D, self.num_bits, st["levels32"])
            xh = st["rev"](self._pad_cols(xs, st["Dp"]), col_bias=st["c32d"])
            out[s:e] = host.to_host(xh)
        return out

    # ---- code-domain search support (FlatQuantizedIndex): not part of the reference surface
    def device_scanner(self, codes: np.ndarray):
        """Codes resident on the device, prepared for the estimator scan on the tensor cores."""
        st = self._state()
        return ops_sq.ERaBitQScanner(host.to_device(np.ascontiguousarray(codes, dtype=np.uint8)), self.D,
                                     self.num_bits, st["levels32"])

    def rotate_queries(self, q_dev: torch.Tensor, metric: str):
        """(q_rot, q_const): L2 -> ((q - c) P, |q - c|^2); IP -> (q P, q.c)."""
        from b200vq import ops
        st = self._state()
        qp = self._pad_cols(q_dev, st["Dp"])
        if metric == "l2":
            qrot = st["fwd"](qp, x_sub=st["c32"])
            qconst = ops.row_sqnorms((q_dev - st["c32d"]).contiguous())
        else:
            qrot = st["fwd"](qp)
            qconst = (q_dev * st["c32d"]).sum(1)
        return qrot, qconst.contiguous()

    def get_compression_ratio(self, X: np.ndarray) -> float:
        D = int(X.shape[1])
        return float(D * 4 / self.code_size)
