"""ExtendedRaBitQuantizer (multi-bit RaBitQ) on the B200.

Mirrors src/haag_vq/methods/extended_rabitq.py:47-204. Model state (centroid,
seeded QR rotation P, sampled Lloyd-Max levels) is data-independent apart from
the mean and is derived with exactly the reference's seeded numpy recipe on the
host (tiny, fit-time only). The per-vector work runs on the device in fp64 like
the reference: residual/normalise (torch elementwise), rotation (fp64 library
GEMM), then the hand-written row tail (level search, rescale factor, MSB-first
bit packing / unpacking).
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
            bounds = 0.5 * (self.levels[:-1] + self.levels[1:])
            if bounds.size == 0:
                bounds = np.zeros(1)
            self._dev = dict(c=host.to_device(self.c), P=host.to_device(self.P),
                             levels=host.to_device(self.levels), bounds=host.to_device(bounds))
        return self._dev

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
        sqrt_d = float(np.sqrt(D))
        def f(xb, s, e):
            r = xb.to(torch.float64) - st["c"]
            nrm = torch.linalg.vector_norm(r, dim=1)
            o = r / torch.clamp(nrm, min=self._eps)[:, None]
            srot = torch.matmul(o, st["P"]) * sqrt_d
            out[s:e] = ops_sq.erabitq_pack(srot, nrm, self.num_bits, st["levels"], st["bounds"])
        host.stream_rows(X, f, batch_rows=host.batch_rows_for(D * 8 * 4, N))
        return host.to_host(out)

    def decompress(self, codes: np.ndarray) -> np.ndarray:
        if self.P is None:
            raise RuntimeError("Quantizer must be fit before decompress().")
        codes = np.ascontiguousarray(codes, dtype=np.uint8)
        st = self._state()
        N = codes.shape[0]
        out = np.empty((N, self.D), dtype=np.float32)
        step = host.batch_rows_for(self.D * 8 * 4, max(N, 1))
        for s in range(0, N, step):
            e = min(N, s + step)
            shat, nrm = ops_sq.erabitq_unpack(host.to_device(codes[s:e]), self.D, self.num_bits, st["levels"])
            xh = torch.matmul(shat, st["P"].t()) * nrm[:, None] + st["c"]
            out[s:e] = host.to_host(xh.to(torch.float32))
        return out

    def get_compression_ratio(self, X: np.ndarray) -> float:
        D = int(X.shape[1])
        return float(D * 4 / self.code_size)
