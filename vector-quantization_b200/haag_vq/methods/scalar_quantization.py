"""ScalarQuantizer on the B200 kernels.

Mirrors src/haag_vq/methods/scalar_quantization.py:6-100: per-dimension min/max
uniform quantiser, 4 / 8 / 16 bits, nibble packing for 4-bit. Results are
bit-identical to the numpy reference for in-range inputs; out-of-range inputs
saturate (the reference's ``astype(uint8)`` is undefined there).
"""
import numpy as np
import torch

from b200vq import host, ops_sq

from .base_quantizer import BaseQuantizer


class ScalarQuantizer(BaseQuantizer):
    def __init__(self, num_bits: int = 8):
        if num_bits not in [4, 8, 16]:
            raise ValueError(f"num_bits must be 4, 8, or 16, got {num_bits}")
        self.num_bits = num_bits
        self.num_levels = 2 ** num_bits
        self.min = None
        self.max = None
        if num_bits <= 8:
            self.dtype = np.uint8
            self.bytes_per_dim = 1
        else:
            self.dtype = np.uint16
            self.bytes_per_dim = 2
        self._dev = None

    def __getstate__(self):
        st = dict(self.__dict__)
        st["_dev"] = None
        return st

    def _params(self):
        if self._dev is None:
            self._dev = (host.to_device(np.asarray(self.min, dtype=np.float32)),
                         host.to_device(np.asarray(self.max, dtype=np.float32)))
        return self._dev

    def fit(self, X):
        X = host.as_f32_2d(X)
        state = [None, None]
        def f(xb, s, e):
            state[0], state[1] = ops_sq.sq_minmax(xb, state[0], state[1])
        host.stream_rows(X, f)
        self._dev = (state[0], state[1])
        self.min = host.to_host(state[0]).copy()
        self.max = host.to_host(state[1]).copy()

    def compress(self, X):
        if self.min is None:
            raise RuntimeError("ScalarQuantizer must be fitted before compress().")
        X = host.as_f32_2d(X)
        n, D = X.shape
        vmin, vmax = self._params()
        shape, dt = ops_sq.sq_code_shape(n, D, self.num_bits)
        out = torch.empty(shape, dtype=dt, device=host.device())
        host.stream_rows(X, lambda xb, s, e: ops_sq.sq_encode(xb, vmin, vmax, self.num_bits, out=out[s:e]))
        codes = host.to_host(out)
        return codes.view(np.uint16) if self.num_bits == 16 else codes

    def decompress(self, codes):
        if self.min is None:
            raise RuntimeError("ScalarQuantizer must be fitted before decompress().")
        codes = np.ascontiguousarray(codes, dtype=self.dtype)
        vmin, vmax = self._params()
        D = len(self.min)
        return host.to_host(ops_sq.sq_decode(host.to_device(codes), vmin, vmax, self.num_bits, D))

    def get_compression_ratio(self, X):
        original_size = X.shape[1] * 4
        if self.num_bits == 4:
            compressed_size = (X.shape[1] + 1) // 2
        else:
            compressed_size = X.shape[1] * self.bytes_per_dim
        return original_size / compressed_size
