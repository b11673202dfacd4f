"""RaBitQuantizer (1 bit / dimension) on the B200 kernels.

Mirrors src/haag_vq/methods/rabit_quantization.py:9-40, a wrapper over
faiss.RaBitQuantizer(D, metric): ``train`` learns nothing (the bare faiss
quantizer has no centroid; only IndexRaBitQ sets one), ``compute_codes`` emits
ceil(D/8) sign bytes + 8 bytes of factors, ``decode`` reconstructs
(bit - 1/2) * 2 * dp_multiplier / sqrt(D).
"""
import numpy as np
import torch

from b200vq import host, ops_sq

from ..utils.faiss_utils import MetricType
from .base_quantizer import BaseQuantizer


class DeviceRaBitQ:
    """Stand-in for faiss.RaBitQuantizer."""

    def __init__(self, d: int, metric_type=MetricType.L2):
        self.d = int(d)
        self.metric_type = metric_type
        self.code_size = (self.d + 7) // 8 + 8
        self.centroid = None       # numpy (d,) or None
        self._centroid_dev = None

    @property
    def _metric(self) -> str:
        return 'ip' if int(self.metric_type) == int(MetricType.INNER_PRODUCT) else 'l2'

    def set_centroid(self, c) -> None:
        self.centroid = None if c is None else np.asarray(c, dtype=np.float32)
        self._centroid_dev = None

    def centroid_dev(self):
        if self.centroid is None:
            return None
        if self._centroid_dev is None:
            self._centroid_dev = host.to_device(self.centroid)
        return self._centroid_dev

    def __getstate__(self):
        st = dict(self.__dict__)
        st["_centroid_dev"] = None
        return st

    def train(self, X) -> None:  # faiss: does nothing
        return None

    def compute_codes(self, X: np.ndarray) -> np.ndarray:
        X = host.as_f32_2d(X)
        if X.shape[1] != self.d:
            raise ValueError(f"compute_codes: got D={X.shape[1]}, expected {self.d}")
        out = torch.empty((X.shape[0], self.code_size), dtype=torch.uint8, device=host.device())
        c = self.centroid_dev()
        def f(xb, s, e):
            out[s:e] = ops_sq.rabitq1_encode(xb, c, self._metric)
        host.stream_rows(X, f)
        return host.to_host(out)

    def decode(self, codes: np.ndarray) -> np.ndarray:
        codes = np.ascontiguousarray(codes, dtype=np.uint8)
        return host.to_host(ops_sq.rabitq1_decode(host.to_device(codes), self.d, self.centroid_dev()))


class RaBitQuantizer(BaseQuantizer):
    def __init__(self, metric_type: MetricType = MetricType.L2):
        self.metric_type = metric_type
        self.rabitq: DeviceRaBitQ = None

    def fit(self, X: np.ndarray):
        N, D = X.shape
        self.rabitq = DeviceRaBitQ(D, self.metric_type)
        self.rabitq.train(X)

    def compress(self, X: np.ndarray) -> np.ndarray:
        if self.rabitq is None:
            raise RuntimeError("RaBitQuantizer must be fitted before compress().")
        return self.rabitq.compute_codes(X)

    def decompress(self, compressed: np.ndarray) -> np.ndarray:
        if self.rabitq is None:
            raise RuntimeError("RaBitQuantizer must be fitted before decompress().")
        return self.rabitq.decode(compressed)

    def get_compression_ratio(self, X: np.ndarray) -> float:
        D = int(X.shape[1])
        return float(D * 4 / int(self.rabitq.code_size))
