"""FlatQuantizedIndex — any BaseQuantizer + brute-force search, on the B200.

Mirrors src/haag_vq/methods/search/flat_quantized_index.py:17-154. ``fit``
compresses the database; ``search_with_scores`` ranks by distance to the
reconstruction. Where the reference decompresses all N vectors on every call and
runs cdist/GEMM + argpartition on the host, this implementation keeps the codes
resident in HBM and
  * PQ / OPQ: scans the codes with per-query ADC look-up tables (the ADC sum
    equals the distance to the reconstruction) and a warp-level top-k;
  * other quantizers: decodes in device batches and runs the exact flat
    GEMM + top-k kernel over the reconstructions.
Same outputs: ids uint32 (nq, k), squared-L2 ascending or inner product
descending, k clipped to N, k <= 0 -> empty arrays.
"""
from __future__ import annotations

import pickle
from pathlib import Path
from typing import Literal, Optional, Tuple

import numpy as np
import torch

from b200vq import host, ops
from b200vq.pq import scan_layout_width

from haag_vq.methods.base_quantizer import BaseQuantizer
from haag_vq.methods.base_search_index import BaseSearchIndex

MAX_K = 128


class FlatQuantizedIndex(BaseSearchIndex):
    def __init__(self, quantizer: BaseQuantizer) -> None:
        self._quantizer = quantizer
        self._codes: Optional[np.ndarray] = None
        self._metric: Literal['l2', 'ip'] = 'l2'
        self._N: int = 0
        self._D: int = 0
        self._codes_dev: Optional[torch.Tensor] = None
        self._recon_dev: Optional[torch.Tensor] = None
        self._codes_rot: Optional[torch.Tensor] = None
        self._erq_scanner = None

    def fit(self, X: np.ndarray, metric: Literal['l2', 'ip'] = 'l2') -> None:
        X = np.ascontiguousarray(X, dtype=np.float32)
        self._metric = metric
        self._N, self._D = X.shape
        self._quantizer.fit(X)
        self._codes = self._quantizer.compress(X)
        self._codes_dev = None
        self._recon_dev = None
        self._codes_rot = None
        self._erq_scanner = None

    def attach_device_codes(self, codes_dev: torch.Tensor, D: int,
                            metric: Literal['l2', 'ip'] = 'l2') -> None:
        """Streaming build hook (the shape of benchmarks/streaming_sweep.py:153-185): adopt
        (N, M) uint8 PQ codes that were encoded batch by batch on the device by the already
        fitted quantizer, without a host round trip. Not part of the reference surface."""
        if self._pq_codec() is None:
            raise RuntimeError("attach_device_codes needs a fitted ProductQuantizer / OPQ")
        self._metric = metric
        self._N, self._D = int(codes_dev.shape[0]), int(D)
        self._codes = None
        self._codes_dev = codes_dev
        self._recon_dev = None
        self._codes_rot = None
        self._erq_scanner = None

    def search(self, Q: np.ndarray, k: int) -> np.ndarray:
        ids, _ = self.search_with_scores(Q, k)
        return ids

    def _pq_codec(self):
        from haag_vq.methods.optimized_product_quantization import OptimizedProductQuantizer
        from haag_vq.methods.product_quantization import ProductQuantizer
        q = self._quantizer
        if isinstance(q, (ProductQuantizer, OptimizedProductQuantizer)) and q.pq is not None:
            return q.pq.codec
        return None

    def _reconstructions(self) -> torch.Tensor:
        """(N, D) fp32 reconstructions resident on the device (decoded once per fit/load)."""
        if self._recon_dev is None:
            self._recon_dev = host.to_device(
                np.ascontiguousarray(self._quantizer.decompress(self._codes), dtype=np.float32))
        return self._recon_dev

    def search_with_scores(self, Q: np.ndarray, k: int) -> Tuple[np.ndarray, np.ndarray]:
        Q = np.ascontiguousarray(Q, dtype=np.float32)
        nq = Q.shape[0]
        k = min(k, self._N)
        if k <= 0:
            return np.empty((nq, 0), dtype=np.uint32), np.empty((nq, 0), dtype=np.float32)
        if Q.shape[1] != self._D:
            raise ValueError(f"query dimension {Q.shape[1]} != index dimension {self._D}")
        q_dev = host.to_device(Q)
        codec = self._pq_codec()
        if k > MAX_K:
            # one scan launch ranks at most 128 rows per query: beyond that the reference's own route (decompress,
            # exact distance to the reconstructions) with the chained exact search of ops.flat_topk
            if self._codes is None:
                raise ValueError(f"k={k} exceeds the device top-k limit ({MAX_K}) of an index built from device codes")
            vals, ids = ops.flat_topk(self._reconstructions(), q_dev, k, self._metric)
        elif codec is not None:
            if self._codes_dev is None:
                # stored codes -> one byte per sub-quantizer index (sub-byte PQ codes are a packed bit stream)
                self._codes_dev = self._quantizer.pq.unpack_codes(
                    host.to_device(np.ascontiguousarray(self._codes, dtype=np.uint8)))
            if codec.ksub == 256 and codec.M <= 96 and (codec.M % 32 == 0 or (codec.M % 16 == 0 and self._N >= 65536)):
                # conflict-free scan layout (b200vq_pq_codes_to_scan_layout), built once per fit/load; M = 16 has its
                # own half-group layout, other M are padded to the group width of 32 with all-zero tables (same
                # scores bit for bit; M % 16 keeps the row-layout kernel's summation order; below 65536 rows the
                # screen does not apply and the row layout is as fast)
                m_pad = scan_layout_width(codec.M)
                if self._codes_rot is None:
                    self._codes_rot = ops.pq_codes_to_scan_layout(codec.pad_codes_for_scan(self._codes_dev))
                vals, ids = codec.search(self._codes_rot, q_dev, k, self._metric, n=self._N, rot=True, m_pad=m_pad)
            else:
                vals, ids = codec.search(self._codes_dev, q_dev, k, self._metric)
        elif self._erabitq_ok():
            # extended RaBitQ: |q - x^|^2 = |q - c|^2 + a^2 sum l^2 - 2 a <l, (q - c) P> (P orthonormal), scored from
            # the packed codes on the tensor cores (csrc/est_tc.cu) - the reconstructions are never materialised
            if self._erq_scanner is None:
                self._erq_scanner = self._quantizer.device_scanner(self._codes)
            qrot, qconst = self._quantizer.rotate_queries(q_dev, self._metric)
            vals, ids = self._erq_scanner.search(qrot, qconst, k, self._metric)
        else:
            vals, ids = ops.flat_topk(self._reconstructions(), q_dev, k, self._metric)
        return host.to_host(ids).view(np.uint32), host.to_host(vals)

    def _erabitq_ok(self) -> bool:
        from haag_vq.methods.extended_rabitq import ExtendedRaBitQuantizer
        q = self._quantizer
        return isinstance(q, ExtendedRaBitQuantizer) and q.P is not None and q.D % 32 == 0 and self._codes is not None

    def memory_footprint(self) -> int:
        if self._codes is not None:
            return int(self._codes.nbytes)
        return int(self._codes_dev.numel()) if self._codes_dev is not None else 0

    def reconstruction_mse(self, X: np.ndarray, sample_ids: Optional[np.ndarray] = None) -> Optional[float]:
        if self._codes is None:
            return None
        if sample_ids is None:
            sample_ids = np.arange(self._N)
        X_hat = self._quantizer.decompress(self._codes[sample_ids]).astype(np.float32)
        return float(np.mean((X[sample_ids].astype(np.float32) - X_hat) ** 2))

    def save(self, path: str | Path) -> None:
        """Pickle of {'quantizer_state', 'codes', 'metric', 'N', 'D'} — the reference's layout;
        a ProductQuantizer is stored as {'type','M','B','flat_centroids','D'}."""
        path = Path(path)
        path.parent.mkdir(parents=True, exist_ok=True)
        from haag_vq.methods.product_quantization import ProductQuantizer as PQ
        if isinstance(self._quantizer, PQ) and self._quantizer.pq is not None:
            quantizer_state = self._quantizer._state()
        else:
            quantizer_state = {'type': 'other', 'obj': self._quantizer}
        codes = self._codes
        if codes is None and self._codes_dev is not None:      # attach_device_codes: bring the codes back for the file
            codes = host.to_host(self._codes_dev)
        with open(path, 'wb') as f:
            pickle.dump({'quantizer_state': quantizer_state, 'codes': codes, 'metric': self._metric,
                         'N': self._N, 'D': self._D}, f)

    def load(self, path: str | Path) -> None:
        with open(path, 'rb') as f:
            state = pickle.load(f)
        qs = state['quantizer_state']
        if qs['type'] == 'ProductQuantizer':
            from haag_vq.methods.product_quantization import ProductQuantizer as PQ
            self._quantizer = PQ._from_state(qs)
        else:
            self._quantizer = qs['obj']
        self._codes = state['codes']
        self._metric = state['metric']
        self._N = state['N']
        self._D = state['D']
        self._codes_dev = None
        self._recon_dev = None
        self._codes_rot = None
        self._erq_scanner = None
