"""IvfQuantizedIndex on the B200: any BaseQuantizer inside a k-means IVF shell.

Mirrors src/haag_vq/methods/search/ivf_quantized_index.py:15-259 (same constructor, fit / search /
search_with_scores / memory_footprint / reconstruction_mse / save / load, same pickle layout).

Fit: coarse k-means on the device (b200vq.ivf.kmeans: the faiss.Kmeans algorithm the reference calls, restated;
tensor-core assignment), every row assigned to its centroid, one residual quantizer per non-empty cluster from
``quantizer_factory`` (fit + compress of ``X[mask] - centroid``, as the reference does). The reconstructions
``decompress(codes_c) + centroid_c`` are then kept resident in HBM in cluster order.
Search: the reference decompresses the probed clusters again for every query in a Python loop; here the probe lists
come from the exact flat search over the centroids and ONE cluster-major device pass (b200vq_ivf_search, csrc/ivf.cu)
scores every probed row exactly against its queries and merges the top-k. Same ids and distances as the reference's
loop on the same centroids and quantizers (exact top-k over the same reconstructions; (distance, id) order).
"""
from __future__ import annotations

import pickle
from pathlib import Path
from typing import Callable, Literal, Optional, Tuple

import numpy as np
import torch

from b200vq import host, ivf, ops

from haag_vq.methods.base_quantizer import BaseQuantizer
from haag_vq.methods.base_search_index import BaseSearchIndex


class IvfQuantizedIndex(BaseSearchIndex):
    def __init__(self, quantizer_factory: Callable[[], BaseQuantizer], K: int = 4096, nprobe: int = 200) -> None:
        self._quantizer_factory = quantizer_factory
        self._K = K
        self._nprobe = nprobe
        self._centroids: Optional[np.ndarray] = None
        self._cluster_quantizers: list[Optional[BaseQuantizer]] = []
        self._cluster_codes: list[np.ndarray] = []
        self._cluster_ids: list[np.ndarray] = []
        self._metric: Literal['l2', 'ip'] = 'l2'
        self._N = 0
        self._D = 0
        self._dev = None      # resident search state (centroids, reconstructions, norms, ids, offsets)

    # ------------------------------------------------------------------ fit
    def fit(self, X: np.ndarray, metric: Literal['l2', 'ip'] = 'l2', centroids: Optional[np.ndarray] = None) -> None:
        """``centroids`` (K, D), optional: skip the k-means and use these (tests pin everything that follows the
        clustering against the oracle this way; the reference always trains)."""
        X = np.ascontiguousarray(X, dtype=np.float32)
        self._metric = metric
        self._N, self._D = X.shape
        spherical = (metric == 'ip')
        xd = host.to_device(X)
        if centroids is None:
            cd = ivf.kmeans(xd, self._K, seed=0, spherical=spherical)
        else:
            cd = host.to_device(np.ascontiguousarray(centroids, dtype=np.float32))
        self._centroids = host.to_host(cd).copy()
        assignments = host.to_host(ivf.CoarseAssigner(cd, spherical)(xd)).astype(np.int32)
        del xd

        order = np.argsort(assignments, kind='stable')
        counts = np.bincount(assignments, minlength=self._K)
        offs = np.concatenate(([0], np.cumsum(counts)))
        self._cluster_quantizers, self._cluster_codes, self._cluster_ids = [], [], []
        for c in range(self._K):
            vid = order[offs[c]:offs[c + 1]].astype(np.uint32)      # ascending ids, like np.where(mask)
            self._cluster_ids.append(vid)
            if vid.size == 0:
                self._cluster_quantizers.append(None)
                self._cluster_codes.append(np.array([], dtype=np.uint8))
                continue
            residuals = X[vid] - self._centroids[c]
            q = self._quantizer_factory()
            q.fit(residuals)
            self._cluster_quantizers.append(q)
            self._cluster_codes.append(q.compress(residuals))
        self._build_device_state()

    def _build_device_state(self) -> None:
        """Reconstructions of all rows in cluster order, resident in HBM (what the reference recomputes per query)."""
        dev = host.device()
        N, D, K = self._N, self._D, self._K
        recon = torch.empty((N, D), dtype=torch.float32, device=dev)
        gid = np.empty(N, dtype=np.uint32)
        row_off = np.zeros(K + 1, dtype=np.int64)
        rowc = np.empty(N, dtype=np.uint32)
        pos = 0
        for c in range(K):
            vid = self._cluster_ids[c]
            q = self._cluster_quantizers[c]
            if q is not None and vid.size:
                r = q.decompress(self._cluster_codes[c]).astype(np.float32) + self._centroids[c]
                recon[pos:pos + vid.size] = host.to_device(np.ascontiguousarray(r))
                gid[pos:pos + vid.size] = vid
                rowc[pos:pos + vid.size] = c
                pos += vid.size
            row_off[c + 1] = pos
        recon = recon[:pos]
        self._dev = {
            'centroids': host.to_device(self._centroids),
            'recon': recon,
            'rnorm': ops.row_sqnorms(recon) if pos else torch.empty(0, dtype=torch.float32, device=dev),
            'gid': host.to_device(gid[:pos]),
            'row_off': torch.from_numpy(row_off).to(dev),
            'rowcluster': host.to_device(rowc[:pos]),
        }

    # --------------------------------------------------------------- search
    def _nearest_centroids(self, Q: np.ndarray) -> np.ndarray:
        """(nq, nprobe) centroid indices sorted by distance."""
        probe = self._probe(host.to_device(np.ascontiguousarray(Q, dtype=np.float32)))
        return host.to_host(probe).astype(np.int64)

    def _probe(self, qd: torch.Tensor) -> torch.Tensor:
        _, ids = ops.flat_topk(self._dev['centroids'], qd, min(self._nprobe, self._K), self._metric)
        return ids

    def search(self, Q: np.ndarray, k: int) -> np.ndarray:
        ids, _ = self.search_with_scores(Q, k)
        return ids

    def search_with_scores(self, Q: np.ndarray, k: int) -> Tuple[np.ndarray, np.ndarray]:
        """Any k: one device pass ranks 128 rows per query, larger k are pages chained by b200vq_ivf_search_after."""
        Q = np.ascontiguousarray(Q, dtype=np.float32)
        nq = Q.shape[0]
        if k <= 0:
            return np.empty((nq, 0), dtype=np.uint32), np.empty((nq, 0), dtype=np.float32)
        if self._dev is None:
            raise RuntimeError("IvfQuantizedIndex must be fitted or loaded before search().")
        fmax = np.finfo(np.float32).max
        dist_sentinel = -fmax if self._metric == 'ip' else fmax
        if nq == 0 or self._dev['recon'].shape[0] == 0:
            return (np.full((nq, k), np.iinfo(np.uint32).max, dtype=np.uint32),
                    np.full((nq, k), dist_sentinel, dtype=np.float32))
        qd = host.to_device(Q)
        d = self._dev
        vals, ids = ivf.ivf_search(d['recon'], d['rnorm'], d['gid'], d['row_off'], self._K, qd, self._probe(qd), k,
                                   self._metric, rowcluster=d['rowcluster'])
        ids_h = host.to_host(ids).view(np.uint32)
        dists = host.to_host(vals)
        if self._metric == 'ip':
            dists = -dists          # the reference ranks and returns -(q . x)
        dists[ids_h == np.iinfo(np.uint32).max] = dist_sentinel
        return ids_h, dists.astype(np.float32)

    # ------------------------------------------------------------- accounting
    def memory_footprint(self) -> int:
        centroid_bytes = self._centroids.nbytes if self._centroids is not None else 0
        return centroid_bytes + sum(c.nbytes for c in self._cluster_codes)

    def reconstruction_mse(self, X: np.ndarray, sample_ids: Optional[np.ndarray] = None) -> Optional[float]:
        if self._centroids is None or self._dev is None:
            return None
        if sample_ids is None:
            sample_ids = np.arange(self._N)
        gid = host.to_host(self._dev['gid']).view(np.uint32)
        where = np.full(self._N, -1, dtype=np.int64)
        where[gid] = np.arange(gid.size)
        sample_ids = np.asarray(sample_ids).astype(np.int64)
        loc = where[sample_ids]
        keep = loc >= 0
        if not keep.any():
            return None
        rows = host.to_host(self._dev['recon'][torch.from_numpy(loc[keep]).to(self._dev['recon'].device)])
        x = np.asarray(X)[sample_ids[keep]].astype(np.float32)
        per_row = np.mean((x - rows) ** 2, axis=1, dtype=np.float32).astype(np.float64)
        return float(per_row.sum() / keep.sum())

    # ------------------------------------------------------------ persistence
    def save(self, path: str | Path) -> None:
        """The reference's pickle layout; the quantizer_factory is not stored."""
        from haag_vq.methods.product_quantization import ProductQuantizer as PQ

        def ser(q):
            if q is None:
                return {'type': 'none'}
            if isinstance(q, PQ) and q.pq is not None:
                return q._state()
            return {'type': 'other', 'obj': q}

        path = Path(path)
        path.parent.mkdir(parents=True, exist_ok=True)
        with open(path, 'wb') as f:
            pickle.dump({
                'K': self._K, 'nprobe': self._nprobe, 'centroids': self._centroids,
                'serialised_quantizers': [ser(q) for q in self._cluster_quantizers],
                'cluster_codes': self._cluster_codes, 'cluster_ids': self._cluster_ids,
                'metric': self._metric, 'N': self._N, 'D': self._D,
            }, f)

    def load(self, path: str | Path) -> None:
        from haag_vq.methods.product_quantization import ProductQuantizer as PQ

        def de(d):
            if d['type'] == 'none':
                return None
            if d['type'] == 'ProductQuantizer':
                return PQ._from_state(d)
            return d['obj']

        with open(path, 'rb') as f:
            state = pickle.load(f)
        self._K = state['K']
        self._nprobe = state['nprobe']
        self._centroids = state['centroids']
        self._cluster_quantizers = [de(d) for d in state['serialised_quantizers']]
        self._cluster_codes = state['cluster_codes']
        self._cluster_ids = state['cluster_ids']
        self._metric = state['metric']
        self._N = state['N']
        self._D = state['D']
        self._build_device_state()
