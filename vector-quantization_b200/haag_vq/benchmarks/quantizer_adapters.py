"""Quantizer protocol + adapter for the benchmark study.

Mirrors src/haag_vq/benchmarks/quantizer_adapters.py:1-59: fit compresses the
whole database once and caches the codes; reconstruct(ids) decodes
``codes[ids]``; code_bytes = stored code bytes + 4 bytes/vector for the exact-norm
side channel. Errors: RuntimeError before fit, ValueError on negative ids; a
failing compress leaves the previous state intact.
"""
from __future__ import annotations

from typing import Optional, Protocol, runtime_checkable

import numpy as np

from haag_vq.methods.base_quantizer import BaseQuantizer

NORM_SIDECHANNEL_BYTES = 4  # float32 exact norm per vector


@runtime_checkable
class Quantizer(Protocol):
    def fit(self, X: np.ndarray) -> None: ...
    def reconstruct(self, ids: np.ndarray) -> np.ndarray: ...
    def code_bytes(self) -> int: ...


class FaissQuantizerAdapter:
    def __init__(self, quantizer: BaseQuantizer) -> None:
        self._q = quantizer
        self._codes: Optional[np.ndarray] = None
        self._n = 0

    def fit(self, X: np.ndarray) -> None:
        X = np.ascontiguousarray(X, dtype=np.float32)
        self._q.fit(X)
        codes = self._q.compress(X)   # may raise; leave prior state intact
        self._codes = codes
        self._n = X.shape[0]

    def reconstruct(self, ids: np.ndarray) -> np.ndarray:
        if self._codes is None:
            raise RuntimeError("FaissQuantizerAdapter.reconstruct() before fit()")
        ids = np.asarray(ids, dtype=np.int64)
        if np.any(ids < 0):
            raise ValueError("reconstruct(): negative ids are not allowed")
        return np.asarray(self._q.decompress(self._codes[ids]), dtype=np.float32)

    def code_bytes(self) -> int:
        if self._codes is None:
            raise RuntimeError("FaissQuantizerAdapter.code_bytes() before fit()")
        return int(self._codes.nbytes) + self._n * NORM_SIDECHANNEL_BYTES


class SaqEngineAdapter:
    """The SAQ engine behind the Quantizer protocol: a single-cluster (K = 1) index whose
    reconstruct() is the engine's decompress(ids).

    Mirrors src/haag_vq/benchmarks/quantizer_adapters.py:62-135 (same constructor keywords, same
    errors, same code_bytes formula). The engine itself (`saq.IVF(n, D, K, cfg).fit / .decompress`)
    is b200vq.saq_engine.SaqEngine on the device: quant_type "CAQ" with the DP allocator (method
    ``saq_paper``) or the greedy allocator on the per-dimension Lloyd MSE table (``greedy=True``, methods ``ours`` /
    ``ours_exact``: external/saq/include/saq/quantization_plan.h:246-266, src/bit_allocator_greedy.cpp:11-116).
    ``derive_codebooks`` / ``exact_codebooks``: the reference's CPU engine (the `saq.IVF` this adapter drives) builds
    per-dimension codebooks in construct() (src/ivf_index.cpp:50-121) but only its CUDA construct path consumes
    them (segment_codebooks is read in src/gpu/gpu_ivf_construct.cpp alone); the CPU encoder stays CAQ's uniform
    grid, so the flags change no code and no reconstruction there, and none here.
    """

    def __init__(self, *, quant_type: str = "CAQ", avg_bits: float = 4.0, greedy: bool = False,
                 derive_codebooks: bool = False, exact_codebooks: bool = False, apply_pca: bool = True,
                 K: int = 1, num_threads: int = 8, seed: int = 0, max_bits: int = 13) -> None:
        if quant_type != "CAQ":
            raise ValueError(f"SaqEngineAdapter: unsupported quant_type {quant_type!r} (CAQ only)")
        self._greedy = bool(greedy)
        self._derive_codebooks, self._exact_codebooks = bool(derive_codebooks), bool(exact_codebooks)
        if K != 1:
            raise NotImplementedError("SaqEngineAdapter on the B200 path is the K = 1 (flat) configuration")
        self._avg_bits = float(avg_bits)
        self._apply_pca = apply_pca
        self._K, self._num_threads, self._seed, self._max_bits = K, num_threads, seed, max_bits
        self._index = None
        self._n = 0
        self._D = 0

    def fit(self, X: np.ndarray) -> None:
        import torch
        from b200vq.saq_engine import SaqEngine
        X = np.ascontiguousarray(X, dtype=np.float32)
        self._n, self._D = X.shape
        eng = SaqEngine(self._avg_bits, apply_pca=self._apply_pca, seed=self._seed,
                        allocator="greedy" if self._greedy else "dp")
        eng.fit(torch.from_numpy(X).cuda())
        self._index = eng

    def reconstruct(self, ids: np.ndarray) -> np.ndarray:
        if self._index is None:
            raise RuntimeError("SaqEngineAdapter.reconstruct() before fit()")
        import torch
        ids = np.ascontiguousarray(ids, dtype=np.uint32)
        out = self._index.reconstruct(torch.from_numpy(ids.astype(np.int64)).cuda())
        return np.asarray(out.cpu().numpy(), dtype=np.float32)

    def code_bytes(self) -> int:
        if self._index is None:
            raise RuntimeError("SaqEngineAdapter.code_bytes() before fit()")
        code = int(self._n * self._D * self._avg_bits / 8.0)
        return code + self._n * NORM_SIDECHANNEL_BYTES
