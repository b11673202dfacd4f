"""Quantizer contract (mirrors src/haag_vq/methods/base_quantizer.py:8-51).

The reference's ``save_codebooks`` export hook (fvecs/ivecs files) is an on-disk
format outside the hot path and is not provided.
"""
from abc import ABC, abstractmethod

import numpy as np


class BaseQuantizer(ABC):
    @abstractmethod
    def fit(self, X: np.ndarray) -> None:
        """Learn quantization parameters from training data of shape (N, D)."""

    @abstractmethod
    def compress(self, X: np.ndarray) -> np.ndarray:
        """Quantize (N, D) data into row-sliceable codes."""

    @abstractmethod
    def decompress(self, codes: np.ndarray) -> np.ndarray:
        """Decode codes back to approximate (N, D) vectors."""
