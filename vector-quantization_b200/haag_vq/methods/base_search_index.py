"""Search-index contract (mirrors src/haag_vq/methods/base_search_index.py:21-89)."""
from __future__ import annotations

from abc import ABC, abstractmethod
from pathlib import Path
from typing import Literal, Optional, Tuple

import numpy as np


class BaseSearchIndex(ABC):
    @abstractmethod
    def fit(self, X: np.ndarray, metric: Literal['l2', 'ip'] = 'l2') -> None:
        """Learn the index from (N, D) vectors."""

    @abstractmethod
    def search(self, Q: np.ndarray, k: int) -> np.ndarray:
        """(nq, k) uint32 ids of the approximate nearest neighbours."""

    @abstractmethod
    def search_with_scores(self, Q: np.ndarray, k: int) -> Tuple[np.ndarray, np.ndarray]:
        """(ids uint32, distances float32); l2 = squared L2 ascending, ip = inner product descending."""

    @abstractmethod
    def memory_footprint(self) -> int:
        """Index memory in bytes (encoded data + auxiliary structures)."""

    @abstractmethod
    def save(self, path: str | Path) -> None:
        """Persist the index."""

    @abstractmethod
    def load(self, path: str | Path) -> None:
        """Restore the index."""

    def reconstruction_mse(self, X: np.ndarray, sample_ids: Optional[np.ndarray] = None) -> Optional[float]:
        return None
