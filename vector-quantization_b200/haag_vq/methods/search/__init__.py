"""Search indexes: the flat scans of the hot path and the k-means IVF shell around any quantizer
(IvfQuantizedIndex). The reference's wrappers of third-party IVF engines
(FaissIvfPqIndex, RaBitQIVFIndex = faiss index_factory strings; SaqIndex = the engine's own IVF) are baselines
outside the path and are not provided."""
from .flat_quantized_index import FlatQuantizedIndex
from .ivf_quantized_index import IvfQuantizedIndex
from .rabitq_index import RaBitQIndex

__all__ = ["FlatQuantizedIndex", "IvfQuantizedIndex", "RaBitQIndex"]
