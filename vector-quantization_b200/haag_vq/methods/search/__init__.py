"""Search indexes on the hot path (flat scan). The reference's IVF shells
(IvfQuantizedIndex, FaissIvfPqIndex, RaBitQIVFIndex, SaqIndex) are out of scope
(IVF is not on the north-star path)."""
from .flat_quantized_index import FlatQuantizedIndex
from .rabitq_index import RaBitQIndex

__all__ = ["FlatQuantizedIndex", "RaBitQIndex"]
