from .base_search_index import BaseSearchIndex
from .search import FlatQuantizedIndex, RaBitQIndex
