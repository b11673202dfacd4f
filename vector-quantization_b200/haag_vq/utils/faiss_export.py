"""fvecs / ivecs I/O, codebook export and codebook queries on the B200 kernels.

Mirrors src/haag_vq/utils/faiss_export.py:46-133 (write_fvecs / write_ivecs / load_fvecs /
load_ivecs), :150-177 (_extract_codebook), :304-347 (export_codebook) and :352-506
(query_codebook). The reference builds faiss.IndexFlatL2 / IndexFlatIP objects over the codebook
(one per PQ chunk) and calls ``index.search``; here the same exact top-k is one call of the flat
top-k kernel per chunk (``b200vq_flat_topk``), so distances are squared L2 ascending or inner
products descending and indices are int64 with the chunk offset added, as faiss returns them.
File formats are byte-identical to the reference's writers: every record is an int32 dimension
followed by ``dim`` float32 / int32 values.
"""
from __future__ import annotations

import os
from pathlib import Path
from typing import Optional, Tuple, Union

import numpy as np

PathLike = Union[str, Path]

METRIC_INNER_PRODUCT = 0   # faiss.METRIC_INNER_PRODUCT
METRIC_L2 = 1              # faiss.METRIC_L2


def _type_name(model) -> str:
    return type(model).__name__


def _is_pq_like(model) -> bool:
    return _type_name(model) == "ProductQuantizer"


def _is_opq(model) -> bool:
    return _type_name(model) == "OptimizedProductQuantizer"


def _is_scalar(model) -> bool:
    return _type_name(model) == "ScalarQuantizer"


def _is_rabit(model) -> bool:
    return _type_name(model) in ("RaBitQuantizer", "ExtendedRaBitQuantizer")


def _is_saq(model) -> bool:
    return _type_name(model) == "SAQ"


def _write_vecs(path: PathLike, vectors: np.ndarray, dtype, what: str) -> Path:
    path = Path(path)
    vectors = np.ascontiguousarray(vectors, dtype=dtype)
    if vectors.ndim != 2:
        raise ValueError(f"{what} expects a 2D array")
    n, d = vectors.shape
    rec = np.empty((n, d + 1), dtype=np.int32)
    rec[:, 0] = d
    rec[:, 1:] = vectors.view(np.int32) if dtype == np.float32 else vectors
    with path.open("wb") as fh:
        fh.write(rec.tobytes())
    try:
        os.chmod(path, 0o666)
    except (OSError, PermissionError):
        pass
    return path


def write_fvecs(path: PathLike, vectors: np.ndarray) -> Path:
    """Write float vectors to a .fvecs file (faiss_export.py:46-66)."""
    return _write_vecs(path, vectors, np.float32, "fvecs")


def write_ivecs(path: PathLike, vectors: np.ndarray) -> Path:
    """Write int vectors to a .ivecs file (faiss_export.py:69-89)."""
    return _write_vecs(path, vectors, np.int32, "ivecs")


def _load_vec_file(path: PathLike, value_dtype) -> np.ndarray:
    path = Path(path)
    if not path.exists():
        raise FileNotFoundError(path)
    raw = path.read_bytes()
    if not raw:
        return np.empty((0, 0), dtype=value_dtype)
    ints = np.frombuffer(raw, dtype=np.int32)
    dim = int(ints[0])
    record = dim + 1
    if dim <= 0:
        raise ValueError(f"Invalid vector dimension ({dim}) in {path}")
    if ints.size % record != 0:
        raise ValueError(f"Corrupt vector file: {path}")
    nvecs = ints.size // record
    table = ints.reshape(nvecs, record)
    if not np.all(table[:, 0] == dim):
        raise ValueError(f"Non-uniform dimensions in {path}")
    data = table[:, 1:]
    if value_dtype == np.float32:
        data = data.view(np.float32)
    elif value_dtype != np.int32:
        raise TypeError(f"Unsupported dtype: {value_dtype}")
    return np.array(data, copy=True)


def load_fvecs(path: PathLike) -> np.ndarray:
    return _load_vec_file(path, np.float32)


def load_ivecs(path: PathLike) -> np.ndarray:
    return _load_vec_file(path, np.int32)


def _extract_codebook(model) -> np.ndarray:
    """2D array of codebook vectors for fvecs export (faiss_export.py:150-177)."""
    if _is_pq_like(model) and hasattr(model, "codebooks"):
        if not model.codebooks:
            raise ValueError("ProductQuantizer has no trained codebooks")
        return np.concatenate([np.asarray(cb, dtype=np.float32) for cb in model.codebooks], axis=0)
    if _is_opq(model):
        if model.pq is None:
            raise ValueError("OptimizedProductQuantizer must be fitted before export")
        flat = np.asarray(model.pq.centroids, dtype=np.float32)
        return flat.reshape(int(model.M) * int(model.pq.ksub), int(model.pq.dsub)).copy()
    if _is_scalar(model):
        if model.min is None or model.max is None:
            raise ValueError("ScalarQuantizer must be fitted before export")
        return np.stack([model.min, model.max]).astype(np.float32)
    if _is_rabit(model) or _is_saq(model):
        raise RuntimeError("This quantizer does not expose a static codebook for export")
    raise TypeError(f"Unsupported quantizer type: {type(model)!r}")


def export_codebook(model, output_dir: PathLike, *, index_key: Optional[str] = None,
                    codes: Optional[np.ndarray] = None, codebook_filename: str = "codebook.fvecs",
                    codes_filename: str = "codes.ivecs") -> dict:
    """Persist a quantizer's codebook (and optional codes) as fvecs / ivecs (faiss_export.py:304-347).

    The reference additionally returns a faiss index object under "index" for non-PQ models; there
    is no faiss here, the codebook matrix itself ("codebook_vectors") is what query_codebook searches."""
    output_path = Path(output_dir)
    output_path.mkdir(parents=True, exist_ok=True)
    try:
        os.chmod(output_path, 0o777)
    except (OSError, PermissionError):
        pass
    codebook = _extract_codebook(model)
    result = {"codebook": write_fvecs(output_path / codebook_filename, codebook), "codebook_vectors": codebook}
    if codes is not None:
        result["codes"] = write_ivecs(output_path / codes_filename, np.asarray(codes, dtype=np.int32))
    return result


def _flat_search(base: np.ndarray, queries: np.ndarray, k: int, metric: int) -> Tuple[np.ndarray, np.ndarray]:
    """faiss.IndexFlatL2 / IndexFlatIP .search on the device: (distances fp32, indices int64)."""
    from b200vq import host, ops
    vals, ids = ops.flat_topk(host.to_device(np.ascontiguousarray(base, dtype=np.float32)),
                              host.to_device(np.ascontiguousarray(queries, dtype=np.float32)), k,
                              "l2" if metric == METRIC_L2 else "ip", impl="simt")
    return host.to_host(vals), host.to_host(ids).view(np.uint32).astype(np.int64)


def _query_product_codebook(queries, model, codebook_vectors, topk, metric):
    """Per-chunk top-k over the PQ / OPQ centroids (faiss_export.py:352-414)."""
    if _is_pq_like(model) and getattr(model, "chunk_dim", None) is not None:
        chunk_dim = int(model.chunk_dim)
        num_chunks = int(getattr(model, "M", getattr(model, "num_chunks")))
        num_clusters = int(2 ** int(model.B))
        transformed = queries
    elif _is_opq(model):
        if model.pq is None or model.opq is None:
            raise ValueError("OptimizedProductQuantizer must be fitted before querying")
        chunk_dim, num_chunks, num_clusters = int(model.pq.dsub), int(model.M), int(model.pq.ksub)
        transformed = model.opq.apply(queries)
    else:
        raise ValueError("ProductQuantizer has no chunk_dim; call fit first")
    expected_rows = num_chunks * num_clusters
    if codebook_vectors.shape != (expected_rows, chunk_dim):
        raise ValueError("ProductQuantizer codebook has unexpected shape; expected "
                         f"({expected_rows}, {chunk_dim}) but received {codebook_vectors.shape}")
    if transformed.shape[1] != chunk_dim * num_chunks:
        raise ValueError("Query dimensionality does not match ProductQuantizer. "
                         f"Expected {chunk_dim * num_chunks} but received {transformed.shape[1]}")
    per_chunk_topk = min(topk, num_clusters)
    if per_chunk_topk <= 0:
        raise ValueError("topk must be positive and <= number of clusters")
    if metric not in (METRIC_L2, METRIC_INNER_PRODUCT):
        raise ValueError("ProductQuantizer queries currently support only METRIC_L2 and METRIC_INNER_PRODUCT")
    qc = transformed.reshape(transformed.shape[0], num_chunks, chunk_dim)
    cb = codebook_vectors.reshape(num_chunks, num_clusters, chunk_dim)
    dist, idx = [], []
    for m in range(num_chunks):
        d, i = _flat_search(cb[m], qc[:, m, :], per_chunk_topk, metric)
        dist.append(d)
        idx.append(i + m * num_clusters)
    return np.concatenate(dist, axis=1), np.concatenate(idx, axis=1)


def query_codebook(queries: np.ndarray, *, model=None, codebook_vectors: Optional[np.ndarray] = None,
                   codebook_path: Optional[PathLike] = None, topk: int = 1, metric: int = METRIC_L2,
                   index_key: Optional[str] = None) -> Tuple[np.ndarray, np.ndarray]:
    """Closest codebook entries to ``queries`` (faiss_export.py:417-506). For PQ / OPQ models topk
    applies per chunk and the outputs are concatenated across chunks."""
    if codebook_vectors is None:
        if codebook_path is None:
            raise ValueError("Provide either codebook_vectors or codebook_path")
        codebook_vectors = load_fvecs(codebook_path)
    codebook_vectors = np.ascontiguousarray(codebook_vectors, dtype=np.float32)
    if codebook_vectors.ndim != 2:
        raise ValueError("Codebook vectors must be 2D")
    queries = np.asarray(queries, dtype=np.float32)
    if queries.ndim == 1:
        queries = queries.reshape(1, -1)
    if queries.ndim != 2:
        raise ValueError("Queries must be a 1D or 2D array")
    if queries.size == 0:
        raise ValueError("No queries provided for search")
    if topk <= 0:
        raise ValueError("topk must be positive")
    topk = int(topk)
    if model is not None and (_is_pq_like(model) or _is_opq(model)):
        return _query_product_codebook(queries, model, codebook_vectors, topk, metric)
    num_entries, dim = codebook_vectors.shape
    if num_entries == 0:
        raise ValueError("Codebook is empty; cannot run queries")
    if queries.shape[1] != dim:
        raise ValueError(f"Query dimensionality ({queries.shape[1]}) does not match codebook ({dim})")
    if metric not in (METRIC_L2, METRIC_INNER_PRODUCT):
        raise ValueError("Provide a quantizer model to build complex FAISS indexes")
    return _flat_search(codebook_vectors, queries, min(topk, num_entries), metric)
