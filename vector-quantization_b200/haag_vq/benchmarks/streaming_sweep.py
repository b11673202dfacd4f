"""Streaming batch compression for corpora that do not fit in memory (MS MARCO 53M shape).

Mirrors src/haag_vq/benchmarks/streaming_sweep.py:41-224: (1) train the quantizer on a subset,
(2) stream the corpus in batches, (3) compress each batch and accumulate the distortion on the fly,
(4) report compression ratio, MSE (mean row SSE, weighted by batch size), vectors and batches.
The reference pulls batches from a HuggingFace streaming dataset and logs to SQLite; both are
outside the hot path, so the corpus is any iterable of (batch, D) float32
arrays (``synthetic_batches`` gives the MS MARCO shape) and the metrics dict is returned.

Device path: for PQ / OPQ / SQ models the batch is uploaded once, encoded, decoded and compared
on the device (``b200vq_pq_sse`` / ``b200vq_sq_err_sum``): one H2D copy of the batch and one
8-byte D2H read per batch, instead of compress -> host codes -> decompress -> host subtraction.
Any other BaseQuantizer goes through compress() / compute_distortion() as the reference does.
Both paths give the same codes; the MSE agrees to fp32 rounding of the per-element differences.
"""
from __future__ import annotations

import uuid
from datetime import datetime
from typing import Callable, Dict, Iterable, Iterator, Optional, Tuple

import numpy as np

from haag_vq.methods.optimized_product_quantization import OptimizedProductQuantizer
from haag_vq.methods.product_quantization import ProductQuantizer
from haag_vq.methods.rabit_quantization import RaBitQuantizer
from haag_vq.methods.scalar_quantization import ScalarQuantizer
from haag_vq.metrics.distortion import compute_distortion
from haag_vq.utils.faiss_utils import MetricType

try:
    from haag_vq.methods.saq import SAQ
except ImportError:  # pragma: no cover
    SAQ = None


def synthetic_batches(n_total: int, D: int, batch_size: int, seed: int = 1234) -> Iterator[np.ndarray]:
    """Unit-norm anisotropic Gaussian rows (the bench's corpus recipe), generated batch by batch."""
    scale = np.linspace(4.0, 0.2, D, dtype=np.float32)
    done = 0
    while done < n_total:
        m = min(batch_size, n_total - done)
        rng = np.random.default_rng(seed + done)
        x = rng.standard_normal((m, D), dtype=np.float32) * scale
        x /= np.maximum(np.linalg.norm(x, axis=1, keepdims=True), 1e-12)
        yield x
        done += m


def build_model(method: str, *, pq_subquantizers: int = 16, pq_bits: int = 8, opq_quantizers: int = 16,
                opq_bits: int = 8, saq_num_bits: int = 4) -> Tuple[object, dict]:
    """Method switch of the reference driver (streaming_sweep.py:106-126)."""
    if method == "pq":
        return ProductQuantizer(M=int(pq_subquantizers), B=int(pq_bits)), {"M": int(pq_subquantizers), "B": int(pq_bits)}
    if method == "opq":
        return (OptimizedProductQuantizer(M=int(opq_quantizers), B=int(opq_bits)),
                {"M": int(opq_quantizers), "B": int(opq_bits)})
    if method == "sq":
        return ScalarQuantizer(), {}
    if method == "saq":
        if SAQ is None:
            raise ValueError("SAQ is not available")
        return SAQ(num_bits=int(saq_num_bits)), {"num_bits": int(saq_num_bits)}
    if method == "rabitq":
        return RaBitQuantizer(metric_type=MetricType.L2), {}
    raise ValueError(f"Unknown method: {method}")


def _device_batch_sse(model) -> Optional[Callable[[np.ndarray], Tuple[np.ndarray, float]]]:
    """batch -> (codes, sum of squared errors) with encode, decode and the error sum on the device,
    for the models whose codec is a pure device kernel; None for everything else."""
    import torch
    from b200vq import host, ops, ops_sq
    from b200vq import pq as dpq

    if isinstance(model, ProductQuantizer) and model.pq is not None:
        codec = model.pq.codec

        def pq_batch(X):
            xd = host.to_device(host.as_f32_2d(X))
            codes = codec.encode(xd)
            return host.to_host(codes), float(ops.pq_sse(xd, codes, codec.centroids).item())
        return pq_batch
    if isinstance(model, OptimizedProductQuantizer) and model.pq is not None and model.opq is not None:
        codec, A = model.pq.codec, model.opq.A_dev

        def opq_batch(X):
            # ||x - R^T c||^2 = ||R x - c||^2 for the orthonormal R: the error is taken in rotated space
            yd = dpq.rotate(host.to_device(host.as_f32_2d(X)), A)
            codes = ops.pq_assign(yd, codec.centroids, codec.cnorm)     # = codec.encode(x): rotate, then assign
            return host.to_host(codes), float(ops.pq_sse(yd, codes, codec.centroids).item())
        return opq_batch
    if isinstance(model, ScalarQuantizer) and model.min is not None:
        vmin, vmax = model._params()
        bits, D = model.num_bits, len(model.min)

        def sq_batch(X):
            xd = host.to_device(host.as_f32_2d(X))
            codes = ops_sq.sq_encode(xd, vmin, vmax, bits)
            sse = float(ops_sq.sq_err_sum(xd, ops_sq.sq_decode(codes, vmin, vmax, bits, D)).item())
            out = host.to_host(codes)
            return (out.view(np.uint16) if bits == 16 else out), sse
        return sq_batch
    return None


def stream_compress(model, batches: Iterable[np.ndarray], *, max_batches: Optional[int] = None,
                    sink: Optional[Callable[[np.ndarray, int], None]] = None, device_path: bool = True) -> Dict[str, float]:
    """Steps 3-4 of the reference driver (streaming_sweep.py:143-203) for a trained ``model``.

    sink(codes, first_row) receives every batch of codes (the reference discards them).
    Returns {"mse", "total_vectors_compressed", "num_batches"}; mse = sum over batches of
    (batch mean row SSE x batch rows) / rows, exactly the reference's weighted average."""
    fused = _device_batch_sse(model) if device_path else None
    total_sse, total, batch_count = 0.0, 0, 0
    for batch in batches:
        batch = np.asarray(batch, dtype=np.float32)
        if batch.shape[0] == 0:
            continue
        if fused is not None:
            codes, sse = fused(batch)
        else:
            codes = model.compress(batch)
            sse = compute_distortion(batch, codes, model) * batch.shape[0]
        if sink is not None:
            sink(codes, total)
        total_sse += sse
        total += batch.shape[0]
        batch_count += 1
        if max_batches and batch_count >= max_batches:
            break
    return {"mse": total_sse / total if total else 0.0, "total_vectors_compressed": total,
            "num_batches": batch_count}


def streaming_sweep(method: str = "pq", dataset: str = "synthetic-msmarco", training_size: int = 1_000_000,
                    batch_size: int = 10_000, max_batches: Optional[int] = None, *,
                    training_vectors: Optional[np.ndarray] = None, batches: Optional[Iterable[np.ndarray]] = None,
                    n_total: int = 53_000_000, D: int = 1024, pq_subquantizers: str = "16", pq_bits: str = "8",
                    opq_quantizers: str = "16", opq_bits: str = "8", saq_num_bits: str = "4",
                    seed: int = 1234) -> Dict[str, object]:
    """Train on a subset, stream-compress the corpus, return the reference's metrics dict plus the
    sweep id and config. ``training_vectors`` / ``batches`` replace the reference's HuggingFace
    stream; left out, a synthetic corpus of ``n_total`` x ``D`` unit-norm rows is generated."""
    sweep_id = f"streaming_{method}_{datetime.now().strftime('%Y%m%d_%H%M%S')}_{uuid.uuid4().hex[:8]}"
    if training_vectors is None:
        training_vectors = np.concatenate(list(synthetic_batches(min(training_size, n_total), D, 65536, seed + 7)))
    training_vectors = np.asarray(training_vectors, dtype=np.float32)[:training_size]
    model, config = build_model(method, pq_subquantizers=int(pq_subquantizers), pq_bits=int(pq_bits),
                                opq_quantizers=int(opq_quantizers), opq_bits=int(opq_bits),
                                saq_num_bits=int(saq_num_bits))
    model.fit(training_vectors)
    if batches is None:
        batches = synthetic_batches(n_total, training_vectors.shape[1], batch_size, seed)
    metrics = stream_compress(model, batches, max_batches=max_batches)
    metrics["compression_ratio"] = model.get_compression_ratio(training_vectors)
    return {"sweep_id": sweep_id, "method": method, "dataset": f"{dataset}-streaming", "config": config,
            "metrics": metrics, "model": model}
