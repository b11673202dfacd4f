"""Registry for the SAQ-study methods outside the faiss family.

Mirrors src/haag_vq/benchmarks/method_registry_saq.py:17-74 for the methods on
the hot path: ``rabitq`` -> ExtendedRaBitQuantizer(num_bits=round(bpd)) and
``saq_paper`` -> SaqEngineAdapter(CAQ, DP allocation) (SURVEY.md §8 rows a9-a11,
a16), ``lvq`` -> LVQQuantizer(num_bits=round(bpd)) (reference :49-52), and the per-dimension family
``rankaware`` (the proposed method: var^0.5-weighted greedy) / ``perdim_mse`` (alpha = 0) ->
RankAwareQuantizer(packing="ffd", codebook="lloyd") (reference :58-65), whose data-fit codebooks come from
b200vq_codebook_lloyd, and ``ours`` / ``ours_exact`` -> SaqEngineAdapter(greedy=True, derive_codebooks=True[,
exact_codebooks=True]) (reference :34-47; see the adapter for what the codebook flags do in the CPU engine).
Still not built: ``rankaware_exact`` / ``perdim_mse_exact`` (optimal-DP 1-D codebooks, build_codebook_exact); they
raise NotImplementedError instead of silently falling back.
"""
from __future__ import annotations

SAQ_METHODS = ("saq_paper", "ours", "ours_exact", "rabitq", "lvq",
               "rankaware", "perdim_mse", "rankaware_exact", "perdim_mse_exact")


def build_saq_quantizer(method: str, bpd: float, D: int):
    b = int(round(bpd))
    if method == "saq_paper":
        import os
        from haag_vq.benchmarks.quantizer_adapters import SaqEngineAdapter
        # VQ_SAQ_APPLY_PCA=0 skips the internal PCA for already-rotated inputs (reference :24-27)
        return SaqEngineAdapter(quant_type="CAQ", avg_bits=bpd, greedy=False, derive_codebooks=False,
                                apply_pca=os.environ.get("VQ_SAQ_APPLY_PCA", "1") != "0")
    if method in ("ours", "ours_exact"):
        import os
        from haag_vq.benchmarks.quantizer_adapters import SaqEngineAdapter
        return SaqEngineAdapter(quant_type="CAQ", avg_bits=bpd, greedy=True, derive_codebooks=True,
                                exact_codebooks=(method == "ours_exact"),
                                apply_pca=os.environ.get("VQ_SAQ_APPLY_PCA", "1") != "0")
    if method == "rabitq":
        from haag_vq.benchmarks.quantizer_adapters import FaissQuantizerAdapter
        from haag_vq.methods.extended_rabitq import ExtendedRaBitQuantizer
        return FaissQuantizerAdapter(ExtendedRaBitQuantizer(num_bits=b))
    if method == "lvq":
        from haag_vq.benchmarks.quantizer_adapters import FaissQuantizerAdapter
        from haag_vq.methods.lvq_quantization import LVQQuantizer
        return FaissQuantizerAdapter(LVQQuantizer(num_bits=b))
    if method in ("rankaware", "perdim_mse"):
        from haag_vq.benchmarks.quantizer_adapters import FaissQuantizerAdapter
        from haag_vq.methods.rank_aware_quantization import RankAwareQuantizer
        return FaissQuantizerAdapter(RankAwareQuantizer(avg_bits=bpd, alpha=0.5 if method == "rankaware" else 0.0,
                                                        packing="ffd", codebook="lloyd"))
    if method in SAQ_METHODS:
        raise NotImplementedError(
            f"method {method!r} is not built on the B200 path yet (SURVEY.md §8f); no fallback is provided")
    raise ValueError(f"Unknown SAQ-study method: {method!r}")
