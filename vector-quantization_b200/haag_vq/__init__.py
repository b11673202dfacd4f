"""haag_vq drop-in: the reference's method-registry / quantizer-adapter surface
(same module paths, class names, argument meaning and error behaviour) hosted on
the B200 kernels of libb200vq.so. numpy in / numpy out; no faiss, no CPU fallback.

Mirrors src/haag_vq of Human-Augment-Analytics/vector-quantization for the
quantize-and-search hot path only; everything else (CLI, data
loaders, plotting, SQLite logging, IVF) is out of scope.
"""
__version__ = "0.3.0+b200"
