"""b200vq — B200-native quantize-and-search kernels behind the haag_vq surface.

Layers:
  _lib      ctypes binding of libb200vq.so (C ABI: include/b200vq.h)
  ops*      torch-CUDA-tensor wrappers around the C ABI (no arithmetic in Python)
  pq        device-resident PQ / OPQ codec (train, encode, decode, ADC search)
  host      numpy <-> device streaming used by the haag_vq drop-in classes
  dist      row-sharded index over torch.distributed (NCCL) with top-k merge

There is no CPU fallback: a missing library or a missing GPU raises.
"""
from ._lib import LIB_PATH, METRIC_IP, METRIC_L2, SIGNATURES, B200VQError, load  # noqa: F401
