"""CPU restatement of RankAwareQuantizer's data path, given trained state. TEST INFRASTRUCTURE ONLY.

Follows src/haag_vq/methods/rank_aware_quantization.py:229-235 (nearest level via searchsorted into the
codebook mid-points), :238-279 (float64 projection, dense MSB-first packing or FFD layout) and :283-320
(decode, back-projection, float32 cast). PINNED by oracle/make_golden_perdim.py against the reference class
(codebook="gaussian", importable with numpy only): identical codes and reconstructions for the committed
tests/golden/perdim_D24.npz.
"""
from __future__ import annotations

import numpy as np

from . import ffd as offd


def project(X, mu, V):
    return (np.asarray(X, dtype=np.float64) - mu) @ V


def indices(Y, bits, cb):
    N, D = Y.shape
    idx = np.zeros((N, D), dtype=np.int64)
    for d in range(D):
        if bits[d]:
            idx[:, d] = np.searchsorted(0.5 * (cb[d][:-1] + cb[d][1:]), Y[:, d])
    return idx


def compress(X, mu, V, bits, cb, packing="dense"):
    bits = np.asarray(bits, dtype=np.int64)
    idx = indices(project(X, mu, V), bits, cb)
    if packing == "ffd":
        bi, bo, nb = offd.ffd_layout(bits)
        return offd.ffd_encode(idx, bits, bi, bo, nb)
    off = np.concatenate(([0], np.cumsum(bits)))
    bitmat = np.zeros((idx.shape[0], int(bits.sum())), dtype=np.uint8)
    for d, b in enumerate(bits):
        if b:
            bitmat[:, off[d]:off[d] + b] = (idx[:, d, None] >> np.arange(b - 1, -1, -1)) & 1
    return np.packbits(bitmat, axis=1)


def decompress(codes, mu, V, bits, cb, packing="dense"):
    bits = np.asarray(bits, dtype=np.int64)
    codes = np.ascontiguousarray(codes, dtype=np.uint8)
    N, D = codes.shape[0], len(bits)
    if packing == "ffd":
        bi, bo, _ = offd.ffd_layout(bits)
        idx = offd.ffd_decode(codes, bits, bi, bo, D)
    else:
        off = np.concatenate(([0], np.cumsum(bits)))
        bitmat = np.unpackbits(codes, axis=1)[:, :int(bits.sum())]
        idx = np.zeros((N, D), dtype=np.int64)
        for d, b in enumerate(bits):
            if b:
                idx[:, d] = bitmat[:, off[d]:off[d] + b].astype(np.int64) @ (1 << np.arange(b - 1, -1, -1))
    Yh = np.zeros((N, D), dtype=np.float64)
    for d, b in enumerate(bits):
        if b:
            Yh[:, d] = cb[d][idx[:, d]]
    return (Yh @ V.T + mu).astype(np.float32)
