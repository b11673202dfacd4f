"""CPU oracle: scalar quantizer (numpy) and the segmented variable-bit SAQ contract.

TEST INFRASTRUCTURE ONLY — see oracle/__init__.py.

``sq_*`` restates src/haag_vq/methods/scalar_quantization.py:37-90 line by line
(PINNED: oracle/make_golden.py checks it against the reference class imported
from the reference tree and commits the vectors under tests/golden/).

``saq_*`` is the specification of the ``SAQ(num_bits | total_bits, allowed_bits,
n_segments)`` CLI contract (src/haag_vq/benchmarks/sweep.py:321-364,418-430,
README.md:153-170). The reference tree holds NO implementation of it
(tests/test_no_placeholder_saq.py:16-18) -> PARITY UNPINNED; the recorded
behaviour of the deleted class (logs/benchmark_runs.db rows 54-55: 8.0x at 4 bits,
4.0x at 8 bits) matches a per-dimension min/max affine quantizer, which is what
is specified here.
"""
from __future__ import annotations

import numpy as np


# ------------------------------------------------------------------ SQ
def sq_fit(X: np.ndarray):
    """scalar_quantization.py:37-39."""
    return X.min(axis=0), X.max(axis=0)


def sq_compress(X: np.ndarray, vmin: np.ndarray, vmax: np.ndarray, num_bits: int) -> np.ndarray:
    """scalar_quantization.py:52-68 (_compress_block)."""
    X = np.asarray(X)
    levels = 2 ** num_bits
    dtype = np.uint8 if num_bits <= 8 else np.uint16
    scaled = (X - vmin) / (vmax - vmin + 1e-8)
    quantized = np.round(scaled * (levels - 1)).astype(dtype)
    if num_bits == 4:
        if X.shape[1] % 2 == 1:
            quantized = np.pad(quantized, ((0, 0), (0, 1)), mode="constant")
        return (quantized[:, 0::2] << 4) | quantized[:, 1::2]
    return quantized


def sq_decompress(codes: np.ndarray, vmin: np.ndarray, vmax: np.ndarray, num_bits: int) -> np.ndarray:
    """scalar_quantization.py:70-90."""
    levels = 2 ** num_bits
    if num_bits == 4:
        hi = (codes >> 4).astype(np.float32)
        lo = (codes & 0x0F).astype(np.float32)
        unpacked = np.empty((codes.shape[0], codes.shape[1] * 2), dtype=np.float32)
        unpacked[:, 0::2] = hi
        unpacked[:, 1::2] = lo
        if len(vmin) % 2 == 1:
            unpacked = unpacked[:, :-1]
        scaled = unpacked / (levels - 1)
    else:
        scaled = codes.astype(np.float32) / (levels - 1)
    return scaled * (vmax - vmin + 1e-8) + vmin


# ------------------------------------------------------------------ SAQ (CLI contract)
def saq_segments(D: int, n_segments: int) -> np.ndarray:
    """Segment boundaries: n_segments contiguous, near-equal slices of the D dims."""
    n_segments = max(1, min(int(n_segments), D))
    return np.linspace(0, D, n_segments + 1).round().astype(np.int64)


def saq_allocate(var: np.ndarray, bounds: np.ndarray, total_bits: int, allowed_bits) -> np.ndarray:
    """Greedy variance-driven allocation of one bit-width per segment.

    Every segment starts at min(allowed); repeatedly upgrade the segment whose
    next allowed width buys the largest distortion decrease per extra bit, with
    the high-resolution model distortion(seg, b) = sum(var_seg) * 4**-b, while
    the per-vector budget sum(len_s * b_s) <= total_bits holds. Ties -> lowest
    segment index. Returns (n_segments,) int32 bit-widths.
    """
    allowed = sorted(set(int(b) for b in allowed_bits))
    nseg = len(bounds) - 1
    lens = np.diff(bounds).astype(np.int64)
    svar = np.array([float(var[bounds[s]:bounds[s + 1]].sum()) for s in range(nseg)], dtype=np.float64)
    level = np.zeros(nseg, dtype=np.int64)  # index into allowed
    used = int((lens * allowed[0]).sum())
    while True:
        best, best_gain = -1, 0.0
        for s in range(nseg):
            if level[s] + 1 >= len(allowed):
                continue
            b0, b1 = allowed[level[s]], allowed[level[s] + 1]
            extra = int(lens[s]) * (b1 - b0)
            if extra <= 0 or used + extra > total_bits:
                continue
            gain = svar[s] * (4.0 ** -b0 - 4.0 ** -b1) / extra
            if gain > best_gain:
                best, best_gain = s, gain
        if best < 0:
            break
        used += int(lens[best]) * (allowed[level[best] + 1] - allowed[level[best]])
        level[best] += 1
    return np.array([allowed[i] for i in level], dtype=np.int32)


def saq_plan(X: np.ndarray, num_bits: int = 8, total_bits: int | None = None,
             allowed_bits=(0, 2, 4, 6, 8), n_segments: int | None = None):
    """Model state of the SAQ contract: per-dim min/max/mean, per-dim bit table."""
    X = np.asarray(X, dtype=np.float32)
    D = X.shape[1]
    vmin, vmax = X.min(axis=0), X.max(axis=0)
    mean = X.mean(axis=0, dtype=np.float64).astype(np.float32)
    if total_bits is None:
        dim_bits = np.full(D, int(num_bits), dtype=np.int32)
    else:
        nseg = n_segments if n_segments is not None else max(1, D // 64)
        bounds = saq_segments(D, nseg)
        var = X.var(axis=0, dtype=np.float64)
        seg_bits = saq_allocate(var, bounds, int(total_bits), allowed_bits)
        dim_bits = np.repeat(seg_bits, np.diff(bounds)).astype(np.int32)
    bit_offset = np.zeros(D + 1, dtype=np.int32)
    np.cumsum(dim_bits, out=bit_offset[1:])
    return dict(vmin=vmin, vmax=vmax, mean=mean, dim_bits=dim_bits, bit_offset=bit_offset,
                code_bytes=max(1, (int(bit_offset[-1]) + 7) // 8))


def saq_compress(X: np.ndarray, plan: dict) -> np.ndarray:
    """Per-dim affine quantiser (same arithmetic as sq_compress with L = 2**b_d,
    saturating) + MSB-first bit string."""
    X = np.asarray(X, dtype=np.float32)
    n, D = X.shape
    bits = plan["dim_bits"]
    lm1 = (np.left_shift(1, bits) - 1).astype(np.float32)
    scaled = (X - plan["vmin"]) / (plan["vmax"] - plan["vmin"] + 1e-8)
    q = np.round(scaled * lm1)
    q = np.minimum(np.maximum(q, 0), lm1).astype(np.uint16)
    total = int(plan["bit_offset"][-1])
    allbits = np.zeros((n, plan["code_bytes"] * 8), dtype=np.uint8)
    for d in range(D):
        b, off = int(bits[d]), int(plan["bit_offset"][d])
        for j in range(b):
            allbits[:, off + j] = (q[:, d] >> (b - 1 - j)) & 1
    assert total <= allbits.shape[1]
    return np.packbits(allbits, axis=1)


def saq_decompress(codes: np.ndarray, plan: dict) -> np.ndarray:
    bits = plan["dim_bits"]
    D = bits.shape[0]
    n = codes.shape[0]
    allbits = np.unpackbits(np.ascontiguousarray(codes, dtype=np.uint8), axis=1)
    out = np.empty((n, D), dtype=np.float32)
    rng = (plan["vmax"] - plan["vmin"] + 1e-8).astype(np.float32)
    for d in range(D):
        b, off = int(bits[d]), int(plan["bit_offset"][d])
        if b == 0:
            out[:, d] = plan["mean"][d]
            continue
        q = np.zeros(n, dtype=np.uint32)
        for j in range(b):
            q = (q << 1) | allbits[:, off + j]
        scaled = q.astype(np.float32) / np.float32((1 << b) - 1)
        out[:, d] = scaled * rng[d] + plan["vmin"][d]
    return out
