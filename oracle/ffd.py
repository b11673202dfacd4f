"""CPU restatement of the reference's FFD byte packing. TEST INFRASTRUCTURE ONLY.

Follows src/haag_vq/methods/ffd_packing.py:25-81 (layout: decreasing widths, width-4 group behind the
width-3 group for 8-bit bytes, first byte with room), :84-98 (encode) and :101-117 (decode). PINNED:
oracle/make_golden_ffd.py runs the reference module on seeded inputs, asserts equality with this file and
commits tests/golden/ffd_kat.npz (including the reference's known answers {4,3,3,2,2,2} -> 2 bytes and
{4,4,3,3,2} -> 2 bytes, tests/test_ffd_packing.py:62-72).
"""
from __future__ import annotations

import numpy as np


def ffd_layout(bits, byte_cap: int = 8):
    b = np.asarray(bits, dtype=np.int64)
    if np.any(b > byte_cap) or np.any(b < 0):
        raise ValueError("bits out of range")
    order = sorted((d for d in range(len(b)) if b[d] > 0), key=lambda d: (-b[d], d))
    if byte_cap == 8 and any(b[d] == 4 for d in order):
        rest = [d for d in order if b[d] != 4]
        cut = next((i for i, d in enumerate(rest) if b[d] <= 2), len(rest))
        order = rest[:cut] + [d for d in order if b[d] == 4] + rest[cut:]
    byte_idx = np.full(len(b), -1, np.int64)
    bit_off = np.full(len(b), -1, np.int64)
    room = []
    for d in order:
        w = int(b[d])
        t = next((i for i, r in enumerate(room) if r >= w), len(room))
        if t == len(room):
            room.append(byte_cap)
        byte_idx[d], bit_off[d] = t, byte_cap - room[t]
        room[t] -= w
    return byte_idx, bit_off, len(room)


def ffd_encode(codes, bits, byte_idx, bit_off, n_bytes, byte_cap: int = 8):
    codes = np.asarray(codes, dtype=np.uint64)
    out = np.zeros((codes.shape[0], n_bytes), dtype=np.uint8)
    for d, w in enumerate(np.asarray(bits, dtype=np.int64)):
        if w:
            out[:, byte_idx[d]] |= (codes[:, d].astype(np.uint8) << int(byte_cap - bit_off[d] - w))
    return out


def ffd_decode(packed, bits, byte_idx, bit_off, D, byte_cap: int = 8):
    packed = np.asarray(packed, dtype=np.uint8)
    codes = np.zeros((packed.shape[0], D), dtype=np.int64)
    for d, w in enumerate(np.asarray(bits, dtype=np.int64)):
        if w:
            codes[:, d] = (packed[:, byte_idx[d]].astype(np.int64) >> int(byte_cap - bit_off[d] - w)) & ((1 << int(w)) - 1)
    return codes
