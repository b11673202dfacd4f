"""First-fit-decreasing packing of per-dimension bit widths into bytes, on the B200.

Mirrors src/haag_vq/methods/ffd_packing.py:14-117 (``byte_cap_minus``, ``ffd_layout``, ``ffd_encode``,
``ffd_decode``): every dimension's b_d-bit code lies wholly inside one byte, so decoding is a byte read,
a shift and a mask. ``ffd_layout`` is host logic (a few hundred dimensions, once per quantizer) and returns
exactly the reference's layout: widths in decreasing order with ties by index, all width-4 dimensions moved
behind the width-3 ones (the "orphaned 4" fix, byte_cap 8 only), each placed in the first byte with room.
``ffd_encode`` / ``ffd_decode`` run in ``b200vq_ffd_pack`` / ``b200vq_ffd_unpack`` (csrc/ffd.cu).
"""
from __future__ import annotations

import heapq
from typing import Tuple

import numpy as np
import torch

from b200vq import _lib, host
from b200vq._lib import check
from b200vq.ops import _stream


def byte_cap_minus(width: int, offset: int, byte_cap: int = 8) -> int:
    """Right shift that aligns a ``width``-bit field at MSB-side ``offset`` with bit 0 (ffd_packing.py:14-22)."""
    return byte_cap - offset - width


def ffd_layout(bits_per_dim: np.ndarray, byte_cap: int = 8) -> Tuple[np.ndarray, np.ndarray, int]:
    """(byte_idx, bit_off, n_bytes) of the reference's first-fit-decreasing layout (ffd_packing.py:25-81)."""
    b = np.asarray(bits_per_dim, dtype=np.int64)
    if np.any(b > byte_cap) or np.any(b < 0):
        raise ValueError(f"bits_per_dim must be in [0, {byte_cap}]")
    D = b.shape[0]
    byte_idx = np.full(D, -1, dtype=np.int64)
    bit_off = np.full(D, -1, dtype=np.int64)
    # placement order: by width, widest first, dimension index inside a width; with 8-bit bytes the
    # width-4 group goes right behind the width-3 group
    widths = list(range(byte_cap, 0, -1))
    if byte_cap == 8:
        widths = [8, 7, 6, 5, 3, 4, 2, 1]
    # first fit = lowest-numbered open byte with enough room: one min-heap of byte numbers per remaining capacity
    open_by_room = [[] for _ in range(byte_cap + 1)]
    n_bytes = 0
    for w in widths:
        for d in np.nonzero(b == w)[0]:
            best_room = -1
            for room in range(w, byte_cap + 1):
                h = open_by_room[room]
                if h and (best_room < 0 or h[0] < open_by_room[best_room][0]):
                    best_room = room
            if best_room < 0:
                target, room = n_bytes, byte_cap
                n_bytes += 1
            else:
                target, room = heapq.heappop(open_by_room[best_room]), best_room
            byte_idx[d] = target
            bit_off[d] = byte_cap - room
            if room - w > 0:
                heapq.heappush(open_by_room[room - w], target)
    return byte_idx, bit_off, int(n_bytes)


def _tables(bits_per_dim, byte_idx, bit_off, n_bytes: int, byte_cap: int):
    b = np.asarray(bits_per_dim, dtype=np.int64)
    bi = np.asarray(byte_idx, dtype=np.int64)
    bo = np.asarray(bit_off, dtype=np.int64)
    D = b.shape[0]
    used = b > 0
    shift = np.zeros(D, dtype=np.uint8)
    shift[used] = (byte_cap - bo[used] - b[used]).astype(np.uint8)
    mask = np.zeros(D, dtype=np.uint8)
    mask[used] = ((1 << b[used]) - 1).astype(np.uint8)
    dim_byte = np.where(used, bi, -1).astype(np.int32)
    byte_dims = np.full((max(n_bytes, 1), 8), -1, dtype=np.int32)
    fill = np.zeros(max(n_bytes, 1), dtype=np.int64)
    for d in np.nonzero(used)[0]:
        byte_dims[bi[d], fill[bi[d]]] = d
        fill[bi[d]] += 1
    return dim_byte, shift, mask, byte_dims


def ffd_encode(codes: np.ndarray, bits_per_dim: np.ndarray, byte_idx: np.ndarray, bit_off: np.ndarray,
               n_bytes: int, byte_cap: int = 8) -> np.ndarray:
    """(N, D) integer codes -> (N, n_bytes) uint8 per the layout (ffd_packing.py:84-98)."""
    codes = np.asarray(codes)
    N, D = codes.shape
    if byte_cap > 8:
        raise ValueError("byte_cap must be <= 8 (fields live inside one uint8)")
    out = torch.zeros((N, n_bytes), dtype=torch.uint8, device=host.device())
    if N and n_bytes:
        _, shift, _, byte_dims = _tables(bits_per_dim, byte_idx, bit_off, n_bytes, byte_cap)
        c8 = host.to_device(np.ascontiguousarray(codes.astype(np.uint64).astype(np.uint8)))
        bd, sh = host.to_device(byte_dims), host.to_device(shift)   # keep the tables alive across the launch
        check(_lib.load().b200vq_ffd_pack(c8.data_ptr(), N, D, bd.data_ptr(), sh.data_ptr(), n_bytes,
                                          out.data_ptr(), _stream()), "ffd_pack")
        torch.cuda.current_stream().synchronize()
    return host.to_host(out)


def ffd_decode(packed: np.ndarray, bits_per_dim: np.ndarray, byte_idx: np.ndarray, bit_off: np.ndarray, D: int,
               byte_cap: int = 8) -> np.ndarray:
    """Inverse of ffd_encode -> (N, D) int64 codes, 0 for zero-width dimensions (ffd_packing.py:101-117)."""
    packed = np.ascontiguousarray(packed, dtype=np.uint8)
    N, n_bytes = packed.shape
    out = torch.zeros((N, D), dtype=torch.uint8, device=host.device())
    if N and D:
        dim_byte, shift, mask, _ = _tables(bits_per_dim, byte_idx, bit_off, n_bytes, byte_cap)
        pd = host.to_device(packed) if n_bytes else out
        db, sh, mk = host.to_device(dim_byte), host.to_device(shift), host.to_device(mask)
        check(_lib.load().b200vq_ffd_unpack(pd.data_ptr(), N, n_bytes, db.data_ptr(), sh.data_ptr(), mk.data_ptr(), D,
                                            out.data_ptr(), _stream()), "ffd_unpack")
        torch.cuda.current_stream().synchronize()
    return host.to_host(out).astype(np.int64)
