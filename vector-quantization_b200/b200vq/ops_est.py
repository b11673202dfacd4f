"""Code-domain estimator scan (csrc/est_tc.cu): top-k of an estimator that is linear in packed
per-dimension codes, scored on the tensor cores straight from the code bytes.

Reference estimators served: the SAQ engine's accurate estimator
(external/saq/include/saq/caq_estimator.h:254-316, saq_estimator.h:27-129) and extended RaBitQ
searched through FlatQuantizedIndex (src/haag_vq/methods/extended_rabitq.py:173-198).
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence

import torch

from . import _lib
from ._lib import B200VQError, check
from .ops import _chk, _stream, _topk_out, metric_id, require_cuda, scratch


class CodeSegments:
    """Host-side description of the stored rows: one entry per segment with bits > 0.
    codes[s]: (n, stride_s) uint8 device tensor (MSB-first bit string of len_s * bits_s bits per row);
    rowscale[s]: (n,) fp32 or None; table[s]: (2^bits,) fp32 value table or None (value = code + base[s])."""

    def __init__(self, lens: Sequence[int], bits: Sequence[int], codes: Sequence[torch.Tensor],
                 rowscale: Sequence[Optional[torch.Tensor]], table: Sequence[Optional[torch.Tensor]],
                 base: Sequence[float]):
        require_cuda()
        ns = len(lens)
        if not (ns == len(bits) == len(codes) == len(rowscale) == len(table) == len(base)) or ns == 0:
            raise B200VQError("CodeSegments: inconsistent segment lists")
        self.n = int(codes[0].shape[0])
        self.D = int(sum(lens))
        self.keep = (list(codes), list(rowscale), list(table))   # the pointers below borrow these tensors
        for c in codes:
            _chk(c, torch.uint8, "codes")
            if c.dim() != 2 or c.shape[0] != self.n or (c.shape[1] > 1 and c.stride(1) != 1):
                raise B200VQError("CodeSegments: codes must be (n, bytes) row-major")
        for t in list(rowscale) + list(table):
            if t is not None and (_chk(t, torch.float32, "rowscale/table").dim() != 1 or not t.is_contiguous()):
                raise B200VQError("CodeSegments: rowscale / table must be contiguous fp32 vectors")
        self.ns = ns
        self.seg_len = (C.c_int * ns)(*[int(v) for v in lens])
        self.seg_bits = (C.c_int * ns)(*[int(v) for v in bits])
        self.seg_codes = (C.c_void_p * ns)(*[c.data_ptr() for c in codes])
        self.seg_stride = (C.c_int64 * ns)(*[int(c.stride(0)) if c.shape[0] > 1 else int(c.shape[1]) for c in codes])
        self.seg_rowscale = (C.c_void_p * ns)(*[(t.data_ptr() if t is not None else None) for t in rowscale])
        self.seg_table = (C.c_void_p * ns)(*[(t.data_ptr() if t is not None else None) for t in table])
        self.seg_base = (C.c_float * ns)(*[float(v) for v in base])
        self.device = codes[0].device

    def _args(self):
        return (self.ns, self.seg_len, self.seg_bits, self.seg_codes, self.seg_stride, self.seg_rowscale,
                self.seg_table, self.seg_base)

    def row_norms(self) -> torch.Tensor:
        """|x_i|^2 of the virtual rows (b200vq_est_row_norms)."""
        out = torch.empty((self.n,), dtype=torch.float32, device=self.device)
        sc = scratch(8192, self.device, "est_tab")
        check(_lib.load().b200vq_est_row_norms(*self._args(), self.n, out.data_ptr(), sc.data_ptr(), sc.numel(),
                                               _stream()), "est_row_norms")
        return out

    def scan_topk(self, row_norm2: torch.Tensor, row_bias: torch.Tensor, queries: torch.Tensor,
                  qconst: Optional[torch.Tensor], k: int, metric):
        """Top-k of C_q + B_i - 2 <x_i, q> (l2, ascending) or C_q + <x_i, q> (ip, descending).
        queries: (nq, D) fp32 concatenated per-segment query coordinates. Returns (vals, ids int32-as-uint32)."""
        _chk(queries, torch.float32, "queries")
        _chk(row_norm2, torch.float32, "row_norm2")
        _chk(row_bias, torch.float32, "row_bias")
        if queries.dim() != 2 or queries.shape[1] != self.D:
            raise B200VQError(f"est_scan_topk: queries must be (nq, {self.D})")
        nq = queries.shape[0]
        if queries.stride(1) != 1 or queries.stride(0) % 4 or queries.data_ptr() % 16:
            ld = (self.D + 3) // 4 * 4
            q2 = torch.zeros((nq, ld), dtype=torch.float32, device=queries.device)
            q2[:, :self.D] = queries
            queries = q2[:, :self.D]
        lib = _lib.load()
        vals, ids = _topk_out(nq, k, queries.device)
        sc = scratch(lib.b200vq_est_scan_scratch_bytes(self.n, nq, k), queries.device, "est_scan")
        check(lib.b200vq_est_scan_topk(*self._args(), row_norm2.data_ptr(), row_bias.data_ptr(), self.n,
                                       queries.data_ptr(), nq, queries.stride(0) if nq > 1 else max(self.D, queries.stride(0)),
                                       qconst.data_ptr() if qconst is not None else None, metric_id(metric), k,
                                       vals.data_ptr(), ids.data_ptr(), sc.data_ptr(), sc.numel(), _stream()),
              "est_scan_topk")
        return vals, ids
