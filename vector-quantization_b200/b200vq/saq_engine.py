"""Host side of the SAQ engine path: `saq.IVF(n, D, K=1, cfg).fit(X, apply_pca, 1, seed, threads)`
and `.decompress(ids)` as driven by SaqEngineAdapter (reference
src/haag_vq/benchmarks/quantizer_adapters.py:113-135), plus the engine's accurate estimator.

Pipeline (external/saq/src/ivf_index.cpp:186-243, saq_quantizer.h:49-83, quantizer.h:56-104):
  PCA (covariance GEMM on the device, fp64 eigh) -> K = 1 "k-means" centroid (a seed row, see
  kmeans1_row) -> DP plan on the PCA variances (b200vq_saq_plan_dp, host) -> per segment:
  rotate by P_s, subtract the rotated centroid, CAQ-encode every row (b200vq_caq_encode_t),
  bit-pack (b200vq_caq_pack_t). Codes stay in HBM segment by segment as MSB-first bit strings
  with three fp32 factors per (row, segment).

Every GEMM runs in libb200vq.so's fp32-accurate tensor-core kernel (csrc/rotate_tc.cu): the PCA
projection and the per-segment rotations are ONE composite matrix Wc = rot . blockdiag(P_s), so
encode is a single GEMM with the mean subtraction, the rotated-centroid subtraction and the
transposed store (the layout b200vq_caq_encode_t consumes) fused in; decode is the same matrix
the other way with the bias fused; search is the code-domain estimator scan on the tensor cores
(csrc/est_tc.cu) straight from the packed codes. There is no CPU fallback and no library GEMM.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Tuple

import numpy as np
import torch

from . import _lib, ops, ops_sq
from ._lib import check
from .ops import _stream, require_cuda

BLK = 64


def pad64(d: int) -> int:
    return (d + BLK - 1) // BLK * BLK


def mt19937_first(seed: int) -> int:
    """First 32-bit output of std::mt19937(seed) (kmeans_eigen.cpp:20)."""
    state = [seed & 0xFFFFFFFF]
    for i in range(1, 624):
        prev = state[-1]
        state.append((1812433253 * (prev ^ (prev >> 30)) + i) & 0xFFFFFFFF)
    y = (state[0] & 0x80000000) | (state[1] & 0x7FFFFFFF)
    x = state[397] ^ (y >> 1) ^ (0x9908B0DF if (y & 1) else 0)
    x ^= x >> 11
    x ^= (x << 7) & 0x9D2C5680
    x ^= (x << 15) & 0xEFC60000
    x ^= x >> 18
    return x & 0xFFFFFFFF


def mt19937_stream(seed: int):
    """Outputs of std::mt19937(seed), one after another."""
    mt = [seed & 0xFFFFFFFF]
    for i in range(1, 624):
        prev = mt[-1]
        mt.append((1812433253 * (prev ^ (prev >> 30)) + i) & 0xFFFFFFFF)
    idx = 624
    while True:
        if idx >= 624:
            for i in range(624):
                y = (mt[i] & 0x80000000) | (mt[(i + 1) % 624] & 0x7FFFFFFF)
                mt[i] = mt[(i + 397) % 624] ^ (y >> 1) ^ (0x9908B0DF if (y & 1) else 0)
            idx = 0
        x = mt[idx]
        idx += 1
        x ^= x >> 11
        x ^= (x << 7) & 0x9D2C5680
        x ^= (x << 15) & 0xEFC60000
        x ^= x >> 18
        yield x & 0xFFFFFFFF


def kmeans1_row(n: int, seed: int) -> int:
    """With K = 1 the engine's Lloyd loop (kmeans_eigen.cpp:40-56) exits before its first update
    (every assignment already is 0), so the centroid is the row drawn by
    uniform_int_distribution<int>(0, n-1) from mt19937(seed): libstdc++'s multiply-shift with rejection
    (a rejected draw takes the generator's next output)."""
    thresh = ((1 << 32) - n) % n
    for u in mt19937_stream(seed):
        prod = u * n
        if (prod & 0xFFFFFFFF) >= thresh:
            return prod >> 32


def plan_dp(variance: np.ndarray, avg_bits: float) -> List[Tuple[int, int]]:
    var = np.ascontiguousarray(variance, dtype=np.float32)
    d_pad = int(var.shape[0])
    cap = d_pad // BLK + 1
    seg_len = np.zeros(cap, dtype=np.int32)
    seg_bits = np.zeros(cap, dtype=np.int32)
    ns = C.c_int(0)
    check(_lib.load().b200vq_saq_plan_dp(var.ctypes.data, d_pad, C.c_float(avg_bits), seg_len.ctypes.data,
                                         seg_bits.ctypes.data, C.addressof(ns)), "saq_plan_dp")
    return [(int(seg_len[i]), int(seg_bits[i])) for i in range(ns.value)]


MAX_QUANT_BIT = 13          # kMaxQuantBit
NUM_BIT_FACTORS = 64        # kNumShortFactors * sizeof(float) * 8


def allocate_greedy(mse_table: np.ndarray, avg_bits: float, dim_padding: int = BLK, max_bits: int = MAX_QUANT_BIT,
                    num_bit_factors: int = NUM_BIT_FACTORS) -> List[Tuple[int, int]]:
    """BitAllocatorGreedy::AllocateJoint (external/saq/src/bit_allocator_greedy.cpp:11-116) driven as in
    quantization_plan.h:246-266: every 64-dimension block starts at 0 bits; the block with the largest marginal MSE
    drop (first index wins ties) is bumped by one bit until the bump would exceed the budget
    avg_bits * d_pad + 64 (bits of all dimensions + 64 factor bits per coded segment, segments = runs of equal
    widths); runs of equal widths are merged into the plan. mse_table: (d_pad, max_bits + 1) host fp32."""
    d_pad = int(mse_table.shape[0])
    nb = d_pad // dim_padding
    total_bits = int(avg_bits * d_pad) + num_bit_factors
    block = np.zeros((nb, max_bits + 1), dtype=np.float64)
    for k in range(dim_padding):                      # s += mse_table(i * pad + k, b), k ascending, in double
        block += mse_table[k::dim_padding].astype(np.float64)[:nb]

    def usage(bits):
        segs = sum(1 for i in range(nb) if bits[i] > 0 and (i == 0 or bits[i - 1] != bits[i]))
        return int(sum(bits)) * dim_padding + segs * num_bit_factors
    bits = [0] * nb
    while True:
        best_i, best_drop = -1, -np.inf
        for i in range(nb):
            if bits[i] < max_bits:
                drop = block[i, bits[i]] - block[i, bits[i] + 1]
                if drop > best_drop:
                    best_drop, best_i = drop, i
        if best_i < 0:
            break
        bits[best_i] += 1
        if usage(bits) > total_bits:
            bits[best_i] -= 1
            break
    plan, i = [], 0
    while i < nb:
        j = i + 1
        while j < nb and bits[j] == bits[i]:
            j += 1
        plan.append(((j - i) * dim_padding, int(bits[i])))
        i = j
    return plan


def equal_segmentation(d_pad: int, avg_bits: float, num_segs: int) -> List[Tuple[int, int]]:
    """quantization_plan.cpp:125-138."""
    plan, d = [], 0
    for i in range(num_segs):
        if d >= d_pad:
            break
        t = pad64((d_pad - d) // (num_segs - i))
        plan.append((t, int(avg_bits)))
        d += t
    return plan


def caq_encode_t(oT: torch.Tensor, bits: int, c_rot: Optional[torch.Tensor] = None, rd_lmt: int = 6,
                 adj_eps: float = 1e-8):
    """oT (len, n) fp32 on the device (row stride >= n). Returns codesT (len, n) uint16 and the five
    (n,) fp32 outputs (l2norm, rescale, error, mse, cos)."""
    require_cuda()
    ln, n = oT.shape
    if oT.stride(1) != 1:
        raise ValueError("caq_encode_t: oT must be row-major (len, n)")
    ldc = (n + 1) // 2 * 2
    codesT = torch.empty((ln, ldc), dtype=torch.int16, device=oT.device)[:, :n]
    outs = [torch.empty((n,), dtype=torch.float32, device=oT.device) for _ in range(5)]
    check(_lib.load().b200vq_caq_encode_t(oT.data_ptr(), n, oT.stride(0), ln,
                                          c_rot.data_ptr() if c_rot is not None else None, bits, rd_lmt,
                                          C.c_float(adj_eps), codesT.data_ptr(), ldc,
                                          *[o.data_ptr() for o in outs], _stream()), "caq_encode_t")
    return codesT, outs


def caq_encode_t_segments(oT: torch.Tensor, plan, rd_lmt: int = 6, adj_eps: float = 1e-8):
    """Every segment of `plan` ((len, bits) pairs, stacked in oT (d_pad, n), already centred) in one launch.
    Returns codesT (d_pad, n) uint16-as-int16 and five (nseg, n) fp32 factor arrays."""
    require_cuda()
    d_pad, n = oT.shape
    ns = len(plan)
    lens = [int(a) for a, _ in plan]
    offs = list(np.cumsum([0] + lens[:-1]))
    if sum(lens) != d_pad or (n > 1 and oT.stride(1) != 1):
        raise ValueError("caq_encode_t_segments: oT must be (sum of segment lengths, n) row-major")
    ldc = (n + 1) // 2 * 2          # even row stride: the encoder brings two codes per 4-byte cp.async
    codesT = torch.empty((d_pad, ldc), dtype=torch.int16, device=oT.device)[:, :n]
    outs = [torch.empty((ns, n), dtype=torch.float32, device=oT.device) for _ in range(5)]
    arr = lambda v: (C.c_int * ns)(*[int(x) for x in v])
    check(_lib.load().b200vq_caq_encode_t_segments(oT.data_ptr(), n, oT.stride(0), ns, arr(offs), arr(lens),
                                                   arr([b for _, b in plan]), rd_lmt, C.c_float(adj_eps),
                                                   codesT.data_ptr(), ldc, *[o.data_ptr() for o in outs], _stream()),
          "caq_encode_t_segments")
    return codesT, outs


def caq_pack_t(codesT: torch.Tensor, bits: int, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    ln, n = codesT.shape
    if out is None:
        out = torch.empty((n, ln * bits // 8), dtype=torch.uint8, device=codesT.device)
    check(_lib.load().b200vq_caq_pack_t(codesT.data_ptr(), n, codesT.stride(0), ln, bits, out.data_ptr(),
                                        _stream()), "caq_pack_t")
    return out


def caq_decode(packed: torch.Tensor, ln: int, bits: int, l2norm: Optional[torch.Tensor],
               ids: Optional[torch.Tensor] = None, mode: int = 0, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """out (optional): an (m, ln) fp32 view with unit column stride (e.g. a column slice of a wider matrix)."""
    n_total = packed.shape[0]
    m = n_total if ids is None else int(ids.shape[0])
    if out is None:
        out = torch.empty((m, ln), dtype=torch.float32, device=packed.device)
    check(_lib.load().b200vq_caq_decode(packed.data_ptr(), n_total, ids.data_ptr() if ids is not None else None, m,
                                        ln, bits, l2norm.data_ptr() if l2norm is not None else None, mode,
                                        out.data_ptr(), out.stride(0) if m > 1 else max(ln, out.stride(0)), _stream()),
          "caq_decode")
    return out


class SaqEngine:
    """Device-resident single-cluster SAQ index (the K = 1 configuration SaqEngineAdapter uses)."""

    def __init__(self, avg_bits: float, apply_pca: bool = True, seed: int = 0, rot_seed: int = 20240607,
                 rd_lmt: int = 6, adj_eps: float = 1e-8, seg_eqseg: int = 0, batch_rows: int = 1 << 17,
                 allocator: str = "dp"):
        if allocator not in ("dp", "greedy"):
            raise ValueError("allocator must be 'dp' or 'greedy'")
        self.allocator = allocator
        self.mse_table = None
        self.avg_bits, self.apply_pca, self.seed, self.rot_seed = float(avg_bits), bool(apply_pca), seed, rot_seed
        self.rd_lmt, self.adj_eps, self.seg_eqseg, self.batch_rows = rd_lmt, adj_eps, seg_eqseg, batch_rows
        self.n = self.D = self.d_pad = 0
        self.plan: List[Tuple[int, int]] = []
        self.segs: list = []
        self._scanner = None

    # ---- training state
    def _pca(self, X: torch.Tensor):
        """Mean (numpy's row-sequential fp32 order), covariance as a Gram matrix on the tensor cores
        (X_c^T X_c = A A^T with A = X_c^T prepared as the GEMM's W operand), fp64 eigh (train time)."""
        n, D = X.shape
        mean = ops_sq.col_mean_seq(X)
        cov = torch.zeros((D, D), dtype=torch.float32, device=X.device)
        for s in range(0, n, self.batch_rows):
            xb = X[s:s + self.batch_rows]
            rows = xb.shape[0]
            xt = torch.zeros((D, (rows + 31) // 32 * 32), dtype=torch.float32, device=X.device)
            xt[:, :rows] = (xb - mean).t()
            cov += ops.RotationTC(xt)(xt)
        evals, evecs = torch.linalg.eigh(cov.double() / float(n - 1))
        return mean, evecs.flip(1).float().contiguous(), evals.flip(0).clamp_min(0).float()

    def _train(self, X: torch.Tensor, state: Optional[dict]):
        dev = X.device
        n, D = X.shape
        if state is not None:
            self.mean = torch.as_tensor(state["mean"], device=dev)
            self.rot = torch.as_tensor(state["rot"], device=dev)
            self.var = torch.as_tensor(state["var"], device=dev)
            self.cent_row = int(state["cent_row"])
            self.plan = [(int(a), int(b)) for a, b in state["plan"]]
            self.P = [torch.as_tensor(p, device=dev) for p in state["P"]]
            return
        if self.apply_pca:
            self.mean, self.rot, self.var = self._pca(X)
        else:
            self.mean = torch.zeros(D, device=dev)
            self.rot = torch.eye(D, device=dev)
            self.var = torch.ones(D, device=dev)
        self.cent_row = kmeans1_row(n, self.seed)
        var_pad = np.zeros(self.d_pad, np.float32)
        var_pad[:D] = self.var.cpu().numpy()
        if self.seg_eqseg > 0:
            self.plan = equal_segmentation(self.d_pad, self.avg_bits, self.seg_eqseg)
        elif self.allocator == "greedy":
            self.mse_table = self._mse_table(X)
            self.plan = allocate_greedy(self.mse_table, self.avg_bits)
        else:
            self.plan = plan_dp(var_pad, self.avg_bits)
        g = torch.Generator(device="cpu").manual_seed(self.rot_seed)
        self.P = []
        for ln, _ in self.plan:   # P = Q^T of a QR of a Gaussian matrix (rotator.h:43-48)
            Q, _r = torch.linalg.qr(torch.randn((ln, ln), generator=g, dtype=torch.float64))
            self.P.append(Q.t().float().contiguous().to(dev))

    def _mse_table(self, X: torch.Tensor) -> np.ndarray:
        """build_mse_table_for_allocation (external/saq/src/quantization_plan.cpp:21-51): per PCA dimension the MSE of
        its 1-D Lloyd codebook (k-means++ seeding, seed 0) at every bit-rate 0..13, on the PCA-projected data
        (every row while n <= recommended_sample_size = 4 096 000). (d_pad, 14) host fp32; padding dimensions
        are constant columns (cost 0)."""
        n, D = X.shape
        if n > 500 * (1 << MAX_QUANT_BIT):
            raise NotImplementedError("greedy allocation samples the column beyond 4 096 000 rows (not built)")
        Dk = (D + 31) // 32 * 32
        W = torch.zeros((self.d_pad, Dk), dtype=torch.float32, device=X.device)      # y = (x - mean) rot: W = rot^T
        W[:D, :D] = (self.rot if self.apply_pca else torch.eye(D, device=X.device)).t()
        mean_k = torch.zeros(Dk, dtype=torch.float32, device=X.device)
        mean_k[:D] = self.mean if self.apply_pca else 0.0
        xp = X
        if D != Dk or X.stride(0) % 4 or X.data_ptr() % 16:
            xp = torch.zeros((n, Dk), dtype=torch.float32, device=X.device)
            xp[:, :D] = X
        colsT = ops.RotationTC(W)(xp, x_sub=mean_k, transpose_out=True).contiguous()   # (d_pad, n)
        pre = ops_sq.codebook_prefix(colsT)
        table = np.zeros((self.d_pad, MAX_QUANT_BIT + 1), dtype=np.float32)
        for b in range(MAX_QUANT_BIT + 1):
            _, _, costs = ops_sq.codebook_lloyd(colsT, b, seed=0, max_bits=max(b, 1), presorted=pre)
            table[:, b] = costs.cpu().numpy()
        return table

    def _compose(self, dev):
        """Wc (Dk, d_pad) = [rot | 0] . blockdiag(P_s): projection and segment rotations as one matrix
        (Dk = D rounded up to the GEMM's K granule, the extra rows are zero). Built with the same GEMM."""
        D, d_pad = self.D, self.d_pad
        self.Dk = (D + 31) // 32 * 32
        rot_pad = torch.zeros((self.Dk, d_pad), dtype=torch.float32, device=dev)
        rot_pad[:D, :D] = self.rot if self.apply_pca else torch.eye(D, device=dev)
        Wc = torch.empty((self.Dk, d_pad), dtype=torch.float32, device=dev)
        off = 0
        for (ln, _), P in zip(self.plan, self.P):
            # rot_pad[:, seg] . P_s  =  x W^T with W = P_s^T
            ops.RotationTC(P.t().contiguous())(rot_pad[:, off:off + ln].contiguous(), out=Wc[:, off:off + ln])
            off += ln
        self.Wc = Wc
        self.enc = ops.RotationTC(Wc.t().contiguous())     # o = (x - mean) Wc        (N = d_pad, K = Dk)
        self.dec = ops.RotationTC(Wc[:D].contiguous())     # x = oa Wc^T + bias       (N = D, K = d_pad)
        self.mean_k = torch.zeros(self.Dk, dtype=torch.float32, device=dev)
        self.mean_k[:D] = self.mean if self.apply_pca else 0.0

    def _pad_k(self, Xb: torch.Tensor) -> torch.Tensor:
        if Xb.shape[1] == self.Dk and Xb.stride(0) % 4 == 0 and Xb.data_ptr() % 16 == 0 and Xb.stride(1) == 1:
            return Xb
        xp = torch.zeros((Xb.shape[0], self.Dk), dtype=torch.float32, device=Xb.device)
        xp[:, :Xb.shape[1]] = Xb
        return xp

    def _project(self, Xb: torch.Tensor, transpose_out: bool = False) -> torch.Tensor:
        """(x - mean) rot blockdiag(P) - c_rot for every row: (rows, d_pad), or its transpose."""
        return self.enc(self._pad_k(Xb), x_sub=self.mean_k, col_bias=self.neg_c_rot, transpose_out=transpose_out)

    # ---- fit
    def fit(self, X: torch.Tensor, state: Optional[dict] = None) -> "SaqEngine":
        require_cuda()
        if X.dtype != torch.float32 or X.dim() != 2 or not X.is_cuda:
            raise ValueError("SaqEngine.fit expects a float32 (n, D) CUDA tensor")
        n, D = X.shape
        self.n, self.D, self.d_pad = n, D, pad64(D)
        self._train(X, state)
        dev = X.device
        self._compose(dev)
        # rotated centroid of every segment = the centroid row through the same composite matrix
        self.neg_c_rot = None
        c_rot_all = self._project(X[self.cent_row:self.cent_row + 1])[0].contiguous()
        self.neg_c_rot = (-c_rot_all).contiguous()
        # decode bias: (c_rot_s P_s^T)[:D] rot^T + mean = c_rot_all Wc^T + mean (ivf_index.cpp:296-374)
        self.dec_bias = self.dec(c_rot_all.unsqueeze(0), col_bias=(self.mean if self.apply_pca else None))[0].contiguous()
        self.segs, off = [], 0
        for (ln, bits), P in zip(self.plan, self.P):
            self.segs.append(dict(len=ln, bits=bits, off=off, c_rot=c_rot_all[off:off + ln],
                                  packed=(torch.empty((n, ln * bits // 8), dtype=torch.uint8, device=dev)
                                          if bits else None),
                                  l2norm=torch.empty(n, device=dev), rescale=torch.empty(n, device=dev),
                                  error=torch.empty(n, device=dev), mse=torch.empty(n, device=dev),
                                  cos=torch.empty(n, device=dev)))
            off += ln
        for s in range(0, n, self.batch_rows):
            e = min(n, s + self.batch_rows)
            oT = self._project(X[s:e], transpose_out=True)          # (d_pad, rows): every segment, centred
            codesT, outs = caq_encode_t_segments(oT, self.plan, self.rd_lmt, self.adj_eps)
            for si, seg in enumerate(self.segs):
                for key, o in zip(("l2norm", "rescale", "error", "mse", "cos"), outs):
                    seg[key][s:e] = o[si]
                if seg["bits"]:
                    caq_pack_t(codesT[seg["off"]:seg["off"] + seg["len"]], seg["bits"], out=seg["packed"][s:e])
        self._scanner = None
        return self

    # ---- queries on the fitted index
    def recon_mse(self) -> float:
        """Mean over rows of sum over segments of recon_mse * len / d_pad (engine metric)."""
        tot = sum((seg["mse"].double().sum().item() * seg["len"]) for seg in self.segs)
        return tot / (self.n * self.d_pad)

    def stored_bytes(self) -> int:
        return sum((seg["packed"].numel() if seg["packed"] is not None else 0) + 12 * self.n for seg in self.segs)

    def reconstruct(self, ids: torch.Tensor) -> torch.Tensor:
        """IVF::decompress (ivf_index.cpp:245-374): ids int64 on the device -> (m, D) fp32."""
        m = int(ids.shape[0])
        oa = torch.zeros((m, self.d_pad), dtype=torch.float32, device=ids.device)
        for seg in self.segs:
            if seg["bits"]:
                caq_decode(seg["packed"], seg["len"], seg["bits"], seg["l2norm"], ids, mode=0,
                           out=oa[:, seg["off"]:seg["off"] + seg["len"]])
        return self.dec(oa, col_bias=self.dec_bias)

    def _scan_state(self):
        """Per-index state of the code-domain scan: the coded segments, rowscale = rescale * Delta,
        B_i = sum_s |o_s|^2 (all segments) and |x_i|^2 of the virtual rows."""
        if self._scanner is None:
            from .ops_est import CodeSegments
            coded = [seg for seg in self.segs if seg["bits"] > 0]
            bias = torch.zeros(self.n, dtype=torch.float32, device=self.segs[0]["l2norm"].device)
            for seg in self.segs:
                bias += seg["l2norm"] * seg["l2norm"]
            if not coded:
                self._scanner = dict(segs=None, bias=bias)
                return self._scanner
            scales = [(seg["rescale"] * (2.0 / (1 << seg["bits"]))).contiguous() for seg in coded]
            cs = CodeSegments([seg["len"] for seg in coded], [seg["bits"] for seg in coded],
                              [seg["packed"] for seg in coded], scales, [None] * len(coded),
                              [-((1 << seg["bits"]) - 1) / 2.0 for seg in coded])
            cols = torch.cat([torch.arange(seg["off"], seg["off"] + seg["len"], device=bias.device) for seg in coded])
            self._scanner = dict(segs=cs, bias=bias, norm2=cs.row_norms(), cols=cols)
        return self._scanner

    def search(self, Q: torch.Tensor, k: int):
        """Top-k by the engine's accurate estimator (caq_estimator.h:254-283, saq_estimator.h:183-190), squared L2:
        sum_s |o_s|^2 + |q_s|^2 - 2 rescale_s (Delta_s <q_s, code> + (-1 + Delta_s / 2) sum(q_s)),
        q_s = Q_pca[:, seg] P_s - c_rot_s, scored from the packed codes on the tensor cores (csrc/est_tc.cu).
        Returns (dist (nq, k) fp32 ascending, ids (nq, k) int32-as-uint32)."""
        sc = self._scan_state()
        q_all = self._project(Q)                                # (nq, d_pad): every segment's query coordinates
        qc = ops.row_sqnorms(q_all.contiguous() if q_all.stride(1) != 1 else q_all)
        if sc["segs"] is None:
            raise ValueError("SaqEngine.search: the plan has no coded segment")
        qk = q_all.index_select(1, sc["cols"]) if sc["cols"].numel() != self.d_pad else q_all
        return sc["segs"].scan_topk(sc["norm2"], sc["bias"], qk, qc, k, "l2")

    def export_state(self) -> dict:
        return dict(mean=self.mean.cpu().numpy(), rot=self.rot.cpu().numpy(), var=self.var.cpu().numpy(),
                    cent_row=self.cent_row, plan=list(self.plan), P=[p.cpu().numpy() for p in self.P],
                    mse_table=self.mse_table)
