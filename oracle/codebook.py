"""CPU restatement (TEST INFRASTRUCTURE ONLY) of the SAQ engine's per-dimension codebook builder and greedy
bit allocator:

  external/saq/src/preprocessing/codebook_builder.cpp:14-27,44-49,64-128   prefix sums, nudge, lloyd_k (+ repair)
  external/saq/src/preprocessing/codebook_builder.cpp:146-188             init_kmeanspp (the default init)
  external/saq/src/preprocessing/codebook_builder.cpp:341-420             build_codebook_lloyd
  external/saq/include/saq/preprocessing/codebook_builder.h:33-43,90-95   LloydOpts defaults, recommended_sample_size
  external/saq/src/bit_allocator_greedy.cpp:11-116                        BitAllocatorGreedy::AllocateJoint
  external/saq/src/quantization_plan.cpp:21-51                            build_mse_table_for_allocation

Pinned: oracle/_ref/libsaq_ref_codebook.so is the reference's own codebook_builder.cpp compiled where it lies
(oracle/Makefile target ref_codebook; glog / Eigen replaced by two shim headers); tests/test_oracle_codebook.py
checks this restatement against it and against tests/golden/codebook_lloyd.npz (generated from it by
oracle/make_golden_codebook.py). The random draws are std::mt19937_64 through libstdc++'s
uniform_int_distribution / uniform_real_distribution (GCC >= 11 algorithms), restated below.
"""
from __future__ import annotations

import math

import numpy as np

MASK64 = (1 << 64) - 1


class MT19937_64:
    """std::mt19937_64."""
    NN, MM = 312, 156
    MATRIX_A, UM, LM = 0xB5026F5AA96619E9, 0xFFFFFFFF80000000, 0x7FFFFFFF

    def __init__(self, seed: int):
        mt = [seed & MASK64]
        for i in range(1, self.NN):
            mt.append((6364136223846793005 * (mt[-1] ^ (mt[-1] >> 62)) + i) & MASK64)
        self.mt, self.idx = mt, self.NN

    def _twist(self):
        mt, NN, MM = self.mt, self.NN, self.MM
        for i in range(NN):
            x = (mt[i] & self.UM) | (mt[(i + 1) % NN] & self.LM)
            mt[i] = mt[(i + MM) % NN] ^ (x >> 1) ^ (self.MATRIX_A if (x & 1) else 0)
        self.idx = 0

    def __call__(self) -> int:
        if self.idx >= self.NN:
            self._twist()
        x = self.mt[self.idx]
        self.idx += 1
        x ^= (x >> 29) & 0x5555555555555555
        x ^= (x << 17) & 0x71D67FFFEDA60000
        x ^= (x << 37) & 0xFFF7EEE000000000
        x ^= x >> 43
        return x & MASK64


def uniform_int(rng, hi: int) -> int:
    """std::uniform_int_distribution<size_t>(0, hi)(rng) for a 64-bit generator (libstdc++ _S_nd, Lemire)."""
    rng_range = hi + 1
    product = rng() * rng_range
    low = product & MASK64
    if low < rng_range:
        threshold = ((1 << 64) - rng_range) % rng_range
        while low < threshold:
            product = rng() * rng_range
            low = product & MASK64
    return product >> 64


def uniform_real(rng, a: float, b: float) -> float:
    """std::uniform_real_distribution<double>(a, b)(rng): generate_canonical<double, 53> * (b - a) + a."""
    ret = float(rng()) / 18446744073709551616.0
    if ret >= 1.0:
        ret = math.nextafter(1.0, 0.0)
    return ret * (b - a) + a


def nudge_strictly_increasing(c: np.ndarray) -> None:
    for i in range(1, len(c)):
        if c[i] <= c[i - 1]:
            c[i] = c[i - 1] + max(1e-9, abs(c[i - 1]) * 1e-6)


class Prefix:
    def __init__(self, s: np.ndarray):
        v = s.astype(np.float64)
        self.ps = np.concatenate(([0.0], np.cumsum(v)))          # sequential, like the reference's loop
        self.psq = np.concatenate(([0.0], np.cumsum(v * v)))
        self.n = len(s)

    def mean(self, a: int, b: int) -> float:        # [a, b] inclusive
        return float((self.ps[b + 1] - self.ps[a]) / float(b - a + 1))

    def sse(self, a: int, b: int) -> float:
        c, sm = float(b - a + 1), float(self.ps[b + 1] - self.ps[a])
        return float((self.psq[b + 1] - self.psq[a]) - sm * sm / c)


def init_kmeanspp(s: np.ndarray, k: int, seed: int) -> np.ndarray:
    n = len(s)
    rng = MT19937_64(seed)
    s64 = s.astype(np.float64)
    centers = [float(s[uniform_int(rng, n - 1)])]
    d2 = (s64 - centers[0]) ** 2
    while len(centers) < k:
        cum = np.cumsum(d2)                       # acc += d2[i], sequential
        total = float(cum[-1])
        if total <= 0.0:
            chosen = uniform_int(rng, n - 1)
        else:
            target = uniform_real(rng, 0.0, total)
            chosen = int(np.searchsorted(cum, target, side="left"))   # first i with acc >= target
            if chosen >= n:
                chosen = n - 1
        cv = float(s[chosen])
        centers.append(cv)
        d2 = np.minimum(d2, (s64 - cv) ** 2)
    c = np.sort(np.array(centers, dtype=np.float64))
    nudge_strictly_increasing(c)
    return c


def init_equal_mass(pf: Prefix, n: int, k: int) -> np.ndarray:
    c = np.empty(k)
    for i in range(k):
        a, b = (i * n) // k, ((i + 1) * n) // k
        if b <= a:
            b = a + 1
        b = min(b, n)
        c[i] = pf.mean(a, b - 1)
    nudge_strictly_increasing(c)
    return c


def lloyd_k(s: np.ndarray, pf: Prefix, k: int, c: np.ndarray, max_iters: int, tol: float) -> float:
    n = len(s)
    tol = float(np.float32(tol))

    def assign():
        mids = (0.5 * (c[:-1] + c[1:])).astype(np.float32)
        bnd = np.empty(k + 1, dtype=np.int64)
        bnd[0], bnd[k] = 0, n
        bnd[1:k] = np.searchsorted(s, mids, side="right")           # upper_bound on the float mid
        for i in range(1, k):
            if bnd[i] < bnd[i - 1]:
                bnd[i] = bnd[i - 1]
        return bnd
    repairs, budget = 0, 2 * k
    for _ in range(max_iters):
        bnd = assign()
        maxmove = 0.0
        for i in range(k):
            a, b = int(bnd[i]), int(bnd[i + 1])
            nc = pf.mean(a, b - 1) if b > a else float(c[i])
            maxmove = max(maxmove, abs(nc - float(c[i])))
            c[i] = nc
        repaired = False
        if repairs < budget:
            empty = next((i for i in range(k) if bnd[i + 1] <= bnd[i]), k)
            if empty < k:
                best, bsse = k, -1.0
                for j in range(k):
                    a, b = int(bnd[j]), int(bnd[j + 1])
                    if b > a + 1:
                        e = pf.sse(a, b - 1)
                        if e > bsse:
                            bsse, best = e, j
                if best < k:
                    a, b = int(bnd[best]), int(bnd[best + 1])
                    m = a + (b - a) // 2
                    c[best] = pf.mean(a, m - 1)
                    c[empty] = pf.mean(m, b - 1)
                    c.sort()
                    nudge_strictly_increasing(c)
                    repairs += 1
                    repaired = True
        if repaired:
            continue
        if maxmove < tol:
            break
    bnd = assign()
    return float(sum(pf.sse(int(bnd[i]), int(bnd[i + 1]) - 1) for i in range(k) if bnd[i + 1] > bnd[i]))


def sample_column(values: np.ndarray, sample_size: int, seed: int) -> np.ndarray:
    """The partial Fisher-Yates draw of build_codebook_lloyd (codebook_builder.cpp:350-364)."""
    n = len(values)
    if not (0 < sample_size < n):
        return np.array(values, dtype=np.float32)
    rng = MT19937_64(seed)
    idx = list(range(n))
    out = np.empty(sample_size, dtype=np.float32)
    for i in range(sample_size):
        j = i + uniform_int(rng, n - 1 - i)
        idx[i], idx[j] = idx[j], idx[i]
        out[i] = values[idx[i]]
    return out


def recommended_sample_size(n: int, max_bits: int) -> int:
    lower = max(200_000, 500 * (1 << max_bits))
    return n if n < lower else lower


def build_codebook_lloyd(values: np.ndarray, max_bits: int, init: str = "kmeanspp", restarts: int = 1,
                         max_iters: int = 50, tol: float = 1e-6, seed: int = 0, sample_size: int = 0,
                         levels=None):
    """Returns (costs float32 [max_bits + 1], codebooks: list of float32 arrays of 2^b entries (level 0: 1)).
    levels: iterable of bit-rates to build (default all); the others are left None (each level is independent)."""
    s = np.sort(sample_column(np.asarray(values, dtype=np.float32), sample_size, seed))
    ns = len(s)
    pf = Prefix(s)
    ndist = 1 + int(np.count_nonzero(s[1:] != s[:-1]))
    costs = np.zeros(max_bits + 1, dtype=np.float32)
    books = [None] * (max_bits + 1)
    costs[0] = np.float32(pf.sse(0, ns - 1) / ns)
    books[0] = np.array([np.float32(pf.mean(0, ns - 1))], dtype=np.float32)
    for bits in range(1, max_bits + 1):
        if levels is not None and bits not in levels:
            continue
        k = 1 << bits
        if k >= ndist:
            books[bits] = s[np.concatenate(([True], s[1:] != s[:-1]))].copy()
            costs[bits] = 0.0
            continue
        best_sse, best_c = math.inf, None
        for r in range(max(1, restarts)):
            c = init_kmeanspp(s, k, seed + r) if init == "kmeanspp" else init_equal_mass(pf, ns, k)
            sse = lloyd_k(s, pf, k, c, max_iters, tol)
            if sse < best_sse:
                best_sse, best_c = sse, c.copy()
        cen = np.unique(np.sort(best_c.astype(np.float32)))
        if len(cen) < k:
            cen = np.concatenate([cen, np.full(k - len(cen), cen[-1], dtype=np.float32)])
        books[bits] = cen
        costs[bits] = np.float32(best_sse / ns)
    return costs, books


def allocate_greedy(mse_table: np.ndarray, dim_padding: int, max_bits: int, total_bits: int, num_bit_factors: int):
    """BitAllocatorGreedy::AllocateJoint -> [(dim_len, bits)] (bit_allocator_greedy.cpp:11-116)."""
    d_pad = mse_table.shape[0]
    nb = d_pad // dim_padding
    block = np.zeros((nb, max_bits + 1))
    for i in range(nb):
        for b in range(max_bits + 1):
            sacc = 0.0
            for kk in range(dim_padding):
                sacc += float(mse_table[i * dim_padding + kk, b])
            block[i, b] = sacc

    def usage(bits):
        segs, i = 0, 0
        while i < nb:
            j = i + 1
            while j < nb and bits[j] == bits[i]:
                j += 1
            if bits[i] > 0:
                segs += 1
            i = j
        return int(sum(bits)) * dim_padding + segs * num_bit_factors
    bits = [0] * nb
    while True:
        best_i, best_drop = -1, -math.inf
        for i in range(nb):
            if bits[i] >= max_bits:
                continue
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
        plan.append(((j - i) * dim_padding, bits[i]))
        i = j
    return plan
