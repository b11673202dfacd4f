// SAQ engine path on the device: CAQ segment encoder, bit packing, decode, and the host-side
// joint segmentation / bit-allocation DP.
//
// Replaces, for one segment of the SAQ plan:
//   external/saq/include/saq/caq_encoder.h:170-218,67-140,220-261   CAQEncoder::encode /
//       code_adjustment / encode_and_fac  (driven by quantizer.h:56-104, one call per row)
//   external/saq/src/ivf_index.cpp:296-338                          per-segment dequantisation
//   external/saq/src/quantization_plan.cpp:144-255                  SaqDataMaker::dynamic_programming
//
// Encoder layout. The reference walks the dimensions of a row SEQUENTIALLY (every +-1 adjustment
// changes the running <o,oa> and |oa|^2 the next test reads), in fp64, for up to 6 sweeps. That
// chain cannot be split across lanes without changing the codes, so one THREAD owns one row and
// the parallelism is over rows (10^6 rows x segments). For that mapping to be coalesced the
// residuals are consumed TRANSPOSED, oT[j][row] (the rotation GEMM that produces them writes
// either orientation at the same cost), and the working codes live in a transposed uint16
// scratch; a second kernel packs them into the row-major MSB-first bit string kept in HBM.
// Every fp32/fp64 operation uses the round-to-nearest intrinsics so no multiply-add is
// contracted: the codes are bit-identical to the CPU order of operations (oracle/caq.c).
#include <vector>
#include <algorithm>
#include <limits>

#include "common.cuh"

namespace b200vq {
namespace caq {

__device__ __forceinline__ double dmul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double dadd(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double dsub(double a, double b) { return __dsub_rn(a, b); }

// One thread = one row of the segment. oT: (len, ld) with element (j, row) at oT[j * ld + row].
__device__ __forceinline__ void
caq_encode_row(const int64_t i, const float* __restrict__ oT, int64_t ld, int len,
               const float* __restrict__ c_rot, int bits, int rd_lmt, float adj_eps,
               uint16_t* __restrict__ codesT, int64_t ldc, float* __restrict__ l2norm,
               float* __restrict__ rescale, float* __restrict__ error, float* __restrict__ mse,
               float* __restrict__ cosv) {
    const float* op = oT + i;
    uint16_t* cp = codesT + i;
    auto O = [&](int j) -> float {
        const float v = __ldg(op + (int64_t)j * ld);
        return c_rot ? __fsub_rn(v, __ldg(c_rot + j)) : v;
    };
    float mx = O(0), mn = mx, fsum = 0.f, fsq = 0.f;
    for (int j = 0; j < len; ++j) {
        const float v = O(j);
        mx = fmaxf(mx, v);
        mn = fminf(mn, v);
        fsum = __fadd_rn(fsum, v);
        fsq = __fadd_rn(fsq, __fmul_rn(v, v));
    }
    const double o_l2sqr = (double)fsq;
    if (bits == 0) {
        for (int j = 0; j < len; ++j) cp[(int64_t)j * ldc] = 0;
        l2norm[i] = (float)sqrt(o_l2sqr);
        rescale[i] = 0.f;
        error[i] = 0.f;
        mse[i] = __fdiv_rn(fsq, (float)len);
        cosv[i] = 0.f;
        return;
    }
    const int code_max = (1 << bits) - 1;
    const double v_mx = (double)((mx > -mn) ? mx : -mn);
    const double v_mi = -v_mx;
    const double delta = dsub(v_mx, v_mi) / (double)(code_max + 1);
    const float fmi = (float)v_mi, fdelta = (float)delta;
    double ip_o_code = 0.0;
    unsigned long long code_l2sqr = 0;
    uint32_t code_sum = 0;
    for (int j = 0; j < len; ++j) {
        const float v = O(j);
        float cf = 0.f;
        if (delta != 0.0) {
            cf = floorf(__fdiv_rn(__fsub_rn(v, fmi), fdelta));
            if (cf > (float)code_max) cf = (float)code_max;
        }
        ip_o_code = dadd(ip_o_code, (double)__fmul_rn(cf, v));
        const int c = (int)cf;
        cp[(int64_t)j * ldc] = (uint16_t)c;
        code_l2sqr += (unsigned long long)(c * c);
        code_sum += (uint32_t)c;
    }
    const double dd = dmul(delta, delta);
    double ip = dadd(dmul(ip_o_code, delta), dmul(dadd(v_mi, dmul(0.5, delta)), (double)fsum));
    double oa2 = dadd(dmul(dd, (double)code_l2sqr),
                      dmul(dadd(dd, dmul(dmul(2.0, delta), v_mi)), (double)code_sum));
    oa2 = dadd(oa2, dmul(dadd(dadd(dmul(0.25, dd), dmul(delta, v_mi)), dmul(v_mi, v_mi)), (double)len));

    if (rd_lmt && oa2 != 0.0) {
        const double re_eps = dmul((double)adj_eps, oa2);
        for (int adj_cnt = 1, round = 1; adj_cnt; round++) {
            adj_cnt = 0;
            for (int j = 0; j < len; ++j) {
                const float of = O(j);
                const int c0 = cp[(int64_t)j * ldc];
                int c = c0, curr = 0;
                double oa = dadd(dmul(dadd((double)c0, 0.5), delta), v_mi);
                const double tmp = dsub(oa2, dmul(oa, oa));
                const double ip_delta = dmul(delta, (double)of);
                while (c < code_max) {
                    const double nq = dadd(oa, delta);
                    const double nl = dadd(tmp, dmul(nq, nq));
                    const double nip = dadd(ip, ip_delta);
                    if (dmul(dadd(dmul(ip, ip), re_eps), nl) >= dmul(dmul(nip, nip), oa2)) break;
                    c++; ip = nip; oa = nq; oa2 = nl; curr++;
                }
                while (c > 0) {
                    const double nq = dsub(oa, delta);
                    const double nl = dadd(tmp, dmul(nq, nq));
                    const double nip = dsub(ip, ip_delta);
                    if (dmul(dadd(dmul(ip, ip), re_eps), nl) >= dmul(dmul(nip, nip), oa2)) break;
                    c--; ip = nip; oa = nq; oa2 = nl; curr++;
                }
                if (c != c0) {
                    cp[(int64_t)j * ldc] = (uint16_t)c;
                    adj_cnt += curr;
                }
            }
            if (rd_lmt && round >= rd_lmt) break;
            double cl2 = 0.0, cip = 0.0;
            for (int j = 0; j < len; ++j) {
                const double q = dadd(dmul(dadd((double)cp[(int64_t)j * ldc], 0.5), delta), v_mi);
                cip = dadd(cip, dmul(q, (double)O(j)));
                cl2 = dadd(cl2, dmul(q, q));
            }
            oa2 = cl2;
            ip = cip;
        }
    }
    const double o_l2norm = sqrt(o_l2sqr);
    double fac_rescale = (ip != 0.0) ? o_l2sqr / ip : 0.0;
    const double fac_error =
        dmul(dmul(o_l2sqr, (double)1.9f), sqrt(dsub(dmul(o_l2sqr, oa2) / dmul(ip, ip), 1.0) / (double)(len - 1)));
    float diff = 0.f;
    for (int j = 0; j < len; ++j) {
        const float oa = __fadd_rn(__fmul_rn((float)cp[(int64_t)j * ldc], (float)delta), (float)v_mi);
        const float d = __fsub_rn(O(j), oa);
        diff = __fadd_rn(diff, __fmul_rn(d, d));
    }
    mse[i] = __fdiv_rn(diff, (float)len);
    cosv[i] = (bits > 1) ? (float)((ip / o_l2norm) / sqrt(oa2)) : 0.f;
    if (v_mx != 0.0) fac_rescale = dmul(fac_rescale, v_mx);
    l2norm[i] = (float)o_l2norm;
    rescale[i] = (float)fac_rescale;
    error[i] = (float)fac_error;
}

// ---------------------------------------------------------------------------------------------
// Warp-tiled form of the same encoder. The per-row arithmetic (order, rounding) is caq_encode_row's, but the
// residuals and working codes of a warp's 32 rows stream through shared memory in tiles of 32 dimensions,
// brought in by cp.async one tile ahead: the sweep is a latency-bound fp64 chain, and with the loads issued
// from inside it (the store of dimension j may alias the load of j + 1 as far as the compiler knows) every
// step used to wait on L2 / DRAM (ncu: 36 % of the samples on the code load alone). The exact re-accumulation of
// <o, oa> and |oa|^2 that follows every sweep is folded into the sweep (dimension j's code is final once its
// step is done, and the sums run over j in the same order), which saves a pass over the segment per round.
constexpr int CW_TILE = 16;
struct CaqWarpTile {
    float o[2][CW_TILE][32];
    uint16_t c[2][CW_TILE][32];
};
__device__ __forceinline__ void cp_async4(void* smem_dst, const void* gsrc, bool valid) {
    const uint32_t d = static_cast<uint32_t>(__cvta_generic_to_shared(smem_dst));
    const int sz = valid ? 4 : 0;
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(d), "l"(gsrc), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// One warp = rows i0 .. i0 + 31 of one segment (i0 % 32 == 0). Lanes past n run on row n - 1's data and store nothing.
__device__ void caq_encode_rows_warp(CaqWarpTile& T, const int64_t i0, const int64_t n, const float* __restrict__ oT,
                                     int64_t ld, int len, const float* __restrict__ c_rot, int bits, int rd_lmt,
                                     float adj_eps, uint16_t* __restrict__ codesT, int64_t ldc,
                                     float* __restrict__ l2norm, float* __restrict__ rescale, float* __restrict__ error,
                                     float* __restrict__ mse, float* __restrict__ cosv) {
    const int lane = threadIdx.x & 31;
    const int64_t i = i0 + lane;
    const bool live = i < n;
    const int ntile = (len + CW_TILE - 1) / CW_TILE;
    const bool pair_codes = ((ldc & 1) == 0) && ((reinterpret_cast<uintptr_t>(codesT) & 3) == 0);
    // tile t of the residuals (and, with_codes, of the working codes) -> buffer t & 1, as one cp.async group
    auto load_tile = [&](int t, bool with_codes) {
        const int buf = t & 1;
        if (t < ntile) {
            const int j0 = t * CW_TILE;
            const int nj = min(CW_TILE, len - j0);
            for (int jj = 0; jj < nj; ++jj) {
                cp_async4(&T.o[buf][jj][lane], oT + (int64_t)(j0 + jj) * ld + (live ? i : n - 1), true);
                if (with_codes && pair_codes && lane < 16)
                    cp_async4(&T.c[buf][jj][2 * lane], codesT + (int64_t)(j0 + jj) * ldc + i0 + 2 * lane,
                              i0 + 2 * lane < ldc);
            }
        }
        cp_async_commit();
    };
    auto finish_tile_load = [&](int t, bool with_codes) {     // codes of an odd-stride scratch: plain loads
        if (with_codes && !pair_codes) {
            const int buf = t & 1, j0 = t * CW_TILE, nj = min(CW_TILE, len - j0);
            for (int jj = 0; jj < nj; ++jj) T.c[buf][jj][lane] = live ? codesT[(int64_t)(j0 + jj) * ldc + i] : (uint16_t)0;
        }
    };
    auto store_codes = [&](int t) {
        const int buf = t & 1, j0 = t * CW_TILE, nj = min(CW_TILE, len - j0);
        if (live)
            for (int jj = 0; jj < nj; ++jj) codesT[(int64_t)(j0 + jj) * ldc + i] = T.c[buf][jj][lane];
    };
    // walk the segment: body(j, o_j, code slot) for j = 0 .. len-1 in order
    auto walk = [&](bool read_codes, bool write_codes, auto&& body) {
        __syncwarp();
        load_tile(0, read_codes);
        for (int t = 0; t < ntile; ++t) {
            load_tile(t + 1, read_codes);
            cp_async_wait<1>();
            __syncwarp();
            finish_tile_load(t, read_codes);
            const int buf = t & 1, j0 = t * CW_TILE, nj = min(CW_TILE, len - j0);
            for (int jj = 0; jj < nj; ++jj) {
                float v = T.o[buf][jj][lane];
                if (c_rot) v = __fsub_rn(v, __ldg(c_rot + j0 + jj));
                body(j0 + jj, v, T.c[buf][jj][lane]);
            }
            if (write_codes) store_codes(t);
            __syncwarp();
        }
        cp_async_wait<0>();
    };

    float mx = 0.f, mn = 0.f, fsum = 0.f, fsq = 0.f;
    walk(false, false, [&](int j, float v, uint16_t&) {
        if (j == 0) { mx = v; mn = v; }
        mx = fmaxf(mx, v);
        mn = fminf(mn, v);
        fsum = __fadd_rn(fsum, v);
        fsq = __fadd_rn(fsq, __fmul_rn(v, v));
    });
    const double o_l2sqr = (double)fsq;
    if (bits == 0) {
        if (live) {
            for (int j = 0; j < len; ++j) codesT[(int64_t)j * ldc + i] = 0;
            l2norm[i] = (float)sqrt(o_l2sqr);
            rescale[i] = 0.f;
            error[i] = 0.f;
            mse[i] = __fdiv_rn(fsq, (float)len);
            cosv[i] = 0.f;
        }
        return;
    }
    const int code_max = (1 << bits) - 1;
    const double v_mx = (double)((mx > -mn) ? mx : -mn);
    const double v_mi = -v_mx;
    const double delta = dsub(v_mx, v_mi) / (double)(code_max + 1);
    const float fmi = (float)v_mi, fdelta = (float)delta;
    double ip_o_code = 0.0;
    unsigned long long code_l2sqr = 0;
    uint32_t code_sum = 0;
    walk(false, true, [&](int, float v, uint16_t& slot) {
        float cf = 0.f;
        if (delta != 0.0) {
            cf = floorf(__fdiv_rn(__fsub_rn(v, fmi), fdelta));
            if (cf > (float)code_max) cf = (float)code_max;
        }
        ip_o_code = dadd(ip_o_code, (double)__fmul_rn(cf, v));
        const int c = (int)cf;
        slot = (uint16_t)c;
        code_l2sqr += (unsigned long long)(c * c);
        code_sum += (uint32_t)c;
    });
    const double dd = dmul(delta, delta);
    double ip = dadd(dmul(ip_o_code, delta), dmul(dadd(v_mi, dmul(0.5, delta)), (double)fsum));
    double oa2 = dadd(dmul(dd, (double)code_l2sqr),
                      dmul(dadd(dd, dmul(dmul(2.0, delta), v_mi)), (double)code_sum));
    oa2 = dadd(oa2, dmul(dadd(dadd(dmul(0.25, dd), dmul(delta, v_mi)), dmul(v_mi, v_mi)), (double)len));

    // lanes leave the adjustment at different rounds; the warp keeps walking (cooperative loads) while any lane is in
    bool adjusting = live && rd_lmt && oa2 != 0.0;
    const double re_eps = dmul((double)adj_eps, oa2);
    for (int round = 1; __any_sync(FULL_MASK, adjusting); ++round) {
        int adj_cnt = 0;
        double cl2 = 0.0, cip = 0.0;
        walk(true, true, [&](int, float of, uint16_t& slot) {
            if (!adjusting) return;
            const int c0 = slot;
            int c = c0, curr = 0;
            double oa = dadd(dmul(dadd((double)c0, 0.5), delta), v_mi);
            const double tmp = dsub(oa2, dmul(oa, oa));
            const double ip_delta = dmul(delta, (double)of);
            while (c < code_max) {
                const double nq = dadd(oa, delta);
                const double nl = dadd(tmp, dmul(nq, nq));
                const double nip = dadd(ip, ip_delta);
                if (dmul(dadd(dmul(ip, ip), re_eps), nl) >= dmul(dmul(nip, nip), oa2)) break;
                c++; ip = nip; oa = nq; oa2 = nl; curr++;
            }
            while (c > 0) {
                const double nq = dsub(oa, delta);
                const double nl = dadd(tmp, dmul(nq, nq));
                const double nip = dsub(ip, ip_delta);
                if (dmul(dadd(dmul(ip, ip), re_eps), nl) >= dmul(dmul(nip, nip), oa2)) break;
                c--; ip = nip; oa = nq; oa2 = nl; curr++;
            }
            if (c != c0) {
                slot = (uint16_t)c;
                adj_cnt += curr;
            }
            // exact re-accumulation over the round's final codes (the reference's separate pass, same order)
            const double q = dadd(dmul(dadd((double)c, 0.5), delta), v_mi);
            cip = dadd(cip, dmul(q, (double)of));
            cl2 = dadd(cl2, dmul(q, q));
        });
        if (adjusting) {
            if (rd_lmt && round >= rd_lmt) {
                adjusting = false;
            } else {
                oa2 = cl2;
                ip = cip;
                if (adj_cnt == 0) adjusting = false;
            }
        }
    }
    const double o_l2norm = sqrt(o_l2sqr);
    double fac_rescale = (ip != 0.0) ? o_l2sqr / ip : 0.0;
    const double fac_error =
        dmul(dmul(o_l2sqr, (double)1.9f), sqrt(dsub(dmul(o_l2sqr, oa2) / dmul(ip, ip), 1.0) / (double)(len - 1)));
    float diff = 0.f;
    walk(true, false, [&](int, float ov, uint16_t& slot) {
        const float oa = __fadd_rn(__fmul_rn((float)slot, (float)delta), (float)v_mi);
        const float d = __fsub_rn(ov, oa);
        diff = __fadd_rn(diff, __fmul_rn(d, d));
    });
    if (live) {
        mse[i] = __fdiv_rn(diff, (float)len);
        cosv[i] = (bits > 1) ? (float)((ip / o_l2norm) / sqrt(oa2)) : 0.f;
        if (v_mx != 0.0) fac_rescale = dmul(fac_rescale, v_mx);
        l2norm[i] = (float)o_l2norm;
        rescale[i] = (float)fac_rescale;
        error[i] = (float)fac_error;
    }
}

__global__ void __launch_bounds__(128)
caq_encode_t_kernel(const float* __restrict__ oT, int64_t n, int64_t ld, int len,
                    const float* __restrict__ c_rot, int bits, int rd_lmt, float adj_eps,
                    uint16_t* __restrict__ codesT, int64_t ldc, float* __restrict__ l2norm,
                    float* __restrict__ rescale, float* __restrict__ error, float* __restrict__ mse,
                    float* __restrict__ cosv) {
    extern __shared__ __align__(16) unsigned char caq_smem[];
    CaqWarpTile& T = reinterpret_cast<CaqWarpTile*>(caq_smem)[threadIdx.x >> 5];
    const int64_t i0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) & ~(int64_t)31;
    if (i0 >= n) return;
    caq_encode_rows_warp(T, i0, n, oT, ld, len, c_rot, bits, rd_lmt, adj_eps, codesT, ldc, l2norm, rescale, error, mse, cosv);
}

// Every segment of the plan in ONE grid (blockIdx.y = segment, longest first): the rows of a segment are a
// latency-bound fp64 chain each, so the segments' chains overlap instead of running one launch after another.
constexpr int CAQ_MAX_SEGS = 64;
struct CaqSegs {
    int nseg;
    int off[CAQ_MAX_SEGS], len[CAQ_MAX_SEGS], bits[CAQ_MAX_SEGS], slot[CAQ_MAX_SEGS];
};
__global__ void __launch_bounds__(128)
caq_encode_t_segments_kernel(const float* __restrict__ oT, int64_t n, int64_t ld, const __grid_constant__ CaqSegs segs,
                             int rd_lmt, float adj_eps, uint16_t* __restrict__ codesT, int64_t ldc,
                             float* __restrict__ l2norm, float* __restrict__ rescale, float* __restrict__ error,
                             float* __restrict__ mse, float* __restrict__ cosv) {
    extern __shared__ __align__(16) unsigned char caq_smem[];
    CaqWarpTile& T = reinterpret_cast<CaqWarpTile*>(caq_smem)[threadIdx.x >> 5];
    const int64_t i0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) & ~(int64_t)31;
    if (i0 >= n) return;
    const int s = blockIdx.y;
    const int off = segs.off[s], slot = segs.slot[s];
    const int64_t fo = (int64_t)slot * n;
    caq_encode_rows_warp(T, i0, n, oT + (int64_t)off * ld, ld, segs.len[s], nullptr, segs.bits[s], rd_lmt, adj_eps,
                         codesT + (int64_t)off * ldc, ldc, l2norm + fo, rescale + fo, error + fo, mse + fo, cosv + fo);
}

// codesT (len, ldc) uint16 -> packed (n, row_bytes): MSB-first bit string, `bits` per dimension.
// Block = 32 rows x a 256-byte slice of the row; reads are coalesced over rows, the tile is
// turned in shared memory and written out row-contiguous.
constexpr int PACK_ROWS = 32, PACK_BYTES = 256;
__global__ void __launch_bounds__(256)
caq_pack_t_kernel(const uint16_t* __restrict__ codesT, int64_t n, int64_t ldc, int len, int bits,
                  uint8_t* __restrict__ packed, int row_bytes) {
    __shared__ uint8_t tile[PACK_ROWS][PACK_BYTES + 4];
    const int64_t row0 = (int64_t)blockIdx.x * PACK_ROWS;
    const int byte0 = blockIdx.y * PACK_BYTES;
    const int r = threadIdx.x & 31;
    const int64_t row = row0 + r;
    for (int kb = threadIdx.x >> 5; kb < PACK_BYTES; kb += 8) {
        const int k = byte0 + kb;
        uint32_t acc = 0;
        if (k < row_bytes && row < n) {
            const int b0 = k * 8, b1 = b0 + 8;
            for (int d = b0 / bits; d < len && d * bits < b1; ++d) {
                const uint32_t q = codesT[(int64_t)d * ldc + row];
                const int shift = b1 - (d * bits + bits);
                acc |= (shift >= 0) ? (q << shift) : (q >> (-shift));
            }
        }
        tile[r][kb] = (uint8_t)acc;
    }
    __syncthreads();
    const int nb = min(PACK_BYTES, row_bytes - byte0);
    for (int e = threadIdx.x; e < PACK_ROWS * nb; e += 256) {
        const int rr = e / nb, kk = e - rr * nb;
        if (row0 + rr < n) packed[(row0 + rr) * (int64_t)row_bytes + byte0 + kk] = tile[rr][kk];
    }
}

__device__ __forceinline__ uint32_t read_code(const uint8_t* __restrict__ rowp, int d, int bits, int row_bytes) {
    const int off = d * bits, byte = off >> 3;
    uint32_t w = (uint32_t)rowp[byte] << 16;
    if (byte + 1 < row_bytes) w |= (uint32_t)rowp[byte + 1] << 8;
    if (byte + 2 < row_bytes) w |= (uint32_t)rowp[byte + 2];
    return (w >> (24 - (off & 7) - bits)) & ((1u << bits) - 1u);
}

// Per-segment dequantisation (ivf_index.cpp:296-338): oa = (code + 0.5) * 2/2^b - 1, rescaled to
// the stored residual norm. One warp per output row. mode 0: scaled oa (reconstruction);
// mode 1: raw integer codes as fp32 (estimator operand).
__global__ void __launch_bounds__(256)
caq_decode_kernel(const uint8_t* __restrict__ packed, int row_bytes, const int64_t* __restrict__ ids, int64_t m,
                  int len, int bits, const float* __restrict__ l2norm, int mode, float* __restrict__ out,
                  int64_t ldo) {
    const int lane = threadIdx.x & 31;
    const int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (w >= m) return;
    const int64_t id = ids ? ids[w] : w;
    const uint8_t* rowp = packed + id * (int64_t)row_bytes;
    float* o = out + w * ldo;
    if (mode == 1) {
        for (int d = lane; d < len; d += 32) o[d] = (float)read_code(rowp, d, bits, row_bytes);
        return;
    }
    const float delta = 2.0f / (float)(1u << bits);
    float ss = 0.f;
    for (int d = lane; d < len; d += 32) {
        const float oa = __fadd_rn(__fmul_rn(__fadd_rn((float)read_code(rowp, d, bits, row_bytes), 0.5f), delta), -1.0f);
        ss = __fmaf_rn(oa, oa, ss);
    }
    for (int s = 16; s > 0; s >>= 1) ss += __shfl_xor_sync(FULL_MASK, ss, s);
    const float nrm = sqrtf(ss);
    const float scale = (nrm > 1e-9f) ? __fdiv_rn(l2norm[id], nrm) : 1.0f;
    for (int d = lane; d < len; d += 32) {
        const float oa = __fadd_rn(__fmul_rn(__fadd_rn((float)read_code(rowp, d, bits, row_bytes), 0.5f), delta), -1.0f);
        o[d] = __fmul_rn(oa, scale);
    }
}

}  // namespace caq
}  // namespace b200vq

using namespace b200vq;
using namespace b200vq::caq;

extern "C" int b200vq_caq_encode_t(const float* oT, int64_t n, int64_t ld, int len, const float* c_rot, int bits,
                                   int rd_lmt, float adj_eps, uint16_t* codesT, int64_t ldc, float* l2norm,
                                   float* rescale, float* error, float* mse, float* cosv, void* stream) {
    VQ_CHECK_ARG(n >= 0 && len > 1 && bits >= 0 && bits <= 13 && ld >= n && ldc >= n && rd_lmt >= 0,
                 "caq_encode_t: bad args");
    VQ_CHECK_ARG(n == 0 || (oT && codesT && l2norm && rescale && error && mse && cosv), "caq_encode_t: null pointer");
    if (n == 0) return B200VQ_OK;
    const int block = 128;
    const size_t smem = (block / 32) * sizeof(caq::CaqWarpTile);
    VQ_CUDA(cudaFuncSetAttribute(caq::caq_encode_t_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    caq_encode_t_kernel<<<(unsigned)((n + block - 1) / block), block, smem, as_stream(stream)>>>(
        oT, n, ld, len, c_rot, bits, rd_lmt, adj_eps, codesT, ldc, l2norm, rescale, error, mse, cosv);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" int b200vq_caq_encode_t_segments(const float* oT, int64_t n, int64_t ld, int nseg, const int* seg_off,
                                            const int* seg_len, const int* seg_bits, int rd_lmt, float adj_eps,
                                            uint16_t* codesT, int64_t ldc, float* l2norm, float* rescale,
                                            float* error, float* mse, float* cosv, void* stream) {
    VQ_CHECK_ARG(n >= 0 && nseg >= 1 && nseg <= CAQ_MAX_SEGS && seg_off && seg_len && seg_bits && ld >= n && ldc >= n &&
                     rd_lmt >= 0,
                 "caq_encode_t_segments: bad args (1 <= nseg <= %d)", CAQ_MAX_SEGS);
    VQ_CHECK_ARG(n == 0 || (oT && codesT && l2norm && rescale && error && mse && cosv), "caq_encode_t_segments: null pointer");
    if (n == 0) return B200VQ_OK;
    CaqSegs segs;
    segs.nseg = nseg;
    std::vector<int> order(nseg);
    for (int s = 0; s < nseg; ++s) {
        VQ_CHECK_ARG(seg_len[s] > 1 && seg_bits[s] >= 0 && seg_bits[s] <= 13 && seg_off[s] >= 0,
                     "caq_encode_t_segments: bad segment %d", s);
        order[s] = s;
    }
    // longest chains first: cost ~ len * (1 + sweeps), sweeps only when bits > 0
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) {
        return (long long)seg_len[a] * (seg_bits[a] ? 8 : 1) > (long long)seg_len[b] * (seg_bits[b] ? 8 : 1);
    });
    for (int k = 0; k < nseg; ++k) {
        const int s = order[k];
        segs.off[k] = seg_off[s]; segs.len[k] = seg_len[s]; segs.bits[k] = seg_bits[s]; segs.slot[k] = s;
    }
    const int block = 128;
    dim3 grid((unsigned)((n + block - 1) / block), (unsigned)nseg);
    const size_t smem = (block / 32) * sizeof(caq::CaqWarpTile);
    VQ_CUDA(cudaFuncSetAttribute(caq::caq_encode_t_segments_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    caq_encode_t_segments_kernel<<<grid, block, smem, as_stream(stream)>>>(oT, n, ld, segs, rd_lmt, adj_eps, codesT, ldc,
                                                                       l2norm, rescale, error, mse, cosv);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" int b200vq_caq_pack_t(const uint16_t* codesT, int64_t n, int64_t ldc, int len, int bits, uint8_t* packed,
                                 void* stream) {
    VQ_CHECK_ARG(n >= 0 && len > 0 && bits >= 1 && bits <= 13 && ldc >= n && (len * bits) % 8 == 0,
                 "caq_pack_t: bad args");
    VQ_CHECK_ARG(n == 0 || (codesT && packed), "caq_pack_t: null pointer");
    if (n == 0) return B200VQ_OK;
    const int row_bytes = len * bits / 8;
    dim3 grid((unsigned)((n + PACK_ROWS - 1) / PACK_ROWS), (unsigned)((row_bytes + PACK_BYTES - 1) / PACK_BYTES));
    caq_pack_t_kernel<<<grid, 256, 0, as_stream(stream)>>>(codesT, n, ldc, len, bits, packed, row_bytes);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" int b200vq_caq_decode(const uint8_t* packed, int64_t n_total, const int64_t* ids, int64_t m, int len,
                                 int bits, const float* l2norm, int mode, float* out, int64_t ldo, void* stream) {
    VQ_CHECK_ARG(m >= 0 && n_total >= 0 && len > 0 && bits >= 1 && bits <= 13 && ldo >= len &&
                     (len * bits) % 8 == 0 && (mode == 0 || mode == 1),
                 "caq_decode: bad args");
    VQ_CHECK_ARG(m == 0 || (packed && out && (mode == 1 || l2norm)), "caq_decode: null pointer");
    if (m == 0) return B200VQ_OK;
    const int64_t threads = m * 32;
    caq_decode_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, as_stream(stream)>>>(
        packed, len * bits / 8, ids, m, len, bits, l2norm, mode, out, ldo);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

// Host-side plan: joint DP over 64-dimension blocks (segment boundaries) and bits per segment,
// cost = sum of segment variance / 2^bits, budget avg_bits * d_pad + 64 bits, 64 factor bits
// charged per quantised segment. Only reachable states are expanded; the table keeps, per
// (segments used, blocks covered, bits used), the best cost and the (previous block, bits) step.
extern "C" int b200vq_saq_plan_dp(const float* variance, int d_pad, float avg_bits, int* seg_len, int* seg_bits,
                                  int* n_segments) {
    VQ_CHECK_ARG(variance && seg_len && seg_bits && n_segments && d_pad > 0 && d_pad % 64 == 0 && avg_bits >= 0,
                 "saq_plan_dp: bad args");
    const size_t kBlk = 64, kFac = 2 * sizeof(float) * 8;
    const int kMaxBits = 13;
    const size_t budget = static_cast<size_t>(avg_bits * d_pad + kFac);
    const size_t nblk = (size_t)d_pad / kBlk;
    const size_t seg_cap = std::max<size_t>(1, avg_bits < 2 ? nblk : nblk / 2);
    const double kInf = std::numeric_limits<double>::max();
    struct Cell { double cost; uint32_t step; };
    const size_t width = budget + 1;
    std::vector<Cell> tab;
    try {
        tab.assign((seg_cap + 2) * (nblk + 1) * width, Cell{kInf, 0u});
    } catch (const std::bad_alloc&) {
        set_error("saq_plan_dp: table of %zu cells does not fit in host memory", (seg_cap + 2) * (nblk + 1) * width);
        return B200VQ_ERR_INVALID;
    }
    auto at = [&](size_t s, size_t i, size_t b) -> Cell& { return tab[(s * (nblk + 1) + i) * width + b]; };
    std::vector<float> blk_var(nblk);
    for (size_t k = 0; k < nblk; ++k) {
        float a = 0.f;
        for (size_t t = 0; t < kBlk; ++t) a += variance[k * kBlk + t];
        blk_var[k] = a;
    }
    size_t best_s = 0, best_i = nblk, best_b = budget;
    at(0, 0, 0) = Cell{0.0, 0u};
    for (size_t s = 0; s <= seg_cap; ++s)
        for (size_t i = 0; i <= nblk; ++i)
            for (size_t used = 0; used <= budget; ++used) {
                const double here = at(s, i, used).cost;
                if (!(here < kInf)) continue;
                if (i == nblk) {
                    if (here * 1.01 < at(best_s, best_i, best_b).cost) { best_s = s; best_i = i; best_b = used; }
                    continue;
                }
                if (s == seg_cap) continue;
                double span_var = 0.0;
                for (size_t j = 1; i + j <= nblk; ++j) {
                    span_var += blk_var[i + j - 1];
                    for (int b = 1; b <= kMaxBits; ++b) {
                        const size_t need = used + (size_t)b * j * kBlk + kFac;
                        if (need > budget) break;
                        const double cand = here + span_var / (double)(1 << b);
                        Cell& dst = at(s + 1, i + j, need);
                        if (dst.cost > cand) dst = Cell{cand, (uint32_t)((i << 4) + (size_t)b)};
                    }
                }
                Cell& tail = at(s + 1, nblk, used);   // leave the remaining blocks unquantised
                if (here + span_var < tail.cost) tail = Cell{here + span_var, (uint32_t)(i << 4)};
            }
    int ns = 0;
    for (size_t s = best_s, i = best_i, b = best_b; i > 0;) {
        const uint32_t step = at(s, i, b).step;
        const size_t prev = step >> 4, bits = step & 0xf, ln = (i - prev) * kBlk;
        seg_len[ns] = (int)ln;
        seg_bits[ns] = (int)bits;
        ++ns;
        --s;
        i = prev;
        if (bits) b -= bits * ln + kFac;
    }
    std::reverse(seg_len, seg_len + ns);
    std::reverse(seg_bits, seg_bits + ns);
    *n_segments = ns;
    return B200VQ_OK;
}
