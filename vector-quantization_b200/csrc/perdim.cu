// Per-dimension codebook quantizer: nearest-level search + variable-width bit packing (sm_100a).
//
// Data path of RankAwareQuantizer (src/haag_vq/methods/rank_aware_quantization.py): `_quantize_dim`
// (:229-235: np.searchsorted of the column into the mid-points of its sorted codebook), the dense
// MSB-first bit string of `_compress_block` (:262-279, np.packbits) or the FFD byte layout (:252-260,
// ffd_packing.py:84-98), and the inverse gathers of `decompress` (:283-320). Both layouts are "a b_d-bit
// field at MSB-first bit position pos_d of the row": pos_d = cumulative widths (dense) or
// 8 byte_idx_d + bit_off_d (FFD, never crossing a byte), so one kernel pair serves both.
// The projected coordinates are fp64 (the reference projects in float64); comparisons are exact, the
// packed bytes are integer work: bit-identical for identical inputs.
// HBM-bound: 8 D bytes in + code_size out per vector (encode), the reverse for decode.
#include "common.cuh"
#include "topk.cuh"

namespace b200vq {

// One warp per row: indices are OR-ed into a zeroed shared-memory copy of the row's bytes, then stored.
__global__ void __launch_bounds__(256) perdim_encode_kernel(const double* __restrict__ y, int64_t n, int64_t ldy, int D,
                                                            const double* __restrict__ bounds,
                                                            const int* __restrict__ bound_off,
                                                            const uint8_t* __restrict__ bits,
                                                            const int* __restrict__ bitpos, int code_size,
                                                            uint8_t* __restrict__ codes) {
    extern __shared__ uint32_t pd_smem[];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int words = (code_size + 1 + 3) >> 2;   // + 1: a field's second byte may be the (unused) byte after the row
    uint32_t* buf = pd_smem + (size_t)w * words;
    const int64_t wid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t row = wid; row < n; row += nw) {
        for (int i = lane; i < words; i += 32) buf[i] = 0u;
        __syncwarp();
        const double* yr = y + row * ldy;
        for (int d = lane; d < D; d += 32) {
            const int b = bits[d];
            if (b == 0) continue;
            const double v = yr[d];
            // np.searchsorted(boundaries, v) (side='left'): number of boundaries strictly below v
            const double* bd = bounds + bound_off[d];
            int lo = 0, hi = (1 << b) - 1;
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                if (bd[mid] < v) lo = mid + 1; else hi = mid;
            }
            const int pos = bitpos[d], byte = pos >> 3;
            const uint32_t v16 = (uint32_t)lo << (16 - b - (pos & 7));   // field inside a big-endian 16-bit window
            const uint32_t hi8 = v16 >> 8, lo8 = v16 & 0xFFu;
            if (hi8) atomicOr(buf + (byte >> 2), hi8 << (8 * (byte & 3)));
            if (lo8) atomicOr(buf + ((byte + 1) >> 2), lo8 << (8 * ((byte + 1) & 3)));
        }
        __syncwarp();
        uint8_t* o = codes + row * (int64_t)code_size;
        const uint8_t* src = reinterpret_cast<const uint8_t*>(buf);
        for (int i = lane; i < code_size; i += 32) o[i] = src[i];
        __syncwarp();
    }
}

// The same encode with the O(n D^2) PCA projection on the tensor cores instead of an fp64 GEMM
// (rank_aware_quantization.py:246-250 projects in float64): y (fp32) = (x - mu32) V32 from rotate_tc.cu with
// |y - (x - mu) V| <= err_rel |x - mu| + err_abs. A coordinate whose y lies within that bound of one of its
// codebook's mid-points is recomputed as the exact fp64 dot product sum_k (x_k - mu_k) VT[d][k] by the whole
// warp (four flagged coordinates per pass), so the level indices are those of the fp64 computation.
template <typename T>
__global__ void __launch_bounds__(256, 4)
perdim_encode_tc_kernel(const float* __restrict__ y, int64_t ldy, const T* __restrict__ x, int64_t ldx,
                        const double* __restrict__ mu, const double* __restrict__ VT, const double* __restrict__ nrm,
                        int64_t n, int D, const double* __restrict__ bounds, const int* __restrict__ bound_off,
                        const uint8_t* __restrict__ bits, const int* __restrict__ bitpos, int code_size,
                        double err_rel, double err_abs, uint8_t* __restrict__ codes,
                        unsigned long long* __restrict__ stats) {
    extern __shared__ uint32_t pd_smem[];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int words = (code_size + 1 + 3) >> 2;
    const int fwords = (D + 1) >> 1;                                  // flagged list: uint16 per dimension
    uint32_t* buf = pd_smem + (size_t)w * (words + fwords);
    uint16_t* flist = reinterpret_cast<uint16_t*>(buf + words);
    const int64_t wid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    unsigned long long nfix_total = 0;
    auto level_of = [&](const double* bd, int b, double v) {
        int lo = 0, hi = (1 << b) - 1;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (bd[mid] < v) lo = mid + 1; else hi = mid;
        }
        return lo;
    };
    auto put = [&](int d, int b, int lo) {
        const int pos = bitpos[d], byte = pos >> 3;
        const uint32_t v16 = (uint32_t)lo << (16 - b - (pos & 7));
        const uint32_t hi8 = v16 >> 8, lo8 = v16 & 0xFFu;
        if (hi8) atomicOr(buf + (byte >> 2), hi8 << (8 * (byte & 3)));
        if (lo8) atomicOr(buf + ((byte + 1) >> 2), lo8 << (8 * ((byte + 1) & 3)));
    };
    for (int64_t row = wid; row < n; row += nw) {
        for (int i = lane; i < words; i += 32) buf[i] = 0u;
        __syncwarp();
        const float* yr = y + row * ldy;
        const double eps = err_rel * nrm[row] + err_abs;
        int nflag = 0;
        for (int d0 = 0; d0 < D; d0 += 32) {
            const int d = d0 + lane;
            bool flag = false;
            if (d < D) {
                const int b = bits[d];
                if (b) {
                    const double v = (double)yr[d];
                    const double* bd = bounds + bound_off[d];
                    const int lo = level_of(bd, b, v), L = 1 << b;
                    flag = (lo > 0 && v - bd[lo - 1] < eps) || (lo < L - 1 && bd[lo] - v < eps);
                    if (!flag) put(d, b, lo);
                }
            }
            const unsigned m = __ballot_sync(FULL_MASK, flag);
            if (flag) flist[nflag + __popc(m & ((1u << lane) - 1u))] = (uint16_t)d;
            nflag += __popc(m);
        }
        __syncwarp();
        const T* xr = x + row * ldx;
        for (int f0 = 0; f0 < nflag; f0 += 4) {
            int dd[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) dd[e] = flist[min(f0 + e, nflag - 1)];
            const double* p0 = VT + (int64_t)dd[0] * D;
            const double* p1 = VT + (int64_t)dd[1] * D;
            const double* p2 = VT + (int64_t)dd[2] * D;
            const double* p3 = VT + (int64_t)dd[3] * D;
            double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
#pragma unroll 4
            for (int k = lane; k < D; k += 32) {
                const double rk = (double)xr[k] - mu[k];
                a0 = fma(rk, __ldg(p0 + k), a0);
                a1 = fma(rk, __ldg(p1 + k), a1);
                a2 = fma(rk, __ldg(p2 + k), a2);
                a3 = fma(rk, __ldg(p3 + k), a3);
            }
            for (int o = 16; o > 0; o >>= 1) {
                a0 += __shfl_xor_sync(FULL_MASK, a0, o);
                a1 += __shfl_xor_sync(FULL_MASK, a1, o);
                a2 += __shfl_xor_sync(FULL_MASK, a2, o);
                a3 += __shfl_xor_sync(FULL_MASK, a3, o);
            }
            if (lane == 0) {
                const double av[4] = {a0, a1, a2, a3};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    if (f0 + e >= nflag) break;
                    const int d = dd[e], b = bits[d];
                    put(d, b, level_of(bounds + bound_off[d], b, av[e]));
                }
            }
        }
        nfix_total += (unsigned long long)nflag;
        __syncwarp();
        uint8_t* o = codes + row * (int64_t)code_size;
        const uint8_t* src = reinterpret_cast<const uint8_t*>(buf);
        for (int i = lane; i < code_size; i += 32) o[i] = src[i];
        __syncwarp();
    }
    if (stats && lane == 0 && nfix_total) atomicAdd(stats, nfix_total);
}

// decode to fp32 (the operand of the inverse projection GEMM)
__global__ void __launch_bounds__(256) perdim_decode_f32_kernel(const uint8_t* __restrict__ codes, int64_t n, int code_size,
                                                                int D, const uint8_t* __restrict__ bits,
                                                                const int* __restrict__ bitpos,
                                                                const float* __restrict__ cb,
                                                                const int* __restrict__ cb_off, float* __restrict__ yhat,
                                                                int64_t ldy) {
    const int lane = threadIdx.x & 31;
    const int64_t wid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t row = wid; row < n; row += nw) {
        const uint8_t* c = codes + row * (int64_t)code_size;
        float* o = yhat + row * ldy;
        for (int d = lane; d < D; d += 32) {
            const int b = bits[d];
            float r = 0.f;
            if (b) {
                const int pos = bitpos[d], byte = pos >> 3;
                const uint32_t v16 = ((uint32_t)c[byte] << 8) | (byte + 1 < code_size ? (uint32_t)c[byte + 1] : 0u);
                const uint32_t idx = (v16 >> (16 - b - (pos & 7))) & ((1u << b) - 1u);
                r = cb[cb_off[d] + idx];
            }
            o[d] = r;
        }
    }
}

__global__ void __launch_bounds__(256) perdim_decode_kernel(const uint8_t* __restrict__ codes, int64_t n, int code_size,
                                                            int D, const uint8_t* __restrict__ bits,
                                                            const int* __restrict__ bitpos,
                                                            const double* __restrict__ cb,
                                                            const int* __restrict__ cb_off, double* __restrict__ yhat,
                                                            int64_t ldy) {
    const int lane = threadIdx.x & 31;
    const int64_t wid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t row = wid; row < n; row += nw) {
        const uint8_t* c = codes + row * (int64_t)code_size;
        double* o = yhat + row * ldy;
        for (int d = lane; d < D; d += 32) {
            const int b = bits[d];
            double r = 0.0;   // zero-width dimensions reconstruct to 0 (the component mean)
            if (b) {
                const int pos = bitpos[d], byte = pos >> 3;
                const uint32_t v16 = ((uint32_t)c[byte] << 8) | (byte + 1 < code_size ? (uint32_t)c[byte + 1] : 0u);
                const uint32_t idx = (v16 >> (16 - b - (pos & 7))) & ((1u << b) - 1u);
                r = cb[cb_off[d] + idx];
            }
            o[d] = r;
        }
    }
}

// ADC scan straight on the packed per-dimension layout (dense or FFD) = the reference's micro-benchmark loop
// bench/ffd_speed.cpp:87-127 (score = sum_d LUT[d][code_d], LUT[d][c] = q_d * cb_d[c]) with a running top-k:
// thread = row, PD_QT queries per pass so that a field is extracted and its level fetched once for all of them;
// per query one fp32 FMA chain in dimension order (the arithmetic of the exact flat search over the decoded rows).
// value(q, i) = C_q + B_i - 2 <y^_i, q> (L2, ascending) or C_q + <y^_i, q> (IP, descending).
constexpr int PD_QT = 4;
struct PdMeta {
    uint16_t pos;      // MSB-first bit position of the field in the row
    uint8_t bits, pad;
    int32_t cb_off;
};
template <int CAP>
__global__ void __launch_bounds__(256)
perdim_scan_kernel(const uint8_t* __restrict__ codes, int64_t n, int code_size, int D, const uint8_t* __restrict__ bits,
                   const int* __restrict__ bitpos, const float* __restrict__ cb, const int* __restrict__ cb_off,
                   const float* __restrict__ row_bias, const float* __restrict__ queries, int nq, int64_t ldq,
                   const float* __restrict__ qconst, int metric, int k, int64_t chunk_rows, uint64_t* __restrict__ part) {
    extern __shared__ __align__(16) unsigned char pds[];
    const int qt = blockIdx.x, chunk = blockIdx.y, nchunks = gridDim.y;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    float* qs = reinterpret_cast<float*>(pds);                                  // [PD_QT][D]
    PdMeta* meta = reinterpret_cast<PdMeta*>(qs + (size_t)PD_QT * D);           // [D]
    uint64_t* wbuf = reinterpret_cast<uint64_t*>(meta + D);                     // [8 warps][PD_QT][CAP]
    const int n2 = next_pow2(8 * k);
    uint64_t* mbuf = wbuf + 8 * PD_QT * CAP;
    for (int i = threadIdx.x; i < PD_QT * D; i += blockDim.x) {
        const int t = i / D, d = i - t * D, q = qt * PD_QT + t;
        qs[i] = q < nq ? queries[(int64_t)q * ldq + d] : 0.f;
    }
    for (int d = threadIdx.x; d < D; d += blockDim.x) {
        PdMeta m;
        m.pos = (uint16_t)bitpos[d]; m.bits = bits[d]; m.pad = 0; m.cb_off = cb_off[d];
        meta[d] = m;
    }
    __syncthreads();
    WarpTopK<CAP> tk[PD_QT];
#pragma unroll
    for (int t = 0; t < PD_QT; ++t) tk[t].init(wbuf + ((size_t)w * PD_QT + t) * CAP, k);
    float cq[PD_QT];
#pragma unroll
    for (int t = 0; t < PD_QT; ++t) cq[t] = (qconst && qt * PD_QT + t < nq) ? qconst[qt * PD_QT + t] : 0.f;
    const int64_t r0 = (int64_t)chunk * chunk_rows;
    int64_t r1 = r0 + chunk_rows;
    if (r1 > n) r1 = n;
    for (int64_t base = r0 + (int64_t)w * 32; base < r1; base += 256) {
        const int64_t row = base + lane;
        const bool valid = row < r1;
        float acc[PD_QT];
#pragma unroll
        for (int t = 0; t < PD_QT; ++t) acc[t] = 0.f;
        if (valid) {
            const uint8_t* c = codes + row * (int64_t)code_size;
            for (int d = 0; d < D; ++d) {
                const PdMeta m = meta[d];
                if (m.bits == 0) continue;
                const int byte = m.pos >> 3;
                const uint32_t v16 = ((uint32_t)c[byte] << 8) | (byte + 1 < code_size ? (uint32_t)c[byte + 1] : 0u);
                const uint32_t idx = (v16 >> (16 - m.bits - (m.pos & 7))) & ((1u << m.bits) - 1u);
                const float val = __ldg(cb + m.cb_off + idx);
#pragma unroll
                for (int t = 0; t < PD_QT; ++t) acc[t] = __fmaf_rn(qs[t * D + d], val, acc[t]);
            }
        }
        const float bi = (valid && row_bias) ? row_bias[row] : 0.f;
#pragma unroll
        for (int t = 0; t < PD_QT; ++t) {
            const float key = metric == B200VQ_METRIC_L2 ? __fmaf_rn(-2.f, acc[t], __fadd_rn(cq[t], bi))
                                                        : -__fadd_rn(cq[t], acc[t]);
            tk[t].offer(valid, key, (uint32_t)row);
        }
    }
#pragma unroll
    for (int t = 0; t < PD_QT; ++t) {
        const int q = qt * PD_QT + t;
        uint64_t* dst = part + ((size_t)(q < nq ? q : 0) * nchunks + chunk) * k;
        // queries past nq (tile padding) write nothing: their lists are emitted into a scratch slot of the tile buffer
        block_emit_topk<CAP>(tk[t], mbuf, n2, k, q < nq ? dst : mbuf + n2);
        __syncthreads();
    }
}

}  // namespace b200vq

using namespace b200vq;

static int perdim_grid(int64_t n) {
    int64_t g = (n * 32 + 255) / 256;
    const int64_t cap = (int64_t)device_sm_count() * 8;
    if (g > cap) g = cap;
    return (int)(g < 1 ? 1 : g);
}

extern "C" int b200vq_perdim_encode(const double* y, int64_t n, int64_t ldy, int D, const double* bounds,
                                    const int* bound_off, const uint8_t* bits, const int* bitpos, int code_size,
                                    uint8_t* codes, void* stream) {
    VQ_CHECK_ARG(n >= 0 && D > 0 && ldy >= D && code_size >= 0, "perdim_encode: bad shape");
    if (n == 0 || code_size == 0) return B200VQ_OK;
    VQ_CHECK_ARG(y && bounds && bound_off && bits && bitpos && codes, "perdim_encode: null pointer");
    const size_t smem = (size_t)8 * ((code_size + 4) / 4) * 4;
    VQ_CHECK_ARG(smem <= 200 * 1024, "perdim_encode: code_size %d too large", code_size);
    VQ_CUDA(cudaFuncSetAttribute(perdim_encode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    perdim_encode_kernel<<<perdim_grid(n), 256, smem, as_stream(stream)>>>(y, n, ldy, D, bounds, bound_off, bits, bitpos,
                                                                         code_size, codes);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" int b200vq_perdim_decode(const uint8_t* codes, int64_t n, int code_size, int D, const uint8_t* bits,
                                    const int* bitpos, const double* cb, const int* cb_off, double* yhat, int64_t ldy,
                                    void* stream) {
    VQ_CHECK_ARG(n >= 0 && D > 0 && ldy >= D && code_size >= 0, "perdim_decode: bad shape");
    if (n == 0) return B200VQ_OK;
    VQ_CHECK_ARG(bits && bitpos && cb && cb_off && yhat && (codes || code_size == 0), "perdim_decode: null pointer");
    perdim_decode_kernel<<<perdim_grid(n), 256, 0, as_stream(stream)>>>(codes, n, code_size, D, bits, bitpos, cb, cb_off,
                                                                      yhat, ldy);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" int b200vq_perdim_encode_tc(const float* y, int64_t ldy, const void* x, int x_is_f64, int64_t ldx,
                                       const double* mu, const double* VT, const double* nrm, int64_t n, int D,
                                       const double* bounds, const int* bound_off, const uint8_t* bits,
                                       const int* bitpos, int code_size, double err_rel, double err_abs, uint8_t* codes,
                                       unsigned long long* stats, void* stream) {
    VQ_CHECK_ARG(n >= 0 && D > 0 && D < 65536 && ldy >= D && ldx >= D && code_size >= 0 && err_rel >= 0.0 && err_abs >= 0.0,
                 "perdim_encode_tc: bad shape");
    if (n == 0 || code_size == 0) return B200VQ_OK;
    VQ_CHECK_ARG(y && x && mu && VT && nrm && bounds && bound_off && bits && bitpos && codes, "perdim_encode_tc: null pointer");
    const size_t smem = (size_t)8 * (((code_size + 4) / 4) + ((D + 1) / 2)) * 4;
    VQ_CHECK_ARG(smem <= 200 * 1024, "perdim_encode_tc: row too large");
    int64_t g = (n + 7) / 8;
    const int64_t cap = (int64_t)device_sm_count() * 8;
    if (g > cap) g = cap;
    if (x_is_f64) {
        VQ_CUDA(cudaFuncSetAttribute(perdim_encode_tc_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        perdim_encode_tc_kernel<double><<<(int)g, 256, smem, as_stream(stream)>>>(
            y, ldy, reinterpret_cast<const double*>(x), ldx, mu, VT, nrm, n, D, bounds, bound_off, bits, bitpos, code_size,
            err_rel, err_abs, codes, stats);
    } else {
        VQ_CUDA(cudaFuncSetAttribute(perdim_encode_tc_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        perdim_encode_tc_kernel<float><<<(int)g, 256, smem, as_stream(stream)>>>(
            y, ldy, reinterpret_cast<const float*>(x), ldx, mu, VT, nrm, n, D, bounds, bound_off, bits, bitpos, code_size,
            err_rel, err_abs, codes, stats);
    }
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" int b200vq_perdim_decode_f32(const uint8_t* codes, int64_t n, int code_size, int D, const uint8_t* bits,
                                        const int* bitpos, const float* cb, const int* cb_off, float* yhat, int64_t ldy,
                                        void* stream) {
    VQ_CHECK_ARG(n >= 0 && D > 0 && ldy >= D && code_size >= 0, "perdim_decode_f32: bad shape");
    if (n == 0) return B200VQ_OK;
    VQ_CHECK_ARG(bits && bitpos && cb && cb_off && yhat && (codes || code_size == 0), "perdim_decode_f32: null pointer");
    perdim_decode_f32_kernel<<<perdim_grid(n), 256, 0, as_stream(stream)>>>(codes, n, code_size, D, bits, bitpos, cb, cb_off,
                                                                          yhat, ldy);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

static void pd_plan(int64_t n, int nq, int64_t* chunk_rows, int* nchunks) {
    const int64_t unit = 2048;
    const int qtiles = (nq + PD_QT - 1) / PD_QT;
    int64_t want = ((int64_t)device_sm_count() * 4 + qtiles - 1) / (qtiles > 0 ? qtiles : 1);
    int64_t maxc = (n + unit - 1) / unit;
    if (maxc < 1) maxc = 1;
    if (want < 1) want = 1;
    if (want > maxc) want = maxc;
    if (want > 65535) want = 65535;
    int64_t rows = (n + want - 1) / want;
    rows = (rows + unit - 1) / unit * unit;
    *chunk_rows = rows;
    *nchunks = (int)((n + rows - 1) / rows);
    if (*nchunks < 1) *nchunks = 1;
}

extern "C" size_t b200vq_perdim_scan_scratch_bytes(int64_t n, int nq, int k) {
    if (n < 0 || nq <= 0 || k <= 0) return 0;
    int64_t cr;
    int nc;
    pd_plan(n, nq, &cr, &nc);
    return (size_t)nq * nc * k * sizeof(uint64_t);
}

template <int CAP>
static int pd_scan_launch(const uint8_t* codes, int64_t n, int code_size, int D, const uint8_t* bits, const int* bitpos,
                          const float* cb, const int* cb_off, const float* row_bias, const float* queries, int nq,
                          int64_t ldq, const float* qconst, int metric, int k, int64_t cr, int nc, uint64_t* part,
                          cudaStream_t st) {
    const int n2 = next_pow2(8 * k);
    const size_t smem = (size_t)PD_QT * D * 4 + (size_t)D * sizeof(PdMeta) + (size_t)(8 * PD_QT * CAP + n2 + k) * 8;
    VQ_CHECK_ARG(smem <= 220 * 1024, "perdim_scan_topk: D = %d needs %zu bytes of shared memory", D, smem);
    VQ_CUDA(cudaFuncSetAttribute(perdim_scan_kernel<CAP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid((unsigned)((nq + PD_QT - 1) / PD_QT), (unsigned)nc);
    perdim_scan_kernel<CAP><<<grid, 256, smem, st>>>(codes, n, code_size, D, bits, bitpos, cb, cb_off, row_bias, queries, nq,
                                                    ldq, qconst, metric, k, cr, part);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" int b200vq_perdim_scan_topk(const uint8_t* codes, int64_t n, int code_size, int D, const uint8_t* bits,
                                       const int* bitpos, const float* cb, const int* cb_off, const float* row_bias,
                                       const float* queries, int nq, int64_t ldq, const float* qconst, int metric, int k,
                                       float* out_vals, uint32_t* out_ids, void* scratch, size_t scratch_bytes,
                                       void* stream) {
    VQ_CHECK_ARG(codes && bits && bitpos && cb && cb_off && queries && out_vals && out_ids, "perdim_scan_topk: null pointer");
    VQ_CHECK_ARG(n >= 0 && n < (1ll << 32) && D > 0 && code_size > 0 && code_size < 8192 && nq >= 0 && ldq >= D,
                 "perdim_scan_topk: bad shape");
    VQ_CHECK_ARG(k > 0 && k <= 128, "perdim_scan_topk: k must be in [1,128], got %d", k);
    VQ_CHECK_ARG(metric == B200VQ_METRIC_L2 || metric == B200VQ_METRIC_IP, "perdim_scan_topk: bad metric");
    if (nq == 0) return B200VQ_OK;
    int64_t cr;
    int nc;
    pd_plan(n, nq, &cr, &nc);
    const size_t need = (size_t)nq * nc * k * sizeof(uint64_t);
    VQ_CHECK_ARG(scratch && scratch_bytes >= need, "perdim_scan_topk: scratch too small (%zu < %zu)", scratch_bytes, need);
    cudaStream_t st = as_stream(stream);
    uint64_t* part = reinterpret_cast<uint64_t*>(scratch);
    int rc = (k <= 32) ? pd_scan_launch<64>(codes, n, code_size, D, bits, bitpos, cb, cb_off, row_bias, queries, nq, ldq,
                                            qconst, metric, k, cr, nc, part, st)
                       : pd_scan_launch<256>(codes, n, code_size, D, bits, bitpos, cb, cb_off, row_bias, queries, nq, ldq,
                                             qconst, metric, k, cr, nc, part, st);
    if (rc != B200VQ_OK) return rc;
    merge_partials_kernel<2048><<<nq, 256, 0, st>>>(part, nc, k, metric == B200VQ_METRIC_IP, out_vals, out_ids);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}
