// Product-quantizer kernels (exact fp32 SIMT paths): centroid norms, nearest-
// centroid assignment, decode, Lloyd accumulation/finalisation, SSE, ADC LUTs.
//
// Reference semantics being reproduced (faiss ProductQuantizer as driven by
// src/haag_vq/methods/product_quantization.py:58-86 of the reference):
//   compute_codes : argmin_k (||x_m||^2 + ||c_mk||^2) - 2 x_m.c_mk  in fp32,
//                   negative distances clamped to 0, first index wins ties
//   decode        : concatenation of the selected centroids
//   train         : per-sub-quantizer Lloyd iterations
#include "common.cuh"

namespace b200vq {

// ------------------------------------------------------- centroid norms ---
__global__ void pq_centroid_norms_kernel(const float* __restrict__ c, int total, int dsub,
                                         float* __restrict__ out) {
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const float* p = c + (size_t)idx * dsub;
    float acc = 0.f;
    for (int i = 0; i < dsub; ++i) acc = __fmaf_rn(p[i], p[i], acc);
    out[idx] = acc;
}

// ----------------------------------------------------- exact assignment ---
// One CTA = R rows x one sub-quantizer. Thread r owns row r and walks the
// centroids in blocks of KB, accumulating KB dot products in registers with a
// sequential fp32 FMA chain over the sub-vector (deterministic order).
template <int KB>
__global__ void pq_assign_simt_kernel(const float* __restrict__ x, int64_t n, int64_t ldx,
                                      const float* __restrict__ centroids,
                                      const float* __restrict__ cnorm, int M, int ksub, int dsub,
                                      uint8_t* __restrict__ codes, float* __restrict__ sqdist) {
    extern __shared__ float smem[];
    const int R = blockDim.x;
    const int tid = threadIdx.x;
    const int m = blockIdx.y;
    const int64_t row0 = (int64_t)blockIdx.x * R;
    float* xs = smem;                          // [dsub][R+1]
    float* cs = xs + (((size_t)dsub * (R + 1) + 3) & ~(size_t)3);   // [dsub][KB], 16-byte aligned

    for (int idx = tid; idx < R * dsub; idx += R) {
        int r = idx / dsub, i = idx - r * dsub;
        int64_t row = row0 + r;
        float v = 0.f;
        if (row < n) v = x[row * ldx + (int64_t)m * dsub + i];
        xs[i * (R + 1) + r] = v;
    }
    __syncthreads();

    float xn = 0.f;
    for (int i = 0; i < dsub; ++i) {
        float v = xs[i * (R + 1) + tid];
        xn = __fmaf_rn(v, v, xn);
    }

    float best = __int_as_float(0x7f800000);
    int bestk = 0;
    const float* cm = centroids + (size_t)m * ksub * dsub;
    const float* cnm = cnorm + (size_t)m * ksub;

    for (int k0 = 0; k0 < ksub; k0 += KB) {
        __syncthreads();
        for (int idx = tid; idx < KB * dsub; idx += R) {
            int kk = idx / dsub, i = idx - kk * dsub;
            float v = 0.f;
            if (k0 + kk < ksub) v = cm[(size_t)(k0 + kk) * dsub + i];
            cs[i * KB + kk] = v;
        }
        __syncthreads();

        float acc[KB];
#pragma unroll
        for (int j = 0; j < KB; ++j) acc[j] = 0.f;
        for (int i = 0; i < dsub; ++i) {
            float xv = xs[i * (R + 1) + tid];
            const float4* c4 = reinterpret_cast<const float4*>(cs + i * KB);
#pragma unroll
            for (int j4 = 0; j4 < KB / 4; ++j4) {
                float4 c = c4[j4];
                acc[4 * j4 + 0] = __fmaf_rn(xv, c.x, acc[4 * j4 + 0]);
                acc[4 * j4 + 1] = __fmaf_rn(xv, c.y, acc[4 * j4 + 1]);
                acc[4 * j4 + 2] = __fmaf_rn(xv, c.z, acc[4 * j4 + 2]);
                acc[4 * j4 + 3] = __fmaf_rn(xv, c.w, acc[4 * j4 + 3]);
            }
        }
#pragma unroll
        for (int j = 0; j < KB; ++j) {
            if (k0 + j < ksub) {
                float d = __fmaf_rn(-2.f, acc[j], __fadd_rn(xn, cnm[k0 + j]));
                d = fmaxf(d, 0.f);
                if (d < best) {
                    best = d;
                    bestk = k0 + j;
                }
            }
        }
    }
    int64_t row = row0 + tid;
    if (row < n) {
        codes[row * M + m] = (uint8_t)bestk;
        if (sqdist) sqdist[row * M + m] = best;
    }
}

// ---------------------------------------------------------------- decode ---
template <bool BY_ID>
__global__ void pq_decode_kernel(const uint8_t* __restrict__ codes, const int64_t* __restrict__ ids,
                                 int64_t n, const float* __restrict__ centroids, int M, int ksub,
                                 int dsub, float* __restrict__ out, int64_t ldo) {
    const int D = M * dsub;
    const int64_t total = n * (int64_t)D;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total;
         e += (int64_t)gridDim.x * blockDim.x) {
        int64_t row = e / D;
        int col = (int)(e - row * D);
        int m = col / dsub, i = col - m * dsub;
        int64_t src = BY_ID ? ids[row] : row;
        int code = codes[src * M + m];
        out[row * ldo + col] = centroids[((size_t)m * ksub + code) * dsub + i];
    }
}

// 128-bit variant: one warp per output row, a lane moves 4 consecutive floats of one sub-vector per step.
// No 64-bit index arithmetic per element: the sub-quantizer of a 16-byte piece is a multiply-high by a
// host-computed reciprocal (exact for the sizes here: piece index < 2^16, dsub / 4 <= 2^14).
template <bool BY_ID>
__global__ void __launch_bounds__(256) pq_decode_v4_kernel(const uint8_t* __restrict__ codes,
                                                           const int64_t* __restrict__ ids, int64_t n,
                                                           const float* __restrict__ centroids, int M, int ksub,
                                                           int dsub, uint32_t inv_dsub4, float* __restrict__ out,
                                                           int64_t ldo) {
    const int D4 = (M * dsub) >> 2;
    const int dsub4 = dsub >> 2;
    const int lane = threadIdx.x & 31;
    const int64_t wid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const float4* c4base = reinterpret_cast<const float4*>(centroids);
    for (int64_t row = wid; row < n; row += nw) {
        const uint8_t* code = codes + (BY_ID ? ids[row] : row) * M;
        float4* o = reinterpret_cast<float4*>(out + row * ldo);
#pragma unroll 4
        for (int c4 = lane; c4 < D4; c4 += 32) {
            const int m = dsub4 == 1 ? c4 : (int)__umulhi((uint32_t)c4, inv_dsub4);   // 2^32 / 1 does not fit the reciprocal
            const int i4 = c4 - m * dsub4;
            o[c4] = __ldg(c4base + ((size_t)m * ksub + code[m]) * dsub4 + i4);
        }
    }
}

// --------------------------------------------------------------- k-means ---
__global__ void pq_kmeans_accumulate_kernel(const float* __restrict__ x, int64_t n, int64_t ldx,
                                            const uint8_t* __restrict__ codes, int M, int ksub,
                                            int dsub, float* __restrict__ sums,
                                            int* __restrict__ counts) {
    const int D = M * dsub;
    const int64_t total = n * (int64_t)D;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total;
         e += (int64_t)gridDim.x * blockDim.x) {
        int64_t row = e / D;
        int col = (int)(e - row * D);
        int m = col / dsub, i = col - m * dsub;
        int code = codes[row * M + m];
        size_t cidx = (size_t)m * ksub + code;
        atomicAdd(&sums[cidx * dsub + i], x[row * ldx + col]);
        if (i == 0) atomicAdd(&counts[cidx], 1);
    }
}

// One CTA per sub-quantizer. Empty clusters are re-seeded by splitting the
// largest cluster (deterministic restatement of faiss' split_clusters, which
// draws the donor at random: Clustering.cpp, EPS = 1/1024).
__global__ void pq_kmeans_finalize_kernel(const float* __restrict__ sums, int* __restrict__ counts,
                                          int ksub, int dsub, float* __restrict__ centroids,
                                          int* __restrict__ n_split) {
    const int m = blockIdx.x;
    const float* s = sums + (size_t)m * ksub * dsub;
    int* cnt = counts + (size_t)m * ksub;
    float* c = centroids + (size_t)m * ksub * dsub;
    for (int idx = threadIdx.x; idx < ksub * dsub; idx += blockDim.x) {
        int k = idx / dsub;
        int h = cnt[k];
        if (h > 0) c[idx] = s[idx] * (1.0f / (float)h);
    }
    __syncthreads();
    __shared__ int s_donor;
    int nsplit = 0;
    for (int ci = 0; ci < ksub; ++ci) {
        if (cnt[ci] != 0) continue;  // uniform: everybody reads the same global value
        if (threadIdx.x == 0) {
            int best = 0, bi = 0;
            for (int k = 0; k < ksub; ++k)
                if (cnt[k] > best) {
                    best = cnt[k];
                    bi = k;
                }
            s_donor = bi;
        }
        __syncthreads();
        const int cj = s_donor;
        const float EPS = 1.f / 1024.f;
        for (int j = threadIdx.x; j < dsub; j += blockDim.x) {
            float v = c[(size_t)cj * dsub + j];
            if ((j & 1) == 0) {
                c[(size_t)ci * dsub + j] = v * (1.f + EPS);
                c[(size_t)cj * dsub + j] = v * (1.f - EPS);
            } else {
                c[(size_t)ci * dsub + j] = v * (1.f - EPS);
                c[(size_t)cj * dsub + j] = v * (1.f + EPS);
            }
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            int h = cnt[cj];
            cnt[ci] = h / 2;
            cnt[cj] = h - h / 2;
            __threadfence_block();
        }
        __syncthreads();
        ++nsplit;
    }
    if (threadIdx.x == 0 && n_split && nsplit) atomicAdd(n_split, nsplit);
}

// One warp per row: a lane walks the row's columns (no 64-bit index arithmetic per element), differences
// in fp32, squares and sums in fp64.
__global__ void __launch_bounds__(256) pq_sse_kernel(const float* __restrict__ x, int64_t n, int64_t ldx,
                                                     const uint8_t* __restrict__ codes,
                                                     const float* __restrict__ centroids, int M, int ksub, int dsub,
                                                     uint32_t inv_dsub4, bool vec, double* __restrict__ out) {
    const int D = M * dsub;
    const int64_t wid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    double acc = 0.0;
    for (int64_t row = wid; row < n; row += nw) {
        const float* xr = x + row * ldx;
        const uint8_t* code = codes + row * M;
        if (vec) {   // 128-bit pieces, sub-quantizer by multiply-high (as pq_decode_v4_kernel)
            const int D4 = D >> 2, dsub4 = dsub >> 2;
            for (int c4 = threadIdx.x & 31; c4 < D4; c4 += 32) {
                const int m = dsub4 == 1 ? c4 : (int)__umulhi((uint32_t)c4, inv_dsub4);
                const float4 a = __ldg(reinterpret_cast<const float4*>(xr) + c4);
                const float4 c = __ldg(reinterpret_cast<const float4*>(centroids) +
                                       ((size_t)m * ksub + code[m]) * dsub4 + (c4 - m * dsub4));
                const float d0 = a.x - c.x, d1 = a.y - c.y, d2 = a.z - c.z, d3 = a.w - c.w;
                acc += (double)d0 * (double)d0;
                acc += (double)d1 * (double)d1;
                acc += (double)d2 * (double)d2;
                acc += (double)d3 * (double)d3;
            }
        } else {
            for (int col = threadIdx.x & 31; col < D; col += 32) {
                const int m = col / dsub, i = col - m * dsub;
                const float d = xr[col] - centroids[((size_t)m * ksub + code[m]) * dsub + i];
                acc += (double)d * (double)d;
            }
        }
    }
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(FULL_MASK, acc, o);
    __shared__ double part[32];
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) part[w] = acc;
    __syncthreads();
    if (w == 0) {
        double v = (lane < (blockDim.x >> 5)) ? part[lane] : 0.0;
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL_MASK, v, o);
        if (lane == 0) atomicAdd(out, v);
    }
}

// --------------------------------------------------------------- ADC LUT ---
// grid (nq, M); thread k computes lut[q, m, k] with a sequential fp32 chain.
__global__ void adc_lut_kernel(const float* __restrict__ queries, int64_t ldq,
                               const float* __restrict__ centroids, int M, int ksub, int dsub,
                               int metric, float* __restrict__ lut) {
    extern __shared__ float qs[];
    const int q = blockIdx.x, m = blockIdx.y;
    for (int i = threadIdx.x; i < dsub; i += blockDim.x)
        qs[i] = queries[(int64_t)q * ldq + (int64_t)m * dsub + i];
    __syncthreads();
    for (int k = threadIdx.x; k < ksub; k += blockDim.x) {
        const float* c = centroids + ((size_t)m * ksub + k) * dsub;
        float acc = 0.f;
        if (metric == B200VQ_METRIC_L2) {
            for (int i = 0; i < dsub; ++i) {
                float d = __fsub_rn(qs[i], c[i]);
                acc = __fmaf_rn(d, d, acc);
            }
        } else {
            for (int i = 0; i < dsub; ++i) acc = __fmaf_rn(qs[i], c[i], acc);
        }
        lut[((size_t)q * M + m) * ksub + k] = acc;
    }
}

// Batched LUT build: grid (M, ceil(nq / 32)), 512 threads. The CTA stages the 256 centroids of its sub-quantizer
// dimension-major ([i][256]) and 32 query sub-vectors dimension-major ([i][32]) once; thread (kq, gq) then owns a
// 4 x 4 register tile (centroids 4 kq .. 4 kq + 3, queries 4 gq .. 4 gq + 3) and runs, for each of its 16 entries,
// the same sequential fp32 chain as adc_lut_kernel. Per dimension a thread issues two 128-bit shared loads for 32
// FP instructions: the one-centroid-per-thread form (1 + 1 loads for 8 FP) was bound by shared-memory wavefronts
// (ncu r2: 77 us at 1000 x 1536, the same with 256 or 512 threads).
constexpr int LUT_QB = 32;
constexpr int LUT_THREADS = 512;
__global__ void __launch_bounds__(LUT_THREADS)
adc_lut_batched_kernel(const float* __restrict__ queries, int nq, int64_t ldq, const float* __restrict__ centroids,
                       int M, int dsub, int metric, float* __restrict__ lut) {
    extern __shared__ __align__(16) float lsm[];
    float* qs = lsm;                              // [dsub][LUT_QB]
    float* cs = lsm + (size_t)dsub * LUT_QB;      // [dsub][256]
    const int m = blockIdx.x, q0 = blockIdx.y * LUT_QB;
    const float* cm = centroids + (size_t)m * 256 * dsub;
    {
        const int k = threadIdx.x & 255, half = threadIdx.x >> 8;
        if ((dsub & 3) == 0 && ((uintptr_t)cm & 15) == 0) {
            // thread (k, h) copies every other quad of centroid k with 128-bit loads (independent, several in
            // flight); the transposed stores of a warp are 32 consecutive words
            const float4* row = reinterpret_cast<const float4*>(cm + (size_t)k * dsub);
#pragma unroll 8
            for (int i4 = half; i4 < (dsub >> 2); i4 += 2) {
                const float4 v = __ldg(row + i4);
                cs[(4 * i4 + 0) * 256 + k] = v.x;
                cs[(4 * i4 + 1) * 256 + k] = v.y;
                cs[(4 * i4 + 2) * 256 + k] = v.z;
                cs[(4 * i4 + 3) * 256 + k] = v.w;
            }
        } else {
            for (int i = half; i < dsub; i += 2) cs[i * 256 + k] = cm[(size_t)k * dsub + i];
        }
    }
    for (int idx = threadIdx.x; idx < LUT_QB * dsub; idx += LUT_THREADS) {
        const int qq = idx / dsub, i = idx - qq * dsub;
        qs[i * LUT_QB + qq] = (q0 + qq < nq) ? queries[(int64_t)(q0 + qq) * ldq + (int64_t)m * dsub + i] : 0.f;
    }
    __syncthreads();
    const int kq = threadIdx.x & 63, gq = threadIdx.x >> 6;
    float acc[4][4];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;
    const float4* cs4 = reinterpret_cast<const float4*>(cs) + kq;
    const float4* qs4 = reinterpret_cast<const float4*>(qs) + gq;
    if (metric == B200VQ_METRIC_L2) {
#pragma unroll 2
        for (int i = 0; i < dsub; ++i) {
            const float4 c4 = cs4[i * 64], q4 = qs4[i * (LUT_QB / 4)];
            const float cv[4] = {c4.x, c4.y, c4.z, c4.w}, qv[4] = {q4.x, q4.y, q4.z, q4.w};
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int b = 0; b < 4; ++b) {
                    const float d = __fsub_rn(qv[a], cv[b]);
                    acc[a][b] = __fmaf_rn(d, d, acc[a][b]);
                }
        }
    } else {
#pragma unroll 2
        for (int i = 0; i < dsub; ++i) {
            const float4 c4 = cs4[i * 64], q4 = qs4[i * (LUT_QB / 4)];
            const float cv[4] = {c4.x, c4.y, c4.z, c4.w}, qv[4] = {q4.x, q4.y, q4.z, q4.w};
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int b = 0; b < 4; ++b) acc[a][b] = __fmaf_rn(qv[a], cv[b], acc[a][b]);
        }
    }
#pragma unroll
    for (int a = 0; a < 4; ++a) {
        const int q = q0 + 4 * gq + a;
        if (q < nq)
            *reinterpret_cast<float4*>(lut + ((size_t)q * M + m) * 256 + 4 * kq) =
                make_float4(acc[a][0], acc[a][1], acc[a][2], acc[a][3]);
    }
}

// faiss packs sub-quantizer indices of nbits != 8 into a little-endian bit stream (PQEncoderGeneric /
// PQDecoderGeneric: index m occupies stream bits [m nbits, (m + 1) nbits), stream bit i = bit i % 8 of byte i / 8).
// One thread per packed byte (pack) / per index (unpack); nbits < 8, so an index spans at most two bytes.
__global__ void pq_pack_bits_kernel(const uint8_t* __restrict__ codes, int64_t n, int M, int nbits, int code_size,
                                    uint8_t* __restrict__ packed) {
    const int64_t total = n * (int64_t)code_size;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const int64_t row = e / code_size;
        const int byte = (int)(e - row * code_size);
        const int b0 = byte * 8, b1 = b0 + 8;
        uint32_t acc = 0;
        for (int m = b0 / nbits; m < M && m * nbits < b1; ++m) {
            const int sh = m * nbits - b0;                 // stream position of the index's bit 0 relative to the byte
            const uint32_t q = codes[row * M + m];
            acc |= (sh >= 0) ? (q << sh) : (q >> (-sh));
        }
        packed[e] = (uint8_t)acc;
    }
}
__global__ void pq_unpack_bits_kernel(const uint8_t* __restrict__ packed, int64_t n, int M, int nbits, int code_size,
                                      uint8_t* __restrict__ codes) {
    const int64_t total = n * (int64_t)M;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const int64_t row = e / M;
        const int m = (int)(e - row * M);
        const int bit = m * nbits, byte = bit >> 3;
        const uint8_t* p = packed + row * code_size;
        const uint32_t w = (uint32_t)p[byte] | (byte + 1 < code_size ? (uint32_t)p[byte + 1] << 8 : 0u);
        codes[e] = (uint8_t)((w >> (bit & 7)) & ((1u << nbits) - 1u));
    }
}

}  // namespace b200vq

using namespace b200vq;

extern "C" int b200vq_pq_pack_bits(const uint8_t* codes, int64_t n, int M, int nbits, uint8_t* packed, void* stream) {
    VQ_CHECK_ARG(n >= 0 && M > 0 && nbits >= 1 && nbits <= 8 && (n == 0 || (codes && packed)), "pq_pack_bits: bad args");
    if (n == 0) return B200VQ_OK;
    const int code_size = (M * nbits + 7) / 8;
    int64_t g = (n * code_size + 255) / 256;
    const int64_t cap = (int64_t)device_sm_count() * 16;
    if (g > cap) g = cap;
    pq_pack_bits_kernel<<<(int)g, 256, 0, as_stream(stream)>>>(codes, n, M, nbits, code_size, packed);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}
extern "C" int b200vq_pq_unpack_bits(const uint8_t* packed, int64_t n, int M, int nbits, uint8_t* codes, void* stream) {
    VQ_CHECK_ARG(n >= 0 && M > 0 && nbits >= 1 && nbits <= 8 && (n == 0 || (codes && packed)), "pq_unpack_bits: bad args");
    if (n == 0) return B200VQ_OK;
    const int code_size = (M * nbits + 7) / 8;
    int64_t g = (n * M + 255) / 256;
    const int64_t cap = (int64_t)device_sm_count() * 16;
    if (g > cap) g = cap;
    pq_unpack_bits_kernel<<<(int)g, 256, 0, as_stream(stream)>>>(packed, n, M, nbits, code_size, codes);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" int b200vq_pq_centroid_norms(const float* centroids, int M, int ksub, int dsub,
                                        float* cnorm, void* stream) {
    VQ_CHECK_ARG(centroids && cnorm && M > 0 && ksub > 0 && dsub > 0, "pq_centroid_norms: bad args");
    int total = M * ksub;
    pq_centroid_norms_kernel<<<(total + 255) / 256, 256, 0, as_stream(stream)>>>(centroids, total,
                                                                                 dsub, cnorm);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" int b200vq_pq_assign_f32(const float* x, int64_t n, int64_t ldx, const float* centroids,
                                    const float* cnorm, int M, int ksub, int dsub, uint8_t* codes,
                                    float* sqdist, void* stream) {
    VQ_CHECK_ARG(x && centroids && cnorm && codes, "pq_assign_f32: null pointer");
    VQ_CHECK_ARG(n >= 0 && M > 0 && dsub > 0 && ksub > 0 && ksub <= 256,
                 "pq_assign_f32: need n>=0, M>0, dsub>0, 0<ksub<=256 (got n=%lld M=%d ksub=%d dsub=%d)",
                 (long long)n, M, ksub, dsub);
    VQ_CHECK_ARG(ldx >= (int64_t)M * dsub, "pq_assign_f32: ldx < D");
    VQ_CHECK_ARG(M <= 65535, "pq_assign_f32: M > 65535");
    if (n == 0) return B200VQ_OK;
    constexpr int KB = 16;
    int R = 128;
    auto need = [&](int r) {
        return ((((size_t)dsub * (r + 1) + 3) & ~(size_t)3) + (size_t)dsub * KB) * sizeof(float);
    };
    while (R > 32 && need(R) > 200 * 1024) R >>= 1;
    VQ_CHECK_ARG(need(R) <= 200 * 1024, "pq_assign_f32: dsub=%d too large for the SIMT path", dsub);
    size_t smem = need(R);
    VQ_CUDA(cudaFuncSetAttribute(pq_assign_simt_kernel<KB>,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid((unsigned)((n + R - 1) / R), (unsigned)M);
    pq_assign_simt_kernel<KB><<<grid, R, smem, as_stream(stream)>>>(x, n, ldx, centroids, cnorm, M,
                                                                    ksub, dsub, codes, sqdist);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

static inline int grid_for(int64_t work, int block, int waves = 8) {
    int64_t g = (work + block - 1) / block;
    int64_t cap = (int64_t)device_sm_count() * waves;
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    return (int)g;
}

// ceil(2^32 / (dsub / 4)): __umulhi(c, inv) == c / (dsub / 4) for c < 2^16
static uint32_t decode_inv(int dsub) { return (uint32_t)(((1ull << 32) + (uint64_t)(dsub >> 2) - 1) / (uint64_t)(dsub >> 2)); }

extern "C" int b200vq_pq_decode(const uint8_t* codes, int64_t n, const float* centroids, int M,
                                int ksub, int dsub, float* out, int64_t ldo, void* stream) {
    VQ_CHECK_ARG((codes || n == 0) && centroids && (out || n == 0) && M > 0 && ksub > 0 && dsub > 0 && n >= 0,
                 "pq_decode: bad args");
    VQ_CHECK_ARG(ldo >= (int64_t)M * dsub, "pq_decode: ldo < D");
    if (n == 0) return B200VQ_OK;
    bool v4 = (dsub % 4 == 0) && (ldo % 4 == 0) && (((uintptr_t)out & 15) == 0) &&
              (((uintptr_t)centroids & 15) == 0);
    if (v4) {
        pq_decode_v4_kernel<false><<<grid_for(n * 32, 256), 256, 0, as_stream(stream)>>>(
            codes, nullptr, n, centroids, M, ksub, dsub, decode_inv(dsub), out, ldo);
    } else {
        int64_t work = n * (int64_t)(M * dsub);
        pq_decode_kernel<false><<<grid_for(work, 256), 256, 0, as_stream(stream)>>>(
            codes, nullptr, n, centroids, M, ksub, dsub, out, ldo);
    }
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" int b200vq_pq_decode_ids(const uint8_t* codes, const int64_t* ids, int64_t n_ids,
                                    const float* centroids, int M, int ksub, int dsub, float* out,
                                    int64_t ldo, void* stream) {
    VQ_CHECK_ARG(codes && (ids || n_ids == 0) && centroids && (out || n_ids == 0) && M > 0 && ksub > 0 && dsub > 0 && n_ids >= 0,
                 "pq_decode_ids: bad args");
    VQ_CHECK_ARG(ldo >= (int64_t)M * dsub, "pq_decode_ids: ldo < D");
    if (n_ids == 0) return B200VQ_OK;
    bool v4 = (dsub % 4 == 0) && (ldo % 4 == 0) && (((uintptr_t)out & 15) == 0) &&
              (((uintptr_t)centroids & 15) == 0);
    if (v4) {
        pq_decode_v4_kernel<true><<<grid_for(n_ids * 32, 256), 256, 0, as_stream(stream)>>>(
            codes, ids, n_ids, centroids, M, ksub, dsub, decode_inv(dsub), out, ldo);
    } else {
        int64_t work = n_ids * (int64_t)(M * dsub);
        pq_decode_kernel<true><<<grid_for(work, 256), 256, 0, as_stream(stream)>>>(
            codes, ids, n_ids, centroids, M, ksub, dsub, out, ldo);
    }
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" int b200vq_pq_kmeans_accumulate(const float* x, int64_t n, int64_t ldx,
                                           const uint8_t* codes, int M, int ksub, int dsub,
                                           float* sums, int* counts, void* stream) {
    VQ_CHECK_ARG((x || n == 0) && (codes || n == 0) && sums && counts && M > 0 && ksub > 0 && dsub > 0 && n >= 0,
                 "pq_kmeans_accumulate: bad args");
    if (n == 0) return B200VQ_OK;
    int64_t work = n * (int64_t)(M * dsub);
    pq_kmeans_accumulate_kernel<<<grid_for(work, 256, 16), 256, 0, as_stream(stream)>>>(
        x, n, ldx, codes, M, ksub, dsub, sums, counts);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" int b200vq_pq_kmeans_finalize(const float* sums, int* counts, int M, int ksub, int dsub,
                                         float* centroids, int* n_split, void* stream) {
    VQ_CHECK_ARG(sums && counts && centroids && M > 0 && ksub > 0 && dsub > 0,
                 "pq_kmeans_finalize: bad args");
    pq_kmeans_finalize_kernel<<<M, 256, 0, as_stream(stream)>>>(sums, counts, ksub, dsub, centroids,
                                                               n_split);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" int b200vq_pq_sse(const float* x, int64_t n, int64_t ldx, const uint8_t* codes,
                             const float* centroids, int M, int ksub, int dsub, double* out,
                             void* stream) {
    VQ_CHECK_ARG((x || n == 0) && (codes || n == 0) && centroids && (out || n == 0) && M > 0 && ksub > 0 && dsub > 0 && n >= 0,
                 "pq_sse: bad args");
    if (n == 0) return B200VQ_OK;
    const bool vec = dsub % 4 == 0 && ldx % 4 == 0 && (((uintptr_t)x | (uintptr_t)centroids) & 15) == 0;
    pq_sse_kernel<<<grid_for(n * 32, 256), 256, 0, as_stream(stream)>>>(x, n, ldx, codes, centroids, M, ksub, dsub,
                                                                        vec ? decode_inv(dsub) : 0u, vec, out);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" int b200vq_adc_lut(const float* queries, int nq, int64_t ldq, const float* centroids,
                              int M, int ksub, int dsub, int metric, float* lut, void* stream) {
    VQ_CHECK_ARG(queries && centroids && lut && nq >= 0 && M > 0 && ksub > 0 && dsub > 0,
                 "adc_lut: bad args");
    VQ_CHECK_ARG(metric == B200VQ_METRIC_L2 || metric == B200VQ_METRIC_IP, "adc_lut: bad metric");
    VQ_CHECK_ARG(M <= 65535, "adc_lut: M > 65535");
    VQ_CHECK_ARG((size_t)dsub * sizeof(float) <= 48 * 1024, "adc_lut: dsub too large");
    if (nq == 0) return B200VQ_OK;
    const size_t bsm = ((size_t)dsub * 256 + (size_t)dsub * LUT_QB) * sizeof(float);
    if (ksub == 256 && nq >= 8 && bsm <= 200 * 1024 && ((uintptr_t)lut & 15) == 0) {
        VQ_CUDA(cudaFuncSetAttribute(adc_lut_batched_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bsm));
        // two CTAs per SM need the largest shared-memory carve-out (dsub = 96: 2 x 111 KB)
        VQ_CUDA(cudaFuncSetAttribute(adc_lut_batched_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                     (int)cudaSharedmemCarveoutMaxShared));
        adc_lut_batched_kernel<<<dim3((unsigned)M, (unsigned)((nq + LUT_QB - 1) / LUT_QB)), LUT_THREADS, bsm, as_stream(stream)>>>(
            queries, nq, ldq, centroids, M, dsub, metric, lut);
        VQ_LAUNCH_CHECK();
        return B200VQ_OK;
    }
    dim3 grid((unsigned)nq, (unsigned)M);
    int block = ksub >= 256 ? 256 : ((ksub + 31) / 32) * 32;
    adc_lut_kernel<<<grid, block, dsub * sizeof(float), as_stream(stream)>>>(
        queries, ldq, centroids, M, ksub, dsub, metric, lut);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}
