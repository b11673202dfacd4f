// RaBitQ kernels.
//
// (1) 1-bit RaBitQ with faiss RaBitQuantizer / IndexRaBitQ semantics, as driven
//     by src/haag_vq/methods/rabit_quantization.py:20-29 and
//     src/haag_vq/methods/search/rabitq_index.py:42-70 of the reference.
//     faiss is a third-party dependency that is not vendored in the reference
//     tree; the arithmetic is restated from faiss >= 1.11
//     (faiss/impl/RaBitQuantizer.cpp: compute_codes_core, decode_core,
//     RaBitDistanceComputerQ::{set_query,distance_to_code}) and Gao & Long 2024.
//       code  = ceil(D/8) sign bytes of (x - c), bit j%8 of byte j/8,
//               then float or_minus_c_l2sqr (minus ||x||^2 for IP), float dp_multiplier
//       query = (q - c) scalar-quantised to qb bits, bit-plane transposed
//       est   = ||o-c||^2 + ||q-c||^2 - 2 dp_multiplier (c1 <q~,o> + c2 popc(o) - c34)
// (2) Extended (multi-bit) RaBitQ row tail for
//     src/haag_vq/methods/extended_rabitq.py:125-198 (fp64): level search,
//     rescale factor t, MSB-first B-bit packing with float32 (nrm, t) trailer.
#include "common.cuh"
#include "topk.cuh"
#include "rabitq_est.cuh"

namespace b200vq {

// ============================================================ 1-bit =======
// One warp per row. Lane l owns dimensions j0 + l; the sign bits of 32
// consecutive dimensions are gathered with one __ballot_sync.
__global__ void rabitq1_encode_kernel(const float* __restrict__ x, int64_t n, int64_t ldx, int D,
                                      const float* __restrict__ centroid, int metric,
                                      uint8_t* __restrict__ codes) {
    const int lane = threadIdx.x & 31;
    const int64_t wid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int nbytes = (D + 7) >> 3;
    const int code_size = nbytes + 8;
    const float inv_d_sqrt = 1.0f / sqrtf((float)D);
    const bool word_rows = (code_size & 3) == 0 && ((reinterpret_cast<uintptr_t>(codes) & 3) == 0);
    for (int64_t row = wid; row < n; row += nw) {
        const float* xr = x + row * ldx;
        uint8_t* code = codes + row * (int64_t)code_size;
        float norm = 0.f, orl2 = 0.f, dp = 0.f;
        // eight 128-byte loads in flight per warp; the per-lane accumulation order (j = lane, lane + 32, ...)
        // is the one the oracle restates, only the loads are batched
        for (int j0 = 0; j0 < D; j0 += 256) {
            float vv[8], cc[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int j = j0 + 32 * u + lane;
                vv[u] = (j < D) ? ld_stream_f32(xr + j) : 0.f;
                cc[u] = (j < D && centroid) ? __ldg(centroid + j) : 0.f;
            }
            unsigned mine = 0;
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int j = j0 + 32 * u + lane;
                if (j0 + 32 * u < D) {       // warp-uniform
                    const float v = vv[u];
                    const float omc = (j < D) ? v - cc[u] : 0.f;
                    norm = __fmaf_rn(omc, omc, norm);
                    orl2 = __fmaf_rn(v, v, orl2);
                    dp += fabsf(omc);
                    const unsigned word = __ballot_sync(FULL_MASK, omc > 0.f);
                    if (lane == u) mine = word;
                }
            }
            // lane u holds the sign word of dimensions [j0 + 32 u, j0 + 32 u + 32)
            const int byte0 = (j0 >> 3) + 4 * lane;
            if (lane < 8 && byte0 < nbytes) {
                if (word_rows && byte0 + 4 <= nbytes) {
                    *reinterpret_cast<uint32_t*>(code + byte0) = mine;
                } else {
#pragma unroll
                    for (int b = 0; b < 4; ++b)
                        if (byte0 + b < nbytes) code[byte0 + b] = (uint8_t)(mine >> (8 * b));
                }
            }
        }
        for (int o = 16; o > 0; o >>= 1) {
            norm += __shfl_xor_sync(FULL_MASK, norm, o);
            orl2 += __shfl_xor_sync(FULL_MASK, orl2, o);
            dp += __shfl_xor_sync(FULL_MASK, dp, o);
        }
        if (lane == 0) {
            const float eps = 1.1920929e-07f;  // std::numeric_limits<float>::epsilon()
            float sq = sqrtf(norm);
            float inv_norm = (fabsf(norm) < eps) ? 1.0f : (1.0f / sq);
            float dpn = dp * inv_norm;
            dpn *= inv_d_sqrt;
            float inv_dp = (fabsf(dpn) < eps) ? 1.0f : (1.0f / dpn);
            float f0 = norm;
            if (metric == B200VQ_METRIC_IP) f0 -= orl2;
            float f1 = inv_dp * sq;
            uint32_t u0 = __float_as_uint(f0), u1 = __float_as_uint(f1);
            uint8_t* f = code + nbytes;
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                f[b] = (uint8_t)(u0 >> (8 * b));
                f[4 + b] = (uint8_t)(u1 >> (8 * b));
            }
        }
    }
}

// The same encoder for 16-byte aligned rows with D % 4 == 0: a lane owns 4 consecutive dimensions of every
// 128-dimension block (one 128-bit load of x and of the centroid), its 4 sign bits are merged with the neighbours'
// by three xor-shuffles into the block's four 32-bit words, and the three sums run on four independent chains.
// (The reference's sums are numpy / faiss reductions in their own order; parity on the two factors is to 1e-6.)
__global__ void __launch_bounds__(256)
rabitq1_encode_vec_kernel(const float* __restrict__ x, int64_t n, int64_t ldx, int D,
                          const float* __restrict__ centroid, int metric, uint8_t* __restrict__ codes) {
    const int lane = threadIdx.x & 31;
    // the warp index through a shuffle: the compiler then knows the row loop is warp-uniform and does not wrap every
    // shuffle inside it in collective warp synchronisation
    const int64_t wid = ((int64_t)blockIdx.x * blockDim.x >> 5) + __shfl_sync(FULL_MASK, (int)(threadIdx.x >> 5), 0);
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int nbytes = (D + 7) >> 3;
    const int code_size = nbytes + 8;
    const float inv_d_sqrt = 1.0f / sqrtf((float)D);
    const bool word_rows = (code_size & 3) == 0 && ((reinterpret_cast<uintptr_t>(codes) & 3) == 0);
    for (int64_t row = wid; row < n; row += nw) {
        const float* xr = x + row * ldx;
        uint8_t* code = codes + row * (int64_t)code_size;
        float nr[4] = {0.f, 0.f, 0.f, 0.f}, o2[4] = {0.f, 0.f, 0.f, 0.f}, ad[4] = {0.f, 0.f, 0.f, 0.f};
        constexpr int U = 4;
        for (int j0 = 0; j0 < D; j0 += 128 * U) {
            // four 128-dimension blocks per step: eight 128-bit loads in flight per lane (2 blocks: 0.70 of HBM, 4: 0.80, 6: 0.69 - registers)
            float4 v[U], c[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int j = j0 + 128 * u + 4 * lane;
                v[u] = j < D ? ld_stream_f4(reinterpret_cast<const float4*>(xr + j)) : make_float4(0.f, 0.f, 0.f, 0.f);
                c[u] = (j < D && centroid) ? __ldg(reinterpret_cast<const float4*>(centroid + j))
                                           : make_float4(0.f, 0.f, 0.f, 0.f);
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                // (a block past D holds zeros: no sign bit, nothing added, its store is masked - no branch around the
                // shuffles, which the compiler would otherwise wrap in collective warp synchronisation)
                const float o0 = v[u].x - c[u].x, o1 = v[u].y - c[u].y, o2_ = v[u].z - c[u].z, o3 = v[u].w - c[u].w;
                nr[0] = __fmaf_rn(o0, o0, nr[0]);
                nr[1] = __fmaf_rn(o1, o1, nr[1]);
                nr[2] = __fmaf_rn(o2_, o2_, nr[2]);
                nr[3] = __fmaf_rn(o3, o3, nr[3]);
                o2[0] = __fmaf_rn(v[u].x, v[u].x, o2[0]);
                o2[1] = __fmaf_rn(v[u].y, v[u].y, o2[1]);
                o2[2] = __fmaf_rn(v[u].z, v[u].z, o2[2]);
                o2[3] = __fmaf_rn(v[u].w, v[u].w, o2[3]);
                ad[0] += fabsf(o0);
                ad[1] += fabsf(o1);
                ad[2] += fabsf(o2_);
                ad[3] += fabsf(o3);
                // bit j % 8 of byte j / 8: nibble -> byte -> half word -> word (held by lanes 0, 8, 16, 24)
                unsigned w = (o0 > 0.f ? 1u : 0u) | (o1 > 0.f ? 2u : 0u) | (o2_ > 0.f ? 4u : 0u) | (o3 > 0.f ? 8u : 0u);
                w |= __shfl_xor_sync(FULL_MASK, w, 1) << 4;
                w |= __shfl_xor_sync(FULL_MASK, w, 2) << 8;
                w |= __shfl_xor_sync(FULL_MASK, w, 4) << 16;
                const int byte0 = ((j0 + 128 * u) >> 3) + (lane >> 1);
                if ((lane & 7) == 0 && byte0 < nbytes) {
                    if (word_rows && byte0 + 4 <= nbytes) {
                        *reinterpret_cast<uint32_t*>(code + byte0) = w;
                    } else {
#pragma unroll
                        for (int b = 0; b < 4; ++b)
                            if (byte0 + b < nbytes) code[byte0 + b] = (uint8_t)(w >> (8 * b));
                    }
                }
            }
        }
        float norm = (nr[0] + nr[1]) + (nr[2] + nr[3]), orl2 = (o2[0] + o2[1]) + (o2[2] + o2[3]);
        float dp = (ad[0] + ad[1]) + (ad[2] + ad[3]);
        for (int o = 16; o > 0; o >>= 1) {
            norm += __shfl_xor_sync(FULL_MASK, norm, o);
            orl2 += __shfl_xor_sync(FULL_MASK, orl2, o);
            dp += __shfl_xor_sync(FULL_MASK, dp, o);
        }
        if (lane == 0) {
            const float eps = 1.1920929e-07f;  // std::numeric_limits<float>::epsilon()
            float sq = sqrtf(norm);
            float inv_norm = (fabsf(norm) < eps) ? 1.0f : (1.0f / sq);
            float dpn = dp * inv_norm;
            dpn *= inv_d_sqrt;
            float inv_dp = (fabsf(dpn) < eps) ? 1.0f : (1.0f / dpn);
            float f0 = norm;
            if (metric == B200VQ_METRIC_IP) f0 -= orl2;
            float f1 = inv_dp * sq;
            uint32_t u0 = __float_as_uint(f0), u1 = __float_as_uint(f1);
            uint8_t* f = code + nbytes;
            if (word_rows && (nbytes & 3) == 0) {
                reinterpret_cast<uint32_t*>(f)[0] = u0;
                reinterpret_cast<uint32_t*>(f)[1] = u1;
            } else {
#pragma unroll
                for (int b = 0; b < 4; ++b) {
                    f[b] = (uint8_t)(u0 >> (8 * b));
                    f[4 + b] = (uint8_t)(u1 >> (8 * b));
                }
            }
        }
    }
}

__device__ __forceinline__ float load_f32_unaligned(const uint8_t* p) {
    uint32_t u = (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) |
                 ((uint32_t)p[3] << 24);
    return __uint_as_float(u);
}

// One warp per row: the two possible magnitudes (bit - 0.5) * dp_multiplier * 2 * inv_d_sqrt are formed
// once per row with faiss' left-to-right products, then every lane selects by sign bit and adds the
// centroid for four consecutive dimensions (128-bit stores when the output rows are aligned).
__global__ void rabitq1_decode_kernel(const uint8_t* __restrict__ codes, int64_t n, int D,
                                      const float* __restrict__ centroid, float* __restrict__ out,
                                      int64_t ldo) {
    const int nbytes = (D + 7) >> 3;
    const int code_size = nbytes + 8;
    const float inv_d_sqrt = 1.0f / sqrtf((float)D);
    const int lane = threadIdx.x & 31;
    const int64_t wid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const bool vec = (D % 4 == 0) && (ldo % 4 == 0) && ((reinterpret_cast<uintptr_t>(out) & 15) == 0) &&
                     (!centroid || (reinterpret_cast<uintptr_t>(centroid) & 15) == 0);
    for (int64_t row = wid; row < n; row += nw) {
        const uint8_t* code = codes + row * (int64_t)code_size;
        const float dpm = load_f32_unaligned(code + nbytes + 4);
        // (bit - 0.5f) * dp_multiplier * 2 * inv_d_sqrt + c[j], left to right like faiss
        const float vpos = __fmul_rn(__fmul_rn(__fmul_rn(0.5f, dpm), 2.f), inv_d_sqrt);
        const float vneg = __fmul_rn(__fmul_rn(__fmul_rn(-0.5f, dpm), 2.f), inv_d_sqrt);
        float* o = out + row * ldo;
        if (vec) {
            for (int j = 4 * lane; j < D; j += 128) {
                const uint32_t nib = ((uint32_t)__ldg(code + (j >> 3)) >> (j & 4)) & 0xFu;
                float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
                if (centroid) c = __ldg(reinterpret_cast<const float4*>(centroid + j));
                float4 r;
                r.x = __fadd_rn((nib & 1u) ? vpos : vneg, c.x);
                r.y = __fadd_rn((nib & 2u) ? vpos : vneg, c.y);
                r.z = __fadd_rn((nib & 4u) ? vpos : vneg, c.z);
                r.w = __fadd_rn((nib & 8u) ? vpos : vneg, c.w);
                *reinterpret_cast<float4*>(o + j) = r;
            }
        } else {
            for (int j = lane; j < D; j += 32) {
                const bool bit = (code[j >> 3] >> (j & 7)) & 1;
                o[j] = __fadd_rn(bit ? vpos : vneg, centroid ? centroid[j] : 0.f);
            }
        }
    }
}

// One CTA (256 threads) per query.
__global__ void rabitq1_prepare_queries_kernel(const float* __restrict__ queries, int64_t ldq, int D,
                                               const float* __restrict__ centroid, int qb,
                                               uint32_t* __restrict__ planes,
                                               float* __restrict__ qfac) {
    extern __shared__ float rq[];  // D residuals, then D quantised ints (as int)
    int* qq = reinterpret_cast<int*>(rq + D);
    __shared__ float red[4][32];
    __shared__ unsigned long long s_sum;
    const int q = blockIdx.x, tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int nwarp = blockDim.x >> 5;
    const float* qr = queries + (int64_t)q * ldq;
    float l2 = 0.f, nrm = 0.f;
    float mn = __int_as_float(0x7f800000), mx = __int_as_float(0xff800000);
    for (int j = tid; j < D; j += blockDim.x) {
        float v = qr[j];
        float r = v - (centroid ? centroid[j] : 0.f);
        rq[j] = r;
        l2 = __fmaf_rn(r, r, l2);
        nrm = __fmaf_rn(v, v, nrm);
        mn = fminf(mn, r);
        mx = fmaxf(mx, r);
    }
    for (int o = 16; o > 0; o >>= 1) {
        l2 += __shfl_xor_sync(FULL_MASK, l2, o);
        nrm += __shfl_xor_sync(FULL_MASK, nrm, o);
        mn = fminf(mn, __shfl_xor_sync(FULL_MASK, mn, o));
        mx = fmaxf(mx, __shfl_xor_sync(FULL_MASK, mx, o));
    }
    if (lane == 0) {
        red[0][w] = l2;
        red[1][w] = nrm;
        red[2][w] = mn;
        red[3][w] = mx;
    }
    if (tid == 0) s_sum = 0ull;
    __syncthreads();
    l2 = 0.f;
    nrm = 0.f;
    mn = __int_as_float(0x7f800000);
    mx = __int_as_float(0xff800000);
    for (int i = 0; i < nwarp; ++i) {
        l2 += red[0][i];
        nrm += red[1][i];
        mn = fminf(mn, red[2][i]);
        mx = fmaxf(mx, red[3][i]);
    }
    const float pow2 = (float)(1 << qb);
    const float delta = (mx - mn) / (pow2 - 1.f);
    const float inv_delta = 1.0f / delta;
    unsigned long long local = 0;
    for (int j = tid; j < D; j += blockDim.x) {
        float vq = (rq[j] - mn) * inv_delta;
        int v = (int)roundf(vq);  // std::round: half away from zero
        local += (unsigned long long)(long long)v;
        qq[j] = min(255, max(0, v));
    }
    for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(FULL_MASK, local, o);
    if (lane == 0) atomicAdd(&s_sum, local);
    __syncthreads();
    // bit-plane transpose with ballots: warp w handles words w, w+nwarp, ...
    const int W = (D + 31) >> 5;
    for (int wd = w; wd < W; wd += nwarp) {
        int j = wd * 32 + lane;
        int v = (j < D) ? qq[j] : 0;
        for (int b = 0; b < qb; ++b) {
            unsigned word = __ballot_sync(FULL_MASK, (v >> b) & 1);
            if (lane == 0) planes[((size_t)q * qb + b) * W + wd] = word;
        }
    }
    if (tid == 0) {
        const float inv_d = 1.0f / sqrtf((float)D);
        float* f = qfac + (size_t)q * 8;
        f[0] = 2.f * delta * inv_d;                                    // c1
        f[1] = 2.f * mn * inv_d;                                       // c2
        f[2] = inv_d * (delta * (float)s_sum + (float)D * mn);         // c34
        f[3] = l2;                                                     // ||q - c||^2
        f[4] = nrm;                                                    // ||q||^2
        f[5] = 0.f;
        f[6] = 0.f;
        f[7] = 0.f;
    }
}

// Scan: grid (nq, nchunks), 256 threads. A warp stages 32 code rows into shared
// memory with coalesced loads (row pitch padded to an odd number of words so
// the per-lane row walk is bank-conflict free), then lane = row.
template <int CAP>
__global__ void __launch_bounds__(256)
    rabitq1_scan_kernel(const uint8_t* __restrict__ codes, int64_t n, int D,
                        const uint32_t* __restrict__ planes, const float* __restrict__ qfac, int qb,
                        int metric, int k, int64_t chunk_rows, uint64_t* __restrict__ part,
                        const uint32_t* __restrict__ gate = nullptr, const float* __restrict__ after_vals = nullptr,
                        const uint32_t* __restrict__ after_ids = nullptr) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int q = blockIdx.x, chunk = blockIdx.y, nchunks = gridDim.y;
    if (gate && gate[q] == 0) return;   // gated launch: only flagged queries (rabitq_tc.cu fallback)
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int W = (D + 31) >> 5;
    const int nbytes = (D + 7) >> 3;
    const int code_size = nbytes + 8;
    const int pitch_w = ((code_size + 3) >> 2) | 1;  // words per staged row, odd
    uint32_t* qpl = reinterpret_cast<uint32_t*>(smem_raw);                  // [qb][W]
    uint32_t* stage = qpl + ((qb * W + 3) & ~3);                            // [8 warps][32][pitch_w]
    uint64_t* wbuf = reinterpret_cast<uint64_t*>(
        (reinterpret_cast<uintptr_t>(stage + 8 * 32 * pitch_w) + 15) & ~(uintptr_t)15);
    const int n2 = next_pow2(8 * k);
    uint64_t* mbuf = wbuf + 8 * CAP;

    for (int i = threadIdx.x; i < qb * W; i += blockDim.x) qpl[i] = planes[(size_t)q * qb * W + i];
    const float c1 = qfac[(size_t)q * 8 + 0], c2 = qfac[(size_t)q * 8 + 1],
                c34 = qfac[(size_t)q * 8 + 2], qc2 = qfac[(size_t)q * 8 + 3],
                qn2 = qfac[(size_t)q * 8 + 4];
    __syncthreads();

    WarpTopK<CAP> tk;
    tk.init(wbuf + w * CAP, k);
    uint64_t floor_key = 0;   // continuation (b200vq_rabitq1_scan_topk_after): only keys after the previous page's last
    if (after_vals) floor_key = make_key(metric == B200VQ_METRIC_IP ? -after_vals[q] : after_vals[q], after_ids[q]);
    uint32_t* st = stage + (size_t)w * 32 * pitch_w;
    uint8_t* st8 = reinterpret_cast<uint8_t*>(st);

    const int64_t r0 = (int64_t)chunk * chunk_rows;
    int64_t r1 = r0 + chunk_rows;
    if (r1 > n) r1 = n;
    for (int64_t base = r0 + (int64_t)w * 32; base < r1; base += 256) {
        int rows = (int)((r1 - base) < 32 ? (r1 - base) : 32);
        // coalesced byte-range copy of `rows` code rows
        const uint8_t* src = codes + base * (int64_t)code_size;
        const int64_t nb = (int64_t)rows * code_size;
        __syncwarp();
        if ((((uintptr_t)src) & 3) == 0 && (code_size & 3) == 0) {
            const uint32_t* s4 = reinterpret_cast<const uint32_t*>(src);
            const int wpr = code_size >> 2;
            for (int i = lane; i < rows * wpr; i += 32) {
                int r = i / wpr, c = i - r * wpr;
                st[r * pitch_w + c] = __ldg(s4 + i);
            }
        } else {
            for (int64_t i = lane; i < nb; i += 32) {
                int r = (int)(i / code_size), c = (int)(i - (int64_t)r * code_size);
                st8[(size_t)r * pitch_w * 4 + c] = src[i];
            }
        }
        __syncwarp();
        const bool valid = lane < rows;
        float est = 0.f;
        if (valid) {
            const uint32_t* cw = st + lane * pitch_w;
            const uint8_t* cb = reinterpret_cast<const uint8_t*>(cw);
            uint32_t dot = 0, sumq = 0;
            for (int wd = 0; wd < W; ++wd) {
                uint32_t c = cw[wd];
                if (wd == W - 1 && (D & 31)) c &= (1u << (D & 31)) - 1u;  // drop factor bytes
                sumq += __popc(c);
                for (int b = 0; b < qb; ++b) dot += (uint32_t)__popc(c & qpl[b * W + wd]) << b;
            }
            float f0 = load_f32_unaligned(cb + nbytes);
            float f1 = load_f32_unaligned(cb + nbytes + 4);
            est = rabitq1_estimate((float)dot, (float)sumq, f0, -2.f * f1, c1, c2, c34, qc2, qn2, metric);
        }
        tk.offer(valid && (!after_vals || make_key(est, (uint32_t)(base + lane)) > floor_key), est,
                 (uint32_t)(base + lane));
    }
    // emit
    tk.finish();
    for (int i = threadIdx.x; i < n2; i += blockDim.x) mbuf[i] = KEY_INF;
    __syncthreads();
    for (int i = lane; i < tk.count; i += 32) mbuf[w * k + i] = tk.buf[i];
    block_bitonic_sort(mbuf, n2);
    uint64_t* dst = part + ((size_t)q * nchunks + chunk) * k;
    for (int j = threadIdx.x; j < k; j += blockDim.x) dst[j] = mbuf[j];
}

// ======================================================= extended =========
// One CTA per row (grid-stride). levels/bounds in shared memory.
// One warp per row (no CTA-wide barriers): level search for four elements per lane in flight, the row's
// indices in a per-warp shared buffer, then one lane per output byte.
__global__ void __launch_bounds__(256) erabitq_pack_kernel(const double* __restrict__ srot, const double* __restrict__ nrm,
                                                           int64_t n, int D, int B, const double* __restrict__ levels,
                                                           const double* __restrict__ bounds, uint8_t* __restrict__ codes) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* lv = reinterpret_cast<double*>(smem_raw);        // [2^B]
    double* bd = lv + (1 << B);                              // [2^B - 1] (padded to 2^B)
    const int L = 1 << B;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int dpad = (D + 15) & ~15;
    uint8_t* idx = reinterpret_cast<uint8_t*>(bd + L) + (size_t)w * dpad;  // [warps][D]
    for (int i = tid; i < L; i += blockDim.x) lv[i] = levels[i];
    for (int i = tid; i < L - 1; i += blockDim.x) bd[i] = bounds[i];
    __syncthreads();
    const int ibytes = (D * B + 7) >> 3;
    const int code_size = ibytes + 8;
    const int64_t wid = ((int64_t)blockIdx.x * blockDim.x + tid) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t row = wid; row < n; row += nw) {
        const double* s = srot + row * (int64_t)D;
        double num = 0.0, den = 0.0;
        for (int d0 = lane; d0 < D; d0 += 128) {
            double v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) v[u] = (d0 + 32 * u < D) ? s[d0 + 32 * u] : 0.0;
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int d = d0 + 32 * u;
                if (d < D) {
                    // np.searchsorted(bounds, v) (side='left'): number of bounds < v
                    int lo = 0, hi = L - 1;
                    while (lo < hi) {
                        const int mid = (lo + hi) >> 1;
                        if (bd[mid] < v[u]) lo = mid + 1;
                        else hi = mid;
                    }
                    idx[d] = (uint8_t)lo;
                    const double sh = lv[lo];
                    num += v[u] * sh;
                    den += sh * sh;
                }
            }
        }
        for (int o = 16; o > 0; o >>= 1) {
            num += __shfl_xor_sync(FULL_MASK, num, o);
            den += __shfl_xor_sync(FULL_MASK, den, o);
        }
        __syncwarp();
        uint8_t* code = codes + row * (int64_t)code_size;
        // MSB-first B-bit packing, one lane per output byte
        for (int byte = lane; byte < ibytes; byte += 32) {
            const int b0 = byte * 8, b1 = b0 + 8;
            uint32_t acc = 0;
            for (int d = b0 / B; d < D && d * B < b1; ++d) {
                const int shift = b1 - (d * B + B);
                const uint32_t qv = idx[d];
                acc |= ((shift >= 0) ? (qv << shift) : (qv >> (-shift))) & 0xffu;
            }
            code[byte] = (uint8_t)acc;
        }
        if (lane == 0) {
            const double t = (den > 1e-12) ? (num / den) : 1.0;
            const uint32_t u0 = __float_as_uint((float)nrm[row]);
            const uint32_t u1 = __float_as_uint((float)t);
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                code[ibytes + b] = (uint8_t)(u0 >> (8 * b));
                code[ibytes + 4 + b] = (uint8_t)(u1 >> (8 * b));
            }
        }
        __syncwarp();
    }
}

// ---------------------------------------------------------------------------------------------
// Extended-RaBitQ encode with the rotation on the tensor cores (b200vq_erabitq_pack_tc).
// The reference computes s = ((x - c) / |x - c|) P sqrt(D) in fp64 (extended_rabitq.py:140-149): an
// O(n D^2) DGEMM. Here the rotation is the fp32-accurate tcgen05 GEMM (rotate_tc.cu, y = (x - c32) P32),
// whose error is bounded by err_rel |x - c| + err_abs, and only the coordinates whose level index could
// change inside that bound (|s - mid-point| < eps) are recomputed as exact fp64 dot products
// sum_k (x_k - c_k) P[k][d] by the whole warp: the level indices are those of the fp64 computation.
// y == nullptr: every coordinate is computed exactly (shapes the GEMM does not cover).
template <typename T>
__global__ void __launch_bounds__(256) resid_norm_kernel(const T* __restrict__ x, int64_t n, int64_t ldx, int D,
                                                         const double* __restrict__ c, double* __restrict__ nrm) {
    const int lane = threadIdx.x & 31;
    const int64_t wid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t row = wid; row < n; row += nw) {
        const T* xr = x + row * ldx;
        double acc = 0.0;
        for (int d = lane; d < D; d += 32) {
            const double r = (double)xr[d] - c[d];
            acc = fma(r, r, acc);
        }
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(FULL_MASK, acc, o);
        if (lane == 0) nrm[row] = sqrt(acc);
    }
}

__device__ __forceinline__ int erabitq_level_of(const double* bd, int L, double v) {
    int lo = 0, hi = L - 1;          // np.searchsorted(bounds, v) (side='left'): number of bounds < v
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (bd[mid] < v) lo = mid + 1;
        else hi = mid;
    }
    return lo;
}

// BT = B at compile time (4 or 8: the level search and the packing unroll), 0 = any B
template <typename T, int BT>
__global__ void __launch_bounds__(256, 4)
erabitq_pack_tc_kernel(const float* __restrict__ y, int64_t ldy, const T* __restrict__ x, int64_t ldx,
                       const double* __restrict__ c, const double* __restrict__ PT, const double* __restrict__ nrm,
                       int64_t n, int D, int B_rt, const double* __restrict__ levels, const double* __restrict__ bounds,
                       double err_rel, double err_abs, uint8_t* __restrict__ codes, unsigned long long* __restrict__ stats,
                       unsigned* __restrict__ bcnt, unsigned long long* __restrict__ bucket, unsigned bcap,
                       double* __restrict__ numden) {
    const int B = BT ? BT : B_rt;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* lv = reinterpret_cast<double*>(smem_raw);        // [2^B]
    double* bd = lv + (1 << B);                              // [2^B - 1] (padded to 2^B)
    const int L = 1 << B;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int dpad = (D + 15) & ~15;
    uint8_t* idx = reinterpret_cast<uint8_t*>(bd + L) + (size_t)w * dpad * 3;  // [warps][D] indices
    uint16_t* flist = reinterpret_cast<uint16_t*>(idx + dpad);                 // [warps][D] flagged coordinates
    // fp32 copy of the bounds for the level search of pass 1 (the exact fp64 search is only needed for coordinates
    // within the error bound of a mid-point, and those are re-evaluated anyway)
    __shared__ float bd32[256];
    __shared__ float bmax_s;
    for (int i = tid; i < L; i += blockDim.x) lv[i] = levels[i];
    for (int i = tid; i < L - 1; i += blockDim.x) {
        bd[i] = bounds[i];
        bd32[i] = (float)bounds[i];
    }
    if (tid == 0) bmax_s = fmaxf(fabsf((float)bounds[0]), fabsf((float)bounds[L - 2])) * 1.000001f;   // sorted bounds
    __syncthreads();
    const float bmax = bmax_s;
    const int ibytes = (D * B + 7) >> 3;
    const int code_size = ibytes + 8;
    const double sqrt_d = sqrt((double)D);
    const int64_t wid = ((int64_t)blockIdx.x * blockDim.x >> 5) + __shfl_sync(FULL_MASK, tid >> 5, 0);
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    // a lane's four coordinates are B / 2 whole bytes for B = 8, 4, 2: stored straight from registers
    const int galign = B == 8 ? 4 : (B == 4 ? 2 : 1);
    const bool direct = (B == 8 || B == 4 || B == 2) && (D & 3) == 0 && (code_size % galign) == 0 &&
                        (reinterpret_cast<uintptr_t>(codes) % galign) == 0;
    unsigned long long nfix_total = 0;
    unsigned ndef_lane = 0;
    for (int64_t row = wid; row < n; row += nw) {
        const double nr = nrm[row];
        const double nsafe = fmax(nr, 1e-12);
        const double sc = sqrt_d / nsafe;
        const double eps_s = (err_rel * nr + err_abs) * sc;
        const float sc32 = (float)sc;
        // fp32 margin: eps_s plus the rounding of s32 = fl(y fl(sc)) (2^-23 relative) and of the fp32 bounds (2^-24)
        const float eps32 = (float)(eps_s * 1.000001);
        const float* yr = y ? y + row * ldy : nullptr;
        uint8_t* code = codes + row * (int64_t)code_size;
        double num = 0.0, den = 0.0;
        int nflag = 0;
        // pass 1: four consecutive coordinates per lane and step (one 128-bit load when the row allows it)
        const bool vec4 = yr && ((reinterpret_cast<uintptr_t>(yr) & 15) == 0);
        // (the next step's 128-bit load is issued before this step's arithmetic: with one load in flight per lane
        // 14 % of the samples waited on it)
        auto load4 = [&](int dbase) {
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (yr && dbase < D) {
                if (vec4 && dbase + 3 < D) {
                    v = __ldg(reinterpret_cast<const float4*>(yr + dbase));
                } else {
                    if (dbase + 0 < D) v.x = __ldg(yr + dbase + 0);
                    if (dbase + 1 < D) v.y = __ldg(yr + dbase + 1);
                    if (dbase + 2 < D) v.z = __ldg(yr + dbase + 2);
                    if (dbase + 3 < D) v.w = __ldg(yr + dbase + 3);
                }
            }
            return v;
        };
        float4 vnext = load4(4 * lane);
        for (int d0 = 0; d0 < D; d0 += 128) {
            const int db = d0 + 4 * lane;
            const float4 vcur = vnext;
            vnext = load4(db + 128);
            const float yv[4] = {vcur.x, vcur.y, vcur.z, vcur.w};
            unsigned fmask = 0, los = 0;      // flag bits and the four indices (8 bits each) of this lane's coordinates
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int d = db + e;
                if (d < D) {
                    if (yr) {
                        const float s32 = __fmul_rn(yv[e], sc32);
                        int lo = 0;                      // number of bounds < s32 (branch-free: L is a power of two)
                        if (BT) {
#pragma unroll
                            for (int step = (1 << (BT ? BT : 1)) >> 1; step > 0; step >>= 1)
                                if (bd32[lo + step - 1] < s32) lo += step;
                        } else {
                            for (int step = L >> 1; step > 0; step >>= 1)
                                if (bd32[lo + step - 1] < s32) lo += step;
                        }
                        const float margin = eps32 + 5e-7f * (fabsf(s32) + bmax);
                        bool flag = (lo > 0 && s32 - bd32[lo - 1] < margin) || (lo < L - 1 && bd32[lo] - s32 < margin);
                        idx[d] = (uint8_t)lo;
                        los |= (unsigned)lo << (8 * e);
                        const double sv = (double)yv[e] * sc;
                        const double sh = lv[lo];
                        num = fma(sv, sh, num);
                        den = fma(sh, sh, den);
                        if (flag && bcnt) {
                            // deferred: erabitq_fixup_kernel re-evaluates the coordinate (bucket per coordinate, so
                            // that a row of P^T is read once for all rows that need it); a full bucket falls back
                            // to the in-kernel fix-up below
                            const unsigned pos = atomicAdd(bcnt + d, 1u);
                            if (pos < bcap) {
                                bucket[(size_t)d * bcap + pos] = ((unsigned long long)row << 8) | (unsigned long long)lo;
                                flag = false;
                                ++ndef_lane;
                            }
                        }
                        if (flag) fmask |= 1u << e;
                    } else {
                        fmask |= 1u << e;
                    }
                }
            }
            if (direct && db < D) {
                if (B == 8) {
                    *reinterpret_cast<uint32_t*>(code + db) = los;
                } else if (B == 4) {
                    const unsigned b0 = ((los & 0xFu) << 4) | ((los >> 8) & 0xFu), b1 = (((los >> 16) & 0xFu) << 4) | (los >> 24);
                    *reinterpret_cast<uint16_t*>(code + (db >> 1)) = (uint16_t)(b0 | (b1 << 8));
                } else {
                    code[db >> 2] = (uint8_t)(((los & 3u) << 6) | (((los >> 8) & 3u) << 4) | (((los >> 16) & 3u) << 2) | (los >> 24));
                }
            }
            if (__any_sync(FULL_MASK, fmask != 0u)) {
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const bool flag = (fmask >> e) & 1u;
                    const unsigned m = __ballot_sync(FULL_MASK, flag);
                    if (flag) flist[nflag + __popc(m & ((1u << lane) - 1u))] = (uint16_t)(db + e);
                    nflag += __popc(m);
                }
            }
        }
        __syncwarp();
        // pass 2: exact fp64 dot products, four flagged coordinates at a time (x - c is loaded once for the four)
        const T* xr = x + row * ldx;
        for (int f0 = 0; f0 < nflag; f0 += 4) {
            int dd[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) dd[e] = flist[min(f0 + e, nflag - 1)];
            const double* pt0 = PT + (int64_t)dd[0] * D;
            const double* pt1 = PT + (int64_t)dd[1] * D;
            const double* pt2 = PT + (int64_t)dd[2] * D;
            const double* pt3 = PT + (int64_t)dd[3] * D;
            double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
#pragma unroll 4
            for (int k = lane; k < D; k += 32) {
                const double rk = (double)xr[k] - c[k];
                a0 = fma(rk, __ldg(pt0 + k), a0);
                a1 = fma(rk, __ldg(pt1 + k), a1);
                a2 = fma(rk, __ldg(pt2 + k), a2);
                a3 = fma(rk, __ldg(pt3 + k), a3);
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
                    const int d = dd[e];
                    const double sv = (av[e] / nsafe) * sqrt_d;
                    const int lo = erabitq_level_of(bd, L, sv);
                    if (yr) {
                        const double so = (double)yr[d] * sc;
                        const double sho = lv[idx[d]];
                        num -= so * sho;
                        den -= sho * sho;
                    }
                    idx[d] = (uint8_t)lo;
                    const double sh = lv[lo];
                    num = fma(sv, sh, num);
                    den = fma(sh, sh, den);
                }
            }
        }
        nfix_total += (unsigned long long)nflag;
        for (int o = 16; o > 0; o >>= 1) {
            num += __shfl_xor_sync(FULL_MASK, num, o);
            den += __shfl_xor_sync(FULL_MASK, den, o);
        }
        __syncwarp();
        if (!(direct && nflag == 0))
        for (int byte = lane; byte < ibytes; byte += 32) {       // MSB-first B-bit packing, one lane per output byte
            const int b0 = byte * 8, b1 = b0 + 8;
            uint32_t acc = 0;
            for (int d = b0 / B; d < D && d * B < b1; ++d) {
                const int shift = b1 - (d * B + B);
                const uint32_t qv = idx[d];
                acc |= ((shift >= 0) ? (qv << shift) : (qv >> (-shift))) & 0xffu;
            }
            code[byte] = (uint8_t)acc;
        }
        if (lane == 0) {
            if (numden) {
                numden[2 * row] = num;
                numden[2 * row + 1] = den;
            }
            const double t = (den > 1e-12) ? (num / den) : 1.0;
            const uint32_t u0 = __float_as_uint((float)nr);
            const uint32_t u1 = __float_as_uint((float)t);
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                code[ibytes + b] = (uint8_t)(u0 >> (8 * b));
                code[ibytes + 4 + b] = (uint8_t)(u1 >> (8 * b));
            }
        }
        __syncwarp();
    }
    for (int o = 16; o > 0; o >>= 1) ndef_lane += __shfl_xor_sync(FULL_MASK, ndef_lane, o);
    nfix_total += ndef_lane;
    if (stats && lane == 0 && nfix_total) atomicAdd(stats, nfix_total);
}

__device__ __forceinline__ uint32_t erabitq_index(const uint8_t* code, int d, int B, int ibytes) {
    int off = d * B;
    int byte = off >> 3;
    uint32_t wv = ((uint32_t)code[byte] << 8) | (byte + 1 < ibytes ? code[byte + 1] : 0u);
    int shift = 16 - ((off & 7) + B);
    return (wv >> shift) & ((1u << B) - 1u);
}

// One warp per row; levels[i] / sqrt(D) is formed once per CTA (the same fp64 quotient the per-element form
// computed 2^-B as often), t and nrm are read once per row.
__global__ void __launch_bounds__(256) erabitq_unpack_kernel(const uint8_t* __restrict__ codes, int64_t n, int D, int B,
                                                             const double* __restrict__ levels, double* __restrict__ shat,
                                                             double* __restrict__ nrm_out) {
    __shared__ double lvs[256];
    const double sqrt_d = sqrt((double)D);
    for (int i = threadIdx.x; i < (1 << B); i += blockDim.x) lvs[i] = levels[i] / sqrt_d;
    __syncthreads();
    const int ibytes = (D * B + 7) >> 3;
    const int code_size = ibytes + 8;
    const int lane = threadIdx.x & 31;
    const int64_t wid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t row = wid; row < n; row += nw) {
        const uint8_t* code = codes + row * (int64_t)code_size;
        const double t = (double)load_f32_unaligned(code + ibytes + 4);
        double* o = shat + row * (int64_t)D;
        for (int d = lane; d < D; d += 32) o[d] = lvs[erabitq_index(code, d, B, ibytes)] * t;
        if (lane == 0) nrm_out[row] = (double)load_f32_unaligned(code + ibytes);
    }
}

// Code-domain scan: thread = row, levels (fp32) and the rotated query in shared
// memory. est = qconst + nrm^2 t^2 sum(l^2)/D - 2 nrm t/sqrt(D) <l[idx], qrot>  (L2)
//        ip  = qc_dot + nrm t/sqrt(D) <l[idx], qrot>                         (IP; qconst = q.c)
// Decode for the tensor-core estimator scan: x_i = a_i * levels[idx] (fp32), a_i = nrm_i t_i / sqrt(D), and
// ||x_i||^2 = a_i^2 * sum levels[idx]^2. With these the estimator of erabitq_scan_kernel,
// qc + a^2 l2 - 2 a <levels[idx], q>, IS the squared L2 distance between x_i and the rotated query (and
// qc + a <levels[idx], q> the inner product), so the exact flat search on tcgen05 (flat_tc.cu) serves it.
// One warp per row; HBM-bound: ceil(D B / 8) + 8 bytes in, 4 D + 4 out.
__global__ void __launch_bounds__(256) erabitq_decode_f32_kernel(const uint8_t* __restrict__ codes, int64_t n, int D, int B,
                                                                 const float* __restrict__ levels,
                                                                 float* __restrict__ x, int64_t ldx,
                                                                 float* __restrict__ xnorm) {
    __shared__ float lv[256];
    for (int i = threadIdx.x; i < (1 << B); i += blockDim.x) lv[i] = levels[i];
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int ibytes = (D * B + 7) >> 3, code_size = ibytes + 8;
    const float inv_sqrt_d = 1.0f / sqrtf((float)D);
    const int64_t wid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t row = wid; row < n; row += nw) {
        const uint8_t* code = codes + row * (int64_t)code_size;
        const float a = load_f32_unaligned(code + ibytes) * load_f32_unaligned(code + ibytes + 4) * inv_sqrt_d;
        float l2 = 0.f;
        float* o = x + row * ldx;
        for (int d = lane; d < D; d += 32) {
            const float l = lv[erabitq_index(code, d, B, ibytes)];
            l2 = __fmaf_rn(l, l, l2);
            o[d] = a * l;
        }
        for (int s = 16; s > 0; s >>= 1) l2 += __shfl_xor_sync(FULL_MASK, l2, s);
        if (lane == 0) xnorm[row] = a * a * l2;
    }
}

// a_i = nrm_i t_i / sqrt(D) (the row scale of the estimator), with the products of erabitq_scan_kernel
__global__ void erabitq_rowscale_kernel(const uint8_t* __restrict__ codes, int64_t n, int D, int B, float* __restrict__ a) {
    const int ibytes = (D * B + 7) >> 3, code_size = ibytes + 8;
    const float inv_sqrt_d = 1.0f / sqrtf((float)D);
    for (int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; row < n; row += (int64_t)gridDim.x * blockDim.x) {
        const uint8_t* code = codes + row * (int64_t)code_size;
        a[row] = __fmul_rn(__fmul_rn(load_f32_unaligned(code + ibytes), load_f32_unaligned(code + ibytes + 4)), inv_sqrt_d);
    }
}

template <int CAP>
__global__ void __launch_bounds__(256)
    erabitq_scan_kernel(const uint8_t* __restrict__ codes, int64_t n, int D, int B,
                        const float* __restrict__ levels, const float* __restrict__ qrot,
                        const float* __restrict__ qconst, int metric, int k, int64_t chunk_rows,
                        uint64_t* __restrict__ part) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int q = blockIdx.x, chunk = blockIdx.y, nchunks = gridDim.y;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int L = 1 << B;
    float* qs = reinterpret_cast<float*>(smem_raw);   // [D]
    float* lv = qs + ((D + 3) & ~3);                  // [L]
    uint64_t* wbuf = reinterpret_cast<uint64_t*>(
        (reinterpret_cast<uintptr_t>(lv + L) + 15) & ~(uintptr_t)15);
    const int n2 = next_pow2(8 * k);
    uint64_t* mbuf = wbuf + 8 * CAP;
    for (int i = threadIdx.x; i < D; i += blockDim.x) qs[i] = qrot[(size_t)q * D + i];
    for (int i = threadIdx.x; i < L; i += blockDim.x) lv[i] = levels[i];
    const float qc = qconst[q];
    __syncthreads();
    const int ibytes = (D * B + 7) >> 3;
    const int code_size = ibytes + 8;
    const float inv_sqrt_d = 1.0f / sqrtf((float)D);

    WarpTopK<CAP> tk;
    tk.init(wbuf + w * CAP, k);
    const int64_t r0 = (int64_t)chunk * chunk_rows;
    int64_t r1 = r0 + chunk_rows;
    if (r1 > n) r1 = n;
    for (int64_t base = r0 + (int64_t)w * 32; base < r1; base += 256) {
        const int64_t row = base + lane;
        const bool valid = row < r1;
        float est = 0.f;
        if (valid) {
            const uint8_t* code = codes + row * (int64_t)code_size;
            float dot = 0.f, l2 = 0.f;
            for (int d = 0; d < D; ++d) {
                float lvl = lv[erabitq_index(code, d, B, ibytes)];
                dot = __fmaf_rn(lvl, qs[d], dot);
                l2 = __fmaf_rn(lvl, lvl, l2);
            }
            float nrm = load_f32_unaligned(code + ibytes);
            float t = load_f32_unaligned(code + ibytes + 4);
            // canonical arithmetic, shared with est_tc.cu (one segment): every operation rounded separately
            const float a = __fmul_rn(__fmul_rn(nrm, t), inv_sqrt_d);
            const float acc = __fmul_rn(a, dot);
            if (metric == B200VQ_METRIC_L2) est = __fmaf_rn(-2.f, acc, __fadd_rn(qc, __fmul_rn(__fmul_rn(a, a), l2)));
            else est = -__fadd_rn(qc, acc);
        }
        tk.offer(valid, est, (uint32_t)row);
    }
    tk.finish();
    for (int i = threadIdx.x; i < n2; i += blockDim.x) mbuf[i] = KEY_INF;
    __syncthreads();
    for (int i = lane; i < tk.count; i += 32) mbuf[w * k + i] = tk.buf[i];
    block_bitonic_sort(mbuf, n2);
    uint64_t* dst = part + ((size_t)q * nchunks + chunk) * k;
    for (int j = threadIdx.x; j < k; j += blockDim.x) dst[j] = mbuf[j];
}

// Chunk planning identical to scan.cu (kept local: 2048-row units).
struct RPlan {
    int64_t chunk_rows;
    int nchunks;
};
static RPlan rplan(int64_t n, int nq) {
    const int64_t unit = 2048;
    int64_t target = (int64_t)device_sm_count() * 8;
    int64_t want = (target + nq - 1) / (nq > 0 ? nq : 1);
    if (want < 1) want = 1;
    int64_t maxc = (n + unit - 1) / unit;
    if (maxc < 1) maxc = 1;
    int64_t minc = (n + 262143) / 262144;
    if (minc < 1) minc = 1;
    int64_t nch = want;
    if (nch > maxc) nch = maxc;
    if (nch < minc) nch = minc;
    if (nch > 65535) nch = 65535;
    int64_t rows = (n + nch - 1) / nch;
    rows = ((rows + unit - 1) / unit) * unit;
    if (rows < unit) rows = unit;
    RPlan p;
    p.chunk_rows = rows;
    p.nchunks = (int)((n + rows - 1) / rows);
    if (p.nchunks < 1) p.nchunks = 1;
    return p;
}


// Deferred fix-ups of erabitq_pack_tc_kernel, one CTA per coordinate d: row d of P^T and the centre sit in shared
// memory (fp64) and every warp walks its share of the bucket: exact fp64 <x_i - c, P[:, d]>, exact level, the code
// field is patched in place (atomic and / or on the containing 32-bit words: neighbouring fields belong to other
// CTAs) and the row's (num, den) sums move from the approximate to the exact contribution.
template <typename T>
__global__ void __launch_bounds__(256)
erabitq_fixup_kernel(const float* __restrict__ y, int64_t ldy, const T* __restrict__ x, int64_t ldx,
                     const double* __restrict__ c, const double* __restrict__ PT, const double* __restrict__ nrm, int D,
                     int B, const double* __restrict__ levels, const double* __restrict__ bounds, uint8_t* __restrict__ codes,
                     const unsigned* __restrict__ bcnt, const unsigned long long* __restrict__ bucket, unsigned bcap,
                     double* __restrict__ numden) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* pt = reinterpret_cast<double*>(smem_raw);     // [D]
    double* cs = pt + D;                                  // [D]
    double* lv = cs + D;                                  // [2^B]
    double* bd = lv + (1 << B);                           // [2^B]
    const int d = blockIdx.x, L = 1 << B;
    unsigned cnt = bcnt[d];
    if (cnt == 0) return;
    if (cnt > bcap) cnt = bcap;
    for (int k = threadIdx.x; k < D; k += blockDim.x) {
        pt[k] = PT[(int64_t)d * D + k];
        cs[k] = c[k];
    }
    for (int i = threadIdx.x; i < L; i += blockDim.x) lv[i] = levels[i];
    for (int i = threadIdx.x; i < L - 1; i += blockDim.x) bd[i] = bounds[i];
    __syncthreads();
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const int ibytes = (D * B + 7) >> 3, code_size = ibytes + 8;
    const double sqrt_d = sqrt((double)D);
    for (unsigned e = w; e < cnt; e += nw) {
        const unsigned long long ent = bucket[(size_t)d * bcap + e];
        const int64_t row = (int64_t)(ent >> 8);
        const int lo_old = (int)(ent & 255ull);
        const T* xr = x + row * ldx;
        double a0 = 0.0, a1 = 0.0;
        int k = lane;
        for (; k + 32 < D; k += 64) {
            a0 = fma((double)xr[k] - cs[k], pt[k], a0);
            a1 = fma((double)xr[k + 32] - cs[k + 32], pt[k + 32], a1);
        }
        if (k < D) a0 = fma((double)xr[k] - cs[k], pt[k], a0);
        a0 += a1;
        for (int o = 16; o > 0; o >>= 1) a0 += __shfl_xor_sync(FULL_MASK, a0, o);
        if (lane == 0) {
            const double nr = nrm[row], nsafe = fmax(nr, 1e-12);
            const double sv = (a0 / nsafe) * sqrt_d;
            const int lo = erabitq_level_of(bd, L, sv);
            const double so = (double)y[row * ldy + d] * (sqrt_d / nsafe);
            const double sho = lv[lo_old], sh = lv[lo];
            atomicAdd(numden + 2 * row, sv * sh - so * sho);
            atomicAdd(numden + 2 * row + 1, sh * sh - sho * sho);
            if (lo != lo_old) {
                // MSB-first B-bit field at bit d B of the row's code: at most two bytes
                uint8_t* code = codes + row * (int64_t)code_size;
                const int bit0 = d * B;
                for (int byte = bit0 >> 3; byte * 8 < bit0 + B; ++byte) {
                    const int shift = (byte * 8 + 8) - (bit0 + B);       // > 0: left shift inside this byte
                    const uint32_t mask = ((shift >= 0) ? (((1u << B) - 1u) << shift) : (((1u << B) - 1u) >> (-shift))) & 0xffu;
                    const uint32_t val = ((shift >= 0) ? ((uint32_t)lo << shift) : ((uint32_t)lo >> (-shift))) & 0xffu;
                    const uintptr_t addr = reinterpret_cast<uintptr_t>(code + byte);
                    unsigned* word = reinterpret_cast<unsigned*>(addr & ~(uintptr_t)3);
                    const int sh8 = 8 * (int)(addr & 3);
                    atomicAnd(word, ~(mask << sh8));
                    atomicOr(word, val << sh8);
                }
            }
        }
    }
}
// t = <s, l> / <l, l> of every row from its final sums
__global__ void erabitq_finalize_kernel(const double* __restrict__ numden, int64_t n, int code_size, int ibytes,
                                        uint8_t* __restrict__ codes) {
    const int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= n) return;
    const double num = numden[2 * row], den = numden[2 * row + 1];
    const uint32_t u1 = __float_as_uint((float)((den > 1e-12) ? (num / den) : 1.0));
    uint8_t* f = codes + row * (int64_t)code_size + ibytes + 4;
#pragma unroll
    for (int b = 0; b < 4; ++b) f[b] = (uint8_t)(u1 >> (8 * b));
}

}  // namespace b200vq

using namespace b200vq;

extern "C" int b200vq_rabitq1_encode(const float* x, int64_t n, int64_t ldx, int D,
                                     const float* centroid, int metric, uint8_t* codes,
                                     void* stream) {
    VQ_CHECK_ARG((x || n == 0) && (codes || n == 0) && n >= 0 && D > 0 && ldx >= D, "rabitq1_encode: bad args");
    VQ_CHECK_ARG(metric == B200VQ_METRIC_L2 || metric == B200VQ_METRIC_IP, "rabitq1_encode: bad metric");
    if (n == 0) return B200VQ_OK;
    int64_t g = (n * 32 + 255) / 256;
    int64_t cap = (int64_t)device_sm_count() * 16;
    if (g > cap) g = cap;
    const bool vec = D % 4 == 0 && ldx % 4 == 0 && ((uintptr_t)x & 15) == 0 && (!centroid || ((uintptr_t)centroid & 15) == 0);
    if (vec)
        rabitq1_encode_vec_kernel<<<(int)g, 256, 0, as_stream(stream)>>>(x, n, ldx, D, centroid, metric, codes);
    else
        rabitq1_encode_kernel<<<(int)g, 256, 0, as_stream(stream)>>>(x, n, ldx, D, centroid, metric, codes);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" int b200vq_rabitq1_decode(const uint8_t* codes, int64_t n, int D, const float* centroid,
                                     float* out, int64_t ldo, void* stream) {
    VQ_CHECK_ARG((codes || n == 0) && (out || n == 0) && n >= 0 && D > 0 && ldo >= D, "rabitq1_decode: bad args");
    if (n == 0) return B200VQ_OK;
    int64_t g = (n * 32 + 255) / 256;
    int64_t cap = (int64_t)device_sm_count() * 16;
    if (g > cap) g = cap;
    rabitq1_decode_kernel<<<(int)g, 256, 0, as_stream(stream)>>>(codes, n, D, centroid, out, ldo);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" int b200vq_rabitq1_prepare_queries(const float* queries, int nq, int64_t ldq, int D,
                                              const float* centroid, int qb, uint32_t* planes,
                                              float* qfac, void* stream) {
    VQ_CHECK_ARG(queries && planes && qfac && nq >= 0 && D > 0 && ldq >= D,
                 "rabitq1_prepare_queries: bad args");
    VQ_CHECK_ARG(qb >= 1 && qb <= 8, "rabitq1_prepare_queries: qb must be in [1,8], got %d", qb);
    VQ_CHECK_ARG((size_t)D * 8 <= 160 * 1024, "rabitq1_prepare_queries: D too large");
    if (nq == 0) return B200VQ_OK;
    size_t smem = (size_t)D * 8;
    VQ_CUDA(cudaFuncSetAttribute(rabitq1_prepare_queries_kernel,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    rabitq1_prepare_queries_kernel<<<nq, 256, smem, as_stream(stream)>>>(queries, ldq, D, centroid,
                                                                         qb, planes, qfac);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" size_t b200vq_rabitq1_scan_scratch_bytes(int64_t n, int nq, int k) {
    if (n < 0 || nq <= 0 || k <= 0) return 0;
    RPlan p = rplan(n, nq);
    return (size_t)nq * p.nchunks * k * sizeof(uint64_t);
}

template <int CAP>
static int rabitq1_scan_launch(const uint8_t* codes, int64_t n, int D, const uint32_t* planes,
                               const float* qfac, int nq, int qb, int metric, int k, uint64_t* part,
                               const RPlan& p, cudaStream_t st, const uint32_t* gate = nullptr,
                               const float* after_vals = nullptr, const uint32_t* after_ids = nullptr) {
    const int W = (D + 31) >> 5;
    const int code_size = ((D + 7) >> 3) + 8;
    const int pitch_w = ((code_size + 3) >> 2) | 1;
    const int n2 = next_pow2(8 * k);
    size_t smem = (size_t)((qb * W + 3) & ~3) * 4 + (size_t)8 * 32 * pitch_w * 4 + 16 +
                  (size_t)(8 * CAP + n2) * sizeof(uint64_t);
    VQ_CHECK_ARG(smem <= 227 * 1024, "rabitq1_scan: D=%d needs %zu B of shared memory", D, smem);
    VQ_CUDA(cudaFuncSetAttribute(rabitq1_scan_kernel<CAP>,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid((unsigned)nq, (unsigned)p.nchunks);
    rabitq1_scan_kernel<CAP><<<grid, 256, smem, st>>>(codes, n, D, planes, qfac, qb, metric, k,
                                                      p.chunk_rows, part, gate, after_vals, after_ids);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

// SIMT scan restricted to the queries with gate[q] != 0 (scratch sized by b200vq_rabitq1_scan_scratch_bytes).
int b200vq::rabitq1_scan_gated(const uint8_t* codes, int64_t n, int D, const uint32_t* planes, const float* qfac,
                               int nq, int qb, int metric, int k, float* out_vals, uint32_t* out_ids, void* scratch,
                               const uint32_t* gate, cudaStream_t st) {
    RPlan p = rplan(n, nq);
    uint64_t* part = reinterpret_cast<uint64_t*>(scratch);
    int rc = (k <= 32)
                 ? rabitq1_scan_launch<128>(codes, n, D, planes, qfac, nq, qb, metric, k, part, p, st, gate)
                 : rabitq1_scan_launch<256>(codes, n, D, planes, qfac, nq, qb, metric, k, part, p, st, gate);
    if (rc != B200VQ_OK) return rc;
    merge_partials_kernel<2048><<<nq, 256, 0, st>>>(part, p.nchunks, k, metric == B200VQ_METRIC_IP, out_vals, out_ids,
                                                    gate, 0u);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" int b200vq_rabitq1_scan_topk(const uint8_t* codes, int64_t n, int D,
                                        const uint32_t* planes, const float* qfac, int nq, int qb,
                                        int metric, int k, float* out_vals, uint32_t* out_ids,
                                        void* scratch, size_t scratch_bytes, void* stream) {
    VQ_CHECK_ARG(codes && planes && qfac && out_vals && out_ids, "rabitq1_scan_topk: null pointer");
    VQ_CHECK_ARG(n >= 0 && n < (1ll << 32) && D > 0 && nq >= 0, "rabitq1_scan_topk: bad shape");
    VQ_CHECK_ARG(qb >= 1 && qb <= 8, "rabitq1_scan_topk: qb must be in [1,8]");
    VQ_CHECK_ARG(k > 0 && k <= 128, "rabitq1_scan_topk: k must be in [1,128], got %d", k);
    VQ_CHECK_ARG(metric == B200VQ_METRIC_L2 || metric == B200VQ_METRIC_IP, "rabitq1_scan_topk: bad metric");
    if (nq == 0) return B200VQ_OK;
    RPlan p = rplan(n, nq);
    size_t need = (size_t)nq * p.nchunks * k * sizeof(uint64_t);
    VQ_CHECK_ARG(scratch && scratch_bytes >= need, "rabitq1_scan_topk: scratch too small");
    cudaStream_t st = as_stream(stream);
    uint64_t* part = reinterpret_cast<uint64_t*>(scratch);
    int rc = (k <= 32)
                 ? rabitq1_scan_launch<128>(codes, n, D, planes, qfac, nq, qb, metric, k, part, p, st)
                 : rabitq1_scan_launch<256>(codes, n, D, planes, qfac, nq, qb, metric, k, part, p, st);
    if (rc != B200VQ_OK) return rc;
    merge_partials_kernel<2048><<<nq, 256, 0, st>>>(part, p.nchunks, k, metric == B200VQ_METRIC_IP,
                                                    out_vals, out_ids);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

// Continuation: the best k rows AFTER (after_vals[q], after_ids[q]) in (estimate, id) order, so that any k is served
// in pages of at most 128 (RaBitQIndex.search with k > 128). Exact SIMT scan; scratch as b200vq_rabitq1_scan_topk.
extern "C" int b200vq_rabitq1_scan_topk_after(const uint8_t* codes, int64_t n, int D, const uint32_t* planes,
                                              const float* qfac, int nq, int qb, int metric, int k,
                                              const float* after_vals, const uint32_t* after_ids, float* out_vals,
                                              uint32_t* out_ids, void* scratch, size_t scratch_bytes, void* stream) {
    VQ_CHECK_ARG(codes && planes && qfac && out_vals && out_ids && after_vals && after_ids,
                 "rabitq1_scan_topk_after: null pointer");
    VQ_CHECK_ARG(n >= 0 && n < (1ll << 32) && D > 0 && nq >= 0, "rabitq1_scan_topk_after: bad shape");
    VQ_CHECK_ARG(qb >= 1 && qb <= 8, "rabitq1_scan_topk_after: qb must be in [1,8]");
    VQ_CHECK_ARG(k > 0 && k <= 128, "rabitq1_scan_topk_after: k must be in [1,128], got %d", k);
    VQ_CHECK_ARG(metric == B200VQ_METRIC_L2 || metric == B200VQ_METRIC_IP, "rabitq1_scan_topk_after: bad metric");
    if (nq == 0) return B200VQ_OK;
    RPlan p = rplan(n, nq);
    size_t need = (size_t)nq * p.nchunks * k * sizeof(uint64_t);
    VQ_CHECK_ARG(scratch && scratch_bytes >= need, "rabitq1_scan_topk_after: scratch too small");
    cudaStream_t st = as_stream(stream);
    uint64_t* part = reinterpret_cast<uint64_t*>(scratch);
    int rc = (k <= 32) ? rabitq1_scan_launch<128>(codes, n, D, planes, qfac, nq, qb, metric, k, part, p, st, nullptr,
                                                  after_vals, after_ids)
                       : rabitq1_scan_launch<256>(codes, n, D, planes, qfac, nq, qb, metric, k, part, p, st, nullptr,
                                                  after_vals, after_ids);
    if (rc != B200VQ_OK) return rc;
    merge_partials_kernel<2048><<<nq, 256, 0, st>>>(part, p.nchunks, k, metric == B200VQ_METRIC_IP, out_vals, out_ids);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" int b200vq_erabitq_pack(const double* srot, const double* nrm, int64_t n, int D, int B,
                                   const double* levels, const double* bounds, uint8_t* codes,
                                   void* stream) {
    VQ_CHECK_ARG((srot || n == 0) && (nrm || n == 0) && levels && bounds && (codes || n == 0) && n >= 0 && D > 0,
                 "erabitq_pack: bad args");
    VQ_CHECK_ARG(B >= 1 && B <= 8, "erabitq_pack: B must be in [1,8], got %d", B);
    if (n == 0) return B200VQ_OK;
    size_t smem = (size_t)(2 << B) * sizeof(double) + (size_t)8 * ((D + 15) & ~15) + 16;   // 8 warps' index buffers
    VQ_CHECK_ARG(smem <= 200 * 1024, "erabitq_pack: D too large");
    VQ_CUDA(cudaFuncSetAttribute(erabitq_pack_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)smem));
    int64_t g = (n + 7) / 8;
    int64_t cap = (int64_t)device_sm_count() * 8;
    if (g > cap) g = cap;
    erabitq_pack_kernel<<<(int)g, 256, smem, as_stream(stream)>>>(srot, nrm, n, D, B, levels, bounds,
                                                                 codes);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" int b200vq_resid_norms_f64(const void* x, int x_is_f64, int64_t n, int64_t ldx, int D, const double* c,
                                      double* nrm, void* stream) {
    VQ_CHECK_ARG((x || n == 0) && c && (nrm || n == 0) && n >= 0 && D > 0 && ldx >= D, "resid_norms_f64: bad args");
    if (n == 0) return B200VQ_OK;
    int64_t g = (n * 32 + 255) / 256;
    const int64_t cap = (int64_t)device_sm_count() * 16;
    if (g > cap) g = cap;
    if (x_is_f64)
        resid_norm_kernel<double><<<(int)g, 256, 0, as_stream(stream)>>>(reinterpret_cast<const double*>(x), n, ldx, D, c, nrm);
    else
        resid_norm_kernel<float><<<(int)g, 256, 0, as_stream(stream)>>>(reinterpret_cast<const float*>(x), n, ldx, D, c, nrm);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

namespace {
struct PackPlan {
    unsigned bcap;
    size_t off_cnt, off_bucket, off_numden, total;
};
PackPlan pack_plan(int64_t n, int D) {
    PackPlan p;
    // a coordinate is flagged with probability ~1.3e-3 (B = 4, D = 1536): four times the expectation, at least 256
    int64_t cap = (n / 192 + 255) & ~(int64_t)255;
    if (cap < 256) cap = 256;
    p.bcap = (unsigned)cap;
    auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
    size_t o = 0;
    p.off_cnt = o;     o = up(o + (size_t)D * 4);
    p.off_bucket = o;  o = up(o + (size_t)D * cap * 8);
    p.off_numden = o;  o = up(o + (size_t)n * 16);
    p.total = o;
    return p;
}
int pack_tc_impl(const float* y, int64_t ldy, const void* x, int x_is_f64, int64_t ldx, const double* c, const double* PT,
                 const double* nrm, int64_t n, int D, int B, const double* levels, const double* bounds, double err_rel,
                 double err_abs, uint8_t* codes, unsigned long long* stats, void* scratch, size_t scratch_bytes,
                 void* stream) {
    VQ_CHECK_ARG((x || n == 0) && c && PT && (nrm || n == 0) && levels && bounds && (codes || n == 0) && n >= 0 &&
                     D > 0 && D < 65536 && ldx >= D && (!y || ldy >= D),
                 "erabitq_pack_tc: bad args");
    VQ_CHECK_ARG(B >= 1 && B <= 8, "erabitq_pack_tc: B must be in [1,8], got %d", B);
    VQ_CHECK_ARG(err_rel >= 0.0 && err_abs >= 0.0, "erabitq_pack_tc: negative error bound");
    if (n == 0) return B200VQ_OK;
    VQ_CHECK_ARG(n < (1ll << 40), "erabitq_pack_tc: too many rows");
    const size_t smem = (size_t)(2 << B) * sizeof(double) + (size_t)8 * 3 * ((D + 15) & ~15) + 16;
    VQ_CHECK_ARG(smem <= 200 * 1024, "erabitq_pack_tc: D too large");
    cudaStream_t st = as_stream(stream);
    // deferred fix-ups need the approximate coordinates (y) and a scratch area; otherwise everything happens in-kernel
    unsigned* bcnt = nullptr;
    unsigned long long* bucket = nullptr;
    double* numden = nullptr;
    PackPlan pp{};
    const size_t fsmem = (size_t)(2 * D + (2 << B)) * sizeof(double);
    if (y && scratch && fsmem <= 200 * 1024) {
        pp = pack_plan(n, D);
        VQ_CHECK_ARG(scratch_bytes >= pp.total && ((uintptr_t)scratch & 255) == 0,
                     "erabitq_pack_tc2: scratch too small or not 256-byte aligned (%zu < %zu)", scratch_bytes, pp.total);
        uint8_t* sb = reinterpret_cast<uint8_t*>(scratch);
        bcnt = reinterpret_cast<unsigned*>(sb + pp.off_cnt);
        bucket = reinterpret_cast<unsigned long long*>(sb + pp.off_bucket);
        numden = reinterpret_cast<double*>(sb + pp.off_numden);
        VQ_CUDA(cudaMemsetAsync(bcnt, 0, (size_t)D * 4, st));
    }
    int64_t g = (n + 7) / 8;
    const int64_t cap = (int64_t)device_sm_count() * 8;
    if (g > cap) g = cap;
#define B200VQ_PACK_LAUNCH(TY, BTV)                                                                                  \
    do {                                                                                                            \
        VQ_CUDA(cudaFuncSetAttribute(erabitq_pack_tc_kernel<TY, BTV>, cudaFuncAttributeMaxDynamicSharedMemorySize,   \
                                     (int)smem));                                                                   \
        erabitq_pack_tc_kernel<TY, BTV><<<(int)g, 256, smem, st>>>(                                                  \
            y, ldy, reinterpret_cast<const TY*>(x), ldx, c, PT, nrm, n, D, B, levels, bounds, err_rel, err_abs,      \
            codes, stats, bcnt, bucket, pp.bcap, numden);                                                           \
    } while (0)
    if (x_is_f64) {
        if (B == 4) B200VQ_PACK_LAUNCH(double, 4);
        else if (B == 8) B200VQ_PACK_LAUNCH(double, 8);
        else B200VQ_PACK_LAUNCH(double, 0);
    } else {
        if (B == 4) B200VQ_PACK_LAUNCH(float, 4);
        else if (B == 8) B200VQ_PACK_LAUNCH(float, 8);
        else B200VQ_PACK_LAUNCH(float, 0);
    }
#undef B200VQ_PACK_LAUNCH
    VQ_LAUNCH_CHECK();
    if (bcnt) {
        const int ibytes = (D * B + 7) >> 3;
        if (x_is_f64) {
            VQ_CUDA(cudaFuncSetAttribute(erabitq_fixup_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fsmem));
            erabitq_fixup_kernel<double><<<D, 256, fsmem, st>>>(y, ldy, reinterpret_cast<const double*>(x), ldx, c, PT, nrm, D,
                                                                B, levels, bounds, codes, bcnt, bucket, pp.bcap, numden);
        } else {
            VQ_CUDA(cudaFuncSetAttribute(erabitq_fixup_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fsmem));
            erabitq_fixup_kernel<float><<<D, 256, fsmem, st>>>(y, ldy, reinterpret_cast<const float*>(x), ldx, c, PT, nrm, D, B,
                                                               levels, bounds, codes, bcnt, bucket, pp.bcap, numden);
        }
        VQ_LAUNCH_CHECK();
        erabitq_finalize_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(numden, n, ibytes + 8, ibytes, codes);
        VQ_LAUNCH_CHECK();
    }
    return B200VQ_OK;
}
}  // namespace

extern "C" int b200vq_erabitq_pack_tc(const float* y, int64_t ldy, const void* x, int x_is_f64, int64_t ldx,
                                      const double* c, const double* PT, const double* nrm, int64_t n, int D, int B,
                                      const double* levels, const double* bounds, double err_rel, double err_abs,
                                      uint8_t* codes, unsigned long long* stats, void* stream) {
    return pack_tc_impl(y, ldy, x, x_is_f64, ldx, c, PT, nrm, n, D, B, levels, bounds, err_rel, err_abs, codes, stats,
                        nullptr, 0, stream);
}

// The same encoder with the fix-ups deferred to a second kernel that visits them coordinate by coordinate (a row of
// P^T is read once for all the rows that need it instead of once per flagged coordinate): same codes, same trailers.
extern "C" size_t b200vq_erabitq_pack_tc2_scratch_bytes(int64_t n, int D) {
    if (n <= 0 || D <= 0) return 0;
    return pack_plan(n, D).total;
}
extern "C" int b200vq_erabitq_pack_tc2(const float* y, int64_t ldy, const void* x, int x_is_f64, int64_t ldx,
                                       const double* c, const double* PT, const double* nrm, int64_t n, int D, int B,
                                       const double* levels, const double* bounds, double err_rel, double err_abs,
                                       uint8_t* codes, unsigned long long* stats, void* scratch, size_t scratch_bytes,
                                       void* stream) {
    VQ_CHECK_ARG(y && scratch, "erabitq_pack_tc2: needs the approximate coordinates and a scratch area");
    return pack_tc_impl(y, ldy, x, x_is_f64, ldx, c, PT, nrm, n, D, B, levels, bounds, err_rel, err_abs, codes, stats,
                        scratch, scratch_bytes, stream);
}

extern "C" int b200vq_erabitq_rowscale(const uint8_t* codes, int64_t n, int D, int B, float* a, void* stream) {
    VQ_CHECK_ARG((codes || n == 0) && (a || n == 0) && n >= 0 && D > 0 && B >= 1 && B <= 8, "erabitq_rowscale: bad args");
    if (n == 0) return B200VQ_OK;
    int64_t g = (n + 255) / 256;
    const int64_t cap = (int64_t)device_sm_count() * 16;
    if (g > cap) g = cap;
    erabitq_rowscale_kernel<<<(int)g, 256, 0, as_stream(stream)>>>(codes, n, D, B, a);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" int b200vq_erabitq_unpack(const uint8_t* codes, int64_t n, int D, int B,
                                     const double* levels, double* shat, double* nrm_out,
                                     void* stream) {
    VQ_CHECK_ARG((codes || n == 0) && levels && (shat || n == 0) && (nrm_out || n == 0) && n >= 0 && D > 0, "erabitq_unpack: bad args");
    VQ_CHECK_ARG(B >= 1 && B <= 8, "erabitq_unpack: B must be in [1,8], got %d", B);
    if (n == 0) return B200VQ_OK;
    int64_t g = (n * 32 + 255) / 256;
    int64_t cap = (int64_t)device_sm_count() * 16;
    if (g > cap) g = cap;
    erabitq_unpack_kernel<<<(int)g, 256, 0, as_stream(stream)>>>(codes, n, D, B, levels, shat,
                                                                 nrm_out);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" int b200vq_erabitq_decode_f32(const uint8_t* codes, int64_t n, int D, int B, const float* levels,
                                         float* x, int64_t ldx, float* xnorm, void* stream) {
    VQ_CHECK_ARG(n >= 0 && D > 0 && B >= 1 && B <= 8 && ldx >= D, "erabitq_decode_f32: bad shape");
    if (n == 0) return B200VQ_OK;
    VQ_CHECK_ARG(codes && levels && x && xnorm, "erabitq_decode_f32: null pointer");
    int64_t g = (n * 32 + 255) / 256;
    const int64_t cap = (int64_t)device_sm_count() * 16;
    if (g > cap) g = cap;
    erabitq_decode_f32_kernel<<<(int)g, 256, 0, as_stream(stream)>>>(codes, n, D, B, levels, x, ldx, xnorm);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" size_t b200vq_erabitq_scan_scratch_bytes(int64_t n, int nq, int k) {
    if (n < 0 || nq <= 0 || k <= 0) return 0;
    RPlan p = rplan(n, nq);
    return (size_t)nq * p.nchunks * k * sizeof(uint64_t);
}

template <int CAP>
static int erabitq_scan_launch(const uint8_t* codes, int64_t n, int D, int B, const float* levels,
                               const float* qrot, const float* qconst, int nq, int metric, int k,
                               uint64_t* part, const RPlan& p, cudaStream_t st) {
    const int n2 = next_pow2(8 * k);
    size_t smem = (size_t)(((D + 3) & ~3) + (1 << B)) * 4 + 16 + (size_t)(8 * CAP + n2) * 8;
    VQ_CHECK_ARG(smem <= 227 * 1024, "erabitq_scan: D too large");
    VQ_CUDA(cudaFuncSetAttribute(erabitq_scan_kernel<CAP>,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid((unsigned)nq, (unsigned)p.nchunks);
    erabitq_scan_kernel<CAP><<<grid, 256, smem, st>>>(codes, n, D, B, levels, qrot, qconst, metric,
                                                      k, p.chunk_rows, part);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" int b200vq_erabitq_scan_topk(const uint8_t* codes, int64_t n, int D, int B,
                                        const float* levels, const float* qrot, const float* qconst,
                                        int nq, int metric, int k, float* out_vals,
                                        uint32_t* out_ids, void* scratch, size_t scratch_bytes,
                                        void* stream) {
    VQ_CHECK_ARG(codes && levels && qrot && qconst && out_vals && out_ids,
                 "erabitq_scan_topk: null pointer");
    VQ_CHECK_ARG(n >= 0 && n < (1ll << 32) && D > 0 && nq >= 0 && B >= 1 && B <= 8,
                 "erabitq_scan_topk: bad shape");
    VQ_CHECK_ARG(k > 0 && k <= 128, "erabitq_scan_topk: k must be in [1,128], got %d", k);
    VQ_CHECK_ARG(metric == B200VQ_METRIC_L2 || metric == B200VQ_METRIC_IP, "erabitq_scan_topk: bad metric");
    if (nq == 0) return B200VQ_OK;
    RPlan p = rplan(n, nq);
    size_t need = (size_t)nq * p.nchunks * k * sizeof(uint64_t);
    VQ_CHECK_ARG(scratch && scratch_bytes >= need, "erabitq_scan_topk: scratch too small");
    cudaStream_t st = as_stream(stream);
    uint64_t* part = reinterpret_cast<uint64_t*>(scratch);
    int rc = (k <= 32) ? erabitq_scan_launch<128>(codes, n, D, B, levels, qrot, qconst, nq, metric,
                                                  k, part, p, st)
                       : erabitq_scan_launch<256>(codes, n, D, B, levels, qrot, qconst, nq, metric,
                                                  k, part, p, st);
    if (rc != B200VQ_OK) return rc;
    merge_partials_kernel<2048><<<nq, 256, 0, st>>>(part, p.nchunks, k, metric == B200VQ_METRIC_IP,
                                                    out_vals, out_ids);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}
