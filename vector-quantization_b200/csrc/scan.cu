// Scan kernels: ADC code scan with per-query shared-memory LUT and warp-level
// running top-k; exact flat search (tiled fp32 product + fused top-k); row
// norms; partial-list merges.
//
// Reference behaviour reproduced:
//   - FlatQuantizedIndex.search_with_scores (src/haag_vq/methods/search/
//     flat_quantized_index.py:45-76): distance to the reconstruction, top-k by
//     ascending squared L2 / descending inner product. The ADC sum over
//     sub-quantizer LUT entries equals the distance to the reconstruction.
//   - bench/ffd_speed.cpp:87-127: score = sum_d LUT[d][code_d].
//   - faiss IndexFlatL2/IP.search (benchmarks/exact_search.py:72-77).
#include <mutex>

#include "common.cuh"
#include "topk.cuh"

namespace b200vq {

constexpr int SCAN_THREADS = 256;
constexpr int SCAN_WARPS = SCAN_THREADS / 32;

// ------------------------------------------------------------ ADC scan ---
// grid (nq, nchunks). Each CTA loads one query's LUT (M*ksub fp32) into shared
// memory and streams `chunk_rows` rows of codes. Thread = row.
template <int CAP, bool LUT_SMEM, int VEC>
__global__ void __launch_bounds__(SCAN_THREADS)
    adc_scan_kernel(const uint8_t* __restrict__ codes, int64_t n, int M, int ksub,
                    const float* __restrict__ lut, int k, int64_t chunk_rows, int negate,
                    uint64_t* __restrict__ part) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int q = blockIdx.x, chunk = blockIdx.y, nchunks = gridDim.y;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const size_t lut_elems = LUT_SMEM ? (size_t)M * ksub : 0;
    float* lut_s = reinterpret_cast<float*>(smem_raw);
    uint64_t* wbuf = reinterpret_cast<uint64_t*>(smem_raw + ((lut_elems * 4 + 15) & ~(size_t)15));
    const int n2 = next_pow2(SCAN_WARPS * k);
    uint64_t* mbuf = wbuf + SCAN_WARPS * CAP;

    const float* lut_q = lut + (size_t)q * M * ksub;
    if (LUT_SMEM) {
        for (size_t i = threadIdx.x; i < lut_elems; i += blockDim.x) lut_s[i] = lut_q[i];
        __syncthreads();
    }
    const float* L = LUT_SMEM ? lut_s : lut_q;

    WarpTopK<CAP> tk;
    tk.init(wbuf + w * CAP, k);

    const int64_t r0 = (int64_t)chunk * chunk_rows;
    int64_t r1 = r0 + chunk_rows;
    if (r1 > n) r1 = n;
    for (int64_t base = r0 + (int64_t)w * 32; base < r1; base += SCAN_THREADS) {
        const int64_t row = base + lane;
        const bool valid = row < r1;
        float acc = 0.f;
        if (valid) {
            const uint8_t* cr = codes + row * M;
            if (VEC == 16) {
                const uint4* c4 = reinterpret_cast<const uint4*>(cr);
                for (int g = 0; g < (M >> 4); ++g) {
                    uint4 v = ld_stream_u4(c4 + g);
                    const float* Lg = L + (size_t)(g << 4) * ksub;
                    uint32_t wds[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
#pragma unroll
                        for (int b = 0; b < 4; ++b) {
                            uint32_t code = (wds[j] >> (8 * b)) & 0xffu;
                            acc += Lg[(size_t)(4 * j + b) * ksub + code];
                        }
                    }
                }
            } else if (VEC == 4) {
                const uint32_t* c1 = reinterpret_cast<const uint32_t*>(cr);
                for (int g = 0; g < (M >> 2); ++g) {
                    uint32_t v = __ldg(c1 + g);
                    const float* Lg = L + (size_t)(g << 2) * ksub;
#pragma unroll
                    for (int b = 0; b < 4; ++b) acc += Lg[(size_t)b * ksub + ((v >> (8 * b)) & 0xffu)];
                }
            } else {
                for (int m = 0; m < M; ++m) acc += L[(size_t)m * ksub + cr[m]];
            }
        }
        tk.offer(valid, negate ? -acc : acc, (uint32_t)row);
    }
    block_emit_topk<CAP>(tk, mbuf, n2, k, part + ((size_t)q * nchunks + chunk) * k);
}

// ------------------------------------------------- conflict-free layout ---
// Scan layout: rows are grouped in blocks of 32; sub-quantizers in groups of
// 32. For block b, group g, lane i, step s: byte codes[b*32+i][g*32 + (i+s)%32]
// lives at (b*G + g)*1024 + (s/16)*512 + i*16 + s%16. In the scan, lane i reads
// LUT column (i+s) at step s, so the 32 lanes of a warp always hit 32 distinct
// banks, and each of its two 128-bit code loads is a contiguous 512-byte request.
__global__ void codes_to_scan_layout_kernel(const uint8_t* __restrict__ codes, int64_t n, int M, int GW,
                                            uint8_t* __restrict__ out) {
    // GW = group width: 32, or 16 for the half group of M = 16 (one 16-byte run per lane, 512-byte blocks;
    // step s of lane i holds sub-quantizer (i + s) % 16)
    const int G = M / GW;
    const int64_t nblk = (n + 31) >> 5;
    const int64_t total = nblk * G * 32 * GW;
    const int sh = GW == 32 ? 5 : 4;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total;
         e += (int64_t)gridDim.x * blockDim.x) {
        int s = (int)(e & (GW - 1));
        int i = (int)((e >> sh) & 31);
        int64_t bg = e >> (sh + 5);
        int g = (int)(bg % G);
        int64_t b = bg / G;
        int64_t row = b * 32 + i;
        uint8_t v = 0;
        if (row < n) v = codes[row * M + g * GW + ((i + s) & (GW - 1))];
        // the 32 bytes of a lane are stored as two 16-byte halves, each half contiguous across the
        // 32 lanes (512 B), so that one 128-bit load per lane is one fully used 512-byte request
        out[(bg << (sh + 5)) + ((s >> 4) << 9) + (i << 4) + (s & 15)] = v;
    }
}

// LUT in shared memory as [g][k][64] floats: entry (k, j) = lut[g*32 + (j&31)][k],
// i.e. every 32-column row is stored twice so that lane i can address column
// i+s with the step s as a static immediate offset.
template <int CAP, int G, int GW = 32>
__global__ void __launch_bounds__(SCAN_THREADS)
    adc_scan_rot_kernel(const uint8_t* __restrict__ codes_rot, int64_t n,
                        const float* __restrict__ lut, int k, int64_t chunk_rows, int negate,
                        uint64_t* __restrict__ part, const uint32_t* __restrict__ gate = nullptr,
                        uint32_t gate_cap = 0) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int KSUB = 256;
    constexpr int M = G * GW;
    constexpr int NQW = GW / 4;          // column quads (= code words) of a group
    constexpr int NV = GW / 16;          // 128-bit code loads of a group
    static_assert(GW == 32 || (GW == 16 && G == 1), "group width 32, or one half group of 16");
    const int q = blockIdx.x, chunk = blockIdx.y, nchunks = gridDim.y;
    // gated launch (fallback of the screened scan): only queries whose candidate list overflowed
    if (gate && gate[q] <= gate_cap) return;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    float* lut_s = reinterpret_cast<float*>(smem_raw);
    constexpr size_t lut_bytes = (size_t)G * KSUB * 64 * 4;
    uint64_t* wbuf = reinterpret_cast<uint64_t*>(smem_raw + lut_bytes);
    const int n2 = next_pow2(SCAN_WARPS * k);
    uint64_t* mbuf = wbuf + SCAN_WARPS * CAP;

    // LUT fill: a thread gathers entry kk of four consecutive sub-quantizers and stores them as one
    // float4 (again at columns j + GW, ... up to column 31 + GW); a quarter warp writes 128 contiguous bytes.
    const float* lut_q = lut + (size_t)q * M * KSUB;
    for (int idx = threadIdx.x; idx < G * KSUB * NQW; idx += SCAN_THREADS) {
        const int jq = idx % NQW, kk = (idx / NQW) & (KSUB - 1), g = idx / (NQW * KSUB);
        const float* src = lut_q + (size_t)(g * GW + 4 * jq) * KSUB + kk;
        float4 v = make_float4(src[0], src[KSUB], src[2 * KSUB], src[3 * KSUB]);
        float4* dst = reinterpret_cast<float4*>(lut_s + ((size_t)g * KSUB + kk) * 64 + 4 * jq);
#pragma unroll
        for (int rep = 0; rep * GW < 32 + GW; ++rep) dst[rep * NQW] = v;
    }
    __syncthreads();

    WarpTopK<CAP> tk;
    tk.init(wbuf + w * CAP, k);

    const unsigned char* lbase = smem_raw;
    const uint32_t lane4 = (uint32_t)lane << 2;
    const int64_t r0 = (int64_t)chunk * chunk_rows;  // multiple of 32
    int64_t r1 = r0 + chunk_rows;
    if (r1 > n) r1 = n;
    // two 32-row blocks per iteration: both blocks' code loads are issued before the first lookup
    constexpr int UNR = 2;
    for (int64_t base = r0 + (int64_t)w * 32; base < r1; base += UNR * SCAN_THREADS) {
        uint4 cv[UNR][NV * G];
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
            // blocks past the chunk end are clamped to a valid block and masked out below
            int64_t bb = base + (int64_t)u * SCAN_THREADS;
            if (bb >= r1) bb = base;
            const uint4* cb = reinterpret_cast<const uint4*>(codes_rot + (bb >> 5) * (int64_t)G * (32 * GW));
#pragma unroll
            for (int g = 0; g < G; ++g)
#pragma unroll
                for (int h = 0; h < NV; ++h) cv[u][NV * g + h] = ld_stream_u4(cb + g * 64 + h * 32 + lane);
        }
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
            const int64_t row = base + (int64_t)u * SCAN_THREADS + lane;
            float acc = 0.f;
#pragma unroll
            for (int g = 0; g < G; ++g) {
                uint32_t wds[NQW];
#pragma unroll
                for (int h = 0; h < NV; ++h) {
                    wds[4 * h + 0] = cv[u][NV * g + h].x; wds[4 * h + 1] = cv[u][NV * g + h].y;
                    wds[4 * h + 2] = cv[u][NV * g + h].z; wds[4 * h + 3] = cv[u][NV * g + h].w;
                }
#pragma unroll
                for (int j = 0; j < NQW; ++j) {
#pragma unroll
                    for (int b = 0; b < 4; ++b) {
                        // address = (code << 8) | (lane << 2), built with one PRMT; the group and step
                        // offsets are immediates of the LDS
                        const uint32_t off = __byte_perm(wds[j], lane4, 0x5504u | ((uint32_t)b << 4));
                        acc += *reinterpret_cast<const float*>(lbase + off + g * (KSUB * 64 * 4) + (4 * j + b) * 4);
                    }
                }
            }
            tk.offer(row < r1, negate ? -acc : acc, (uint32_t)row);
        }
    }
    block_emit_topk<CAP>(tk, mbuf, n2, k, part + ((size_t)q * nchunks + chunk) * k);
}

}  // namespace b200vq
#include "scan_screen.cuh"
namespace b200vq {

// ------------------------------------------------------------ row norms ---
// One warp per row, lanes stride the row; fp32 partial chains + shuffle tree.
__global__ void row_sqnorms_kernel(const float* __restrict__ x, int64_t n, int64_t ldx, int D,
                                   float* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t wid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t row = wid; row < n; row += nw) {
        const float* p = x + row * ldx;
        float acc = 0.f;
        for (int i = lane; i < D; i += 32) acc = __fmaf_rn(p[i], p[i], acc);
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(FULL_MASK, acc, o);
        if (lane == 0) out[row] = acc;
    }
}

// ---------------------------------------------------------- exact flat ---
constexpr int FT_Q = 64;    // queries per CTA tile
constexpr int FT_N = 128;   // rows per CTA tile
constexpr int FT_K = 16;    // reduction step
constexpr int FT_THREADS = 256;

template <int CAP>
__global__ void __launch_bounds__(FT_THREADS)
    flat_topk_kernel(const float* __restrict__ x, int64_t n, int64_t ldx, int D,
                     const float* __restrict__ xnorm, const float* __restrict__ row_scale,
                     const float* __restrict__ queries, int nq, int64_t ldq, int metric, int k,
                     int64_t chunk_rows, uint64_t* __restrict__ part, const uint32_t* __restrict__ gate = nullptr,
                     const float* __restrict__ after_vals = nullptr, const uint32_t* __restrict__ after_ids = nullptr) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    if (gate) {   // gated launch (fallback of the tensor-core search): skip tiles without a flagged query
        int any = 0;
        const int qq = blockIdx.x * FT_Q + (threadIdx.x & (FT_Q - 1));
        if (qq < nq && gate[qq] != 0) any = 1;
        if (!__syncthreads_or(any)) return;
    }
    float* Qs = reinterpret_cast<float*>(smem_raw);            // [FT_K][FT_Q+4]
    float* Xs = Qs + FT_K * (FT_Q + 4);                        // [FT_K][FT_N+4]
    float* ds = Xs + FT_K * (FT_N + 4);                        // [FT_Q][FT_N+1]
    float* qn = ds + FT_Q * (FT_N + 1);                        // [FT_Q]
    float* s_thr = qn + FT_Q;                                  // [FT_Q]
    int* s_cnt = reinterpret_cast<int*>(s_thr + FT_Q);         // [FT_Q]
    uint64_t* lists = reinterpret_cast<uint64_t*>(
        (reinterpret_cast<uintptr_t>(s_cnt + FT_Q) + 15) & ~(uintptr_t)15);  // [FT_Q][CAP]
    uint64_t* s_after = lists + (size_t)FT_Q * CAP;            // [FT_Q] continuation floor (see b200vq_flat_topk_after)

    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int ty = tid >> 4, tx = tid & 15;
    const int q0 = blockIdx.x * FT_Q;
    if (after_vals && tid < FT_Q && q0 + tid < nq) {
        const float av = after_vals[q0 + tid];
        s_after[tid] = make_key(metric == B200VQ_METRIC_IP ? -av : av, after_ids[q0 + tid]);
    }
    const int chunk = blockIdx.y, nchunks = gridDim.y;
    const int64_t r0 = (int64_t)chunk * chunk_rows;
    int64_t r1 = r0 + chunk_rows;
    if (r1 > n) r1 = n;

    // query norms (L2 only) and list state
    for (int ql = w; ql < FT_Q; ql += FT_THREADS / 32) {
        float acc = 0.f;
        if (q0 + ql < nq) {
            const float* p = queries + (int64_t)(q0 + ql) * ldq;
            for (int i = lane; i < D; i += 32) acc = __fmaf_rn(p[i], p[i], acc);
        }
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(FULL_MASK, acc, o);
        if (lane == 0) {
            qn[ql] = acc;
            s_thr[ql] = __int_as_float(0x7f800000);
            s_cnt[ql] = 0;
        }
    }
    __syncthreads();

    for (int64_t t0 = r0; t0 < r1; t0 += FT_N) {
        float acc[4][8];
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 8; ++b) acc[a][b] = 0.f;

        for (int k0 = 0; k0 < D; k0 += FT_K) {
            __syncthreads();
            for (int idx = tid; idx < FT_Q * FT_K; idx += FT_THREADS) {
                int r = idx / FT_K, kk = idx - r * FT_K;
                float v = 0.f;
                if (q0 + r < nq && k0 + kk < D) v = queries[(int64_t)(q0 + r) * ldq + k0 + kk];
                Qs[kk * (FT_Q + 4) + r] = v;
            }
            for (int idx = tid; idx < FT_N * FT_K; idx += FT_THREADS) {
                int r = idx / FT_K, kk = idx - r * FT_K;
                int64_t row = t0 + r;
                float v = 0.f;
                if (row < r1 && k0 + kk < D) {
                    v = x[row * ldx + k0 + kk];
                    if (row_scale) v = __fmul_rn(v, row_scale[row]);
                }
                Xs[kk * (FT_N + 4) + r] = v;
            }
            __syncthreads();
#pragma unroll
            for (int kk = 0; kk < FT_K; ++kk) {
                float4 qa = *reinterpret_cast<const float4*>(Qs + kk * (FT_Q + 4) + ty * 4);
                float4 xa = *reinterpret_cast<const float4*>(Xs + kk * (FT_N + 4) + tx * 4);
                float4 xb = *reinterpret_cast<const float4*>(Xs + kk * (FT_N + 4) + 64 + tx * 4);
                float qv[4] = {qa.x, qa.y, qa.z, qa.w};
                float xv[8] = {xa.x, xa.y, xa.z, xa.w, xb.x, xb.y, xb.z, xb.w};
#pragma unroll
                for (int a = 0; a < 4; ++a)
#pragma unroll
                    for (int b = 0; b < 8; ++b) acc[a][b] = __fmaf_rn(qv[a], xv[b], acc[a][b]);
            }
        }
        // scores -> shared tile
#pragma unroll
        for (int a = 0; a < 4; ++a) {
            int ql = ty * 4 + a;
#pragma unroll
            for (int b = 0; b < 8; ++b) {
                int rl = (b < 4) ? (tx * 4 + b) : (64 + tx * 4 + (b - 4));
                int64_t row = t0 + rl;
                float v;
                if (metric == B200VQ_METRIC_L2) {
                    float xn = (row < r1) ? xnorm[row] : 0.f;
                    v = __fmaf_rn(-2.f, acc[a][b], __fadd_rn(qn[ql], xn));
                    v = fmaxf(v, 0.f);
                } else {
                    v = -acc[a][b];
                }
                ds[ql * (FT_N + 1) + rl] = v;
            }
        }
        __syncthreads();
        // per-query running top-k: warp w owns queries w*8 .. w*8+7
        for (int qq = 0; qq < FT_Q / (FT_THREADS / 32); ++qq) {
            int ql = w * (FT_Q / (FT_THREADS / 32)) + qq;
            if (q0 + ql >= nq) break;
            WarpTopK<CAP> tk;
            tk.buf = lists + (size_t)ql * CAP;
            tk.k = k;
            tk.count = s_cnt[ql];
            tk.thr = s_thr[ql];
            for (int c = 0; c < FT_N / 32; ++c) {
                int rl = c * 32 + lane;
                int64_t row = t0 + rl;
                const float dv = ds[ql * (FT_N + 1) + rl];
                bool ok = row < r1;
                if (after_vals) ok = ok && make_key(dv, (uint32_t)row) > s_after[ql];
                tk.offer(ok, dv, (uint32_t)row);
            }
            if (lane == 0) {
                s_cnt[ql] = tk.count;
                s_thr[ql] = tk.thr;
            }
            __syncwarp();
        }
    }
    __syncthreads();
    for (int qq = 0; qq < FT_Q / (FT_THREADS / 32); ++qq) {
        int ql = w * (FT_Q / (FT_THREADS / 32)) + qq;
        if (q0 + ql >= nq) break;
        WarpTopK<CAP> tk;
        tk.buf = lists + (size_t)ql * CAP;
        tk.k = k;
        tk.count = s_cnt[ql];
        tk.thr = s_thr[ql];
        tk.finish();
        uint64_t* dst = part + ((size_t)(q0 + ql) * nchunks + chunk) * k;
        for (int j = lane; j < k; j += 32) dst[j] = (j < tk.count) ? tk.buf[j] : KEY_INF;
    }
}

// ---------------------------------------------- merge with 64-bit ids ---
// One CTA per query; P*k <= MAXE entries sorted by (value, id) lexicographic.
template <int MAXE>
__global__ void topk_merge64_kernel(const float* __restrict__ in_vals,
                                    const int64_t* __restrict__ in_ids, int P, int k, int negate,
                                    float* __restrict__ out_vals, int64_t* __restrict__ out_ids) {
    __shared__ uint32_t sv[MAXE];
    __shared__ int64_t si[MAXE];
    const int q = blockIdx.x;
    const int total = P * k;
    const int n2 = next_pow2(total);
    for (int i = threadIdx.x; i < n2; i += blockDim.x) {
        uint32_t kv = 0xffffffffu;
        int64_t id = 0x7fffffffffffffffll;
        if (i < total) {
            int64_t gid = in_ids[(size_t)q * total + i];
            if (gid >= 0) {
                float v = in_vals[(size_t)q * total + i];
                kv = float_to_ordered(negate ? -v : v);
                id = gid;
            }
        }
        sv[i] = kv;
        si[i] = id;
    }
    for (int size = 2; size <= n2; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            __syncthreads();
            for (int t = threadIdx.x; t < (n2 >> 1); t += blockDim.x) {
                int lo = 2 * t - (t & (stride - 1));
                int hi = lo + stride;
                bool up = ((lo & size) == 0);
                uint32_t av = sv[lo], bv = sv[hi];
                int64_t ai = si[lo], bi = si[hi];
                bool gt = (av > bv) || (av == bv && ai > bi);
                if (gt == up) {
                    sv[lo] = bv;
                    sv[hi] = av;
                    si[lo] = bi;
                    si[hi] = ai;
                }
            }
        }
    }
    __syncthreads();
    for (int j = threadIdx.x; j < k; j += blockDim.x) {
        float v;
        int64_t id;
        if (j < total && si[j] != 0x7fffffffffffffffll) {
            v = ordered_to_float(sv[j]);
            if (negate) v = -v;
            id = si[j];
        } else {
            v = negate ? -__int_as_float(0x7f800000) : __int_as_float(0x7f800000);
            id = -1;
        }
        out_vals[(size_t)q * k + j] = v;
        out_ids[(size_t)q * k + j] = id;
    }
}

// ------------------------------------------------------- host helpers ---
struct ScanPlan {
    int64_t chunk_rows;
    int nchunks;
};

// Chunks are multiples of 2048 rows; enough CTAs for ~4 waves when n allows.
static ScanPlan plan_scan(int64_t n, int nq_ctas) {
    const int64_t unit = 2048;
    int64_t target_ctas = (int64_t)device_sm_count() * 8;
    // every CTA stages its query's tables first (tens of KB): with many query CTAs one chunk each is enough
    int64_t want_chunks = nq_ctas * 2 >= target_ctas ? 1 : (target_ctas + nq_ctas - 1) / (nq_ctas > 0 ? nq_ctas : 1);
    if (want_chunks < 1) want_chunks = 1;
    int64_t max_chunks = (n + unit - 1) / unit;
    if (max_chunks < 1) max_chunks = 1;
    int64_t min_chunks = (n + 65536 * 4 - 1) / (65536 * 4);  // cap chunk at 256K rows
    if (min_chunks < 1) min_chunks = 1;
    int64_t nch = want_chunks;
    if (nch > max_chunks) nch = max_chunks;
    if (nch < min_chunks) nch = min_chunks;
    if (nch > 65535) nch = 65535;
    int64_t rows = (n + nch - 1) / nch;
    rows = ((rows + unit - 1) / unit) * unit;
    if (rows < unit) rows = unit;
    ScanPlan p;
    p.chunk_rows = rows;
    p.nchunks = (int)((n + rows - 1) / rows);
    if (p.nchunks < 1) p.nchunks = 1;
    return p;
}

static size_t partial_bytes(int64_t n, int nq, int k, int nq_ctas) {
    ScanPlan p = plan_scan(n, nq_ctas);
    return (size_t)nq * p.nchunks * k * sizeof(uint64_t);
}

template <int TILE>
static int launch_merge(const uint64_t* part, int nq, int nparts, int k, int negate,
                        float* out_vals, uint32_t* out_ids, cudaStream_t st, const uint32_t* gate = nullptr,
                        uint32_t gate_cap = 0) {
    // the kernel sorts TILE keys per pass whatever the number of partial entries: a few partial lists (small n) take
    // the 256-key instantiation (an order of magnitude fewer compare-exchange steps)
    const int64_t total = (int64_t)nparts * k;
    if (TILE > 256 && k <= 64 && total + k <= 256)
        merge_partials_kernel<256><<<nq, 256, 0, st>>>(part, nparts, k, negate, out_vals, out_ids, gate, gate_cap);
    else if (TILE > 512 && k <= 128 && total + k <= 512)
        merge_partials_kernel<512><<<nq, 256, 0, st>>>(part, nparts, k, negate, out_vals, out_ids, gate, gate_cap);
    else
        merge_partials_kernel<TILE><<<nq, 256, 0, st>>>(part, nparts, k, negate, out_vals, out_ids, gate, gate_cap);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

}  // namespace b200vq

using namespace b200vq;

extern "C" size_t b200vq_adc_scan_scratch_bytes(int64_t n, int nq, int k) {
    if (n < 0 || nq <= 0 || k <= 0) return 0;
    return partial_bytes(n, nq, k, nq);
}

template <int CAP>
static int adc_scan_launch(const uint8_t* codes, int64_t n, int M, int ksub, const float* lut,
                           int nq, int k, int negate, uint64_t* part, const ScanPlan& p,
                           cudaStream_t st) {
    const int n2 = next_pow2(SCAN_WARPS * k);
    size_t lut_bytes = ((size_t)M * ksub * 4 + 15) & ~(size_t)15;
    size_t tail = (size_t)(SCAN_WARPS * CAP + n2) * sizeof(uint64_t);
    bool lut_smem = lut_bytes + tail <= 200 * 1024;
    size_t smem = (lut_smem ? lut_bytes : 0) + tail;
    int vec = (M % 16 == 0 && ((uintptr_t)codes & 15) == 0) ? 16
              : (M % 4 == 0 && ((uintptr_t)codes & 3) == 0) ? 4 : 1;
    dim3 grid((unsigned)nq, (unsigned)p.nchunks);
#define LAUNCH_ADC(LS, V)                                                                      \
    do {                                                                                       \
        VQ_CUDA(cudaFuncSetAttribute(adc_scan_kernel<CAP, LS, V>,                              \
                                     cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
        adc_scan_kernel<CAP, LS, V><<<grid, SCAN_THREADS, smem, st>>>(                         \
            codes, n, M, ksub, lut, k, p.chunk_rows, negate, part);                            \
    } while (0)
    if (lut_smem) {
        if (vec == 16) LAUNCH_ADC(true, 16);
        else if (vec == 4) LAUNCH_ADC(true, 4);
        else LAUNCH_ADC(true, 1);
    } else {
        if (vec == 16) LAUNCH_ADC(false, 16);
        else if (vec == 4) LAUNCH_ADC(false, 4);
        else LAUNCH_ADC(false, 1);
    }
#undef LAUNCH_ADC
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" int b200vq_adc_scan_topk(const uint8_t* codes, int64_t n, int M, int ksub,
                                    const float* lut, int nq, int metric, int k, float* out_vals,
                                    uint32_t* out_ids, void* scratch, size_t scratch_bytes,
                                    void* stream) {
    VQ_CHECK_ARG(codes && lut && out_vals && out_ids, "adc_scan_topk: null pointer");
    VQ_CHECK_ARG(n >= 0 && n < (1ll << 32) && M > 0 && ksub > 0 && ksub <= 256 && nq >= 0,
                 "adc_scan_topk: bad shape n=%lld M=%d ksub=%d nq=%d", (long long)n, M, ksub, nq);
    VQ_CHECK_ARG(k > 0 && k <= 128, "adc_scan_topk: k must be in [1,128], got %d", k);
    VQ_CHECK_ARG(metric == B200VQ_METRIC_L2 || metric == B200VQ_METRIC_IP, "adc_scan_topk: bad metric");
    VQ_CHECK_ARG(nq <= 2147483647 / 1, "adc_scan_topk: nq too large");
    if (nq == 0) return B200VQ_OK;
    size_t need = partial_bytes(n, nq, k, nq);
    VQ_CHECK_ARG(scratch && scratch_bytes >= need, "adc_scan_topk: scratch too small (%zu < %zu)",
                 scratch_bytes, need);
    cudaStream_t st = as_stream(stream);
    ScanPlan p = plan_scan(n, nq);
    uint64_t* part = reinterpret_cast<uint64_t*>(scratch);
    int negate = metric == B200VQ_METRIC_IP;
    int rc = (k <= 32) ? adc_scan_launch<128>(codes, n, M, ksub, lut, nq, k, negate, part, p, st)
                       : adc_scan_launch<256>(codes, n, M, ksub, lut, nq, k, negate, part, p, st);
    if (rc != B200VQ_OK) return rc;
    return launch_merge<2048>(part, nq, p.nchunks, k, negate, out_vals, out_ids, st);
}

extern "C" int b200vq_pq_codes_to_scan_layout(const uint8_t* codes, int64_t n, int M, uint8_t* out,
                                              void* stream) {
    VQ_CHECK_ARG((codes || n == 0) && (out || n == 0) && n >= 0 && M > 0 && (M % 32 == 0 || M == 16),
                 "pq_codes_to_scan_layout: need M %% 32 == 0 or M == 16 (M=%d)", M);
    if (n == 0) return B200VQ_OK;
    int64_t total = ((n + 31) / 32) * 32 * (int64_t)M;
    int64_t g = (total + 255) / 256;
    int64_t cap = (int64_t)device_sm_count() * 16;
    if (g > cap) g = cap;
    codes_to_scan_layout_kernel<<<(int)g, 256, 0, as_stream(stream)>>>(codes, n, M, M == 16 ? 16 : 32, out);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

template <int CAP, int G, int GW = 32>
static int adc_scan_rot_launch_g(const uint8_t* codes_rot, int64_t n, const float* lut, int nq, int k,
                                 int negate, uint64_t* part, const ScanPlan& p, cudaStream_t st,
                                 const uint32_t* gate = nullptr, uint32_t gate_cap = 0) {
    const int n2 = next_pow2(SCAN_WARPS * k);
    size_t smem = (size_t)G * 256 * 64 * 4 + (size_t)(SCAN_WARPS * CAP + n2) * sizeof(uint64_t);
    VQ_CHECK_ARG(smem <= 227 * 1024, "adc_scan_topk_rot: M=%d needs %zu B of shared memory", G * GW, smem);
    VQ_CUDA(cudaFuncSetAttribute(adc_scan_rot_kernel<CAP, G, GW>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)smem));
    dim3 grid((unsigned)nq, (unsigned)p.nchunks);
    adc_scan_rot_kernel<CAP, G, GW><<<grid, SCAN_THREADS, smem, st>>>(codes_rot, n, lut, k, p.chunk_rows, negate,
                                                                      part, gate, gate_cap);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

template <int CAP>
static int adc_scan_rot_launch(const uint8_t* codes_rot, int64_t n, int M, const float* lut, int nq,
                               int k, int negate, uint64_t* part, const ScanPlan& p,
                               cudaStream_t st, const uint32_t* gate = nullptr, uint32_t gate_cap = 0) {
    if (M == 16) return adc_scan_rot_launch_g<CAP, 1, 16>(codes_rot, n, lut, nq, k, negate, part, p, st, gate, gate_cap);
    switch (M / 32) {
        case 1: return adc_scan_rot_launch_g<CAP, 1>(codes_rot, n, lut, nq, k, negate, part, p, st, gate, gate_cap);
        case 2: return adc_scan_rot_launch_g<CAP, 2>(codes_rot, n, lut, nq, k, negate, part, p, st, gate, gate_cap);
        default: return adc_scan_rot_launch_g<CAP, 3>(codes_rot, n, lut, nq, k, negate, part, p, st, gate, gate_cap);
    }
}

extern "C" int b200vq_adc_scan_topk_rot(const uint8_t* codes_rot, int64_t n, int M, int ksub,
                                        const float* lut, int nq, int metric, int k,
                                        float* out_vals, uint32_t* out_ids, void* scratch,
                                        size_t scratch_bytes, void* stream) {
    VQ_CHECK_ARG(codes_rot && lut && out_vals && out_ids, "adc_scan_topk_rot: null pointer");
    VQ_CHECK_ARG(ksub == 256 && M > 0 && (M % 32 == 0 || M == 16) && M <= 96,
                 "adc_scan_topk_rot: needs ksub == 256 and M in {16,32,64,96} (M=%d ksub=%d)", M, ksub);
    VQ_CHECK_ARG(n >= 0 && n < (1ll << 32) && nq >= 0, "adc_scan_topk_rot: bad n/nq");
    VQ_CHECK_ARG(k > 0 && k <= 128, "adc_scan_topk_rot: k must be in [1,128], got %d", k);
    VQ_CHECK_ARG(((uintptr_t)codes_rot & 15) == 0, "adc_scan_topk_rot: codes must be 16B aligned");
    if (nq == 0) return B200VQ_OK;
    size_t need = partial_bytes(n, nq, k, nq);
    VQ_CHECK_ARG(scratch && scratch_bytes >= need, "adc_scan_topk_rot: scratch too small");
    cudaStream_t st = as_stream(stream);
    ScanPlan p = plan_scan(n, nq);
    uint64_t* part = reinterpret_cast<uint64_t*>(scratch);
    int negate = metric == B200VQ_METRIC_IP;
    int rc = (k <= 32) ? adc_scan_rot_launch<128>(codes_rot, n, M, lut, nq, k, negate, part, p, st)
                       : adc_scan_rot_launch<256>(codes_rot, n, M, lut, nq, k, negate, part, p, st);
    if (rc != B200VQ_OK) return rc;
    return launch_merge<2048>(part, nq, p.nchunks, k, negate, out_vals, out_ids, st);
}

// ---- screened scan: sample threshold -> packed 6-bit tables -> screen -> exact finish (+ gated fallback)
struct ScreenPlan {
    int64_t n_sample;
    int ngroups;
    uint32_t cap;
    size_t off_vals, off_ids, off_part, off_qlut, off_thr, off_count, off_cand, total;
};
static ScreenPlan plan_screen(int64_t n, int M, int nq, int k) {
    ScreenPlan s;
    // threshold sample: an eighth of the rows, between 8192 and SCREEN_SAMPLE (a smaller sample only loosens
    // the thresholds; the result stays exact)
    // (shards of at most 262 144 rows: a sixteenth, at most SAMPLE_KTH_MAX rows, scored and radix-selected by
    // adc_sample_kth_kernel - the general path's fixed costs would dominate the whole search)
    int64_t ns = (n / 16) & ~(int64_t)31;
    if (ns < 4096) ns = 4096;
    if (ns > SAMPLE_KTH_MAX) ns = SAMPLE_KTH_MAX;
    // (about k n / ns rows survive a threshold taken from ns rows - several times more when the 6-bit bound is loose:
    // the small sample only while that estimate stays below a sixteenth of the candidate lists,
    // otherwise large k or wide-range tables would overflow them into the exact fallback)
    if (!(n <= (1 << 18) && (M == 16 || M == 32) && (int64_t)k * n <= 256 * ns)) {
        ns = (n / 8) & ~(int64_t)31;
        if (ns < 8192) ns = 8192;
        if (ns > SCREEN_SAMPLE) ns = SCREEN_SAMPLE;
    }
    s.n_sample = ns;
    s.ngroups = (nq + 3) / 4;
    int64_t cap = n / 256;
    if (cap < 4096) cap = 4096;
    if (cap > (1 << 20)) cap = 1 << 20;
    s.cap = (uint32_t)cap;
    auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
    size_t o = 0;
    s.off_vals = o; o = up(o + (size_t)nq * k * 4);
    s.off_ids = o;  o = up(o + (size_t)nq * k * 4);
    size_t part = partial_bytes(n, nq, k, nq);   // fallback partials; the sample's (smaller) alias them
    size_t part_s = partial_bytes(s.n_sample, nq, k, nq);
    s.off_part = o; o = up(o + (part > part_s ? part : part_s));
    s.off_qlut = o; o = up(o + (size_t)s.ngroups * M * 256 * 4);
    s.off_thr = o;  o = up(o + (size_t)s.ngroups * 4 * 4);
    s.off_count = o; o = up(o + (size_t)s.ngroups * 4 * 4);
    s.off_cand = o; o = up(o + (size_t)s.ngroups * 4 * s.cap * 4);
    s.total = o;
    return s;
}
static bool screen_applies(int64_t n, int nq, int k) { return n >= 65536 && nq >= 4 && k <= 64; }

extern "C" size_t b200vq_adc_scan_screen_scratch_bytes(int64_t n, int M, int nq, int k) {
    if (n < 0 || nq <= 0 || k <= 0 || M <= 0) return 0;
    if (!screen_applies(n, nq, k)) return partial_bytes(n, nq, k, nq);
    return plan_screen(n, M, nq, k).total;
}

template <int G, bool ABS, int GW = 32>
static int screen_launch_g(const uint8_t* codes_rot, int64_t n, const uint32_t* qlut4, const int* thr,
                           const ScanPlan& p, int ngroups, uint32_t* cand, uint32_t* count, uint32_t cap,
                           cudaStream_t st) {
    // ABS: up to 64 KiB of padding in front of the table (it starts at a multiple of 64 KiB of the shared window)
    const size_t smem = (size_t)G * 256 * 64 * 4 + (ABS ? 65536 : 0) + (size_t)4 * SCREEN_LCAP * 4;
    VQ_CUDA(cudaFuncSetAttribute(adc_screen_kernel<G, ABS, GW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid((unsigned)ngroups, (unsigned)p.nchunks);
    adc_screen_kernel<G, ABS, GW><<<grid, ABS ? 1024 : SCAN_THREADS, smem, st>>>(codes_rot, n, qlut4, thr, p.chunk_rows,
                                                                            cand, count, cap);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

template <int CAP>
static int screen_finish_launch(const uint8_t* codes_rot, int M, const float* lut, const uint32_t* cand,
                                const uint32_t* count, uint32_t cap, int nq, int k, int negate, float* out_vals,
                                uint32_t* out_ids, cudaStream_t st) {
    const int n2 = next_pow2(SCAN_WARPS * k);
    size_t smem = (size_t)(SCAN_WARPS * CAP + n2) * sizeof(uint64_t);
    // lists up to half the capacity (at most FINISH_SEL_MAX rows) take the select path: keys in shared memory
    uint32_t sel_max = cap / 2 < (uint32_t)FINISH_SEL_MAX ? cap / 2 : (uint32_t)FINISH_SEL_MAX;
    if (smem < (size_t)sel_max * sizeof(uint64_t)) smem = (size_t)sel_max * sizeof(uint64_t);
    smem += (size_t)M * 1024;      // the query's table
    VQ_CUDA(cudaFuncSetAttribute(screen_finish_kernel<CAP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    screen_finish_kernel<CAP><<<nq, SCAN_THREADS, smem, st>>>(codes_rot, M, lut, cand, count, cap, k, negate, sel_max,
                                                             out_vals, out_ids);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" int b200vq_adc_scan_topk_screen(const uint8_t* codes_rot, int64_t n, int M, int ksub, const float* lut,
                                           int nq, int metric, int k, float* out_vals, uint32_t* out_ids,
                                           void* scratch, size_t scratch_bytes, void* stream) {
    VQ_CHECK_ARG(codes_rot && lut && out_vals && out_ids, "adc_scan_topk_screen: null pointer");
    VQ_CHECK_ARG(ksub == 256 && M > 0 && (M % 32 == 0 || M == 16) && M <= 96,
                 "adc_scan_topk_screen: needs ksub == 256 and M in {16,32,64,96} (M=%d ksub=%d)", M, ksub);
    VQ_CHECK_ARG(n >= 0 && n < (1ll << 32) && nq >= 0, "adc_scan_topk_screen: bad n/nq");
    VQ_CHECK_ARG(k > 0 && k <= 128, "adc_scan_topk_screen: k must be in [1,128], got %d", k);
    if (nq == 0) return B200VQ_OK;
    if (!screen_applies(n, nq, k))   // small problems: the exact kernel is already launch-bound
        return b200vq_adc_scan_topk_rot(codes_rot, n, M, ksub, lut, nq, metric, k, out_vals, out_ids, scratch,
                                        scratch_bytes, stream);
    const ScreenPlan sp = plan_screen(n, M, nq, k);
    VQ_CHECK_ARG(scratch && scratch_bytes >= sp.total, "adc_scan_topk_screen: scratch too small");
    cudaStream_t st = as_stream(stream);
    uint8_t* base = reinterpret_cast<uint8_t*>(scratch);
    float* s_vals = reinterpret_cast<float*>(base + sp.off_vals);
    uint32_t* s_ids = reinterpret_cast<uint32_t*>(base + sp.off_ids);
    uint64_t* part = reinterpret_cast<uint64_t*>(base + sp.off_part);
    uint32_t* qlut4 = reinterpret_cast<uint32_t*>(base + sp.off_qlut);
    int* thr = reinterpret_cast<int*>(base + sp.off_thr);
    uint32_t* count = reinterpret_cast<uint32_t*>(base + sp.off_count);
    uint32_t* cand = reinterpret_cast<uint32_t*>(base + sp.off_cand);
    const int negate = metric == B200VQ_METRIC_IP;
    // 1. threshold sample: exact k-th best score of the first n_sample rows
    if (sp.n_sample <= SAMPLE_KTH_MAX && (M == 16 || M == 32)) {
        const size_t smem = (size_t)256 * 64 * 4 + (size_t)sp.n_sample * 4;
        if (M == 16) {
            VQ_CUDA(cudaFuncSetAttribute(adc_sample_kth_kernel<1, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            adc_sample_kth_kernel<1, 16><<<nq, SAMPLE_KTH_THREADS, smem, st>>>(codes_rot, (int)sp.n_sample, lut, k, negate, s_vals);
        } else {
            VQ_CUDA(cudaFuncSetAttribute(adc_sample_kth_kernel<1, 32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            adc_sample_kth_kernel<1, 32><<<nq, SAMPLE_KTH_THREADS, smem, st>>>(codes_rot, (int)sp.n_sample, lut, k, negate, s_vals);
        }
        VQ_LAUNCH_CHECK();
    } else {
        ScanPlan p = plan_scan(sp.n_sample, nq);
        int rc = (k <= 32) ? adc_scan_rot_launch<128>(codes_rot, sp.n_sample, M, lut, nq, k, negate, part, p, st)
                           : adc_scan_rot_launch<256>(codes_rot, sp.n_sample, M, lut, nq, k, negate, part, p, st);
        if (rc != B200VQ_OK) return rc;
        rc = launch_merge<2048>(part, nq, p.nchunks, k, negate, s_vals, s_ids, st);
        if (rc != B200VQ_OK) return rc;
    }
    // 2. packed 6-bit tables + integer thresholds
    VQ_CUDA(cudaMemsetAsync(qlut4, 0, (size_t)sp.ngroups * M * 256 * 4, st));
    VQ_CUDA(cudaMemsetAsync(thr, 0xFF, (size_t)sp.ngroups * 4 * 4, st));      // -1: unused query slots never pass
    VQ_CUDA(cudaMemsetAsync(count, 0, (size_t)sp.ngroups * 4 * 4, st));
    screen_prepare_kernel<<<nq, 256, 0, st>>>(lut, nq, M, s_vals, k, negate, reinterpret_cast<uint8_t*>(qlut4), thr);
    VQ_LAUNCH_CHECK();
    // 3. screen every row, four queries per lookup
    {
        ScanPlan p = plan_scan(n, sp.ngroups);
        int rc;
        if (M == 16) rc = screen_launch_g<1, true, 16>(codes_rot, n, qlut4, thr, p, sp.ngroups, cand, count, sp.cap, st);
        else switch (M / 32) {
            case 1: rc = screen_launch_g<1, true>(codes_rot, n, qlut4, thr, p, sp.ngroups, cand, count, sp.cap, st); break;
            case 2: rc = screen_launch_g<2, true>(codes_rot, n, qlut4, thr, p, sp.ngroups, cand, count, sp.cap, st); break;
            default: rc = screen_launch_g<3, false>(codes_rot, n, qlut4, thr, p, sp.ngroups, cand, count, sp.cap, st); break;
        }
        if (rc != B200VQ_OK) return rc;
    }
    // 4. exact scores of the survivors + top-k
    {
        int rc = (k <= 32) ? screen_finish_launch<128>(codes_rot, M, lut, cand, count, sp.cap, nq, k, negate,
                                                       out_vals, out_ids, st)
                           : screen_finish_launch<256>(codes_rot, M, lut, cand, count, sp.cap, nq, k, negate,
                                                       out_vals, out_ids, st);
        if (rc != B200VQ_OK) return rc;
    }
    // 5. queries whose list overflowed: exact scan, gated per query (CTAs of the others exit at once)
    {
        ScanPlan p = plan_scan(n, nq);
        int rc = (k <= 32) ? adc_scan_rot_launch<128>(codes_rot, n, M, lut, nq, k, negate, part, p, st, count, sp.cap)
                           : adc_scan_rot_launch<256>(codes_rot, n, M, lut, nq, k, negate, part, p, st, count, sp.cap);
        if (rc != B200VQ_OK) return rc;
        return launch_merge<2048>(part, nq, p.nchunks, k, negate, out_vals, out_ids, st, count, sp.cap);
    }
}

extern "C" int b200vq_row_sqnorms(const float* x, int64_t n, int64_t ldx, int D, float* out,
                                  void* stream) {
    VQ_CHECK_ARG((x || n == 0) && (out || n == 0) && n >= 0 && D > 0 && ldx >= D, "row_sqnorms: bad args");
    if (n == 0) return B200VQ_OK;
    int64_t g = (n * 32 + 255) / 256;
    int64_t cap = (int64_t)device_sm_count() * 16;
    if (g > cap) g = cap;
    row_sqnorms_kernel<<<(int)g, 256, 0, as_stream(stream)>>>(x, n, ldx, D, out);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" size_t b200vq_flat_topk_scratch_bytes(int64_t n, int nq, int k) {
    if (n < 0 || nq <= 0 || k <= 0) return 0;
    return partial_bytes(n, nq, k, (nq + FT_Q - 1) / FT_Q);
}

template <int CAP>
static int flat_launch(const float* x, int64_t n, int64_t ldx, int D, const float* xnorm,
                       const float* row_scale, const float* queries, int nq, int64_t ldq,
                       int metric, int k, uint64_t* part, const ScanPlan& p, cudaStream_t st,
                       const uint32_t* gate = nullptr, const float* after_vals = nullptr,
                       const uint32_t* after_ids = nullptr) {
    size_t smem = (size_t)(FT_K * (FT_Q + 4) + FT_K * (FT_N + 4) + FT_Q * (FT_N + 1) + 3 * FT_Q) * 4 +
                  16 + (size_t)FT_Q * (CAP + 1) * sizeof(uint64_t);
    VQ_CUDA(cudaFuncSetAttribute(flat_topk_kernel<CAP>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)smem));
    dim3 grid((unsigned)((nq + FT_Q - 1) / FT_Q), (unsigned)p.nchunks);
    flat_topk_kernel<CAP><<<grid, FT_THREADS, smem, st>>>(x, n, ldx, D, xnorm, row_scale, queries,
                                                          nq, ldq, metric, k, p.chunk_rows, part, gate, after_vals,
                                                          after_ids);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

// SIMT exact search restricted to the queries with gate[q] != 0 (internal: used by flat_tc.cu)
namespace b200vq {
int flat_topk_gated(const float* x, int64_t n, int64_t ldx, int D, const float* xnorm, const float* row_scale,
                    const float* queries, int nq, int64_t ldq, int metric, int k, float* out_vals, uint32_t* out_ids,
                    void* scratch, const uint32_t* gate, cudaStream_t st) {
    int nqc = (nq + FT_Q - 1) / FT_Q;
    ScanPlan p = plan_scan(n, nqc);
    uint64_t* part = reinterpret_cast<uint64_t*>(scratch);
    int rc = (k <= 32) ? flat_launch<64>(x, n, ldx, D, xnorm, row_scale, queries, nq, ldq, metric, k, part, p, st, gate)
                       : flat_launch<256>(x, n, ldx, D, xnorm, row_scale, queries, nq, ldq, metric, k, part, p, st, gate);
    if (rc != B200VQ_OK) return rc;
    return launch_merge<2048>(part, nq, p.nchunks, k, metric == B200VQ_METRIC_IP, out_vals, out_ids, st, gate, 0);
}
}  // namespace b200vq

extern "C" int b200vq_flat_topk(const float* x, int64_t n, int64_t ldx, int D, const float* xnorm,
                                const float* row_scale, const float* queries, int nq, int64_t ldq,
                                int metric, int k, float* out_vals, uint32_t* out_ids,
                                void* scratch, size_t scratch_bytes, void* stream) {
    VQ_CHECK_ARG(x && queries && out_vals && out_ids, "flat_topk: null pointer");
    VQ_CHECK_ARG(n >= 0 && n < (1ll << 32) && D > 0 && ldx >= D && ldq >= D && nq >= 0,
                 "flat_topk: bad shape");
    VQ_CHECK_ARG(k > 0 && k <= 128, "flat_topk: k must be in [1,128], got %d", k);
    VQ_CHECK_ARG(metric == B200VQ_METRIC_L2 || metric == B200VQ_METRIC_IP, "flat_topk: bad metric");
    VQ_CHECK_ARG(metric == B200VQ_METRIC_IP || xnorm, "flat_topk: L2 needs xnorm");
    if (nq == 0) return B200VQ_OK;
    int nqc = (nq + FT_Q - 1) / FT_Q;
    size_t need = partial_bytes(n, nq, k, nqc);
    VQ_CHECK_ARG(scratch && scratch_bytes >= need, "flat_topk: scratch too small (%zu < %zu)",
                 scratch_bytes, need);
    cudaStream_t st = as_stream(stream);
    ScanPlan p = plan_scan(n, nqc);
    uint64_t* part = reinterpret_cast<uint64_t*>(scratch);
    int rc = (k <= 32) ? flat_launch<64>(x, n, ldx, D, xnorm, row_scale, queries, nq, ldq, metric, k,
                                         part, p, st)
                       : flat_launch<256>(x, n, ldx, D, xnorm, row_scale, queries, nq, ldq, metric,
                                          k, part, p, st);
    if (rc != B200VQ_OK) return rc;
    return launch_merge<2048>(part, nq, p.nchunks, k, metric == B200VQ_METRIC_IP, out_vals, out_ids,
                              st);
}

// Continuation of an exact search: the best k rows AFTER (after_vals[q], after_ids[q]) in the (key, id) order of
// b200vq_flat_topk, i.e. its ranks r+1 .. r+k when called with the rank-r entry of every query. Callers chain calls
// to lift the k <= 128 limit of one launch (haag_vq uses it for IVF probe lists longer than 128).
extern "C" int b200vq_flat_topk_after(const float* x, int64_t n, int64_t ldx, int D, const float* xnorm,
                                      const float* row_scale, const float* queries, int nq, int64_t ldq, int metric,
                                      int k, const float* after_vals, const uint32_t* after_ids, float* out_vals,
                                      uint32_t* out_ids, void* scratch, size_t scratch_bytes, void* stream) {
    VQ_CHECK_ARG(x && queries && out_vals && out_ids && after_vals && after_ids, "flat_topk_after: null pointer");
    VQ_CHECK_ARG(n >= 0 && n < (1ll << 32) && D > 0 && ldx >= D && ldq >= D && nq >= 0, "flat_topk_after: bad shape");
    VQ_CHECK_ARG(k > 0 && k <= 128, "flat_topk_after: k must be in [1,128], got %d", k);
    VQ_CHECK_ARG(metric == B200VQ_METRIC_L2 || metric == B200VQ_METRIC_IP, "flat_topk_after: bad metric");
    VQ_CHECK_ARG(metric == B200VQ_METRIC_IP || xnorm, "flat_topk_after: L2 needs xnorm");
    if (nq == 0) return B200VQ_OK;
    int nqc = (nq + FT_Q - 1) / FT_Q;
    size_t need = partial_bytes(n, nq, k, nqc);
    VQ_CHECK_ARG(scratch && scratch_bytes >= need, "flat_topk_after: scratch too small (%zu < %zu)", scratch_bytes, need);
    cudaStream_t st = as_stream(stream);
    ScanPlan p = plan_scan(n, nqc);
    uint64_t* part = reinterpret_cast<uint64_t*>(scratch);
    int rc = (k <= 32) ? flat_launch<64>(x, n, ldx, D, xnorm, row_scale, queries, nq, ldq, metric, k, part, p, st,
                                         nullptr, after_vals, after_ids)
                       : flat_launch<256>(x, n, ldx, D, xnorm, row_scale, queries, nq, ldq, metric, k, part, p, st,
                                          nullptr, after_vals, after_ids);
    if (rc != B200VQ_OK) return rc;
    return launch_merge<2048>(part, nq, p.nchunks, k, metric == B200VQ_METRIC_IP, out_vals, out_ids, st);
}

extern "C" int b200vq_topk_merge(const float* in_vals, const int64_t* in_ids, int nq, int P, int k,
                                 int metric, float* out_vals, int64_t* out_ids, void* stream) {
    VQ_CHECK_ARG(in_vals && in_ids && out_vals && out_ids && nq >= 0 && P > 0 && k > 0,
                 "topk_merge: bad args");
    VQ_CHECK_ARG((int64_t)P * k <= 4096, "topk_merge: P*k must be <= 4096 (got %lld)",
                 (long long)P * k);
    if (nq == 0) return B200VQ_OK;
    topk_merge64_kernel<4096><<<nq, 256, 0, as_stream(stream)>>>(
        in_vals, in_ids, P, k, metric == B200VQ_METRIC_IP, out_vals, out_ids);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

// ------------------------------------------- allocation-free ADC scan ---
// b200vq_adc_scan_topk_static: b200vq_adc_scan_topk for callers that cannot allocate (CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc). The per-(query, row chunk)
// partial top-k lists live in a file-scope device array instead of a caller-owned buffer. The
// array is part of the module image: every GPU that loads libb200vq.so holds its 8 MiB until exit.
// No counterpart in the reference.
//
// In host code the NAME of a __device__ array is only a host-side stand-in for the symbol; the
// device address is obtained once per device (that device current) with cudaGetSymbolAddress and
// cached. Every call shares the array and finds the last call's partial lists in it: nothing has
// to start at zero (each of the nq * nchunks * k slots a call reads is written by that call's
// scan CTAs first), but two calls that overlap on the GPU (different streams) race on it.
static __device__ __align__(256) unsigned char g_adc_static_scratch[B200VQ_ADC_STATIC_SCRATCH_BYTES];

static int adc_static_scratch_address(void** out) {
    constexpr int MAX_DEV = 64;
    static std::mutex mu;
    static void* addr[MAX_DEV] = {};
    int dev = -1;
    VQ_CUDA(cudaGetDevice(&dev));
    VQ_CHECK_ARG(dev >= 0 && dev < MAX_DEV, "adc_scan_topk_static: device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[dev]) {
        void* p = nullptr;
        VQ_CUDA(cudaGetSymbolAddress(&p, g_adc_static_scratch));
        VQ_CHECK_ARG(p != nullptr && ((uintptr_t)p & 255) == 0, "adc_scan_topk_static: bad symbol address");
        addr[dev] = p;
    }
    *out = addr[dev];
    return B200VQ_OK;
}

extern "C" int b200vq_adc_scan_topk_static(const uint8_t* codes, int64_t n, int M, int ksub, const float* lut,
                                           int nq, int metric, int k, float* out_vals, uint32_t* out_ids,
                                           void* stream) {
    VQ_CHECK_ARG(codes && lut && out_vals && out_ids, "adc_scan_topk_static: null pointer");
    VQ_CHECK_ARG(n >= 0 && n < (1ll << 32) && M > 0 && ksub > 0 && ksub <= 256 && nq >= 0,
                 "adc_scan_topk_static: bad shape n=%lld M=%d ksub=%d nq=%d", (long long)n, M, ksub, nq);
    VQ_CHECK_ARG(k > 0 && k <= 128, "adc_scan_topk_static: k must be in [1,128], got %d", k);
    VQ_CHECK_ARG(metric == B200VQ_METRIC_L2 || metric == B200VQ_METRIC_IP, "adc_scan_topk_static: bad metric");
    if (nq == 0) return B200VQ_OK;
    size_t need = partial_bytes(n, nq, k, nq);
    VQ_CHECK_ARG(need <= (size_t)B200VQ_ADC_STATIC_SCRATCH_BYTES,
                 "adc_scan_topk_static: n=%lld nq=%d k=%d needs %zu bytes of partial lists, the built-in scratch "
                 "holds %zu; use b200vq_adc_scan_topk with caller-owned scratch",
                 (long long)n, nq, k, need, (size_t)B200VQ_ADC_STATIC_SCRATCH_BYTES);
    void* scratch = nullptr;
    int rc = adc_static_scratch_address(&scratch);
    if (rc != B200VQ_OK) return rc;
    cudaStream_t st = as_stream(stream);
    ScanPlan p = plan_scan(n, nq);
    uint64_t* part = reinterpret_cast<uint64_t*>(scratch);
    int negate = metric == B200VQ_METRIC_IP;
    rc = (k <= 32) ? adc_scan_launch<128>(codes, n, M, ksub, lut, nq, k, negate, part, p, st)
                   : adc_scan_launch<256>(codes, n, M, ksub, lut, nq, k, negate, part, p, st);
    if (rc != B200VQ_OK) return rc;
    return launch_merge<2048>(part, nq, p.nchunks, k, negate, out_vals, out_ids, st);
}
