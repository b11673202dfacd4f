// Screened ADC scan: four queries per shared-memory lookup, exact results.
//
// The exact scan (adc_scan_rot_kernel) performs one 4-byte table lookup per (code byte, query)
// and sits at the L1/shared-memory wavefront ceiling (ncu: 94.8 % of the data-pipe peak). Here
// the table entries of FOUR queries are quantised to 6 bits and packed into one 32-bit word per
// (sub-quantizer, code), so that one lookup serves four queries and four looked-up words are
// added before they are unpacked (no carries: 4 * 63 < 256). The integer sum S is a rigorous
// LOWER BOUND of the exact score,
//     score(row) >= sum_m lo[m] + S(row) * step,   u[m][c] = min(63, floor((T[m][c] - lo[m]) / step)),
// so a row can only belong to the top-k if S <= thr, where thr is derived from tau0 = the exact
// k-th best score of a 65 536-row sample (an upper bound of the final k-th best). Surviving row
// ids (about 0.06 % of the rows on the benchmark data) are appended to a per-query list; a
// second kernel evaluates them with the exact fp32 table and selects the top-k in the usual
// (value, id) order. A query whose list overflows is re-done by the exact kernel (gated launch),
// so the result is always the exact scan's. Same contract as FlatQuantizedIndex.search_with_scores
// (src/haag_vq/methods/search/flat_quantized_index.py:45-76) and bench/ffd_speed.cpp:87-127.
#pragma once
#include "common.cuh"
#include "topk.cuh"

namespace b200vq {

constexpr int SCREEN_QMAX = 63;          // 6-bit entries
constexpr int SCREEN_SAMPLE = 65536;     // rows of the threshold sample

// One CTA (256 threads) per query: lo[m] = min_c T'[m][c] (T' = -T for inner product), the step
// from the threshold budget, the integer threshold, and this query's byte lane of the packed
// table qlut4[(q/4)][m][c] (uint32 = 4 queries).
__global__ void __launch_bounds__(256)
screen_prepare_kernel(const float* __restrict__ lut, int nq, int M, const float* __restrict__ sample_vals, int k,
                      int negate, uint8_t* __restrict__ qlut4, int* __restrict__ thr) {
    __shared__ float red[8], redx[8];
    __shared__ float lo_s[96];
    __shared__ float amax_s[96];
    __shared__ int active_s;
    __shared__ float step_s;
    const int q = blockIdx.x, t = threadIdx.x, lane = t & 31, w = t >> 5;
    const float* T = lut + (size_t)q * M * 256;
    const float sgn = negate ? -1.f : 1.f;
    if (t == 0) active_s = 0;
    for (int m = 0; m < M; ++m) {
        float v = sgn * T[m * 256 + t], vx = v;
        for (int s = 16; s > 0; s >>= 1) {
            v = fminf(v, __shfl_xor_sync(FULL_MASK, v, s));
            vx = fmaxf(vx, __shfl_xor_sync(FULL_MASK, vx, s));
        }
        if (lane == 0) {
            red[w] = v;
            redx[w] = vx;
        }
        __syncthreads();
        if (t == 0) {
            float mn = red[0], mx = redx[0];
            for (int i = 1; i < 8; ++i) {
                mn = fminf(mn, red[i]);
                mx = fmaxf(mx, redx[i]);
            }
            lo_s[m] = mn;
            amax_s[m] = fmaxf(fabsf(mn), fabsf(mx));
            // constant tables (the all-zero tables of sub-quantizers added by padding M to the scan layout's
            // group width) always contribute exactly lo: they take no share of the budget
            if (mx > mn) ++active_s;
        }
        __syncthreads();
    }
    if (t == 0) {
        double sum_lo = 0.0, sum_abs = 0.0, sum_amax = 0.0;
        for (int m = 0; m < M; ++m) { sum_lo += lo_s[m]; sum_abs += fabs((double)lo_s[m]); sum_amax += (double)amax_s[m]; }
        const float tv = sample_vals[(size_t)q * k + (k - 1)];
        int th;
        float step;
        if (!(fabsf(tv) < 3.0e38f)) {      // fewer than k sample rows: no threshold, everything passes
            step = 1.0f;
            th = 65535;
        } else {
            const double tau = (double)(sgn * tv);
            // slack for the fp32 summation error of the sample's and the candidates' exact scores: an M-term fp32 sum
            // is off by at most (M - 1) 2^-24 sum_m |T[m][c]| <= (M - 1) 2^-24 sum_m max|T[m]| (mixed-sign inner-product
            // tables included), once for the threshold row and once for the candidate
            const double worst = 2.0 * (double)M * 5.9604645e-8 * 1.01 * sum_amax;
            double margin = 7.63e-6 * (fabs(tau) + sum_abs) + 1e-30;
            if (worst > margin) margin = worst;
            double budget = tau - sum_lo;
            if (budget < 0) budget = 0;
            // clip level of one table entry = 6x the mean budget per (non-constant) sub-quantizer of a threshold row
            double st = budget * 6.0 / ((double)(active_s > 0 ? active_s : 1) * SCREEN_QMAX);
            const double floor_st = (fabs(tau) + sum_abs + 1e-30) * 1e-9;
            if (st < floor_st) st = floor_st;
            step = (float)st;
            double tq = floor((budget + margin) / (double)step) + 1.0;
            th = tq > 65535.0 ? 65535 : (int)tq;
        }
        step_s = step;
        thr[q] = th;
    }
    __syncthreads();
    const float step = step_s;
    uint8_t* dst = qlut4 + ((size_t)(q >> 2) * M * 256) * 4 + (q & 3);
    for (int m = 0; m < M; ++m) {
        // u <= floor of the real quotient: the (1 - 2^-21) factor outweighs the two fp32 roundings
        float x = __fdiv_rn(__fsub_rn(sgn * T[m * 256 + t], lo_s[m]), step) * 0.99999952316284f;
        float u = floorf(x);
        u = fminf(fmaxf(u, 0.f), (float)SCREEN_QMAX);
        dst[((size_t)m * 256 + t) * 4] = (uint8_t)u;
    }
}

// ld.shared with a compile-time immediate offset from a 32-bit shared-window address
template <int IMM>
__device__ __forceinline__ uint32_t lds_abs(uint32_t addr) {
    uint32_t v;
    asm("ld.shared.u32 %0, [%1+%2];" : "=r"(v) : "r"(addr), "n"(IMM));
    return v;
}
// Lookups of code word W (4 code bytes = 4 consecutive steps) of the 8*G words of a lane, unrolled at
// compile time so that the group / step offsets are LDS immediates.
template <int G, int W>
struct ScreenWords {
    static __device__ __forceinline__ void run(const uint32_t (&wds)[8 * G], uint32_t lane4, uint32_t& lo16,
                                               uint32_t& hi16) {
        constexpr int g = W / 8, j = W % 8;
        constexpr int imm = g * (256 * 64 * 4) + (4 * j) * 4;
        uint32_t t4 = lds_abs<imm>(__byte_perm(wds[W], lane4, 0x7604u));
        t4 += lds_abs<imm + 4>(__byte_perm(wds[W], lane4, 0x7614u));
        t4 += lds_abs<imm + 8>(__byte_perm(wds[W], lane4, 0x7624u));
        t4 += lds_abs<imm + 12>(__byte_perm(wds[W], lane4, 0x7634u));
        lo16 += t4 & 0x00FF00FFu;
        hi16 += (t4 >> 8) & 0x00FF00FFu;
        ScreenWords<G, W + 1>::run(wds, lane4, lo16, hi16);
    }
};
template <int G>
struct ScreenWords<G, 8 * G> {
    static __device__ __forceinline__ void run(const uint32_t (&)[8 * G], uint32_t, uint32_t&, uint32_t&) {}
};

// grid (query groups of 4, row chunks). Table in shared memory as [g][code][64] uint32 (columns
// stored twice, as in adc_scan_rot_kernel). Survivors are appended to cand[q * cap + pos].
// ABS = true: the table starts at an ABSOLUTE shared-window address that is a multiple of 64 KiB,
// so the PRMT that merges (code << 8) with (lane << 2) also supplies the upper address bytes and
// its result IS the LDS address (no per-lookup base add). That costs up to 64 KiB of padding, so
// the CTA is 1024 threads (one per SM) instead of three CTAs of 256.
template <int G, bool ABS>
__global__ void __launch_bounds__(ABS ? 1024 : SCAN_THREADS)
adc_screen_kernel(const uint8_t* __restrict__ codes_rot, int64_t n, const uint32_t* __restrict__ qlut4,
                  const int* __restrict__ thr, int64_t chunk_rows, uint32_t* __restrict__ cand,
                  uint32_t* __restrict__ count, uint32_t cap) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int KSUB = 256;
    constexpr int M = G * 32;
    constexpr int THREADS = ABS ? 1024 : SCAN_THREADS;
    const int grp = blockIdx.x, chunk = blockIdx.y;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t base32 = (uint32_t)__cvta_generic_to_shared(smem_raw);
    const uint32_t tab32 = ABS ? ((base32 + 0xFFFFu) & ~0xFFFFu) : base32;
    uint32_t* tab = reinterpret_cast<uint32_t*>(smem_raw + (tab32 - base32));
    const uint32_t* src_q = qlut4 + (size_t)grp * M * KSUB;
    for (int idx = threadIdx.x; idx < G * KSUB * 8; idx += THREADS) {
        const int jq = idx & 7, kk = (idx >> 3) & (KSUB - 1), g = idx >> 11;
        const uint32_t* src = src_q + (size_t)(g * 32 + 4 * jq) * KSUB + kk;
        uint4 v = make_uint4(src[0], src[KSUB], src[2 * KSUB], src[3 * KSUB]);
        uint4* dst = reinterpret_cast<uint4*>(tab + ((size_t)g * KSUB + kk) * 64 + 4 * jq);
        dst[0] = v;
        dst[8] = v;
    }
    const int t0 = thr[grp * 4 + 0], t1 = thr[grp * 4 + 1], t2 = thr[grp * 4 + 2], t3 = thr[grp * 4 + 3];
    __syncthreads();

    // second PRMT operand: byte 0 = lane * 4, bytes 2..3 = upper bytes of the table address (ABS)
    const uint32_t lane4 = ((uint32_t)lane << 2) | (ABS ? (tab32 & 0xFFFF0000u) : 0u);
    const int64_t r0 = (int64_t)chunk * chunk_rows;  // multiple of 32
    int64_t r1 = r0 + chunk_rows;
    if (r1 > n) r1 = n;
    constexpr int UNR = 2;
    for (int64_t base = r0 + (int64_t)w * 32; base < r1; base += UNR * THREADS) {
        uint4 cv[UNR][2 * G];
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
            int64_t bb = base + (int64_t)u * THREADS;
            if (bb >= r1) bb = base;
            const uint4* cb = reinterpret_cast<const uint4*>(codes_rot + (bb >> 5) * (int64_t)G * 1024);
#pragma unroll
            for (int g = 0; g < G; ++g) {
                cv[u][2 * g] = ld_stream_u4(cb + g * 64 + lane);
                cv[u][2 * g + 1] = ld_stream_u4(cb + g * 64 + 32 + lane);
            }
        }
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
            const int64_t row = base + (int64_t)u * THREADS + lane;
            uint32_t lo16 = 0, hi16 = 0;   // 16-bit fields: (query 0, query 2) and (query 1, query 3)
            if (ABS) {
                uint32_t wds[8 * G];
#pragma unroll
                for (int g = 0; g < G; ++g) {
                    wds[8 * g + 0] = cv[u][2 * g].x; wds[8 * g + 1] = cv[u][2 * g].y;
                    wds[8 * g + 2] = cv[u][2 * g].z; wds[8 * g + 3] = cv[u][2 * g].w;
                    wds[8 * g + 4] = cv[u][2 * g + 1].x; wds[8 * g + 5] = cv[u][2 * g + 1].y;
                    wds[8 * g + 6] = cv[u][2 * g + 1].z; wds[8 * g + 7] = cv[u][2 * g + 1].w;
                }
                ScreenWords<G, 0>::run(wds, lane4, lo16, hi16);
            } else {
#pragma unroll
                for (int g = 0; g < G; ++g) {
                    const uint32_t wds[8] = {cv[u][2 * g].x, cv[u][2 * g].y, cv[u][2 * g].z, cv[u][2 * g].w,
                                             cv[u][2 * g + 1].x, cv[u][2 * g + 1].y, cv[u][2 * g + 1].z,
                                             cv[u][2 * g + 1].w};
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        uint32_t t4 = 0;
#pragma unroll
                        for (int b = 0; b < 4; ++b) {
                            const uint32_t off = __byte_perm(wds[j], lane4, 0x5504u | ((uint32_t)b << 4));
                            t4 += *reinterpret_cast<const uint32_t*>(smem_raw + off + g * (KSUB * 64 * 4) + (4 * j + b) * 4);
                        }
                        lo16 += t4 & 0x00FF00FFu;
                        hi16 += (t4 >> 8) & 0x00FF00FFu;
                    }
                }
            }
            const bool live = row < r1;
            const bool p0 = live && (int)(lo16 & 0xFFFFu) <= t0;
            const bool p1 = live && (int)(hi16 & 0xFFFFu) <= t1;
            const bool p2 = live && (int)(lo16 >> 16) <= t2;
            const bool p3 = live && (int)(hi16 >> 16) <= t3;
            if (__any_sync(FULL_MASK, p0 | p1 | p2 | p3)) {
                const bool ps[4] = {p0, p1, p2, p3};
#pragma unroll
                for (int qq = 0; qq < 4; ++qq) {
                    const unsigned mk = __ballot_sync(FULL_MASK, ps[qq]);
                    if (mk) {
                        uint32_t pos = 0;
                        if (lane == 0) pos = atomicAdd(count + grp * 4 + qq, (uint32_t)__popc(mk));
                        pos = __shfl_sync(FULL_MASK, pos, 0) + __popc(mk & ((1u << lane) - 1u));
                        if (ps[qq] && pos < cap) cand[(size_t)(grp * 4 + qq) * cap + pos] = (uint32_t)row;
                    }
                }
            }
        }
    }
}

// One CTA per query: exact fp32 score of every listed row (sub-quantizers in increasing order,
// codes read back from the scan layout), running top-k, final (value, id)-ordered k results.
template <int CAP>
__global__ void __launch_bounds__(SCAN_THREADS)
screen_finish_kernel(const uint8_t* __restrict__ codes_rot, int M, const float* __restrict__ lut,
                     const uint32_t* __restrict__ cand, const uint32_t* __restrict__ count, uint32_t cap, int k,
                     int negate, float* __restrict__ out_vals, uint32_t* __restrict__ out_ids) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int q = blockIdx.x;
    const uint32_t cnt = count[q];
    if (cnt > cap) return;     // overflow: the gated exact kernels produce this query's result
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int G = M >> 5;
    uint64_t* wbuf = reinterpret_cast<uint64_t*>(smem_raw);
    const int n2 = next_pow2(SCAN_WARPS * k);
    uint64_t* mbuf = wbuf + SCAN_WARPS * CAP;
    __shared__ uint64_t best[128];
    WarpTopK<CAP> tk;
    tk.init(wbuf + w * CAP, k);
    const float* T = lut + (size_t)q * M * 256;
    const uint32_t* list = cand + (size_t)q * cap;
    const uint32_t rounds = (cnt + SCAN_THREADS - 1) / SCAN_THREADS;
    for (uint32_t it = 0; it < rounds; ++it) {
        const uint32_t e = it * SCAN_THREADS + threadIdx.x;
        const bool valid = e < cnt;
        float acc = 0.f;
        uint32_t row = 0;
        if (valid) {
            row = list[e];
            const uint32_t i = row & 31u;
            const uint8_t* blk = codes_rot + (size_t)(row >> 5) * G * 1024 + (i << 4);
            for (int m = 0; m < M; ++m) {
                const uint32_t s = ((uint32_t)m - i) & 31u;
                const uint8_t c = blk[(size_t)(m >> 5) * 1024 + ((s >> 4) << 9) + (s & 15u)];
                acc += __ldg(T + m * 256 + c);
            }
        }
        tk.offer(valid, negate ? -acc : acc, row);
    }
    block_emit_topk<CAP>(tk, mbuf, n2, k, best);
    __syncthreads();
    for (int j = threadIdx.x; j < k; j += blockDim.x) {
        const uint64_t key = best[j];
        float v;
        uint32_t id;
        if (key == KEY_INF) {
            v = __int_as_float(0x7f800000);
            id = 0xffffffffu;
        } else {
            v = key_value(key);
            id = key_id(key);
        }
        if (negate) v = -v;
        out_vals[(size_t)q * k + j] = v;
        out_ids[(size_t)q * k + j] = id;
    }
}

}  // namespace b200vq
