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
constexpr int SCREEN_LCAP = 512;         // survivors per (CTA, query) buffered in shared memory

// One CTA (256 threads) per query: lo[m] = min_c T'[m][c] (T' = -T for inner product), the step
// from the threshold budget, the integer threshold, and this query's byte lane of the packed
// table qlut4[(q/4)][m][c] (uint32 = 4 queries).
__global__ void __launch_bounds__(256)
screen_prepare_kernel(const float* __restrict__ lut, int nq, int M, const float* __restrict__ sample_vals, int k,
                      int negate, uint8_t* __restrict__ qlut4, int* __restrict__ thr) {
    __shared__ float lo_s[96];
    __shared__ float amax_s[96];
    __shared__ int active_s;
    __shared__ float step_s;
    const int q = blockIdx.x, t = threadIdx.x, lane = t & 31, w = t >> 5;
    const float* T = lut + (size_t)q * M * 256;
    const float sgn = negate ? -1.f : 1.f;
    if (t == 0) active_s = 0;
    __syncthreads();
    // warp w reduces the tables of sub-quantizers w, w + 8, ...: eight entries per lane, min / max by shuffles
    for (int m = w; m < M; m += 8) {
        float v = sgn * T[m * 256 + lane], vx = v;
#pragma unroll
        for (int j = 1; j < 8; ++j) {
            const float e = sgn * T[m * 256 + j * 32 + lane];
            v = fminf(v, e);
            vx = fmaxf(vx, e);
        }
        for (int s = 16; s > 0; s >>= 1) {
            v = fminf(v, __shfl_xor_sync(FULL_MASK, v, s));
            vx = fmaxf(vx, __shfl_xor_sync(FULL_MASK, vx, s));
        }
        if (lane == 0) {
            lo_s[m] = v;
            amax_s[m] = fmaxf(fabsf(v), fabsf(vx));
            // constant tables (the all-zero tables of sub-quantizers added by padding M to the scan layout's
            // group width) always contribute exactly lo: they take no share of the budget
            if (vx > v) atomicAdd(&active_s, 1);
        }
    }
    __syncthreads();
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

// Threshold sample of a SMALL shard (n_sample <= SAMPLE_KTH_MAX rows): one CTA per query scores the first n_sample rows
// exactly (same tables in shared memory and the same lookups as adc_scan_rot_kernel), keeps the scores in shared memory
// and radix-selects the k-th smallest (four 8-bit passes over the order-preserving integer image of the floats).
// Only that value is needed (screen_prepare_kernel reads sample_vals[q * k + k - 1]); the running top-k lists, their
// bitonic compactions and the merge launch of the general path were 60 % of a 100 000-row search (ncu, r2).
constexpr int SAMPLE_KTH_MAX = 8192;
constexpr int SAMPLE_KTH_THREADS = 512;
template <int G, int GW>
__global__ void __launch_bounds__(SAMPLE_KTH_THREADS)
adc_sample_kth_kernel(const uint8_t* __restrict__ codes_rot, int n_sample, const float* __restrict__ lut, int k,
                      int negate, float* __restrict__ sample_vals) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int KSUB = 256;
    constexpr int M = G * GW;
    constexpr int NQW = GW / 4;
    constexpr int NV = GW / 16;
    const int q = blockIdx.x;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    float* lut_s = reinterpret_cast<float*>(smem_raw);
    constexpr size_t lut_bytes = (size_t)G * KSUB * 64 * 4;
    uint32_t* keys = reinterpret_cast<uint32_t*>(smem_raw + lut_bytes);     // [n_sample] ordered-integer scores
    __shared__ uint32_t hist[256];
    __shared__ uint32_t sel_prefix, sel_k;

    const float* lut_q = lut + (size_t)q * M * KSUB;
    for (int idx = threadIdx.x; idx < G * KSUB * NQW; idx += SAMPLE_KTH_THREADS) {
        const int jq = idx % NQW, kk = (idx / NQW) & (KSUB - 1), g = idx / (NQW * KSUB);
        const float* src = lut_q + (size_t)(g * GW + 4 * jq) * KSUB + kk;
        float4 v = make_float4(src[0], src[KSUB], src[2 * KSUB], src[3 * KSUB]);
        float4* dst = reinterpret_cast<float4*>(lut_s + ((size_t)g * KSUB + kk) * 64 + 4 * jq);
#pragma unroll
        for (int rep = 0; rep * GW < 32 + GW; ++rep) dst[rep * NQW] = v;
    }
    __syncthreads();

    const unsigned char* lbase = smem_raw;
    const uint32_t lane4 = (uint32_t)lane << 2;
    constexpr int UNR = 4;
    for (int base = w * 32; base < n_sample; base += UNR * SAMPLE_KTH_THREADS) {
        uint4 cv[UNR][NV * G];
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
            int bb = base + u * SAMPLE_KTH_THREADS;
            if (bb >= n_sample) bb = base;
            const uint4* cb = reinterpret_cast<const uint4*>(codes_rot + (size_t)(bb >> 5) * G * (32 * GW));
#pragma unroll
            for (int g = 0; g < G; ++g)
#pragma unroll
                for (int h = 0; h < NV; ++h) cv[u][NV * g + h] = ld_stream_u4(cb + g * 64 + h * 32 + lane);
        }
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
            const int row = base + u * SAMPLE_KTH_THREADS + lane;
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
                for (int j = 0; j < NQW; ++j)
#pragma unroll
                    for (int b = 0; b < 4; ++b) {
                        const uint32_t off = __byte_perm(wds[j], lane4, 0x5504u | ((uint32_t)b << 4));
                        acc += *reinterpret_cast<const float*>(lbase + off + g * (KSUB * 64 * 4) + (4 * j + b) * 4);
                    }
            }
            if (row < n_sample) keys[row] = float_to_ordered(negate ? -acc : acc);
        }
    }
    if (threadIdx.x == 0) {
        sel_prefix = 0;
        sel_k = (uint32_t)k;
    }
    __syncthreads();
    // radix select: after pass p the top 8 (p + 1) bits of the k-th smallest key are known
    for (int pass = 0; pass < 4; ++pass) {
        const int shift = 24 - 8 * pass;
        if (threadIdx.x < 256) hist[threadIdx.x] = 0;
        __syncthreads();
        const uint32_t prefix = sel_prefix, kk = sel_k;
        for (int i = threadIdx.x; i < n_sample; i += SAMPLE_KTH_THREADS) {
            const uint32_t key = keys[i];
            if (pass == 0 || ((key ^ prefix) >> (shift + 8)) == 0) atomicAdd(&hist[(key >> shift) & 255u], 1u);
        }
        __syncthreads();
        if (w == 0) {
            uint32_t c[8], tot = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                c[j] = hist[lane * 8 + j];
                tot += c[j];
            }
            uint32_t incl = tot;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(FULL_MASK, incl, o);
                if (lane >= o) incl += t;
            }
            const uint32_t excl = incl - tot;
            if (excl < kk && kk <= incl) {           // exactly one lane: its 8 bins hold the k-th key
                uint32_t run = excl;
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    if (run < kk && kk <= run + c[j]) {
                        sel_prefix = prefix | ((uint32_t)(lane * 8 + j) << shift);
                        sel_k = kk - run;
                    }
                    run += c[j];
                }
            }
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const float v = ordered_to_float(sel_prefix);
        sample_vals[(size_t)q * k + (k - 1)] = negate ? -v : v;
    }
}

// ld.shared with a compile-time immediate offset from a 32-bit shared-window address
template <int IMM>
__device__ __forceinline__ uint32_t lds_abs(uint32_t addr) {
    uint32_t v;
    asm("ld.shared.u32 %0, [%1+%2];" : "=r"(v) : "r"(addr), "n"(IMM));
    return v;
}
__device__ __forceinline__ uint32_t imad_acc(uint32_t a, uint32_t one, uint32_t acc) {
    uint32_t r;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(one), "r"(acc));
    return r;
}
// Lookups of code word W (4 code bytes = 4 consecutive steps) of the 8*G words of a lane, unrolled at
// compile time so that the group / step offsets are LDS immediates.
template <int NW, int W>
struct ScreenWords {
    static __device__ __forceinline__ void run(const uint32_t (&wds)[NW], uint32_t lane4, uint32_t& lo16,
                                               uint32_t& hi16, uint32_t one) {
        constexpr int g = W / 8, j = W % 8;
        constexpr int imm = g * (256 * 64 * 4) + (4 * j) * 4;
        uint32_t t4 = lds_abs<imm>(__byte_perm(wds[W], lane4, 0x7604u));
        t4 += lds_abs<imm + 4>(__byte_perm(wds[W], lane4, 0x7614u));
        t4 = imad_acc(lds_abs<imm + 8>(__byte_perm(wds[W], lane4, 0x7624u)), one, t4);
        t4 = imad_acc(lds_abs<imm + 12>(__byte_perm(wds[W], lane4, 0x7634u)), one, t4);
        // The kernel is bound by the integer ALU pipe (PRMT, LOP3, SHF, IADD3: 2.5 ALU instructions per look-up in the
        // first form, "math pipe throttle" the top stall), so: (1) adds as IMAD (x * 1 + acc, `one` is a runtime 1) run on
        // the FMA pipe; (2) (t4 >> 8) & 0x00FF00FF is one PRMT; (3) the even-byte sums are not masked out per word:
        // `lo16` accumulates the whole words (mod 2^32) and the caller subtracts hi16 << 8 once per row, which leaves
        // sum b0 + 2^16 sum b2 exactly. C5 scan: 10.37 -> 10.08 (1) -> 9.68 (2) ms; IMAD.HI for the shifted add: 10.83
        lo16 = imad_acc(t4, one, lo16);
        hi16 = imad_acc(__byte_perm(t4, 0u, 0x4341u), one, hi16);
        ScreenWords<NW, W + 1>::run(wds, lane4, lo16, hi16, one);
    }
};
template <int NW>
struct ScreenWords<NW, NW> {
    static __device__ __forceinline__ void run(const uint32_t (&)[NW], uint32_t, uint32_t&, uint32_t&, uint32_t) {}
};

// grid (query groups of 4, row chunks). Table in shared memory as [g][code][64] uint32 (columns
// stored twice, as in adc_scan_rot_kernel). Survivors are appended to cand[q * cap + pos].
// ABS = true: the table starts at an ABSOLUTE shared-window address that is a multiple of 64 KiB,
// so the PRMT that merges (code << 8) with (lane << 2) also supplies the upper address bytes and
// its result IS the LDS address (no per-lookup base add). That costs up to 64 KiB of padding, so
// the CTA is 1024 threads (one per SM) instead of three CTAs of 256.
// GW = 16: the half group of M = 16 (one 128-bit code load and 16 lookups per row; table columns j = m, m + 16, m + 32).
template <int G, bool ABS, int GW = 32>
__global__ void __launch_bounds__(ABS ? 1024 : SCAN_THREADS)
adc_screen_kernel(const uint8_t* __restrict__ codes_rot, int64_t n, const uint32_t* __restrict__ qlut4,
                  const int* __restrict__ thr, int64_t chunk_rows, uint32_t* __restrict__ cand,
                  uint32_t* __restrict__ count, uint32_t cap) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int KSUB = 256;
    constexpr int M = G * GW;
    constexpr int NQW = GW / 4;          // column quads (= code words) of a group
    constexpr int NV = GW / 16;          // 128-bit code loads of a group
    static_assert(GW == 32 || (GW == 16 && G == 1), "group width 32, or one half group of 16");
    constexpr int THREADS = ABS ? 1024 : SCAN_THREADS;
    const int grp = blockIdx.x, chunk = blockIdx.y;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t base32 = (uint32_t)__cvta_generic_to_shared(smem_raw);
    const uint32_t tab32 = ABS ? ((base32 + 0xFFFFu) & ~0xFFFFu) : base32;
    uint32_t* tab = reinterpret_cast<uint32_t*>(smem_raw + (tab32 - base32));
    const uint32_t* src_q = qlut4 + (size_t)grp * M * KSUB;
    for (int idx = threadIdx.x; idx < G * KSUB * NQW; idx += THREADS) {
        const int jq = idx % NQW, kk = (idx / NQW) & (KSUB - 1), g = idx / (NQW * KSUB);
        const uint32_t* src = src_q + (size_t)(g * GW + 4 * jq) * KSUB + kk;
        uint4 v = make_uint4(src[0], src[KSUB], src[2 * KSUB], src[3 * KSUB]);
        uint4* dst = reinterpret_cast<uint4*>(tab + ((size_t)g * KSUB + kk) * 64 + 4 * jq);
#pragma unroll
        for (int rep = 0; rep * GW < 32 + GW; ++rep) dst[rep * NQW] = v;
    }
    const int t0 = thr[grp * 4 + 0], t1 = thr[grp * 4 + 1], t2 = thr[grp * 4 + 2], t3 = thr[grp * 4 + 3];
    const uint32_t one = cap ? 1u : 0u;      // runtime 1 (see ScreenWords)
    // survivors are collected per CTA in shared memory (SCREEN_LCAP per query) and appended to the global lists with
    // one atomic per query at the end: a global atomicAdd with a return value inside the scan loop stalled the warp
    // for a memory round trip whenever one of its 128 (row, query) pairs survived (40 % of the iterations at C1)
    __shared__ uint32_t s_cnt[4], s_base[4];
    uint32_t* lbuf = tab + (size_t)G * KSUB * 64;          // [4][SCREEN_LCAP]
    if (threadIdx.x < 4) s_cnt[threadIdx.x] = 0u;
    __syncthreads();

    // second PRMT operand: byte 0 = lane * 4, bytes 2..3 = upper bytes of the table address (ABS)
    const uint32_t lane4 = ((uint32_t)lane << 2) | (ABS ? (tab32 & 0xFFFF0000u) : 0u);
    const int64_t r0 = (int64_t)chunk * chunk_rows;  // multiple of 32
    int64_t r1 = r0 + chunk_rows;
    if (r1 > n) r1 = n;
    constexpr int UNR = GW == 32 ? 2 : 4;
    for (int64_t base = r0 + (int64_t)w * 32; base < r1; base += UNR * THREADS) {
        uint4 cv[UNR][NV * G];
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
            int64_t bb = base + (int64_t)u * THREADS;
            if (bb >= r1) bb = base;
            const uint4* cb = reinterpret_cast<const uint4*>(codes_rot + (bb >> 5) * (int64_t)G * (32 * GW));
#pragma unroll
            for (int g = 0; g < G; ++g)
#pragma unroll
                for (int h = 0; h < NV; ++h) cv[u][NV * g + h] = ld_stream_u4(cb + g * 64 + h * 32 + lane);
        }
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
            const int64_t row = base + (int64_t)u * THREADS + lane;
            uint32_t lo16 = 0, hi16 = 0;   // 16-bit fields: (query 0, query 2) and (query 1, query 3)
            if (ABS) {
                uint32_t wds[NQW * G];
#pragma unroll
                for (int g = 0; g < G; ++g)
#pragma unroll
                    for (int h = 0; h < NV; ++h) {
                        wds[NQW * g + 4 * h + 0] = cv[u][NV * g + h].x; wds[NQW * g + 4 * h + 1] = cv[u][NV * g + h].y;
                        wds[NQW * g + 4 * h + 2] = cv[u][NV * g + h].z; wds[NQW * g + 4 * h + 3] = cv[u][NV * g + h].w;
                    }
                ScreenWords<NQW * G, 0>::run(wds, lane4, lo16, hi16, one);
                lo16 -= hi16 << 8;        // whole-word sums -> even-byte sums (see ScreenWords)
            } else {
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
                        uint32_t t4 = 0;
#pragma unroll
                        for (int b = 0; b < 4; ++b) {
                            const uint32_t off = __byte_perm(wds[j], lane4, 0x5504u | ((uint32_t)b << 4));
                            t4 += *reinterpret_cast<const uint32_t*>(smem_raw + off + g * (KSUB * 64 * 4) + (4 * j + b) * 4);
                        }
                        lo16 = imad_acc(t4, one, lo16);                                    // whole words (see ScreenWords)
                        hi16 = imad_acc(__byte_perm(t4, 0u, 0x4341u), one, hi16);
                    }
                }
                lo16 -= hi16 << 8;
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
                        if (lane == 0) pos = atomicAdd(&s_cnt[qq], (uint32_t)__popc(mk));
                        pos = __shfl_sync(FULL_MASK, pos, 0) + __popc(mk & ((1u << lane) - 1u));
                        if (ps[qq]) {
                            if (pos < (uint32_t)SCREEN_LCAP) {
                                lbuf[qq * SCREEN_LCAP + pos] = (uint32_t)row;
                            } else {        // local buffer full (dense survivors): straight to the global list
                                const uint32_t gp = atomicAdd(count + grp * 4 + qq, 1u);
                                if (gp < cap) cand[(size_t)(grp * 4 + qq) * cap + gp] = (uint32_t)row;
                            }
                        }
                    }
                }
            }
        }
    }
    __syncthreads();
    if (threadIdx.x < 4) {
        const uint32_t nloc = min(s_cnt[threadIdx.x], (uint32_t)SCREEN_LCAP);
        s_base[threadIdx.x] = nloc ? atomicAdd(count + grp * 4 + threadIdx.x, nloc) : 0u;
    }
    __syncthreads();
#pragma unroll
    for (int qq = 0; qq < 4; ++qq) {
        const uint32_t nloc = min(s_cnt[qq], (uint32_t)SCREEN_LCAP), gb = s_base[qq];
        for (uint32_t i2 = threadIdx.x; i2 < nloc; i2 += THREADS)
            if (gb + i2 < cap) cand[(size_t)(grp * 4 + qq) * cap + gb + i2] = lbuf[qq * SCREEN_LCAP + i2];
    }
}

// Rotate the 4 * NW bytes held in w[0 .. NW) left by r byte positions towards lower indices: byte j of the result is
// byte (j + r) % (4 NW) of the input (a barrel shifter: funnel shifts for r % 4, conditional word moves for r / 4).
template <int NW>
__device__ __forceinline__ void rotate_bytes_down(uint32_t (&w)[NW], uint32_t r) {
    const uint32_t bits = (r & 3u) * 8u;
    uint32_t t[NW];
#pragma unroll
    for (int j = 0; j < NW; ++j) t[j] = __funnelshift_r(w[j], w[(j + 1) % NW], bits);
#pragma unroll
    for (int sh = 1; sh < NW; sh <<= 1) {
        const bool on = ((r >> 2) & (uint32_t)sh) != 0;
        uint32_t u[NW];
#pragma unroll
        for (int j = 0; j < NW; ++j) u[j] = on ? t[(j + sh) % NW] : t[j];
#pragma unroll
        for (int j = 0; j < NW; ++j) t[j] = u[j];
    }
#pragma unroll
    for (int j = 0; j < NW; ++j) w[j] = t[j];
}

// Exact fp32 score of one row from the scan layout: the row's codes are one (M = 16) or two 128-bit loads per group,
// rotated so that byte m is sub-quantizer m (lane i stores sub-quantizer (i + s) % GW at step s); the table entries
// come from the query's table T (shared memory) and are summed in increasing m, as the row kernel does.
__device__ __forceinline__ float screen_exact_score(const uint8_t* __restrict__ codes_rot, int M, const float* T,
                                                    uint32_t row) {
    const uint32_t i = row & 31u;
    float acc = 0.f;
    if (M == 16) {      // half group: 512-byte blocks
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(codes_rot + (size_t)(row >> 5) * 512 + (i << 4)));
        uint32_t w[4] = {v.x, v.y, v.z, v.w};
        rotate_bytes_down<4>(w, (16u - (i & 15u)) & 15u);     // byte m <- step (m - i) % 16
#pragma unroll
        for (int m = 0; m < 16; ++m) acc += T[m * 256 + ((w[m >> 2] >> ((m & 3) * 8)) & 0xFFu)];
        return acc;
    }
    const int G = M >> 5;
    const uint8_t* blk = codes_rot + (size_t)(row >> 5) * G * 1024 + (i << 4);
    for (int g = 0; g < G; ++g) {
        const uint4 v0 = __ldg(reinterpret_cast<const uint4*>(blk + (size_t)g * 1024));
        const uint4 v1 = __ldg(reinterpret_cast<const uint4*>(blk + (size_t)g * 1024 + 512));
        uint32_t w[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
        rotate_bytes_down<8>(w, (32u - i) & 31u);             // byte m <- step (m - i) % 32
        const float* Tg = T + (size_t)g * 32 * 256;
#pragma unroll
        for (int m = 0; m < 32; ++m) acc += Tg[m * 256 + ((w[m >> 2] >> ((m & 3) * 8)) & 0xFFu)];
    }
    return acc;
}

// One CTA per query: exact fp32 score of every listed row, final (value, id)-ordered k results. Lists of at most
// sel_max rows (the usual case) are scored into shared memory as (value, id) keys; the k-th smallest key is found by
// a radix select over the 64-bit keys (8-bit digits, most significant first, stopping as soon as the digit's bin is
// taken whole), the k keys up to it are compacted and only those are sorted. A full bitonic sort moved every key
// through shared memory at every stage (65 % of the kernel at 400 rows per query); longer lists go through the warps'
// running top-k.
constexpr int FINISH_SEL_MAX = 8192;
template <int CAP>
__global__ void __launch_bounds__(SCAN_THREADS)
screen_finish_kernel(const uint8_t* __restrict__ codes_rot, int M, const float* __restrict__ lut,
                     const uint32_t* __restrict__ cand, const uint32_t* __restrict__ count, uint32_t cap, int k,
                     int negate, uint32_t sel_max, float* __restrict__ out_vals, uint32_t* __restrict__ out_ids) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int q = blockIdx.x;
    const uint32_t cnt = count[q];
    if (cnt > cap) return;     // overflow: the gated exact kernels produce this query's result
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    // the query's table in shared memory: M KB read once, coalesced (gathering it from L2 entry by entry cost a
    // 32-byte sector per look-up: 13x the bytes at 400 rows per query)
    float* T = reinterpret_cast<float*>(smem_raw);
    {
        const float4* src = reinterpret_cast<const float4*>(lut + (size_t)q * M * 256);
        float4* dst = reinterpret_cast<float4*>(T);
        for (int idx = threadIdx.x; idx < M * 64; idx += SCAN_THREADS) dst[idx] = __ldg(src + idx);
    }
    __syncthreads();
    unsigned char* smem_top = smem_raw + (size_t)M * 1024;
    const uint32_t* list = cand + (size_t)q * cap;
    const uint64_t* best;
    __shared__ uint64_t best_s[128];
    if (cnt <= sel_max) {
        uint64_t* keys = reinterpret_cast<uint64_t*>(smem_top);
        __shared__ uint32_t hist[256];
        __shared__ uint64_t sel_prefix;
        __shared__ uint32_t sel_k, sel_done, out_n;
        // two list entries per thread and pass: the gathers of both rows are in flight together
        for (uint32_t e = threadIdx.x; e < cnt; e += 2 * SCAN_THREADS) {
            const uint32_t e1 = e + SCAN_THREADS;
            const bool v1 = e1 < cnt;
            const uint32_t row0 = list[e], row1 = v1 ? list[e1] : 0u;
            const float acc0 = screen_exact_score(codes_rot, M, T, row0);
            float acc1 = 0.f;
            if (v1) acc1 = screen_exact_score(codes_rot, M, T, row1);
            keys[e] = make_key(negate ? -acc0 : acc0, row0);
            if (v1) keys[e1] = make_key(negate ? -acc1 : acc1, row1);
        }
        const uint32_t want = cnt < (uint32_t)k ? cnt : (uint32_t)k;
        if (threadIdx.x == 0) {
            sel_prefix = 0ull;
            sel_k = want;
            sel_done = want == cnt ? 1u : 0u;      // everything is kept: no selection needed
            out_n = 0u;
        }
        for (int i = threadIdx.x; i < 128; i += SCAN_THREADS) best_s[i] = KEY_INF;
        __syncthreads();
        if (sel_done) {
            if (threadIdx.x == 0) sel_prefix = KEY_INF;
        } else {
            for (int pass = 0; pass < 8; ++pass) {
                const int shift = 56 - 8 * pass;
                hist[threadIdx.x] = 0;
                __syncthreads();
                const uint64_t prefix = sel_prefix;
                const uint32_t kk = sel_k;
                for (uint32_t e = threadIdx.x; e < cnt; e += SCAN_THREADS) {
                    const uint64_t key = keys[e];
                    if (pass == 0 || ((key ^ prefix) >> (shift + 8)) == 0ull)
                        atomicAdd(&hist[(uint32_t)(key >> shift) & 255u], 1u);
                }
                __syncthreads();
                if (w == 0) {
                    uint32_t c[8], tot = 0;
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        c[j] = hist[lane * 8 + j];
                        tot += c[j];
                    }
                    uint32_t incl = tot;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) {
                        const uint32_t t = __shfl_up_sync(FULL_MASK, incl, o);
                        if (lane >= o) incl += t;
                    }
                    const uint32_t excl = incl - tot;
                    if (excl < kk && kk <= incl) {           // exactly one lane: its 8 bins hold the k-th key
                        uint32_t run = excl;
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            if (run < kk && kk <= run + c[j]) {
                                uint64_t np = prefix | ((uint64_t)(lane * 8 + j) << shift);
                                if (kk - run == c[j] || pass == 7) {       // the bin is taken whole: done
                                    np |= shift ? ((1ull << shift) - 1ull) : 0ull;
                                    sel_done = 1u;
                                }
                                sel_prefix = np;
                                sel_k = kk - run;
                            }
                            run += c[j];
                        }
                    }
                }
                __syncthreads();
                if (sel_done) break;
            }
        }
        __syncthreads();
        {
            const uint64_t kth = sel_prefix;           // every key <= kth belongs to the result: exactly `want` keys
            for (uint32_t e = threadIdx.x; e < cnt; e += SCAN_THREADS) {
                const uint64_t key = keys[e];
                if (key <= kth) {
                    const uint32_t pos = atomicAdd(&out_n, 1u);
                    if (pos < 128u) best_s[pos] = key;
                }
            }
        }
        block_bitonic_sort(best_s, next_pow2(k < 2 ? 2 : k));
        best = best_s;
    } else {
        uint64_t* wbuf = reinterpret_cast<uint64_t*>(smem_top);
        const int n2 = next_pow2(SCAN_WARPS * k);
        uint64_t* mbuf = wbuf + SCAN_WARPS * CAP;
        WarpTopK<CAP> tk;
        tk.init(wbuf + w * CAP, k);
        const uint32_t rounds = (cnt + SCAN_THREADS - 1) / SCAN_THREADS;
        for (uint32_t it = 0; it < rounds; ++it) {
            const uint32_t e = it * SCAN_THREADS + threadIdx.x;
            const bool valid = e < cnt;
            float acc = 0.f;
            uint32_t row = 0;
            if (valid) {
                row = list[e];
                acc = screen_exact_score(codes_rot, M, T, row);
            }
            tk.offer(valid, negate ? -acc : acc, row);
        }
        block_emit_topk<CAP>(tk, mbuf, n2, k, best_s);
        __syncthreads();
        best = best_s;
    }
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
