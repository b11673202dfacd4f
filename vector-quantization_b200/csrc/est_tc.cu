// Code-domain estimator scan + top-k on the sm_100a tensor cores, straight from the packed codes.
//
// Serves the two estimators of the reference that score a query against bit-packed per-dimension codes:
//   * the SAQ engine's accurate estimator (external/saq/include/saq/caq_estimator.h:254-316 with the
//     per-segment combination of saq_estimator.h:27-129,183-190; GPU twin src/gpu/gpu_search.cu:37-99):
//       dist = sum_s |o_s|^2 + |q_s|^2 - 2 rescale_s (ip_s Delta_s + (-1 + Delta_s / 2) sum(q_s))
//     which, with u = code - (2^b - 1) / 2, is  sum_s |o_s|^2 + |q|^2 - 2 <x_i, q>,  x_is = rescale_is Delta_s u_is;
//   * extended RaBitQ searched through FlatQuantizedIndex (src/haag_vq/methods/extended_rabitq.py:173-198 decoded,
//     then methods/search/flat_quantized_index.py:57-68): |q - x^|^2 = qc + a^2 sum(l^2) - 2 a <l, q_rot>,
//     a = nrm t / sqrt(D), l = levels[idx]  (b200vq_erabitq_scan_topk computes the same form).
// Both are  key(q, i) = C_q + B_i - 2 <x_i, q>  (L2 form) or  -(C_q + <x_i, q>)  (inner-product form) over a
// VIRTUAL row x_i = concat_s rowscale_is * value_s(code), value_s = a 2^b-entry table or code + base.
//
// The decoded n x D matrix never exists in HBM (round 1 wrote it: 6.1 GB per million codes at D = 1536).
// Per (128-row, 256-query) unit and 32-dimension chunk, converter warps read 4 b bytes per row from the
// packed MSB-first bit string, expand them to 32 fp32 values and store them at the SWIZZLE_128B positions of
// the K-major A tile; TMA brings the matching query tile; one thread issues tcgen05.mma kind::tf32
// (128 x 256 x 8) into one of two TMEM accumulators; eight epilogue warps test the rigorous lower bound
// a P + b_i <= T_q (flat_tc.cu's error model: both operands truncated to TF32) and append survivors to
// per-query lists. est_finish_kernel scores survivors with the CANONICAL fp32 arithmetic (one sequential
// FMA chain per segment in dimension order, then fma(rowscale, ip, acc) in segment order - for one segment
// exactly b200vq_erabitq_scan_topk's arithmetic), merges them into the running top-k and tightens T_q for
// the next, three times larger, row stage. A query whose list overflows is redone exactly over all rows
// (gated launch of the finish kernel on the identity list), so the result never depends on the screen.
#include "tc_common.cuh"
#include "topk.cuh"

namespace b200vq {
namespace est {
using namespace tc;

constexpr int TILE_ROWS = 128;
constexpr int TILE_Q = 256;
constexpr int CH = 32;                       // K chunk: 32 floats = one 128-byte swizzle row
constexpr int NST = 4;
constexpr int CONV_WARPS = 8;                // two groups of four warps: group g converts the chunks with (index & 1) == g
constexpr int EPI_WARPS = 8;                 // 2 accumulator stages x 4 TMEM lane quarters
constexpr int THREADS = 64 + 32 * (CONV_WARPS + EPI_WARPS);
constexpr uint32_t A_BYTES = TILE_ROWS * CH * 4;
constexpr uint32_t B_BYTES = TILE_Q * CH * 4;
constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
constexpr uint32_t BAR_OFF = NST * STAGE_BYTES;
constexpr uint32_t THR_OFF = BAR_OFF + 256;
constexpr uint32_t MAP_OFF = THR_OFF + 2 * TILE_Q * 4;        // chunk -> segment (uint8[MAX_CHUNKS])
constexpr int MAX_CHUNKS = 512;                                // D <= 16384
constexpr uint32_t TAB_OFF = MAP_OFF + MAX_CHUNKS;
constexpr int MAX_SEG = 32;
constexpr int MAX_TAB = 2048;                                  // table floats over all segments
constexpr int MAX_BITS = 13;
enum { B_FULL = 0, B_CONV = NST, B_EMPTY = 2 * NST, B_TFULL = 3 * NST, B_TEMPTY = 3 * NST + 2, B_TMEM = 3 * NST + 4 };

struct Seg {
    const uint8_t* codes;     // (n, stride) packed MSB-first bit string of this segment
    const float* rowscale;    // (n) or null (= 1)
    int64_t stride;           // bytes between rows
    int chunk0, nchunk;       // position / length in 32-dimension chunks of the concatenated row
    int bits;                 // 1..13
    int tab_off;              // offset of the 2^bits value table in `tables`, or -1: value = code + base
    float base;
};
struct SegTable {
    int nseg;
    Seg s[MAX_SEG];
};

struct Params {
    const float* norm2;       // (n) |x_i|^2 of the virtual rows (bounds the screen error)
    const float* bias;        // (n) B_i
    const float* thr;         // (ntq * 256) T_q, NaN for padding queries
    const float* qstats;      // [0] = max_q |q|^2, [1] = max_q |C_q|
    const float* tables;      // concatenated value tables
    uint32_t* cand;           // (nq, cap)
    float* cand_lb;           // (nq, cap) the survivors' lower bounds
    uint32_t* count;          // (nq)
    uint32_t cap;
    int dense;                // stage 0: every row is a survivor, slot = row - r0 (no counters, no atomics)
    int64_t r0, r1;           // rows of this stage
    int nq, nchunk, ntq, ntab;
    int64_t ntr;
    int metric;
    float ebound;
};

// ------------------------------------------------------------------ decoding ---
// One chunk = 32 fields of B bits = B 32-bit words of the big-endian bit stream.
template <int B>
__device__ __forceinline__ void load_words_raw(const uint8_t* p, uint32_t (&w)[MAX_BITS]) {
    const uint32_t* q = reinterpret_cast<const uint32_t*>(p);
#pragma unroll
    for (int i = 0; i < B; ++i) w[i] = __ldg(q + i);   // no use of the value here: the load stays in flight
}
template <int B>
__device__ __forceinline__ void swap_words(uint32_t (&w)[MAX_BITS]) {
#pragma unroll
    for (int i = 0; i < B; ++i) w[i] = __byte_perm(w[i], 0, 0x0123);   // big-endian bit stream
}
template <int B>
__device__ __forceinline__ void load_words(const uint8_t* p, uint32_t (&w)[MAX_BITS]) {
    load_words_raw<B>(p, w);
    swap_words<B>(w);
}
template <int B>
__device__ __forceinline__ void zero_words(uint32_t (&w)[MAX_BITS]) {
#pragma unroll
    for (int i = 0; i < B; ++i) w[i] = 0;
}
template <int B>
__device__ __forceinline__ uint32_t field(const uint32_t (&w)[MAX_BITS], int j) {
    const int bit = j * B, wi = bit >> 5, off = bit & 31;
    if (off + B <= 32) return (w[wi] >> (32 - off - B)) & ((1u << B) - 1u);
    return __funnelshift_l(w[wi + 1 < B ? wi + 1 : wi], w[wi], off) >> (32 - B);
}
template <int B, bool TAB>
__device__ __forceinline__ float value(const uint32_t (&w)[MAX_BITS], int j, const float* tab, float base) {
    const uint32_t c = field<B>(w, j);
    return TAB ? tab[c] : __fadd_rn((float)c, base);
}

// A-tile row of one chunk: 32 values * rowscale at the SWIZZLE_128B positions of row r.
template <int B, bool TAB>
__device__ __forceinline__ void convert_chunk(const uint32_t (&w)[MAX_BITS], float rs, const float* tab, float base,
                                              uint8_t* a_row, int r7) {
#pragma unroll
    for (int j4 = 0; j4 < 8; ++j4) {
        float4 o;
        o.x = __fmul_rn(rs, value<B, TAB>(w, 4 * j4 + 0, tab, base));
        o.y = __fmul_rn(rs, value<B, TAB>(w, 4 * j4 + 1, tab, base));
        o.z = __fmul_rn(rs, value<B, TAB>(w, 4 * j4 + 2, tab, base));
        o.w = __fmul_rn(rs, value<B, TAB>(w, 4 * j4 + 3, tab, base));
        *reinterpret_cast<float4*>(a_row + ((j4 ^ r7) << 4)) = o;
    }
}
// canonical chain: ip = fma(q_d, value_d, ip), d ascending
template <int B, bool TAB>
__device__ __forceinline__ float chain_chunk(const uint32_t (&w)[MAX_BITS], const float* tab, float base, const float* q,
                                             float ip) {
#pragma unroll
    for (int j = 0; j < 32; ++j) ip = __fmaf_rn(q[j], value<B, TAB>(w, j, tab, base), ip);
    return ip;
}
template <int B, bool TAB>
__device__ __forceinline__ float sq_chunk(const uint32_t (&w)[MAX_BITS], const float* tab, float base, float l2) {
#pragma unroll
    for (int j = 0; j < 32; ++j) {
        const float v = value<B, TAB>(w, j, tab, base);
        l2 = __fmaf_rn(v, v, l2);
    }
    return l2;
}

#define EST_SWITCH_BITS(bits, STMT)                 \
    switch (bits) {                                 \
        case 1: { constexpr int BB = 1; STMT; } break;   \
        case 2: { constexpr int BB = 2; STMT; } break;   \
        case 3: { constexpr int BB = 3; STMT; } break;   \
        case 4: { constexpr int BB = 4; STMT; } break;   \
        case 5: { constexpr int BB = 5; STMT; } break;   \
        case 6: { constexpr int BB = 6; STMT; } break;   \
        case 7: { constexpr int BB = 7; STMT; } break;   \
        case 8: { constexpr int BB = 8; STMT; } break;   \
        case 9: { constexpr int BB = 9; STMT; } break;   \
        case 10: { constexpr int BB = 10; STMT; } break; \
        case 11: { constexpr int BB = 11; STMT; } break; \
        case 12: { constexpr int BB = 12; STMT; } break; \
        default: { constexpr int BB = 13; STMT; } break; \
    }

// ------------------------------------------------------------- screen kernel ---
// FB > 0: every row is ONE segment of FB-bit codes with a value table (extended RaBitQ): the converters run
// without per-chunk dispatch and keep two chunks of code words in flight. FB == 0: any segment list.
template <int FB>
__global__ void __launch_bounds__(THREADS, 1)
est_screen_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ SegTable segs, const Params p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;
    uint8_t* gbase = smem_raw + (base - raw);
    const uint32_t bar0 = base + BAR_OFF;
    auto bar = [&](int slot) { return bar0 + 8u * (uint32_t)slot; };
    volatile uint32_t* tmem_ptr_smem = reinterpret_cast<volatile uint32_t*>(gbase + BAR_OFF + 8 * B_TMEM);
    float* thr_s = reinterpret_cast<float*>(gbase + THR_OFF);    // [2][TILE_Q]
    uint8_t* seg_of = gbase + MAP_OFF;                           // [nchunk]
    float* tab_s = reinterpret_cast<float*>(gbase + TAB_OFF);    // [ntab]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t nunits = p.ntr * p.ntq;

    for (int i = threadIdx.x; i < p.ntab; i += THREADS) tab_s[i] = p.tables[i];
    for (int i = threadIdx.x; i < p.nchunk; i += THREADS) {
        int s = 0;
        while (s + 1 < segs.nseg && i >= segs.s[s + 1].chunk0) ++s;
        seg_of[i] = (uint8_t)s;
    }
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&map_q);
        for (int i = 0; i < NST; ++i) {
            mbar_init(bar(B_FULL + i), 1);
            mbar_init(bar(B_CONV + i), CONV_WARPS / 2);   // one arrival per warp of the group that owns the chunk
            mbar_init(bar(B_EMPTY + i), 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(bar(B_TFULL + i), 1);
            mbar_init(bar(B_TEMPTY + i), 4);
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(bar(B_TMEM), 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;

    if (warp == 0) {
        // ===================================================== TMA producer (query tiles) ===
        if (lane == 0) {
            uint32_t st = 0, phase = 0;
            for (int64_t unit = blockIdx.x; unit < nunits; unit += gridDim.x) {
                const int u = (int)(unit % p.ntq);
                for (int c = 0; c < p.nchunk; ++c) {
                    mbar_wait_backoff(bar(B_EMPTY + st), phase ^ 1, 256);
                    mbar_expect_tx(bar(B_FULL + st), B_BYTES);
                    tma_load_2d(base + st * STAGE_BYTES + A_BYTES, &map_q, bar(B_FULL + st), c * CH, u * TILE_Q);
                    if (++st == NST) { st = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ======================================================= MMA issuer ===
        if (lane == 0) {
            uint32_t st = 0, phase = 0, acc = 0, t_phase = 0;
            for (int64_t unit = blockIdx.x; unit < nunits; unit += gridDim.x) {
                mbar_wait_backoff(bar(B_TEMPTY + acc), t_phase ^ 1, 32);
                tc_fence_after();
                const uint32_t d = tmem_base + acc * 256u;
                uint32_t accum = 0;
                for (int c = 0; c < p.nchunk; ++c) {
                    mbar_wait_backoff(bar(B_FULL + st), phase, 32);
                    mbar_wait_backoff(bar(B_CONV + st), phase, 32);
                    tc_fence_after();
                    const uint32_t a_addr = base + st * STAGE_BYTES, b_addr = a_addr + A_BYTES;
#pragma unroll
                    for (int kk = 0; kk < CH / 8; ++kk) {
                        umma_tf32(d, make_kmajor_desc(a_addr + kk * 32, 1024, 2), make_kmajor_desc(b_addr + kk * 32, 1024, 2),
                                  IDESC_TF32_128x256, accum);
                        accum = 1;
                    }
                    umma_commit(bar(B_EMPTY + st));
                    if (++st == NST) { st = 0; phase ^= 1; }
                }
                umma_commit(bar(B_TFULL + acc));
                if (++acc == 2) { acc = 0; t_phase ^= 1; }
            }
        }
    } else if (warp < 2 + CONV_WARPS) {
        // ======================================================= converters ===
        // thread = tile row; group g = (warp - 2) / 4 converts the chunks whose running index has parity g.
        // The words of the group's NEXT chunk are loaded before the current one is converted.
        const int cw = warp - 2;
        const int grp = cw >> 2;
        const int r = (cw & 3) * 32 + lane;
        const int r7 = r & 7;
        uint8_t* a_row0 = gbase + (uint32_t)(r >> 3) * 1024u + (uint32_t)r7 * 128u;
        // (unit, row tile, chunk) cursors advance without divisions: a unit step adds gridDim.x units
        const int dq = (int)(gridDim.x / (unsigned)p.ntq), dr = (int)(gridDim.x % (unsigned)p.ntq);
        struct Cursor {
            int64_t unit, t;
            int u, c;
        };
        auto advance = [&](Cursor& k, int by) {
            k.c += by;
            while (k.c >= p.nchunk) {
                k.c -= p.nchunk;
                k.unit += gridDim.x;
                k.t += dq;
                k.u += dr;
                if (k.u >= p.ntq) { k.u -= p.ntq; ++k.t; }
            }
        };
        Cursor kc, kf;      // chunk being converted / chunk being fetched
        kc.unit = blockIdx.x; kc.t = blockIdx.x / p.ntq; kc.u = (int)(blockIdx.x % p.ntq); kc.c = 0;
        advance(kc, grp);
        kf = kc;
        if constexpr (FB > 0) {
            const uint8_t* codes0 = segs.s[0].codes;
            const float* rowscale0 = segs.s[0].rowscale;
            const int64_t stride0 = segs.s[0].stride;
            uint32_t w0[MAX_BITS], w1[MAX_BITS], w2[MAX_BITS];
            float s0 = 0.f, s1 = 0.f, s2 = 0.f;
            // raw words only: the byte swap happens at conversion time, so nothing waits on a load here
            auto fetch1 = [&](const Cursor& k, uint32_t (&w)[MAX_BITS], float& rs) {
                const int64_t row = p.r0 + k.t * TILE_ROWS + r;
                if (k.unit < nunits && row < p.r1) {
                    load_words_raw<FB>(codes0 + row * stride0 + (int64_t)k.c * (4 * FB), w);
                    rs = rowscale0 ? __ldg(rowscale0 + row) : 1.f;
                } else {
                    zero_words<FB>(w);
                    rs = 0.f;
                }
            };
            fetch1(kf, w1, s1);
            advance(kf, 2);
            fetch1(kf, w2, s2);
            for (uint32_t g = (uint32_t)grp; kc.unit < nunits; g += 2) {
#pragma unroll
                for (int i = 0; i < FB; ++i) { w0[i] = w1[i]; w1[i] = w2[i]; }
                s0 = s1; s1 = s2;
                advance(kf, 2);
                fetch1(kf, w2, s2);
                const uint32_t st = g & (NST - 1), phase = (g / NST) & 1u;
                mbar_wait(bar(B_EMPTY + st), phase ^ 1);
                swap_words<FB>(w0);
                convert_chunk<FB, true>(w0, s0, tab_s, 0.f, a_row0 + st * STAGE_BYTES, r7);
                fence_proxy_async();
                __syncwarp();
                if (lane == 0) mbar_arrive(bar(B_CONV + st));
                advance(kc, 2);
            }
        } else {
            uint32_t cur[MAX_BITS], nxt[MAX_BITS];
            float rs_cur = 0.f, rs_nxt = 0.f;
            auto fetch = [&](const Cursor& k, uint32_t (&w)[MAX_BITS], float& rs) {
                const int64_t row = p.r0 + k.t * TILE_ROWS + r;
                const Seg& sg = segs.s[seg_of[k.c]];
                if (k.unit < nunits && row < p.r1) {
                    const uint8_t* src = sg.codes + row * sg.stride + (int64_t)(k.c - sg.chunk0) * (4 * sg.bits);
                    EST_SWITCH_BITS(sg.bits, load_words_raw<BB>(src, w));
                    rs = sg.rowscale ? __ldg(sg.rowscale + row) : 1.f;
                } else {
                    EST_SWITCH_BITS(sg.bits, zero_words<BB>(w));
                    rs = 0.f;
                }
            };
            fetch(kf, nxt, rs_nxt);
            for (uint32_t g = (uint32_t)grp; kc.unit < nunits; g += 2) {
#pragma unroll
                for (int i = 0; i < MAX_BITS; ++i) cur[i] = nxt[i];
                rs_cur = rs_nxt;
                advance(kf, 2);
                fetch(kf, nxt, rs_nxt);
                const uint32_t st = g & (NST - 1), phase = (g / NST) & 1u;
                const Seg& sg = segs.s[seg_of[kc.c]];
                mbar_wait(bar(B_EMPTY + st), phase ^ 1);
                uint8_t* a_row = a_row0 + st * STAGE_BYTES;
                if (sg.tab_off >= 0) {
                    const float* tab = tab_s + sg.tab_off;
                    EST_SWITCH_BITS(sg.bits, (swap_words<(BB <= 8 ? BB : 8)>(cur),
                                              convert_chunk<(BB <= 8 ? BB : 8), true>(cur, rs_cur, tab, 0.f, a_row, r7)));
                } else {
                    EST_SWITCH_BITS(sg.bits, (swap_words<BB>(cur),
                                              convert_chunk<BB, false>(cur, rs_cur, nullptr, sg.base, a_row, r7)));
                }
                fence_proxy_async();
                __syncwarp();
                if (lane == 0) mbar_arrive(bar(B_CONV + st));
                advance(kc, 2);
            }
        }
    } else {
        // ========================================================= epilogue ===
        const int ew = warp - 2 - CONV_WARPS;                 // 0..7
        const uint32_t acc = (uint32_t)ew >> 2;               // accumulator stage served by this warp
        const int q4 = warp & 3;                              // TMEM lane quarter this warp may read
        const uint32_t lane_addr = tmem_base + ((uint32_t)(q4 * 32) << 16) + acc * 256u;
        float* ts = thr_s + acc * TILE_Q;
        const int gt = (ew & 3) * 32 + lane;                  // 0..127 inside the group
        const uint32_t grp_bar = 1u + acc;
        const float qn_max = p.qstats[0], qc_max = p.qstats[1], q_max = sqrtf(qn_max) * 1.0000005f;
        uint32_t t_phase = 0;
        int64_t unit = blockIdx.x;
        for (uint32_t it = 0; unit < nunits; unit += gridDim.x, ++it) {
            if ((it & 1u) != acc) continue;
            const int64_t t = unit / p.ntq;
            const int u = (int)(unit - t * p.ntq);
            const int64_t row = p.r0 + t * TILE_ROWS + q4 * 32 + lane;
            const bool live = row < p.r1;
            float a = 0.f, b = 0.f;
            if (live) {
                const float xn = p.norm2[row];
                const float bi = p.bias[row];
                const float nrm = sqrtf(xn) * 1.0000005f;
                const float E = p.ebound * nrm * q_max + 2.4e-7f * (xn + fabsf(bi) + qn_max + qc_max);
                if (p.metric == B200VQ_METRIC_L2) {
                    a = -2.f * 1.0009765625f;
                    b = bi - 2.f * E;
                } else {
                    a = -1.0009765625f;
                    b = -E;
                }
            }
            asm volatile("bar.sync %0, 128;" ::"r"(grp_bar) : "memory");   // previous unit's thresholds consumed
            ts[gt] = p.thr[(size_t)u * TILE_Q + gt];
            ts[gt + 128] = p.thr[(size_t)u * TILE_Q + gt + 128];
            asm volatile("bar.sync %0, 128;" ::"r"(grp_bar) : "memory");
            mbar_wait(bar(B_TFULL + acc), t_phase);
            tc_fence_after();
            float v[32];
#pragma unroll 1
            for (int cb = 0; cb < 8; ++cb) {
                tmem_ld32(lane_addr + cb * 32, v);
                const float4* t4 = reinterpret_cast<const float4*>(ts + cb * 32);
#pragma unroll
                for (int j4 = 0; j4 < 8; ++j4) {
                    const float4 tt = t4[j4];
                    const float th[4] = {tt.x, tt.y, tt.z, tt.w};
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float lb = __fmaf_rn(a, v[4 * j4 + e], b);
                        if (p.dense) {
                            const int q = u * TILE_Q + cb * 32 + 4 * j4 + e;
                            if (live && q < p.nq) {
                                p.cand[(size_t)q * p.cap + (uint32_t)(row - p.r0)] = (uint32_t)row;
                                p.cand_lb[(size_t)q * p.cap + (uint32_t)(row - p.r0)] = lb;
                            }
                        } else if (live && lb <= th[e]) {
                            const int q = u * TILE_Q + cb * 32 + 4 * j4 + e;
                            const uint32_t pos = atomicAdd(p.count + q, 1u);
                            if (pos < p.cap) {
                                p.cand[(size_t)q * p.cap + pos] = (uint32_t)row;
                                p.cand_lb[(size_t)q * p.cap + pos] = lb;
                            }
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar(B_TEMPTY + acc));
            t_phase ^= 1;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// ------------------------------------------------------------ canonical score ---
// acc = sum_s fma(rowscale_s, ip_s, acc), ip_s the sequential chain of segment s; q points at the
// query's concatenated coordinates (shared memory), tab_s at the staged tables.
__device__ __forceinline__ float canonical_acc(const SegTable& segs, const float* tab_s, const float* q, int64_t row) {
    float acc = 0.f;
    uint32_t w[MAX_BITS];
    for (int s = 0; s < segs.nseg; ++s) {
        const Seg& sg = segs.s[s];
        const uint8_t* src = sg.codes + row * sg.stride;
        const float* qs = q + sg.chunk0 * CH;
        float ip = 0.f;
        if (sg.tab_off >= 0) {
            const float* tab = tab_s + sg.tab_off;
            for (int c = 0; c < sg.nchunk; ++c) {
                EST_SWITCH_BITS(sg.bits, (load_words<(BB <= 8 ? BB : 8)>(src + (int64_t)c * 4 * sg.bits, w),
                                          ip = chain_chunk<(BB <= 8 ? BB : 8), true>(w, tab, 0.f, qs + c * CH, ip)));
            }
        } else {
            for (int c = 0; c < sg.nchunk; ++c) {
                EST_SWITCH_BITS(sg.bits, (load_words<BB>(src + (int64_t)c * 4 * sg.bits, w),
                                          ip = chain_chunk<BB, false>(w, nullptr, sg.base, qs + c * CH, ip)));
            }
        }
        const float rs = sg.rowscale ? __ldg(sg.rowscale + row) : 1.f;
        acc = __fmaf_rn(rs, ip, acc);
    }
    return acc;
}
__device__ __forceinline__ float canonical_key(float acc, float cq, float bi, int metric) {
    if (metric == B200VQ_METRIC_L2) return __fmaf_rn(-2.f, acc, __fadd_rn(cq, bi));
    return -__fadd_rn(cq, acc);
}

// |x_i|^2 = sum_s fma(rowscale^2, sum_d value^2, acc): thread per row.
__global__ void __launch_bounds__(256) est_rownorm_kernel(const __grid_constant__ SegTable segs, const float* __restrict__ tables,
                                                          int ntab, int64_t n, float* __restrict__ norm2) {
    __shared__ float tab_s[MAX_TAB];
    for (int i = threadIdx.x; i < ntab; i += blockDim.x) tab_s[i] = tables[i];
    __syncthreads();
    uint32_t w[MAX_BITS];
    for (int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; row < n; row += (int64_t)gridDim.x * blockDim.x) {
        float acc = 0.f;
        for (int s = 0; s < segs.nseg; ++s) {
            const Seg& sg = segs.s[s];
            const uint8_t* src = sg.codes + row * sg.stride;
            float l2 = 0.f;
            if (sg.tab_off >= 0) {
                const float* tab = tab_s + sg.tab_off;
                for (int c = 0; c < sg.nchunk; ++c) {
                    EST_SWITCH_BITS(sg.bits, (load_words<(BB <= 8 ? BB : 8)>(src + (int64_t)c * 4 * sg.bits, w),
                                              l2 = sq_chunk<(BB <= 8 ? BB : 8), true>(w, tab, 0.f, l2)));
                }
            } else {
                for (int c = 0; c < sg.nchunk; ++c) {
                    EST_SWITCH_BITS(sg.bits, (load_words<BB>(src + (int64_t)c * 4 * sg.bits, w),
                                              l2 = sq_chunk<BB, false>(w, nullptr, sg.base, l2)));
                }
            }
            const float rs = sg.rowscale ? __ldg(sg.rowscale + row) : 1.f;
            acc = __fmaf_rn(__fmul_rn(rs, rs), l2, acc);
        }
        norm2[row] = acc;
    }
}

// max_q |q|^2 (lane-strided chain + tree; only bounds the screen error) and max_q |C_q|
__global__ void est_qstats_kernel(const float* __restrict__ queries, int nq, int64_t ldq, int D, const float* __restrict__ qconst,
                                  float* __restrict__ qstats, float* __restrict__ thr) {
    const int lane = threadIdx.x & 31;
    const int q = (int)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    if (q >= nq) return;
    const float* pq = queries + (int64_t)q * ldq;
    float acc = 0.f;
    for (int i = lane; i < D; i += 32) acc = __fmaf_rn(pq[i], pq[i], acc);
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(FULL_MASK, acc, o);
    if (lane == 0) {
        atomicMax(reinterpret_cast<int*>(qstats), __float_as_int(acc));   // non-negative floats order as ints
        atomicMax(reinterpret_cast<int*>(qstats) + 1, __float_as_int(fabsf(qconst ? qconst[q] : 0.f)));
        thr[q] = __int_as_float(0x7f800000);    // first stage: every row passes (padding queries keep their NaN)
    }
}

// Candidate pool between the stages (the scheme of flat_tc.cu: flat_pool_kernel): no code row is decoded again until
// the one exact pass at the end. UB = LB + 4E (L2) / LB + 2E (IP) with the screen's own E; U = k-th smallest UB over
// pool and new survivors bounds the final k-th best key from above, the pool keeps the rows with LB <= U and U is the
// next stage's threshold. One CTA per query; pool arrays double-buffered by stage parity.
constexpr int POOL_CAP = 1024;
constexpr int POOL_MAX_ENTRIES = 8192;
__global__ void __launch_bounds__(256)
est_pool_kernel(const float* __restrict__ norm2, const float* __restrict__ bias, const float* __restrict__ qstats,
                float ebound, int metric, int k, const uint32_t* __restrict__ cand, const float* __restrict__ cand_lb,
                float* __restrict__ cand_ub, uint32_t* __restrict__ count, uint32_t cap,
                const uint32_t* __restrict__ prow_in, const float* __restrict__ plb_in, const float* __restrict__ pub_in,
                uint32_t* __restrict__ prow_out, float* __restrict__ plb_out, float* __restrict__ pub_out,
                uint32_t* __restrict__ pool_cnt, float* __restrict__ thr, uint32_t* __restrict__ flags,
                uint32_t dense_count) {
    extern __shared__ __align__(16) unsigned char pool_smem[];
    uint32_t* keys = reinterpret_cast<uint32_t*>(pool_smem);
    __shared__ uint32_t s_np, s_hist[256], s_sel[2];
    const int q = blockIdx.x;
    uint32_t cnt = dense_count ? dense_count : count[q];
    bool overflow = false;
    if (cnt > cap) {
        cnt = cap;
        overflow = true;
    }
    const uint32_t pc = pool_cnt[q];
    uint32_t total = pc + cnt;
    if (total > POOL_MAX_ENTRIES) {
        total = POOL_MAX_ENTRIES;
        overflow = true;
    }
    const int n2 = next_pow2((int)(total < 2 ? 2 : total));
    const float qn_max = qstats[0], qc_max = qstats[1], q_max = sqrtf(qn_max) * 1.0000005f;
    const uint32_t* list = cand + (size_t)q * cap;
    const float* llb = cand_lb + (size_t)q * cap;
    float* lub = cand_ub + (size_t)q * cap;
    const size_t pq = (size_t)q * POOL_CAP;
    if (threadIdx.x == 0) s_np = 0;
    for (uint32_t e = threadIdx.x; e < total; e += blockDim.x) {
        float ubv;
        if (e < pc) {
            ubv = pub_in[pq + e];
        } else {
            const uint32_t row = list[e - pc];
            const float xn = norm2[row];
            const float slack = 2.4e-7f * (xn + fabsf(bias[row]) + qn_max + qc_max);
            const float E = ebound * (sqrtf(xn) * 1.0000005f) * q_max + slack;     // est_screen_kernel's bound
            const float ub = llb[e - pc] + (metric == B200VQ_METRIC_L2 ? 4.f : 2.f) * E + slack;
            lub[e - pc] = ub;
            ubv = ub;
        }
        keys[e] = float_to_ordered(ubv);
    }
    __syncthreads();
    // U = the k-th smallest upper bound: a radix select (only the value is needed; the full bitonic sort of up to
    // 8192 keys was 0.1 ms per stage)
    float U = __int_as_float(0x7f800000);
    if ((uint32_t)k <= total) U = ordered_to_float(block_kth_smallest(keys, (int)total, (uint32_t)k, s_hist, s_sel));
    __syncthreads();
    for (uint32_t e = threadIdx.x; e < total; e += blockDim.x) {
        uint32_t row;
        float lb, ub;
        if (e < pc) {
            row = prow_in[pq + e];
            lb = plb_in[pq + e];
            ub = pub_in[pq + e];
        } else {
            row = list[e - pc];
            lb = llb[e - pc];
            ub = lub[e - pc];
        }
        if (lb <= U) {
            const uint32_t pos = atomicAdd(&s_np, 1u);
            if (pos < POOL_CAP) {
                prow_out[pq + pos] = row;
                plb_out[pq + pos] = lb;
                pub_out[pq + pos] = ub;
            }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t np = s_np;
        if (np > POOL_CAP) {
            np = POOL_CAP;
            overflow = true;
        }
        pool_cnt[q] = np;
        thr[q] = U;
        count[q] = 0;
        if (overflow) flags[q] = 1;     // redone exactly by the gated launch
    }
}

constexpr int FIN_THREADS = 256;
// One CTA per query: canonical keys of the listed rows, merged with the running list.
// seed_rows > 0: the "list" is rows [0, seed_rows). gate != null: only queries with gate[q] != 0 run, and they
// start from an empty running list (the exact redo of an overflowed query).
template <int CAP>
__global__ void __launch_bounds__(FIN_THREADS)
est_finish_kernel(const __grid_constant__ SegTable segs, const float* __restrict__ tables, int ntab, int D,
                  const float* __restrict__ bias, const float* __restrict__ queries, int64_t ldq,
                  const float* __restrict__ qconst, int metric, int k, const uint32_t* __restrict__ cand,
                  uint32_t* __restrict__ count, uint32_t cap, uint64_t* __restrict__ running, float* __restrict__ thr,
                  uint32_t* __restrict__ flags, int last, float* __restrict__ out_vals, uint32_t* __restrict__ out_ids,
                  int64_t seed_rows, const uint32_t* __restrict__ gate) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ uint64_t best[128];
    const int q = blockIdx.x;
    if (gate && !gate[q]) return;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    float* qs = reinterpret_cast<float*>(smem_raw);                       // [D]
    float* tab_s = qs + D;                                                // [ntab]
    uint64_t* wbuf = reinterpret_cast<uint64_t*>(smem_raw + ((((size_t)D + ntab) * 4 + 15) & ~(size_t)15));
    const int n2 = next_pow2((FIN_THREADS / 32) * k);
    uint64_t* mbuf = wbuf + (FIN_THREADS / 32) * CAP;
    const float* pq = queries + (int64_t)q * ldq;
    for (int i = threadIdx.x; i < D; i += FIN_THREADS) qs[i] = pq[i];
    for (int i = threadIdx.x; i < ntab; i += FIN_THREADS) tab_s[i] = tables[i];
    __syncthreads();
    const float cq = qconst ? qconst[q] : 0.f;
    int64_t cnt = seed_rows ? seed_rows : (int64_t)count[q];
    if (!seed_rows && cnt > (int64_t)cap) {
        if (threadIdx.x == 0) flags[q] = 1;     // redone exactly by the gated launch
        cnt = cap;
    }
    WarpTopK<CAP> tk;
    tk.init(wbuf + w * CAP, k);
    if (!gate) {
        for (int e0 = 0; e0 < k; e0 += FIN_THREADS) {
            const int e = e0 + threadIdx.x;
            uint64_t key = KEY_INF;
            if (e < k) key = running[(size_t)q * k + e];
            tk.offer(key != KEY_INF, key_value(key), key_id(key));
        }
    }
    const uint32_t* list = cand + (size_t)q * cap;
    const int64_t ngroups = (cnt + 31) / 32;
    for (int64_t g = w; g < ngroups; g += FIN_THREADS / 32) {
        const int64_t e = g * 32 + lane;
        const bool valid = e < cnt;
        float key = 0.f;
        uint32_t row = 0;
        if (valid) {
            row = seed_rows ? (uint32_t)e : list[e];
            key = canonical_key(canonical_acc(segs, tab_s, qs, (int64_t)row), cq, bias[row], metric);
        }
        tk.offer(valid, key, row);
    }
    block_emit_topk<CAP>(tk, mbuf, n2, k, best);
    __syncthreads();
    for (int j = threadIdx.x; j < k; j += FIN_THREADS) {
        const uint64_t kk = best[j];
        running[(size_t)q * k + j] = kk;
        if (last) {
            float v;
            uint32_t id;
            if (kk == KEY_INF) {
                v = __int_as_float(0x7f800000);
                id = 0xffffffffu;
            } else {
                v = key_value(kk);
                id = key_id(kk);
            }
            if (metric == B200VQ_METRIC_IP) v = -v;
            out_vals[(size_t)q * k + j] = v;
            out_ids[(size_t)q * k + j] = id;
        }
    }
    if (threadIdx.x == 0) {
        const uint64_t kk = best[k - 1];
        float t = __int_as_float(0x7f800000);
        // the screen tests a bound of (key - C_q) for L2 and of (key + C_q) = -<x, q> for inner product
        if (kk != KEY_INF) t = (metric == B200VQ_METRIC_IP) ? key_value(kk) + cq : key_value(kk) - cq;
        thr[q] = t;
        count[q] = 0;
    }
}

}  // namespace est
}  // namespace b200vq

using namespace b200vq;
using namespace b200vq::est;

namespace {
constexpr int64_t STAGE_GROWTH = 8;     // as flat_tc.cu: a stage costs a screen launch and a 30 us pool kernel
constexpr int64_t FIRST_ROWS = 2048;    // stage 0: infinite threshold, dense slots
constexpr int64_t MIN_SCREEN_ROWS = 16384;   // below this the exact kernel alone scores every row
struct EstPlan {
    uint32_t cap;
    int ntq;
    size_t off_running, off_thr, off_count, off_flags, off_qstats, off_tables, off_cand, off_cand_lb, off_cand_ub, off_pool,
        off_pool_cnt, total;
};
EstPlan plan_est(int nq, int k) {
    EstPlan f;
    f.ntq = (nq + TILE_Q - 1) / TILE_Q;
    int64_t cap = 64 * (int64_t)k;
    if (cap < 4096) cap = 4096;
    f.cap = (uint32_t)cap;
    auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
    size_t o = 0;
    f.off_running = o; o = up(o + (size_t)nq * k * 8);
    f.off_thr = o;     o = up(o + (size_t)f.ntq * TILE_Q * 4);
    f.off_count = o;   o = up(o + (size_t)nq * 4);
    f.off_flags = o;   o = up(o + (size_t)nq * 4);
    f.off_qstats = o;  o = up(o + 16);
    f.off_tables = o;  o = up(o + (size_t)MAX_TAB * 4);
    f.off_cand = o;    o = up(o + (size_t)nq * f.cap * 4);
    f.off_cand_lb = o; o = up(o + (size_t)nq * f.cap * 4);
    f.off_cand_ub = o; o = up(o + (size_t)nq * f.cap * 4);
    f.off_pool = o;    o = up(o + (size_t)nq * POOL_CAP * 4 * 6);     // rows, lb, ub: two generations
    f.off_pool_cnt = o; o = up(o + (size_t)nq * 4);
    f.total = o;
    return f;
}

// Host description -> SegTable; the value tables are gathered into `tables_dev` (device, MAX_TAB floats).
int build_segs(int nseg, const int* seg_len, const int* seg_bits, const uint8_t* const* seg_codes,
               const int64_t* seg_stride, const float* const* seg_rowscale, const float* const* seg_table,
               const float* seg_base, float* tables_dev, cudaStream_t st, SegTable* out, int* D_out, int* ntab_out) {
    VQ_CHECK_ARG(nseg >= 1 && nseg <= MAX_SEG, "est_scan: 1 <= nseg <= %d (got %d)", MAX_SEG, nseg);
    VQ_CHECK_ARG(seg_len && seg_bits && seg_codes && seg_stride, "est_scan: null segment description");
    int chunk = 0, ntab = 0;
    out->nseg = nseg;
    for (int s = 0; s < nseg; ++s) {
        Seg& g = out->s[s];
        VQ_CHECK_ARG(seg_len[s] > 0 && seg_len[s] % CH == 0, "est_scan: segment %d length %d is not a multiple of 32", s,
                     seg_len[s]);
        VQ_CHECK_ARG(seg_bits[s] >= 1 && seg_bits[s] <= MAX_BITS, "est_scan: segment %d has %d bits (1..13)", s,
                     seg_bits[s]);
        VQ_CHECK_ARG(seg_codes[s] && ((uintptr_t)seg_codes[s] & 3) == 0 && seg_stride[s] % 4 == 0 &&
                         seg_stride[s] >= (int64_t)seg_len[s] * seg_bits[s] / 8,
                     "est_scan: segment %d codes must be 4-byte aligned with a stride that is a multiple of 4", s);
        g.codes = seg_codes[s];
        g.rowscale = seg_rowscale ? seg_rowscale[s] : nullptr;
        g.stride = seg_stride[s];
        g.chunk0 = chunk;
        g.nchunk = seg_len[s] / CH;
        g.bits = seg_bits[s];
        g.base = seg_base ? seg_base[s] : 0.f;
        g.tab_off = -1;
        if (seg_table && seg_table[s]) {
            VQ_CHECK_ARG(seg_bits[s] <= 8, "est_scan: value tables need bits <= 8");
            const int L = 1 << seg_bits[s];
            VQ_CHECK_ARG(ntab + L <= MAX_TAB, "est_scan: value tables exceed %d entries", MAX_TAB);
            g.tab_off = ntab;
            if (tables_dev)
                VQ_CUDA(cudaMemcpyAsync(tables_dev + ntab, seg_table[s], (size_t)L * 4, cudaMemcpyDeviceToDevice, st));
            ntab += L;
        }
        chunk += g.nchunk;
    }
    VQ_CHECK_ARG(chunk <= MAX_CHUNKS, "est_scan: more than %d dimensions", MAX_CHUNKS * CH);
    *D_out = chunk * CH;
    *ntab_out = ntab;
    return B200VQ_OK;
}

template <int CAP>
int est_finish_launch(const SegTable& segs, const float* tables, int ntab, int D, const float* bias, const float* queries,
                      int nq, int64_t ldq, const float* qconst, int metric, int k, const uint32_t* cand, uint32_t* count,
                      uint32_t cap, uint64_t* running, float* thr, uint32_t* flags, int last, float* out_vals,
                      uint32_t* out_ids, cudaStream_t st, int64_t seed_rows, const uint32_t* gate) {
    const int n2 = next_pow2((FIN_THREADS / 32) * k);
    const size_t smem = ((((size_t)D + ntab) * 4 + 15) & ~(size_t)15) + (size_t)((FIN_THREADS / 32) * CAP + n2) * 8;
    VQ_CHECK_ARG(smem <= 200 * 1024, "est_scan: D too large for the finish kernel");
    VQ_CUDA(cudaFuncSetAttribute(est_finish_kernel<CAP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    est_finish_kernel<CAP><<<nq, FIN_THREADS, smem, st>>>(segs, tables, ntab, D, bias, queries, ldq, qconst, metric, k, cand,
                                                        count, cap, running, thr, flags, last, out_vals, out_ids,
                                                        seed_rows, gate);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}
}  // namespace

extern "C" int b200vq_est_row_norms(int nseg, const int* seg_len, const int* seg_bits, const uint8_t* const* seg_codes,
                                    const int64_t* seg_stride, const float* const* seg_rowscale,
                                    const float* const* seg_table, const float* seg_base, int64_t n, float* norm2,
                                    void* scratch, size_t scratch_bytes, void* stream) {
    VQ_CHECK_ARG(n >= 0 && (norm2 || n == 0), "est_row_norms: bad args");
    VQ_CHECK_ARG(scratch && scratch_bytes >= (size_t)MAX_TAB * 4, "est_row_norms: scratch too small (%zu < %zu)",
                 scratch_bytes, (size_t)MAX_TAB * 4);
    cudaStream_t st = as_stream(stream);
    SegTable segs;
    int D = 0, ntab = 0;
    float* tables = reinterpret_cast<float*>(scratch);
    int rc = build_segs(nseg, seg_len, seg_bits, seg_codes, seg_stride, seg_rowscale, seg_table, seg_base, tables, st,
                        &segs, &D, &ntab);
    if (rc != B200VQ_OK || n == 0) return rc;
    int64_t grid = (n + 255) / 256;
    const int64_t cap = (int64_t)device_sm_count() * 8;
    if (grid > cap) grid = cap;
    est_rownorm_kernel<<<(unsigned)grid, 256, 0, st>>>(segs, tables, ntab, n, norm2);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" size_t b200vq_est_scan_scratch_bytes(int64_t n, int nq, int k) {
    if (n < 0 || nq <= 0 || k <= 0) return 0;
    return plan_est(nq, k).total;
}

extern "C" int b200vq_est_scan_topk(int nseg, const int* seg_len, const int* seg_bits, const uint8_t* const* seg_codes,
                                    const int64_t* seg_stride, const float* const* seg_rowscale,
                                    const float* const* seg_table, const float* seg_base, const float* row_norm2,
                                    const float* row_bias, int64_t n, const float* queries, int nq, int64_t ldq,
                                    const float* qconst, int metric, int k, float* out_vals, uint32_t* out_ids,
                                    void* scratch, size_t scratch_bytes, void* stream) {
    VQ_CHECK_ARG(queries && out_vals && out_ids && row_norm2 && row_bias, "est_scan_topk: null pointer");
    VQ_CHECK_ARG(n >= 0 && n < (1ll << 31) && nq >= 0, "est_scan_topk: bad shape");
    VQ_CHECK_ARG(k > 0 && k <= 128, "est_scan_topk: k must be in [1,128], got %d", k);
    VQ_CHECK_ARG(metric == B200VQ_METRIC_L2 || metric == B200VQ_METRIC_IP, "est_scan_topk: bad metric");
    if (nq == 0) return B200VQ_OK;
    const EstPlan fp = plan_est(nq, k);
    VQ_CHECK_ARG(scratch && scratch_bytes >= fp.total, "est_scan_topk: scratch too small (%zu < %zu)", scratch_bytes,
                 fp.total);
    cudaStream_t st = as_stream(stream);
    uint8_t* sb = reinterpret_cast<uint8_t*>(scratch);
    uint64_t* running = reinterpret_cast<uint64_t*>(sb + fp.off_running);
    float* thr = reinterpret_cast<float*>(sb + fp.off_thr);
    uint32_t* count = reinterpret_cast<uint32_t*>(sb + fp.off_count);
    uint32_t* flags = reinterpret_cast<uint32_t*>(sb + fp.off_flags);
    float* qstats = reinterpret_cast<float*>(sb + fp.off_qstats);
    float* tables = reinterpret_cast<float*>(sb + fp.off_tables);
    uint32_t* cand = reinterpret_cast<uint32_t*>(sb + fp.off_cand);
    SegTable segs;
    int D = 0, ntab = 0;
    int rc = build_segs(nseg, seg_len, seg_bits, seg_codes, seg_stride, seg_rowscale, seg_table, seg_base, tables, st,
                        &segs, &D, &ntab);
    if (rc != B200VQ_OK) return rc;
    VQ_CHECK_ARG(ldq >= D && ldq % 4 == 0 && ((uintptr_t)queries & 15) == 0,
                 "est_scan_topk: queries must be 16-byte aligned (nq, %d) rows with ldq %% 4 == 0", D);

    VQ_CUDA(cudaMemsetAsync(running, 0xFF, (size_t)nq * k * 8, st));        // KEY_INF
    VQ_CUDA(cudaMemsetAsync(flags, 0, (size_t)nq * 4, st));
    VQ_CUDA(cudaMemsetAsync(count, 0, (size_t)nq * 4, st));
#define EST_FINISH(LAST, SEED, GATE)                                                                                   \
    ((k <= 32) ? est_finish_launch<64>(segs, tables, ntab, D, row_bias, queries, nq, ldq, qconst, metric, k, cand,     \
                                       count, fp.cap, running, thr, flags, LAST, out_vals, out_ids, st, SEED, GATE)    \
               : est_finish_launch<256>(segs, tables, ntab, D, row_bias, queries, nq, ldq, qconst, metric, k, cand,    \
                                        count, fp.cap, running, thr, flags, LAST, out_vals, out_ids, st, SEED, GATE))
    if (n < MIN_SCREEN_ROWS) return EST_FINISH(1, n, nullptr);   // n == 0: empty lists -> +inf / 0xffffffff

    float* cand_lb = reinterpret_cast<float*>(sb + fp.off_cand_lb);
    float* cand_ub = reinterpret_cast<float*>(sb + fp.off_cand_ub);
    uint32_t* pool_cnt = reinterpret_cast<uint32_t*>(sb + fp.off_pool_cnt);
    const size_t pstride = (size_t)nq * POOL_CAP;
    uint32_t* prow[2] = {reinterpret_cast<uint32_t*>(sb + fp.off_pool), reinterpret_cast<uint32_t*>(sb + fp.off_pool) + pstride};
    float* plb[2] = {reinterpret_cast<float*>(sb + fp.off_pool) + 2 * pstride, reinterpret_cast<float*>(sb + fp.off_pool) + 3 * pstride};
    float* pub[2] = {reinterpret_cast<float*>(sb + fp.off_pool) + 4 * pstride, reinterpret_cast<float*>(sb + fp.off_pool) + 5 * pstride};
    VQ_CUDA(cudaMemsetAsync(qstats, 0, 16, st));
    VQ_CUDA(cudaMemsetAsync(thr, 0xFF, (size_t)fp.ntq * TILE_Q * 4, st));   // NaN: padding queries never pass
    VQ_CUDA(cudaMemsetAsync(pool_cnt, 0, (size_t)nq * 4, st));
    est_qstats_kernel<<<(unsigned)(((int64_t)nq * 32 + 255) / 256), 256, 0, st>>>(queries, nq, ldq, D, qconst, qstats, thr);
    VQ_LAUNCH_CHECK();

    CUtensorMap map_q;
    if ((rc = encode_2d(&map_q, queries, (uint64_t)nq, (uint64_t)D, (uint64_t)ldq, TILE_Q, CH, CU_TENSOR_MAP_SWIZZLE_128B)))
        return rc;
    const uint32_t smem = TAB_OFF + (uint32_t)ntab * 4 + 1024;
    typedef void (*ScreenFn)(const CUtensorMap, const SegTable, const Params);
    ScreenFn screen = est_screen_kernel<0>;
    if (segs.nseg == 1 && segs.s[0].tab_off == 0) {   // one table segment (extended RaBitQ): width-specialised converters
        static const ScreenFn fixed[9] = {nullptr, est_screen_kernel<1>, est_screen_kernel<2>, est_screen_kernel<3>,
                                          est_screen_kernel<4>, est_screen_kernel<5>, est_screen_kernel<6>,
                                          est_screen_kernel<7>, est_screen_kernel<8>};
        screen = fixed[segs.s[0].bits];
    }
    VQ_CUDA(cudaFuncSetAttribute(screen, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const size_t pool_smem_bytes = (size_t)POOL_MAX_ENTRIES * 8;
    VQ_CUDA(cudaFuncSetAttribute(est_pool_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pool_smem_bytes));
    Params p;
    p.norm2 = row_norm2; p.bias = row_bias; p.thr = thr; p.qstats = qstats; p.tables = tables; p.cand = cand;
    p.cand_lb = cand_lb;
    p.count = count; p.cap = fp.cap; p.nq = nq; p.nchunk = D / CH; p.ntq = fp.ntq; p.ntab = ntab; p.metric = metric;
    p.ebound = 0.0009765625f * 1.02f + 3.0f * (float)D * 5.9604645e-8f;
    const int sms = device_sm_count();
    int gen = 0;
    for (int64_t r = 0; r < n;) {
        int64_t r1 = r == 0 ? FIRST_ROWS : r * STAGE_GROWTH;
        if (r1 > n || n - r1 < r1 / 4) r1 = n;  // fold a short tail into the last stage
        p.r0 = r; p.r1 = r1;
        p.dense = r == 0;
        p.ntr = (r1 - r + TILE_ROWS - 1) / TILE_ROWS;
        const int64_t units = p.ntr * p.ntq;
        screen<<<(unsigned)(units < sms ? units : sms), THREADS, smem, st>>>(map_q, segs, p);
        VQ_LAUNCH_CHECK();
        est_pool_kernel<<<nq, 256, pool_smem_bytes, st>>>(row_norm2, row_bias, qstats, p.ebound, metric, k, cand, cand_lb,
                                                          cand_ub, count, fp.cap, prow[gen], plb[gen], pub[gen],
                                                          prow[gen ^ 1], plb[gen ^ 1], pub[gen ^ 1], pool_cnt, thr, flags,
                                                          p.dense ? (uint32_t)(r1 - r) : 0u);
        VQ_LAUNCH_CHECK();
        gen ^= 1;
        r = r1;
    }
    // canonical keys of the pool rows, top-k in (key, id) order
    rc = (k <= 32) ? est_finish_launch<64>(segs, tables, ntab, D, row_bias, queries, nq, ldq, qconst, metric, k, prow[gen],
                                           pool_cnt, (uint32_t)POOL_CAP, running, thr, flags, 1, out_vals, out_ids, st, 0,
                                           nullptr)
                   : est_finish_launch<256>(segs, tables, ntab, D, row_bias, queries, nq, ldq, qconst, metric, k, prow[gen],
                                            pool_cnt, (uint32_t)POOL_CAP, running, thr, flags, 1, out_vals, out_ids, st, 0,
                                            nullptr);
    if (rc != B200VQ_OK) return rc;
    // queries whose survivor list overflowed in some stage: exact scan of all rows, gated per query
    return EST_FINISH(1, n, flags);
#undef EST_FINISH
}
