// 1-bit RaBitQ scan + top-k on the sm_100a tensor cores.
//
// Replaces faiss IndexRaBitQ.search as driven by the reference at
// src/haag_vq/methods/search/rabitq_index.py:62-70 (qb-bit scalar-quantised query residual against
// the sign codes, unbiased estimator, top-k); same results as b200vq_rabitq1_scan_topk (rabitq.cu).
//
// The estimator's only (row, query) term is the integer  dot = sum_b 2^b popc(code & plane_b)
//   = sum_d bit_d(code) * qv_d,   qv_d in [0, 2^qb)  the quantised query value,
// i.e. a GEMM of the n x D sign-bit matrix with the nq x D matrix of small integers. For qb <= 4 both
// factors are exact in FP8 E4M3 (0/1 and 0..15) and every partial sum is an integer below 2^24, so
// tcgen05.mma kind::f8f6f4 with fp32 accumulation returns `dot` EXACTLY: the tensor pipe replaces
// D/32 * qb AND+POPC pairs per (row, query) of the SIMT kernel.
//
// Per (128-row, 256-query) unit: eight converter warps expand the sign bits of a 128-dimension chunk
// (16 bytes per row, read straight from the code rows in global memory, prefetched one chunk ahead)
// into a K-major SWIZZLE_128B tile of E4M3 0x00 / 0x38 bytes; TMA brings the matching 256 x 128-byte
// tile of query values; one thread issues four tcgen05.mma (128 x 256 x 32) per chunk into one of two
// TMEM accumulators; eight epilogue warps (accumulator stage x TMEM lane quarter, lane = row) turn
// dot into a conservative form of the estimate with three FMAs per value and test it against the
// query's current threshold. Survivors (dot, row) go to per-query lists; rq_finish_kernel evaluates
// the canonical estimate (rabitq1_estimate, shared with the SIMT kernel) for them, merges them into
// the running top-k and tightens the thresholds for the next, twice as large, row stage. A query
// whose list overflows is redone by the SIMT kernel (gated launch), so results never depend on the
// screen. Row side information (f0, -2 f1, popc) is extracted once per call by rq_rowinfo_kernel.
#include "tc_common.cuh"
#include "topk.cuh"
#include "rabitq_est.cuh"

namespace b200vq {
namespace rqtc {
using namespace tc;

constexpr int TILE_ROWS = 128;
constexpr int TILE_Q = 256;
constexpr int CHB = 128;                     // K chunk: 128 dimensions = 128 bytes = one swizzle row
constexpr int NST = 4;
constexpr int CONV_WARPS = 8, EPI_WARPS = 16;
constexpr int THREADS = 64 + 32 * (CONV_WARPS + EPI_WARPS);
constexpr uint32_t A_BYTES = TILE_ROWS * CHB;      // 16 KB
constexpr uint32_t B_BYTES = TILE_Q * CHB;         // 32 KB
constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
constexpr uint32_t BAR_OFF = NST * STAGE_BYTES;
constexpr uint32_t QC_OFF = BAR_OFF + 256;
constexpr uint32_t SMEM_BYTES = QC_OFF + 2 * TILE_Q * 16 + 1024;   // + alignment slack
enum { B_FULL = 0, B_CONV = NST, B_EMPTY = 2 * NST, B_TFULL = 3 * NST, B_TEMPTY = 3 * NST + 2, B_TMEM = 3 * NST + 4 };
constexpr int SEED_ROWS = 1024;
constexpr int64_t MIN_ROWS = 32768;

// kind::f8f6f4, A = B = E4M3 (format 0), fp32 accumulate, K-major, M = 128, N = 256
constexpr uint32_t IDESC_E4M3_128x256 = (1u << 4) | ((256u >> 3) << 17) | ((128u >> 4) << 24);

__device__ __forceinline__ void umma_f8(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                        uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}

struct Params {
    const uint8_t* codes;     // (n, code_size)
    const float4* rowinfo;    // (n) {f0, -2 f1, popc, 0}
    const float4* qscreen;    // (ntq * 256) {c1, c2, -c34, T'}: T' = NaN for padding queries
    uint64_t* cand;           // (nq, cap) survivors: dot (fp32 bits) << 32 | row
    uint32_t* count;          // (nq)
    uint32_t cap;
    int64_t r0, r1;           // rows of this stage
    int code_size, nchunk, ntq;
    int D, seed;              // seed: first stage, thresholds are +inf
    unsigned long long* dbg;  // optional (B200VQ_RQ_DEBUG): MMA-thread wait cycles [t_empty, b_full, conv, total]
    int64_t ntr;
};

// ------------------------------------------------------------ row side info ---
// One warp per row: popc of the sign bits (lane-strided words), the two trailing factors, and the
// maximum of |a| = |2 f1| over all rows (bounds the rounding slack of the screen).
__global__ void rq_rowinfo_kernel(const uint8_t* __restrict__ codes, int64_t n, int D, float4* __restrict__ rowinfo,
                                  float* __restrict__ amax) {
    const int lane = threadIdx.x & 31;
    const int64_t wid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int nbytes = D >> 3, code_size = nbytes + 8;       // D % 64 == 0 on this path: rows are 8-byte aligned
    float am = 0.f;
    for (int64_t row = wid; row < n; row += nw) {
        const uint32_t* cw = reinterpret_cast<const uint32_t*>(codes + row * (int64_t)code_size);
        uint32_t pc = 0;
        for (int w = lane; w < (nbytes >> 2); w += 32) pc += __popc(__ldg(cw + w));
        for (int o = 16; o > 0; o >>= 1) pc += __shfl_xor_sync(FULL_MASK, pc, o);
        if (lane == 0) {
            const float f0 = __uint_as_float(__ldg(cw + (nbytes >> 2)));
            const float f1 = __uint_as_float(__ldg(cw + (nbytes >> 2) + 1));
            const float a = -2.f * f1;
            rowinfo[row] = make_float4(f0, a, (float)pc, 0.f);
            am = fmaxf(am, fabsf(a));     // NaN factors are dropped here; such rows never pass the screen
        }
    }
    if (lane == 0 && am > 0.f) atomicMax(reinterpret_cast<int*>(amax), __float_as_int(am));
}

// ------------------------------------------------------------- query matrix ---
// grid nq_pad, 128 threads: B row q = the quantised query values as E4M3 bytes (Dp = D rounded up to
// 128 columns, zero padded; rows >= nq are zero), and the initial screen constants.
__constant__ uint8_t E4M3_OF_INT[16] = {0x00, 0x38, 0x40, 0x44, 0x48, 0x4A, 0x4C, 0x4E,
                                        0x50, 0x51, 0x52, 0x53, 0x54, 0x55, 0x56, 0x57};
__global__ void rq_prep_kernel(const uint32_t* __restrict__ planes, const float* __restrict__ qfac, int nq, int D,
                               int Dp, int qb, uint8_t* __restrict__ bmat, float4* __restrict__ qscreen) {
    const int q = blockIdx.x, W = (D + 31) >> 5;
    uint32_t* row = reinterpret_cast<uint32_t*>(bmat + (size_t)q * Dp);
    for (int g = threadIdx.x; g < (Dp >> 2); g += blockDim.x) {     // four dimensions per thread and step
        uint32_t out = 0;
        if (q < nq && 4 * g < D) {
            const int wd = g >> 3, sh = (g & 7) * 4;
            uint32_t v[4] = {0, 0, 0, 0};
            for (int b = 0; b < qb; ++b) {
                const uint32_t nib = (planes[((size_t)q * qb + b) * W + wd] >> sh) & 0xFu;
#pragma unroll
                for (int e = 0; e < 4; ++e) v[e] |= ((nib >> e) & 1u) << b;
            }
#pragma unroll
            for (int e = 0; e < 4; ++e)
                if (4 * g + e < D) out |= (uint32_t)E4M3_OF_INT[v[e]] << (8 * e);
        }
        row[g] = out;
    }
    if (threadIdx.x == 0) {
        float4 s = make_float4(0.f, 0.f, 0.f, __int_as_float(0x7fc00000));    // padding query: NaN never passes
        if (q < nq) s = make_float4(qfac[(size_t)q * 8], qfac[(size_t)q * 8 + 1], -qfac[(size_t)q * 8 + 2],
                                    __int_as_float(0x7f800000));               // first stage: everything passes
        qscreen[q] = s;
    }
}

// --------------------------------------------------------------- main kernel ---
__global__ void __launch_bounds__(THREADS, 1)
rq_screen_kernel(const __grid_constant__ CUtensorMap map_b, const Params p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;
    uint8_t* gbase = smem_raw + (base - raw);
    const uint32_t bar0 = base + BAR_OFF;
    auto bar = [&](int slot) { return bar0 + 8u * (uint32_t)slot; };
    volatile uint32_t* tmem_ptr_smem = reinterpret_cast<volatile uint32_t*>(gbase + BAR_OFF + 8 * B_TMEM);
    float4* qc_s = reinterpret_cast<float4*>(gbase + QC_OFF);     // [2][TILE_Q]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t nunits = p.ntr * p.ntq;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&map_b);
        for (int i = 0; i < NST; ++i) {
            mbar_init(bar(B_FULL + i), 1);
            mbar_init(bar(B_CONV + i), CONV_WARPS);
            mbar_init(bar(B_EMPTY + i), 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(bar(B_TFULL + i), 1);
            mbar_init(bar(B_TEMPTY + i), 8);
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
                    tma_load_2d(base + st * STAGE_BYTES + A_BYTES, &map_b, bar(B_FULL + st), c * CHB, u * TILE_Q);
                    if (++st == NST) { st = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ======================================================= MMA issuer ===
        if (lane == 0) {
            uint32_t st = 0, phase = 0, acc = 0, t_phase = 0;
            long long w_t = 0, w_b = 0, w_c = 0;
            const long long t_begin = clock64();
            for (int64_t unit = blockIdx.x; unit < nunits; unit += gridDim.x) {
                { const long long c0 = clock64(); mbar_wait_backoff(bar(B_TEMPTY + acc), t_phase ^ 1, 32); w_t += clock64() - c0; }
                tc_fence_after();
                const uint32_t d = tmem_base + acc * 256u;
                uint32_t accum = 0;
                for (int c = 0; c < p.nchunk; ++c) {
                    { const long long c0 = clock64(); mbar_wait_backoff(bar(B_FULL + st), phase, 32); w_b += clock64() - c0; }
                    { const long long c0 = clock64(); mbar_wait_backoff(bar(B_CONV + st), phase, 32); w_c += clock64() - c0; }
                    tc_fence_after();
                    const uint32_t a_addr = base + st * STAGE_BYTES, b_addr = a_addr + A_BYTES;
#pragma unroll
                    for (int kk = 0; kk < CHB / 32; ++kk) {
                        umma_f8(d, make_kmajor_desc(a_addr + kk * 32, 1024, 2), make_kmajor_desc(b_addr + kk * 32, 1024, 2),
                                IDESC_E4M3_128x256, accum);
                        accum = 1;
                    }
                    umma_commit(bar(B_EMPTY + st));
                    if (++st == NST) { st = 0; phase ^= 1; }
                }
                umma_commit(bar(B_TFULL + acc));
                if (++acc == 2) { acc = 0; t_phase ^= 1; }
            }
            if (p.dbg) {
                atomicAdd(p.dbg + 0, (unsigned long long)w_t);
                atomicAdd(p.dbg + 1, (unsigned long long)w_b);
                atomicAdd(p.dbg + 2, (unsigned long long)w_c);
                atomicAdd(p.dbg + 3, (unsigned long long)(clock64() - t_begin));
            }
        }
    } else if (warp < 2 + CONV_WARPS) {
        // ======================================================= converters ===
        // thread = (tile row, 8-byte half of the chunk): 64 sign bits -> 64 E4M3 bytes (0x38 where the bit is
        // set) at the SWIZZLE_128B position of the row; the next chunk's bytes are in flight meanwhile
        const int cth = threadIdx.x - 64;
        const int r = cth & 127, half = cth >> 7;
        const uint32_t row_off = (uint32_t)(r >> 3) * 1024u + (uint32_t)(r & 7) * 128u;
        const int half_chunks = p.D >> 6;                     // 8-byte groups of sign bits in a row
        // The code bytes come straight from global memory (L2), PD chunks ahead of their use: a prefetch
        // cursor walks the same (unit, chunk) sequence as the main loop, across unit boundaries.
        constexpr int PD = 4;
        int64_t k_unit = blockIdx.x;
        int k_c = 0;
        const uint2* k_src = nullptr;
        bool k_live = false;
        auto k_open = [&]() {
            k_live = false;
            if (k_unit < nunits) {
                const int64_t row = p.r0 + (k_unit / p.ntq) * TILE_ROWS + r;
                k_live = row < p.r1;
                k_src = reinterpret_cast<const uint2*>(p.codes + row * (int64_t)p.code_size) + half;
            }
        };
        auto k_fetch = [&]() {
            uint2 v = make_uint2(0, 0);
            if (k_unit < nunits) {
                if (k_live && 2 * k_c + half < half_chunks) v = __ldg(k_src + 2 * k_c);
                if (++k_c == p.nchunk) {
                    k_c = 0;
                    k_unit += gridDim.x;
                    k_open();
                }
            }
            return v;
        };
        k_open();
        uint2 ring[PD];
#pragma unroll
        for (int i = 0; i < PD; ++i) ring[i] = k_fetch();
        uint32_t st = 0, phase = 0;
        for (int64_t unit = blockIdx.x; unit < nunits; unit += gridDim.x) {
            for (int c = 0; c < p.nchunk; ++c) {
                const uint2 cur = ring[0];
#pragma unroll
                for (int i = 0; i + 1 < PD; ++i) ring[i] = ring[i + 1];
                ring[PD - 1] = k_fetch();
                mbar_wait(bar(B_EMPTY + st), phase ^ 1);
                uint8_t* a_tile = gbase + st * STAGE_BYTES + row_off;
                const uint32_t w[2] = {cur.x, cur.y};
#pragma unroll
                for (int j = 0; j < 4; ++j) {                 // 16 dimensions = 16 bits -> one 16-byte unit
                    const uint32_t bits = (w[j >> 1] >> ((j & 1) * 16)) & 0xFFFFu;
                    uint4 o;
                    o.x = (((bits & 0xFu) * 0x00204081u) & 0x01010101u) * 0x38u;
                    o.y = ((((bits >> 4) & 0xFu) * 0x00204081u) & 0x01010101u) * 0x38u;
                    o.z = ((((bits >> 8) & 0xFu) * 0x00204081u) & 0x01010101u) * 0x38u;
                    o.w = (((bits >> 12) * 0x00204081u) & 0x01010101u) * 0x38u;
                    *reinterpret_cast<uint4*>(a_tile + (((half * 4 + j) ^ (r & 7)) << 4)) = o;
                }
                fence_proxy_async();
                __syncwarp();
                if (lane == 0) mbar_arrive(bar(B_CONV + st));
                if (++st == NST) { st = 0; phase ^= 1; }
            }
        }
    } else {
        // ========================================================= epilogue ===
        const int ew = warp - 2 - CONV_WARPS;                 // 0..15
        const uint32_t acc = (uint32_t)ew >> 3;               // accumulator stage served by this warp
        const int q4 = warp & 3;                              // TMEM lane quarter this warp may read
        const int ch = (ew >> 2) & 1;                         // column half of the 256-query tile
        const uint32_t lane_addr = tmem_base + ((uint32_t)(q4 * 32) << 16) + acc * 256u + (uint32_t)ch * 128u;
        float4* qs = qc_s + acc * TILE_Q;
        const int gt = (ew & 7) * 32 + lane;                  // 0..255 inside the stage group
        const uint32_t grp_bar = 1u + acc;
        uint32_t t_phase = 0;
        int64_t unit = blockIdx.x + (int64_t)acc * gridDim.x;  // this group takes every second unit of the CTA
        float4 myq = make_float4(0.f, 0.f, 0.f, 0.f);
        if (unit < nunits) myq = p.qscreen[(size_t)(unit % p.ntq) * TILE_Q + gt];
        for (; unit < nunits; unit += 2 * (int64_t)gridDim.x) {
            const int64_t t = unit / p.ntq;
            const int u = (int)(unit - t * p.ntq);
            const int64_t row = p.r0 + t * TILE_ROWS + q4 * 32 + lane;
            const bool live = row < p.r1;
            float4 ri = make_float4(0.f, 0.f, 0.f, 0.f);
            if (live) ri = __ldg(p.rowinfo + row);
            asm volatile("bar.sync %0, 256;" ::"r"(grp_bar) : "memory");   // previous unit's constants consumed
            qs[gt] = myq;
            asm volatile("bar.sync %0, 256;" ::"r"(grp_bar) : "memory");
            // the group's next unit: its constants travel while this one is processed
            const int64_t nu = unit + 2 * (int64_t)gridDim.x;
            if (nu < nunits) myq = p.qscreen[(size_t)(nu % p.ntq) * TILE_Q + gt];
            mbar_wait(bar(B_TFULL + acc), t_phase);
            tc_fence_after();
            if (p.seed) {
                // first stage (thresholds +inf): every row passes, so the appends are aggregated per warp
#pragma unroll 1
                for (int cb = 0; cb < 4; ++cb) {
                    float v[32];
                    tmem_ld32(lane_addr + cb * 32, v);
#pragma unroll 1
                    for (int j = 0; j < 32; ++j) {
                        const int q = u * TILE_Q + ch * 128 + cb * 32 + j;
                        const float4 c = qs[ch * 128 + cb * 32 + j];
                        const bool hit = live && !(c.w != c.w);           // padding queries carry NaN
                        const unsigned m = __ballot_sync(FULL_MASK, hit);
                        if (m == 0) continue;
                        uint32_t pos0 = 0;
                        if (lane == 0) pos0 = atomicAdd(p.count + q, (uint32_t)__popc(m));
                        pos0 = __shfl_sync(FULL_MASK, pos0, 0);
                        const uint32_t pos = pos0 + __popc(m & ((1u << lane) - 1u));
                        float vj = 0.f;
#pragma unroll
                        for (int e = 0; e < 32; ++e) vj = (e == j) ? v[e] : vj;
                        if (hit && pos < p.cap)
                            p.cand[(size_t)q * p.cap + pos] = ((uint64_t)__float_as_uint(vj) << 32) | (uint32_t)row;
                    }
                }
            } else {
#pragma unroll 1
            for (int cb = 0; cb < 4; ++cb) {
                float v[32];
                tmem_ld32(lane_addr + cb * 32, v);
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const float4 c = qs[ch * 128 + cb * 32 + j];          // {c1, c2, -c34, T'}
                    const float wq = __fmaf_rn(c.y, ri.z, c.z);
                    const float inner = __fmaf_rn(c.x, v[j], wq);
                    const float e = __fmaf_rn(ri.y, inner, ri.x);
                    if (live && e <= c.w) {
                        const int q = u * TILE_Q + ch * 128 + cb * 32 + j;
                        const uint32_t pos = atomicAdd(p.count + q, 1u);
                        if (pos < p.cap)
                            p.cand[(size_t)q * p.cap + pos] = ((uint64_t)__float_as_uint(v[j]) << 32) | (uint32_t)row;
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

// -------------------------------------------------------------------- finish ---
// One CTA per query: canonical estimate of every survivor (dot, row), merge with the running top-k
// (ordering (value, id)), new screen threshold, list reset. A list that overflowed flags the query.
constexpr int FIN_THREADS = 256;
template <int CAP>
__global__ void __launch_bounds__(FIN_THREADS)
rq_finish_kernel(const float4* __restrict__ rowinfo, const float* __restrict__ qfac, const float* __restrict__ amax,
                 int D, int qb, int metric, int k, const uint64_t* __restrict__ cand, uint32_t* __restrict__ count,
                 uint32_t cap, uint64_t* __restrict__ running, float4* __restrict__ qscreen,
                 uint32_t* __restrict__ flags, int last, float* __restrict__ out_vals, uint32_t* __restrict__ out_ids) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ uint64_t best[128];
    const int q = blockIdx.x;
    const int w = threadIdx.x >> 5;
    uint64_t* wbuf = reinterpret_cast<uint64_t*>(smem_raw);
    const int n2 = next_pow2((FIN_THREADS / 32) * k);
    uint64_t* mbuf = wbuf + (FIN_THREADS / 32) * CAP;
    const float c1 = qfac[(size_t)q * 8 + 0], c2 = qfac[(size_t)q * 8 + 1], c34 = qfac[(size_t)q * 8 + 2],
                qc2 = qfac[(size_t)q * 8 + 3], qn2 = qfac[(size_t)q * 8 + 4];
    uint32_t cnt = count[q];
    if (cnt > cap) {
        if (threadIdx.x == 0) flags[q] = 1;     // redone by the gated SIMT launch
        cnt = cap;
    }
    WarpTopK<CAP> tk;
    tk.init(wbuf + w * CAP, k);
    for (int e0 = 0; e0 < k; e0 += FIN_THREADS) {
        const int e = e0 + threadIdx.x;
        uint64_t key = KEY_INF;
        if (e < k) key = running[(size_t)q * k + e];
        tk.offer(key != KEY_INF, key_value(key), key_id(key));
    }
    const uint64_t* list = cand + (size_t)q * cap;
    const uint32_t rounds = (cnt + FIN_THREADS - 1) / FIN_THREADS;
    for (uint32_t itr = 0; itr < rounds; ++itr) {
        const uint32_t e = itr * FIN_THREADS + threadIdx.x;
        const bool valid = e < cnt;
        float est = 0.f;
        uint32_t row = 0;
        if (valid) {
            const uint64_t ent = list[e];
            row = (uint32_t)ent;
            const float4 ri = __ldg(rowinfo + row);
            est = rabitq1_estimate(__uint_as_float((uint32_t)(ent >> 32)), ri.z, ri.x, ri.y, c1, c2, c34, qc2, qn2, metric);
        }
        tk.offer(valid, est, row);
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
        // screen test: fma(a, fma(c1, dot, fma(c2, s, -c34)), f0) <= T'. It differs from the canonical
        // estimate by the association of the sums: a few ulps of the magnitudes involved, covered by slack.
        const uint64_t kk = best[k - 1];
        float tp = __int_as_float(0x7f800000);
        if (kk != KEY_INF) {
            const float T = key_value(kk);
            const float lim = metric == B200VQ_METRIC_L2 ? T - qc2 : 2.f * T + qn2 - qc2;
            const float mag = fabsf(c1) * (float)((1 << qb) - 1) * (float)D + fabsf(c2) * (float)D + fabsf(c34);
            const float slack = 9.5367432e-7f * (2.f * fabsf(T) + fabsf(qn2) + 2.f * fabsf(qc2) + amax[0] * mag);
            tp = lim + slack + 1e-30f;
        }
        qscreen[q].w = tp;
        count[q] = 0;
    }
}

}  // namespace rqtc
}  // namespace b200vq

using namespace b200vq;
using namespace b200vq::rqtc;

namespace {
struct RqPlan {
    uint32_t cap;
    int ntq, Dp;
    size_t off_simt, off_rowinfo, off_bmat, off_qscreen, off_running, off_count, off_flags, off_amax, off_cand, total;
};
size_t al256(size_t x) { return (x + 255) & ~(size_t)255; }
RqPlan plan_rq(int64_t n, int D, int nq, int k) {
    RqPlan f;
    int64_t cap = 64 * (int64_t)k;
    if (cap < 4096) cap = 4096;
    f.cap = (uint32_t)cap;
    f.ntq = (nq + TILE_Q - 1) / TILE_Q;
    f.Dp = (D + CHB - 1) / CHB * CHB;
    size_t o = 0;
    f.off_simt = o;     o += al256(b200vq_rabitq1_scan_scratch_bytes(n, nq, k));
    f.off_rowinfo = o;  o += al256((size_t)(n > 0 ? n : 1) * 16);
    f.off_bmat = o;     o += al256((size_t)f.ntq * TILE_Q * f.Dp);
    f.off_qscreen = o;  o += al256((size_t)f.ntq * TILE_Q * 16);
    f.off_running = o;  o += al256((size_t)nq * k * 8);
    f.off_count = o;    o += al256((size_t)nq * 4);
    f.off_flags = o;    o += al256((size_t)nq * 4);
    f.off_amax = o;     o += 256;
    f.off_cand = o;     o += al256((size_t)nq * f.cap * 8);
    f.total = o;
    return f;
}
bool rq_tc_applies(const uint8_t* codes, int64_t n, int D, int nq, int qb, int k) {
    return n >= MIN_ROWS && D % 64 == 0 && D >= 128 && qb >= 1 && qb <= 4 && nq >= 1 && k <= 128 &&
           (reinterpret_cast<uintptr_t>(codes) & 15) == 0 && n < (1ll << 32);
}

int encode_u8_2d(CUtensorMap* map, const void* ptr, uint64_t rows, uint64_t cols, uint32_t box_rows, uint32_t box_cols) {
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) {
        set_error("rabitq1_scan_tc: cuTensorMapEncodeTiled is not available from the driver");
        return B200VQ_ERR_CUDA;
    }
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {cols};
    cuuint32_t box[2] = {box_cols, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("rabitq1_scan_tc: cuTensorMapEncodeTiled failed (%d)", (int)r);
        return B200VQ_ERR_CUDA;
    }
    return B200VQ_OK;
}

template <int CAP>
int rq_finish_launch(const float4* rowinfo, const float* qfac, const float* amax, int D, int qb, int metric, int k,
                     const uint64_t* cand, uint32_t* count, uint32_t cap, uint64_t* running, float4* qscreen,
                     uint32_t* flags, int last, float* out_vals, uint32_t* out_ids, int nq, cudaStream_t st) {
    const int n2 = next_pow2((FIN_THREADS / 32) * k);
    const size_t smem = (size_t)((FIN_THREADS / 32) * CAP + n2) * 8;
    VQ_CUDA(cudaFuncSetAttribute(rq_finish_kernel<CAP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    rq_finish_kernel<CAP><<<nq, FIN_THREADS, smem, st>>>(rowinfo, qfac, amax, D, qb, metric, k, cand, count, cap, running,
                                                        qscreen, flags, last, out_vals, out_ids);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}
}  // namespace

extern "C" size_t b200vq_rabitq1_scan_tc_scratch_bytes(int64_t n, int D, int nq, int k) {
    if (n < 0 || nq <= 0 || k <= 0 || D <= 0) return 0;
    return plan_rq(n, D, nq, k).total;
}

extern "C" int b200vq_rabitq1_scan_topk_tc(const uint8_t* codes, int64_t n, int D, const uint32_t* planes,
                                           const float* qfac, int nq, int qb, int metric, int k, float* out_vals,
                                           uint32_t* out_ids, void* scratch, size_t scratch_bytes, void* stream) {
    VQ_CHECK_ARG(codes && planes && qfac && out_vals && out_ids, "rabitq1_scan_topk_tc: null pointer");
    VQ_CHECK_ARG(n >= 0 && n < (1ll << 32) && D > 0 && nq >= 0, "rabitq1_scan_topk_tc: bad shape");
    VQ_CHECK_ARG(qb >= 1 && qb <= 8, "rabitq1_scan_topk_tc: qb must be in [1,8]");
    VQ_CHECK_ARG(k > 0 && k <= 128, "rabitq1_scan_topk_tc: k must be in [1,128], got %d", k);
    VQ_CHECK_ARG(metric == B200VQ_METRIC_L2 || metric == B200VQ_METRIC_IP, "rabitq1_scan_topk_tc: bad metric");
    if (nq == 0) return B200VQ_OK;
    const RqPlan fp = plan_rq(n, D, nq, k);
    VQ_CHECK_ARG(scratch && scratch_bytes >= fp.total, "rabitq1_scan_topk_tc: scratch too small (%zu < %zu)",
                 scratch_bytes, fp.total);
    uint8_t* sb = reinterpret_cast<uint8_t*>(scratch);
    if (!rq_tc_applies(codes, n, D, nq, qb, k))      // shapes the tensor-core path does not cover
        return b200vq_rabitq1_scan_topk(codes, n, D, planes, qfac, nq, qb, metric, k, out_vals, out_ids,
                                        sb + fp.off_simt, fp.off_rowinfo - fp.off_simt, stream);
    cudaStream_t st = as_stream(stream);
    float4* rowinfo = reinterpret_cast<float4*>(sb + fp.off_rowinfo);
    uint8_t* bmat = sb + fp.off_bmat;
    float4* qscreen = reinterpret_cast<float4*>(sb + fp.off_qscreen);
    uint64_t* running = reinterpret_cast<uint64_t*>(sb + fp.off_running);
    uint32_t* count = reinterpret_cast<uint32_t*>(sb + fp.off_count);
    uint32_t* flags = reinterpret_cast<uint32_t*>(sb + fp.off_flags);
    float* amax = reinterpret_cast<float*>(sb + fp.off_amax);
    uint64_t* cand = reinterpret_cast<uint64_t*>(sb + fp.off_cand);
    const int sms = device_sm_count();

    VQ_CUDA(cudaMemsetAsync(amax, 0, 256, st));
    VQ_CUDA(cudaMemsetAsync(running, 0xFF, (size_t)nq * k * 8, st));     // KEY_INF
    VQ_CUDA(cudaMemsetAsync(count, 0, (size_t)nq * 4, st));
    VQ_CUDA(cudaMemsetAsync(flags, 0, (size_t)nq * 4, st));
    rq_rowinfo_kernel<<<sms * 8, 256, 0, st>>>(codes, n, D, rowinfo, amax);
    VQ_LAUNCH_CHECK();
    rq_prep_kernel<<<fp.ntq * TILE_Q, 128, 0, st>>>(planes, qfac, nq, D, fp.Dp, qb, bmat, qscreen);
    VQ_LAUNCH_CHECK();

    CUtensorMap map_b;
    int rc;
    if ((rc = encode_u8_2d(&map_b, bmat, (uint64_t)fp.ntq * TILE_Q, (uint64_t)fp.Dp, TILE_Q, CHB))) return rc;
    VQ_CUDA(cudaFuncSetAttribute(rq_screen_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
    Params p;
    p.codes = codes; p.rowinfo = rowinfo; p.qscreen = qscreen; p.cand = cand; p.count = count; p.cap = fp.cap;
    p.code_size = D / 8 + 8; p.nchunk = fp.Dp / CHB; p.ntq = fp.ntq; p.D = D;
    p.dbg = nullptr;
#ifdef B200VQ_DEBUG_BUILD   // wait-cycle counters of the MMA thread (tools/ builds only)
    const bool debug = getenv("B200VQ_RQ_DEBUG") != nullptr;
#else
    const bool debug = false;
#endif
    unsigned long long* dbg_dev = reinterpret_cast<unsigned long long*>(amax + 8);   // inside the 256-byte amax slot
    for (int64_t r = 0; r < n;) {
        int64_t r1 = r == 0 ? SEED_ROWS : r * 4;   // x4: about 3k expected survivors per query and stage
        if (r1 > n || n - r1 < r1 / 2) r1 = n;     // fold a short tail into the last stage
        p.r0 = r; p.r1 = r1; p.seed = r == 0;
        p.ntr = (r1 - r + TILE_ROWS - 1) / TILE_ROWS;
        const int64_t units = p.ntr * p.ntq;
        p.dbg = (debug && r1 == n) ? dbg_dev : nullptr;
        rq_screen_kernel<<<(unsigned)(units < sms ? units : sms), THREADS, SMEM_BYTES, st>>>(map_b, p);
        VQ_LAUNCH_CHECK();
        if (p.dbg) {
            unsigned long long h[4];
            VQ_CUDA(cudaMemcpyAsync(h, dbg_dev, sizeof(h), cudaMemcpyDeviceToHost, st));
            VQ_CUDA(cudaStreamSynchronize(st));
            fprintf(stderr, "rq_screen last stage: MMA-thread cycles summed over CTAs: t_empty %llu, b_full %llu, conv %llu, total %llu\n",
                    h[0], h[1], h[2], h[3]);
            VQ_CUDA(cudaMemsetAsync(dbg_dev, 0, 32, st));
        }
        const int last = r1 == n;
        rc = (k <= 32) ? rq_finish_launch<64>(rowinfo, qfac, amax, D, qb, metric, k, cand, count, fp.cap, running, qscreen,
                                              flags, last, out_vals, out_ids, nq, st)
                       : rq_finish_launch<256>(rowinfo, qfac, amax, D, qb, metric, k, cand, count, fp.cap, running, qscreen,
                                               flags, last, out_vals, out_ids, nq, st);
        if (rc != B200VQ_OK) return rc;
        r = r1;
    }
    // queries whose list overflowed in some stage: exact SIMT scan over all rows (CTAs of other queries exit at once)
    return rabitq1_scan_gated(codes, n, D, planes, qfac, nq, qb, metric, k, out_vals, out_ids, sb + fp.off_simt, flags, st);
}
