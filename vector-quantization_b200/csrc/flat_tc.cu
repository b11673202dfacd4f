// Exact flat search (ground truth) on the sm_100a tensor cores.
//
// Replaces faiss IndexFlatL2 / IndexFlatIP.search as driven by the reference at
// src/haag_vq/benchmarks/exact_search.py:44-51,72-77, search_bench.py:42-56 and
// data/datasets.py:21-34; same contract and the SAME results as the exact fp32 SIMT kernel
// b200vq_flat_topk (scan.cu): that kernel's arithmetic (sequential fp32 FMA chain over d,
// (|q|^2 + |x|^2) - 2 q.x clamped at 0, (value, id) order) is what the final scores are
// computed with. The tensor cores only SCREEN:
//
//   P[i][q] = x_i . q   with tcgen05.mma kind::tf32 (operands truncated to 10 mantissa bits),
//   128 rows x 256 queries per accumulator, K = D streamed through a 4-stage TMA ring.
//   |P (1 + 2^-10) - x_i.q| <= E_i = (2^-10 * 1.02 + 3 D 2^-24) |x_i| max_q|q|  (truncation is
//   one-sided, hence the centring factor; the D 2^-24 terms cover both fp32 accumulations),
//   so LB = a_i P + b_i is a rigorous lower bound of the exact key of (row i, query q) and the
//   row can only enter the top-k if LB <= T_q. The same bound gives UB = LB + 4 E_i (L2) / LB + 2 E_i (IP).
//
// Rows are processed in geometric stages (2048 rows, then x8 per stage). Each stage screens its rows against T_q and
// appends the survivors WITH their lower bounds to per-query lists; flat_pool_kernel merges them into the query's
// candidate pool = every row seen so far with LB <= U_q, where U_q = the k-th smallest UB (an upper bound of the final
// k-th best key), and U_q becomes T_q of the next stage. No row is re-read between the stages (round 1 scored every
// survivor exactly after every stage: k ln(growth) rows of 6 KB per query and stage, 25 % of the search at 1M rows and
// 75 % at 100K). After the last stage flat_finish_kernel scores the pool rows - k plus the few whose [LB, UB]
// straddles U_q - exactly, once, and ranks them in (key, id) order. A query whose list or pool overflows (thousands
// of rows within the TF32 uncertainty of its k-th best) is flagged and redone by the SIMT kernel (gated launch), so
// the result never depends on the screen. x is read from HBM once (the query tiles of a row tile are consumed back to
// back, out of L2).
#include "tc_common.cuh"
#include "topk.cuh"

namespace b200vq {
namespace ftc {
using namespace tc;

constexpr int TILE_ROWS = 128;
constexpr int TILE_Q = 256;
constexpr int CH = 32;                       // K chunk: 32 floats = one 128-byte swizzle row
constexpr int NST = 4;
constexpr int EPI_WARPS = 8;                 // 2 accumulator stages x 4 TMEM lane quarters
constexpr int THREADS = 64 + 32 * EPI_WARPS;
constexpr uint32_t A_BYTES = TILE_ROWS * CH * 4;
constexpr uint32_t B_BYTES = TILE_Q * CH * 4;
constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
constexpr uint32_t BAR_OFF = NST * STAGE_BYTES;
constexpr uint32_t THR_OFF = BAR_OFF + 256;
constexpr uint32_t SMEM_BYTES = THR_OFF + 2 * TILE_Q * 4 + 1024;   // + alignment slack
enum { B_FULL = 0, B_EMPTY = NST, B_TFULL = 2 * NST, B_TEMPTY = 2 * NST + 2, B_TMEM = 2 * NST + 4 };

struct Params {
    const float* xnorm;       // (n) squared norms of the raw rows
    const float* row_scale;   // (n) or null
    const float* thr;         // (ntq * 256) T_q, NaN for padding queries
    const float* qstats;      // [0] = max_q |q|^2
    uint32_t* cand;           // (nq, cap)
    float* cand_lb;           // (nq, cap) the survivors' lower bounds
    uint32_t* count;          // (nq)
    uint32_t cap;
    int64_t r0, r1;           // rows of this stage
    int nq, nchunk, ntq;      // queries, D / 32, query tiles
    int64_t ntr;              // row tiles of this stage
    int metric;
    float ebound;             // 2^-10 * 1.02 + 3 D 2^-24
    int dense;                // stage 0: every row is a survivor, slot = row - r0 (no counters, no atomics)
    const uint32_t* rowcluster;   // IVF mode (b200vq_ivf_search_tc): cluster of every row, or null
    const uint32_t* probe_bits;   // IVF mode: (nq, kwords) bitmap of the clusters each query probes
    int kwords;
};

__global__ void __launch_bounds__(THREADS, 1)
flat_screen_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_q, const Params p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;
    uint8_t* gbase = smem_raw + (base - raw);
    const uint32_t bar0 = base + BAR_OFF;
    auto bar = [&](int slot) { return bar0 + 8u * (uint32_t)slot; };
    volatile uint32_t* tmem_ptr_smem = reinterpret_cast<volatile uint32_t*>(gbase + BAR_OFF + 8 * B_TMEM);
    float* thr_s = reinterpret_cast<float*>(gbase + THR_OFF);    // [2][TILE_Q]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t nunits = p.ntr * p.ntq;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&map_x);
        tma_prefetch_desc(&map_q);
        for (int i = 0; i < NST; ++i) {
            mbar_init(bar(B_FULL + i), 1);
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
        if (lane == 0) {
            uint32_t st = 0, phase = 0;
            for (int64_t unit = blockIdx.x; unit < nunits; unit += gridDim.x) {
                const int64_t t = unit / p.ntq;
                const int u = (int)(unit - t * p.ntq);
                const int row0 = (int)(p.r0 + t * TILE_ROWS);
                for (int c = 0; c < p.nchunk; ++c) {
                    mbar_wait_backoff(bar(B_EMPTY + st), phase ^ 1);
                    mbar_expect_tx(bar(B_FULL + st), STAGE_BYTES);
                    tma_load_2d(base + st * STAGE_BYTES, &map_x, bar(B_FULL + st), c * CH, row0);
                    tma_load_2d(base + st * STAGE_BYTES + A_BYTES, &map_q, bar(B_FULL + st), c * CH, u * TILE_Q);
                    if (++st == NST) { st = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            uint32_t st = 0, phase = 0, acc = 0, t_phase = 0;
            for (int64_t unit = blockIdx.x; unit < nunits; unit += gridDim.x) {
                mbar_wait_backoff(bar(B_TEMPTY + acc), t_phase ^ 1);
                tc_fence_after();
                const uint32_t d = tmem_base + acc * 256u;
                uint32_t accum = 0;
                for (int c = 0; c < p.nchunk; ++c) {
                    mbar_wait_backoff(bar(B_FULL + st), phase);
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
    } else {
        const uint32_t acc = (uint32_t)(warp - 2) >> 2;      // accumulator stage served by this warp
        const int q4 = warp & 3;                              // TMEM lane quarter this warp may read
        const uint32_t lane_addr = tmem_base + ((uint32_t)(q4 * 32) << 16) + acc * 256u;
        float* ts = thr_s + acc * TILE_Q;
        const int gt = (warp - 2 - (int)acc * 4) * 32 + lane;  // 0..127 inside the group
        const uint32_t grp_bar = 1u + acc;
        const float qn_max = p.qstats[0], q_max = sqrtf(qn_max) * 1.0000005f;
        uint32_t t_phase = 0;
        int64_t unit = blockIdx.x;
        for (uint32_t it = 0; unit < nunits; unit += gridDim.x, ++it) {
            if ((it & 1u) != acc) continue;
            const int64_t t = unit / p.ntq;
            const int u = (int)(unit - t * p.ntq);
            const int64_t row = p.r0 + t * TILE_ROWS + q4 * 32 + lane;
            const bool live = row < p.r1;
            float a = 0.f, b = 0.f;
            const uint32_t rc = (live && p.rowcluster) ? p.rowcluster[row] : 0u;
            if (live) {
                const float xn = p.xnorm[row];
                const float rs = p.row_scale ? p.row_scale[row] : 1.f;
                const float nrm = sqrtf(xn) * fabsf(rs) * 1.0000005f;
                const float E = p.ebound * nrm * q_max + 2.4e-7f * (xn * rs * rs + qn_max);
                if (p.metric == B200VQ_METRIC_L2) {
                    a = -2.f * 1.0009765625f * rs;
                    b = xn - 2.f * E;
                } else {
                    a = -1.0009765625f * rs;
                    b = -E;
                }
            }
            asm volatile("bar.sync %0, 128;" ::"r"(grp_bar) : "memory");   // previous unit's thresholds consumed
            ts[gt] = p.thr[(size_t)u * TILE_Q + gt];
            ts[gt + 128] = p.thr[(size_t)u * TILE_Q + gt + 128];
            asm volatile("bar.sync %0, 128;" ::"r"(grp_bar) : "memory");
            mbar_wait(bar(B_TFULL + acc), t_phase);
            tc_fence_after();
            float v[64];
#pragma unroll 1
            for (int cb = 0; cb < 4; ++cb) {
                tmem_ld64(lane_addr + cb * 64, v);
                const float4* t4 = reinterpret_cast<const float4*>(ts + cb * 64);
#pragma unroll
                for (int j4 = 0; j4 < 16; ++j4) {
                    const float4 tt = t4[j4];
                    const float th[4] = {tt.x, tt.y, tt.z, tt.w};
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float lb = __fmaf_rn(a, v[4 * j4 + e], b);
                        if (p.dense) {
                            const int q = u * TILE_Q + cb * 64 + 4 * j4 + e;
                            if (live && q < p.nq) {
                                p.cand[(size_t)q * p.cap + (uint32_t)(row - p.r0)] = (uint32_t)row;
                                p.cand_lb[(size_t)q * p.cap + (uint32_t)(row - p.r0)] = lb;
                            }
                        } else if (live && lb <= th[e]) {
                            const int q = u * TILE_Q + cb * 64 + 4 * j4 + e;
                            // IVF mode: the row only counts for queries that probe its cluster (tested on the few
                            // survivors, not on every accumulator value)
                            if (p.probe_bits && !((p.probe_bits[(size_t)q * p.kwords + (rc >> 5)] >> (rc & 31u)) & 1u)) continue;
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

// |q|^2 of every query with the SIMT kernel's arithmetic (lane-strided FMA chain + xor tree), and their maximum.
__global__ void flat_qstats_kernel(const float* __restrict__ queries, int nq, int64_t ldq, int D, float* __restrict__ qn,
                                   float* __restrict__ qstats, float* __restrict__ thr,
                                   const float* __restrict__ init_vals, int metric) {
    const int lane = threadIdx.x & 31;
    const int q = (int)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    if (q >= nq) return;
    const float* pq = queries + (int64_t)q * ldq;
    float acc = 0.f;
    for (int i = lane; i < D; i += 32) acc = __fmaf_rn(pq[i], pq[i], acc);
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(FULL_MASK, acc, o);
    if (lane == 0) {
        qn[q] = acc;
        // first stage: every row passes (padding queries keep their NaN), unless the caller knows an upper bound of
        // the k-th best value already (IVF mode: the exact k-th best of the nearest cluster)
        float t = __int_as_float(0x7f800000);
        if (init_vals && init_vals[q] < t && init_vals[q] > -t) t = metric == B200VQ_METRIC_L2 ? init_vals[q] - acc : -init_vals[q];
        thr[q] = t;
        atomicMax(reinterpret_cast<int*>(qstats), __float_as_int(acc));   // non-negative floats order as ints
    }
}

// ---------------------------------------------------------------- candidate pool ---
// Between the stages NO row is re-read. Every survivor of a screen carries its rigorous lower bound lb of the exact
// key; the same error bound gives ub = lb + 4E (L2) / lb + 2E (IP) >= exact key. Per query the pool keeps every row
// seen so far that can still be in the top-k: with U = the k-th smallest ub (an upper bound of the final k-th best
// key), exactly the rows with lb <= U. U is the next stage's screening threshold. After the last stage
// flat_finish_kernel scores the pool rows (k plus the few rows whose interval straddles U) exactly, once.
// One CTA per query; pool arrays are double-buffered by stage parity. A pool or entry overflow (thousands of rows
// within the TF32 uncertainty of the k-th best: duplicated rows) flags the query for the gated exact SIMT search.
constexpr int POOL_CAP = 1024;
constexpr int POOL_MAX_ENTRIES = 8192;
__global__ void __launch_bounds__(256)
flat_pool_kernel(const float* __restrict__ xnorm, const float* __restrict__ row_scale, const float* __restrict__ qstats,
                 float ebound, int metric, int k, const uint32_t* __restrict__ cand, const float* __restrict__ cand_lb,
                 float* __restrict__ cand_ub, uint32_t* __restrict__ count, uint32_t cap,
                 const uint32_t* __restrict__ prow_in, const float* __restrict__ plb_in, const float* __restrict__ pub_in,
                 uint32_t* __restrict__ prow_out, float* __restrict__ plb_out, float* __restrict__ pub_out,
                 uint32_t* __restrict__ pool_cnt, float* __restrict__ thr, uint32_t* __restrict__ flags,
                 uint32_t dense_count) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint32_t* keys = reinterpret_cast<uint32_t*>(smem_raw);
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
    const float qn_max = qstats[0], q_max = sqrtf(qn_max) * 1.0000005f;
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
            const float xn = xnorm[row];
            const float rs = row_scale ? row_scale[row] : 1.f;
            const float nrm = sqrtf(xn) * fabsf(rs) * 1.0000005f;
            const float slack = 2.4e-7f * (xn * rs * rs + qn_max);
            const float E = ebound * nrm * q_max + slack;      // the screen's own bound (flat_screen_kernel)
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
        if (overflow) flags[q] = 1;     // redone by the gated SIMT launch
    }
}

constexpr int FIN_THREADS = 256;
constexpr int FIN_TILE_LD = 36;             // padded tile row (floats): conflict-free 128-bit stores and loads
// One CTA per query: exact keys of the listed rows (the SIMT kernel's arithmetic), merged with the running list.
template <int CAP>
__global__ void __launch_bounds__(FIN_THREADS)
flat_finish_kernel(const float* __restrict__ x, int64_t ldx, int D, const float* __restrict__ xnorm,
                   const float* __restrict__ row_scale, const float* __restrict__ queries, int64_t ldq, int metric, int k,
                   const uint32_t* __restrict__ cand, uint32_t* __restrict__ count, uint32_t cap,
                   uint64_t* __restrict__ running, float* __restrict__ thr, uint32_t* __restrict__ flags, int last,
                   float* __restrict__ out_vals, uint32_t* __restrict__ out_ids, uint32_t seed_rows,
                   const uint32_t* __restrict__ idmap = nullptr) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ uint64_t best[128];
    __shared__ float qn_s;
    const int q = blockIdx.x;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    float* qs = reinterpret_cast<float*>(smem_raw);                       // [D]
    uint64_t* wbuf = reinterpret_cast<uint64_t*>(smem_raw + (((size_t)D * 4 + 15) & ~(size_t)15));
    const int n2 = next_pow2((FIN_THREADS / 32) * k);
    uint64_t* mbuf = wbuf + (FIN_THREADS / 32) * CAP;
    const float* pq = queries + (int64_t)q * ldq;
    for (int i = threadIdx.x; i < D; i += FIN_THREADS) qs[i] = pq[i];
    if (w == 0) {
        float acc = 0.f;
        for (int i = lane; i < D; i += 32) acc = __fmaf_rn(pq[i], pq[i], acc);
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(FULL_MASK, acc, o);
        if (lane == 0) qn_s = acc;
    }
    __syncthreads();
    const float qn = qn_s;
    // seed_rows > 0: the "list" is rows [0, seed_rows) (first stage: exact top-k of the first rows)
    uint32_t cnt = seed_rows ? seed_rows : count[q];
    if (!seed_rows && cnt > cap) {
        if (threadIdx.x == 0) flags[q] = 1;     // redone by the gated SIMT launch
        cnt = cap;
    }
    WarpTopK<CAP> tk;
    tk.init(wbuf + w * CAP, k);
    // the running list first
    for (int e0 = 0; e0 < k; e0 += FIN_THREADS) {
        const int e = e0 + threadIdx.x;
        uint64_t key = KEY_INF;
        if (e < k) key = running[(size_t)q * k + e];
        tk.offer(key != KEY_INF, key_value(key), key_id(key));
    }
    const uint32_t* list = cand + (size_t)q * cap;
    // One warp per group of 32 candidates. The rows are fetched with coalesced 128-byte
    // requests (8 lanes x float4 per row chunk, four rows per instruction) into a padded
    // warp-private tile and each lane then runs the FMA chain of ITS candidate over the tile
    // row: the chain is the SIMT kernel's (d = 0 .. D-1, one accumulator), only the loads
    // are shared. (A lane-per-row gather moved 32-byte sectors for 16-byte loads and sat on
    // the L2 sector rate.)
    float* tile = reinterpret_cast<float*>(mbuf + n2) + w * (32 * FIN_TILE_LD);
    const uint32_t ngroups = (cnt + 31) / 32;
    const int sub = lane >> 3, l8 = lane & 7;
    const int nchunk = D / 32;
    for (uint32_t g = w; g < ngroups; g += FIN_THREADS / 32) {
        const uint32_t e = g * 32 + lane;
        const bool valid = e < cnt;
        const uint32_t row = valid ? (seed_rows ? e : list[e]) : 0u;
        const float rs = (row_scale && valid) ? row_scale[row] : 1.f;
        const float4* src[8];
        bool ok[8];
        float4 cur[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const uint32_t rj = __shfl_sync(FULL_MASK, row, 4 * j + sub);
            ok[j] = g * 32 + 4 * j + sub < cnt;
            src[j] = reinterpret_cast<const float4*>(x + (int64_t)rj * ldx) + l8;
            cur[j] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (ok[j]) cur[j] = __ldg(src[j]);
        }
        float acc = 0.f;
        for (int c = 0; c < nchunk; ++c) {
            __syncwarp();
#pragma unroll
            for (int j = 0; j < 8; ++j)
                *reinterpret_cast<float4*>(tile + (4 * j + sub) * FIN_TILE_LD + 4 * l8) = cur[j];
            if (c + 1 < nchunk) {
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    if (ok[j]) cur[j] = __ldg(src[j] + (c + 1) * 8);
            }
            __syncwarp();
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                float4 xv = *reinterpret_cast<const float4*>(tile + lane * FIN_TILE_LD + 4 * i);
                if (row_scale) {
                    xv.x = __fmul_rn(xv.x, rs); xv.y = __fmul_rn(xv.y, rs);
                    xv.z = __fmul_rn(xv.z, rs); xv.w = __fmul_rn(xv.w, rs);
                }
                const float4 qv = *reinterpret_cast<const float4*>(qs + c * 32 + 4 * i);
                acc = __fmaf_rn(qv.x, xv.x, acc);
                acc = __fmaf_rn(qv.y, xv.y, acc);
                acc = __fmaf_rn(qv.z, xv.z, acc);
                acc = __fmaf_rn(qv.w, xv.w, acc);
            }
        }
        float key;
        if (metric == B200VQ_METRIC_L2) key = fmaxf(__fmaf_rn(-2.f, acc, __fadd_rn(qn, valid ? xnorm[row] : 0.f)), 0.f);
        else key = -acc;
        tk.offer(valid, key, (idmap && valid) ? idmap[row] : row);    // IVF mode: ranked and reported by global id
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
        if (kk != KEY_INF) t = (metric == B200VQ_METRIC_IP) ? key_value(kk) : key_value(kk) - qn;
        thr[q] = t;
        count[q] = 0;
    }
}

}  // namespace ftc
}  // namespace b200vq

using namespace b200vq;
using namespace b200vq::ftc;

namespace {
constexpr int64_t FLAT_STAGE_GROWTH = 8;   // rows seen after a stage / rows seen before it. With the candidate pool a stage costs a
                                          // screen launch + a 30 us pool kernel and k ln(growth) appended survivors per query:
                                          // 3 -> 5.14 / 1.69 ms, 8 -> 4.64 / 1.43 ms (1M x 1536 k = 10 / 100K x 1536 k = 100, 1000 queries)
constexpr int64_t FIRST_ROWS = 2048;  // stage 0 (infinite threshold: every row enters the pool kernel, dense slots, no atomics)
constexpr int64_t MIN_ROWS = 1024;    // below this the SIMT kernel alone is as fast (one dense stage covers up to 2048 rows)
struct FlatPlan {
    uint32_t cap;
    int ntq;
    size_t off_simt, off_running, off_thr, off_count, off_flags, off_qn, off_qstats, off_cand, off_cand_lb, off_cand_ub,
        off_pool, off_pool_cnt, total;
};
FlatPlan plan_flat(int64_t n, int nq, int k) {
    FlatPlan f;
    f.ntq = (nq + TILE_Q - 1) / TILE_Q;
    int64_t cap = 64 * (int64_t)k;
    if (cap < 4096) cap = 4096;
    f.cap = (uint32_t)cap;
    auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
    size_t o = 0;
    f.off_simt = o;    o = up(o + b200vq_flat_topk_scratch_bytes(n, nq, k));
    f.off_running = o; o = up(o + (size_t)nq * k * 8);
    f.off_thr = o;     o = up(o + (size_t)f.ntq * TILE_Q * 4);
    f.off_count = o;   o = up(o + (size_t)nq * 4);
    f.off_flags = o;   o = up(o + (size_t)nq * 4);
    f.off_qn = o;      o = up(o + (size_t)nq * 4);
    f.off_qstats = o;  o = up(o + 16);
    f.off_cand = o;    o = up(o + (size_t)nq * f.cap * 4);
    f.off_cand_lb = o; o = up(o + (size_t)nq * f.cap * 4);
    f.off_cand_ub = o; o = up(o + (size_t)nq * f.cap * 4);
    f.off_pool = o;    o = up(o + (size_t)nq * POOL_CAP * 4 * 6);     // rows, lb, ub: two generations
    f.off_pool_cnt = o; o = up(o + (size_t)nq * 4);
    f.total = o;
    return f;
}
bool tc_applies(const float* x, int64_t n, int64_t ldx, int D, const float* xnorm, const float* row_scale,
                const float* queries, int nq, int64_t ldq, int metric, int k) {
    return n >= MIN_ROWS && n < (1ll << 31) && D % 32 == 0 && D <= 8192 && ldx % 4 == 0 && ldq % 4 == 0 &&
           ((uintptr_t)x & 15) == 0 && ((uintptr_t)queries & 15) == 0 && xnorm != nullptr && nq >= 1 && k <= 128 &&
           !(metric == B200VQ_METRIC_L2 && row_scale != nullptr);
}
}  // namespace

extern "C" size_t b200vq_flat_topk_tc_scratch_bytes(int64_t n, int nq, int k) {
    if (n < 0 || nq <= 0 || k <= 0) return 0;
    return plan_flat(n, nq, k).total;
}

template <int CAP>
static int finish_launch(const float* x, int64_t ldx, int D, const float* xnorm, const float* row_scale,
                         const float* queries, int nq, int64_t ldq, int metric, int k, const uint32_t* cand,
                         uint32_t* count, uint32_t cap, uint64_t* running, float* thr, uint32_t* flags, int last,
                         float* out_vals, uint32_t* out_ids, cudaStream_t st, uint32_t seed_rows = 0,
                         const uint32_t* idmap = nullptr) {
    const int n2 = next_pow2((FIN_THREADS / 32) * k);
    const size_t smem = (((size_t)D * 4 + 15) & ~(size_t)15) + (size_t)((FIN_THREADS / 32) * CAP + n2) * 8 +
                        (size_t)(FIN_THREADS / 32) * 32 * FIN_TILE_LD * 4;
    VQ_CUDA(cudaFuncSetAttribute(flat_finish_kernel<CAP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    flat_finish_kernel<CAP><<<nq, FIN_THREADS, smem, st>>>(x, ldx, D, xnorm, row_scale, queries, ldq, metric, k, cand,
                                                         count, cap, running, thr, flags, last, out_vals, out_ids,
                                                         seed_rows, idmap);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

// IVF mode of the same search (b200vq_ivf_search_tc): rows = cluster-sorted reconstructions, a row only counts for the
// queries whose probe list holds its cluster, ids are mapped to global ids, flagged queries are reported to the caller.
struct IvfMode {
    const uint32_t* rowcluster;
    const uint32_t* probe;     // (nq, nprobe)
    int nprobe, K;
    const uint32_t* gid;
    uint32_t* flags_out;       // (nq)
    uint32_t* bits;            // (nq, kwords) scratch
    const float* init_vals;    // (nq) or null: an upper bound of every query's k-th best value
};
__global__ void ivf_bits_kernel(const uint32_t* __restrict__ probe, int64_t total, int nprobe, int K, int kwords,
                                uint32_t* __restrict__ bits) {
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const uint32_t c = probe[e];
        if (c < (uint32_t)K) atomicOr(bits + (e / nprobe) * kwords + (c >> 5), 1u << (c & 31u));
    }
}

static int flat_tc_run(const float* x, int64_t n, int64_t ldx, int D, const float* xnorm,
                       const float* row_scale, const float* queries, int nq, int64_t ldq, int metric,
                       int k, float* out_vals, uint32_t* out_ids, void* scratch, size_t scratch_bytes,
                       void* stream, const IvfMode* ivf) {
    VQ_CHECK_ARG(x && queries && out_vals && out_ids, "flat_topk_tc: null pointer");
    VQ_CHECK_ARG(n >= 0 && n < (1ll << 32) && D > 0 && ldx >= D && ldq >= D && nq >= 0, "flat_topk_tc: bad shape");
    VQ_CHECK_ARG(k > 0 && k <= 128, "flat_topk_tc: k must be in [1,128], got %d", k);
    VQ_CHECK_ARG(metric == B200VQ_METRIC_L2 || metric == B200VQ_METRIC_IP, "flat_topk_tc: bad metric");
    VQ_CHECK_ARG(metric == B200VQ_METRIC_IP || xnorm, "flat_topk_tc: L2 needs xnorm");
    if (nq == 0) return B200VQ_OK;
    const FlatPlan fp = plan_flat(n, nq, k);
    VQ_CHECK_ARG(scratch && scratch_bytes >= fp.total, "flat_topk_tc: scratch too small (%zu < %zu)", scratch_bytes,
                 fp.total);
    uint8_t* sb = reinterpret_cast<uint8_t*>(scratch);
    if (ivf) {
        VQ_CHECK_ARG(tc_applies(x, n, ldx, D, xnorm, row_scale, queries, nq, ldq, metric, k),
                     "ivf_search_tc: shape not covered by the tensor-core screen (n >= 65536, D %% 32 == 0, aligned rows)");
    } else if (!tc_applies(x, n, ldx, D, xnorm, row_scale, queries, nq, ldq, metric, k))   // shapes the screen does not cover
        return b200vq_flat_topk(x, n, ldx, D, xnorm, row_scale, queries, nq, ldq, metric, k, out_vals, out_ids,
                                sb + fp.off_simt, fp.off_running - fp.off_simt, stream);
    cudaStream_t st = as_stream(stream);
    uint64_t* running = reinterpret_cast<uint64_t*>(sb + fp.off_running);
    float* thr = reinterpret_cast<float*>(sb + fp.off_thr);
    uint32_t* count = reinterpret_cast<uint32_t*>(sb + fp.off_count);
    uint32_t* flags = reinterpret_cast<uint32_t*>(sb + fp.off_flags);
    float* qn = reinterpret_cast<float*>(sb + fp.off_qn);
    float* qstats = reinterpret_cast<float*>(sb + fp.off_qstats);
    uint32_t* cand = reinterpret_cast<uint32_t*>(sb + fp.off_cand);

    float* cand_lb = reinterpret_cast<float*>(sb + fp.off_cand_lb);
    float* cand_ub = reinterpret_cast<float*>(sb + fp.off_cand_ub);
    uint32_t* pool_cnt = reinterpret_cast<uint32_t*>(sb + fp.off_pool_cnt);
    const size_t pstride = (size_t)nq * POOL_CAP;
    uint32_t* prow[2] = {reinterpret_cast<uint32_t*>(sb + fp.off_pool), reinterpret_cast<uint32_t*>(sb + fp.off_pool) + pstride};
    float* plb[2] = {reinterpret_cast<float*>(sb + fp.off_pool) + 2 * pstride, reinterpret_cast<float*>(sb + fp.off_pool) + 3 * pstride};
    float* pub[2] = {reinterpret_cast<float*>(sb + fp.off_pool) + 4 * pstride, reinterpret_cast<float*>(sb + fp.off_pool) + 5 * pstride};

    // query norms and their maximum; thresholds +inf (NaN for the padding queries: they never pass)
    VQ_CUDA(cudaMemsetAsync(qstats, 0, 16, st));
    VQ_CUDA(cudaMemsetAsync(thr, 0xFF, (size_t)fp.ntq * TILE_Q * 4, st));
    VQ_CUDA(cudaMemsetAsync(running, 0xFF, (size_t)nq * k * 8, st));        // KEY_INF
    VQ_CUDA(cudaMemsetAsync(flags, 0, (size_t)nq * 4, st));
    VQ_CUDA(cudaMemsetAsync(count, 0, (size_t)nq * 4, st));
    VQ_CUDA(cudaMemsetAsync(pool_cnt, 0, (size_t)nq * 4, st));
    flat_qstats_kernel<<<(unsigned)(((int64_t)nq * 32 + 255) / 256), 256, 0, st>>>(queries, nq, ldq, D, qn, qstats, thr,
                                                                                   ivf ? ivf->init_vals : nullptr, metric);
    VQ_LAUNCH_CHECK();

    CUtensorMap map_x, map_q;
    int rc;
    if ((rc = encode_2d(&map_x, x, (uint64_t)n, (uint64_t)D, (uint64_t)ldx, TILE_ROWS, CH, CU_TENSOR_MAP_SWIZZLE_128B)))
        return rc;
    if ((rc = encode_2d(&map_q, queries, (uint64_t)nq, (uint64_t)D, (uint64_t)ldq, TILE_Q, CH, CU_TENSOR_MAP_SWIZZLE_128B)))
        return rc;
    VQ_CUDA(cudaFuncSetAttribute(flat_screen_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
    const size_t pool_smem = (size_t)POOL_MAX_ENTRIES * 8;
    VQ_CUDA(cudaFuncSetAttribute(flat_pool_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pool_smem));
    Params p;
    p.xnorm = xnorm; p.row_scale = row_scale; p.thr = thr; p.qstats = qstats; p.cand = cand; p.cand_lb = cand_lb;
    p.count = count;
    p.cap = fp.cap; p.nq = nq; p.nchunk = D / CH; p.ntq = fp.ntq; p.metric = metric;
    p.ebound = 0.0009765625f * 1.02f + 3.0f * (float)D * 5.9604645e-8f;
    p.rowcluster = nullptr; p.probe_bits = nullptr; p.kwords = 0;
    if (ivf) {
        p.kwords = (ivf->K + 31) / 32;
        VQ_CUDA(cudaMemsetAsync(ivf->bits, 0, (size_t)nq * p.kwords * 4, st));
        const int64_t total = (int64_t)nq * ivf->nprobe;
        ivf_bits_kernel<<<(unsigned)((total + 255) / 256 < 4096 ? (total + 255) / 256 : 4096), 256, 0, st>>>(
            ivf->probe, total, ivf->nprobe, ivf->K, p.kwords, ivf->bits);
        VQ_LAUNCH_CHECK();
        p.rowcluster = ivf->rowcluster;
        p.probe_bits = ivf->bits;
    }
    const int sms = device_sm_count();
    // IVF mode: a query admits every probed row until k of them have been seen, so the stages grow slowly (x3): a
    // stage of S rows appends nprobe / K * S rows to such a query's list (4096 slots)
    const int64_t growth = ivf ? 3 : FLAT_STAGE_GROWTH;
    int gen = 0;
    // Stage 0 covers rows [0, FIRST_ROWS) with an infinite threshold (all of them enter the pool kernel, which keeps
    // the k + few that matter); later stages grow x3. No row is re-read until the single exact pass at the end.
    for (int64_t r = 0; r < n;) {
        int64_t r1 = r == 0 ? FIRST_ROWS : r * growth;
        if (r1 > n || n - r1 < r1 / 4) r1 = n;  // fold a short tail into the last stage
        p.r0 = r; p.r1 = r1;
        p.dense = r == 0 && !ivf;
        p.ntr = (r1 - r + TILE_ROWS - 1) / TILE_ROWS;
        const int64_t units = p.ntr * p.ntq;
        flat_screen_kernel<<<(unsigned)(units < sms ? units : sms), THREADS, SMEM_BYTES, st>>>(map_x, map_q, p);
        VQ_LAUNCH_CHECK();
        flat_pool_kernel<<<nq, 256, pool_smem, st>>>(xnorm, row_scale, qstats, p.ebound, metric, k, cand, cand_lb, cand_ub,
                                                     count, fp.cap, prow[gen], plb[gen], pub[gen], prow[gen ^ 1],
                                                     plb[gen ^ 1], pub[gen ^ 1], pool_cnt, thr, flags,
                                                     p.dense ? (uint32_t)(r1 - r) : 0u);
        VQ_LAUNCH_CHECK();
        gen ^= 1;
        r = r1;
    }
    // exact keys of the pool rows, top-k in (key, id) order
    const uint32_t* idmap = ivf ? ivf->gid : nullptr;
    rc = (k <= 32) ? finish_launch<64>(x, ldx, D, xnorm, row_scale, queries, nq, ldq, metric, k, prow[gen], pool_cnt,
                                       (uint32_t)POOL_CAP, running, thr, flags, 1, out_vals, out_ids, st, 0, idmap)
                   : finish_launch<256>(x, ldx, D, xnorm, row_scale, queries, nq, ldq, metric, k, prow[gen], pool_cnt,
                                        (uint32_t)POOL_CAP, running, thr, flags, 1, out_vals, out_ids, st, 0, idmap);
    if (rc != B200VQ_OK) return rc;
    if (ivf) {   // the caller redoes flagged queries with the exact cluster-major kernel (b200vq_ivf_search)
        VQ_CUDA(cudaMemcpyAsync(ivf->flags_out, flags, (size_t)nq * 4, cudaMemcpyDeviceToDevice, st));
        return B200VQ_OK;
    }
    // queries whose survivor list overflowed in some stage: exact SIMT search, gated per query
    return flat_topk_gated(x, n, ldx, D, xnorm, row_scale, queries, nq, ldq, metric, k, out_vals, out_ids,
                           sb + fp.off_simt, flags, st);
}

extern "C" int b200vq_flat_topk_tc(const float* x, int64_t n, int64_t ldx, int D, const float* xnorm,
                                   const float* row_scale, const float* queries, int nq, int64_t ldq, int metric,
                                   int k, float* out_vals, uint32_t* out_ids, void* scratch, size_t scratch_bytes,
                                   void* stream) {
    return flat_tc_run(x, n, ldx, D, xnorm, row_scale, queries, nq, ldq, metric, k, out_vals, out_ids, scratch,
                       scratch_bytes, stream, nullptr);
}

extern "C" size_t b200vq_ivf_search_tc_scratch_bytes(int64_t n, int nq, int K, int k) {
    if (n < 0 || nq <= 0 || k <= 0 || K <= 0) return 0;
    return plan_flat(n, nq, k).total + 256 + (size_t)nq * ((K + 31) / 32) * 4;
}

extern "C" int b200vq_ivf_search_tc(const float* recon, int64_t n, int64_t ldr, int D, const float* rnorm,
                                    const uint32_t* gid, const uint32_t* rowcluster, int K, const float* queries, int nq,
                                    int64_t ldq, const uint32_t* probe, int nprobe, int metric, int k, float* out_vals,
                                    uint32_t* out_ids, uint32_t* flags_out, const float* init_vals, void* scratch,
                                    size_t scratch_bytes, void* stream) {
    VQ_CHECK_ARG(gid && rowcluster && probe && flags_out && rnorm, "ivf_search_tc: null pointer");
    VQ_CHECK_ARG(K > 0 && nprobe > 0 && nq >= 0, "ivf_search_tc: bad shape");
    if (nq == 0) return B200VQ_OK;
    const size_t base = plan_flat(n, nq, k).total;
    const size_t need = b200vq_ivf_search_tc_scratch_bytes(n, nq, K, k);
    VQ_CHECK_ARG(scratch && scratch_bytes >= need, "ivf_search_tc: scratch too small (%zu < %zu)", scratch_bytes, need);
    IvfMode m;
    m.rowcluster = rowcluster; m.probe = probe; m.nprobe = nprobe; m.K = K; m.gid = gid; m.flags_out = flags_out;
    m.init_vals = init_vals;
    m.bits = reinterpret_cast<uint32_t*>(reinterpret_cast<uint8_t*>(scratch) + ((base + 255) & ~(size_t)255));
    return flat_tc_run(recon, n, ldr, D, rnorm, nullptr, queries, nq, ldq, metric, k, out_vals, out_ids, scratch, base,
                       stream, &m);
}
