// PQ / OPQ nearest-centroid assignment on the sm_100a tensor cores.
//
// Replaces faiss ProductQuantizer.compute_codes as driven by the reference at
// src/haag_vq/methods/product_quantization.py:76-80 and
// src/haag_vq/methods/optimized_product_quantization.py:30-31; same contract and
// the same codes as the exact SIMT kernel b200vq_pq_assign_f32 (pq.cu).
//
// Per (128-row tile, sub-quantizer m) one tcgen05.mma chain computes
//     v[r][j] = x_m[r] . (-2 c_mj)  +  ||c_mj||^2          (kind::tf32, fp32 accumulate)
// into a 128 x 256 TMEM accumulator (||c||^2 enters through one extra K=8 slice:
// A = [1 1 1 0..], B = three-way TF32 split of the fp32 norm, so the epilogue needs
// no per-value add). v differs from the true distance by the row constant ||x_m||^2.
//
// TF32 truncates both operands to 10 mantissa bits, so the argmin of v is only a
// SCREEN: |v - (1-2^-10) v_true| <= E with E = 2^-9 * 1.02 * |x_m| * max_j|c_mj| + slack
// (the norm term is pre-scaled by 1-2^-10 to centre the one-sided truncation error).
// The epilogue (one TMEM lane = one row, 128 columns per warp) makes two passes over the accumulator:
//   pass 1   row minimum with 3-input FMNMX                  (0.5 ALU op / value); the two warps of a
//            pair exchange their minima and 2E through shared memory (64-thread named barrier)
//   pass 2   per value FSETP + predicated IMAD-immediate on four independent chains:
//            acc += 1 | j << 8 | j*j << 17   (count, sum j, sum j^2 of the candidates v <= min + 2E)
// which yields the candidate count and, for 1 or 2 candidates, their indices with no branch per
// value. One candidate: it IS the exact argmin. Two: the lower index is stored as the provisional
// code, the other in the byte side array cand2, and pair_fixup_kernel evaluates both with the SIMT
// kernel's exact fp32 arithmetic (sequential FMA chain, (||x||^2 + ||c||^2) - 2 x.c, clamp at 0,
// lower index wins ties). Three or more, or none (NaN rows) (0.3 % of pairs on the benchmark data):
// marker (255, 255); pair_fixup_kernel appends the pair to a per-sub-quantizer list and
// fixup_kernel runs the exact 256-way scan for it. The result is bit-identical to
// b200vq_pq_assign_f32 whenever E bounds the screen error. Nothing in the main kernel waits on
// global memory.
//
// Data movement: persistent CTAs, one per SM. A CTA keeps the (-2 c) tiles of KS consecutive
// sub-quantizers resident in shared memory (TMA, hardware swizzle) and streams 128-row slices of
// x (only the KS*dsub columns it needs) through an NST-deep TMA ring, so x is read from HBM
// exactly once and the centroids are not re-read per tile. Warp roles: 16 = TMA producer,
// 17 = MMA issuer (owns TMEM), 0..15 = epilogue (accumulator stage x TMEM lane quarter x column
// half). The two 256-column accumulators double-buffer MMA against the epilogue.
// Launch sequence: memset(cand2) -> prepare_kernel -> assign_tc_kernel -> pair_fixup_kernel ->
// fixup_kernel, all on the caller's stream.
#include "tc_common.cuh"

namespace b200vq {
namespace tc {

constexpr int TILE_ROWS = 128;
constexpr int KSUB = 256;
constexpr int EPI_WARPS = 16;   // 2 accumulator stages x 4 TMEM lane quarters x 2 column halves
constexpr int THREADS = 64 + 32 * EPI_WARPS;
// Warps 0-15 = epilogue, 16 = TMA producer, 17 = MMA issuer: the warp scheduler favours high warp
// ids, so the two single-thread roles are never queued behind the epilogue's arithmetic.
constexpr int SVC_WARP = EPI_WARPS;
constexpr int MAX_STAGES = 4;
constexpr uint32_t SMEM_BUDGET = 220 * 1024;

// Candidate bookkeeping of the epilogue's second pass, two instructions per value and no branch:
// FSETP (ALU pipe) + predicated IMAD-immediate (FMA pipe):  if (v <= thr) acc += K(J), with
//   K(J) = 1 | J << 8 | J*J << 17   ->  bits [0,8) count, [8,17) sum of J, [17,32) sum of J^2
// (J < 128 is the column index inside the warp's half; the upper fields are exact whenever the
// count is <= 2, the only case that reads them). `one` is a runtime 1 so the add stays an IMAD.
template <int J>
__device__ __forceinline__ void flag_accumulate(float v, float thr, uint32_t one, uint32_t& acc) {
    constexpr uint32_t K = 1u | ((uint32_t)J << 8) | ((uint32_t)(J * J) << 17);
    asm("{\n\t.reg .pred p;\n\tsetp.le.f32 p, %1, %2;\n\t@p mad.lo.u32 %0, %3, %4, %0;\n\t}"
        : "+r"(acc)
        : "f"(v), "f"(thr), "r"(one), "n"(K));
}
template <int CB>
__device__ __forceinline__ void flag_chunk(const float (&v)[64], float thr, uint32_t one, uint32_t (&ia)[4]) {
    // four independent accumulator chains: the IMADs of one chain are 4 cycles apart
#define B200VQ_FA(i) flag_accumulate<CB * 64 + (i)>(v[i], thr, one, ia[(i) & 3]);
#define B200VQ_FA8(b) B200VQ_FA(b) B200VQ_FA(b + 1) B200VQ_FA(b + 2) B200VQ_FA(b + 3) \
                      B200VQ_FA(b + 4) B200VQ_FA(b + 5) B200VQ_FA(b + 6) B200VQ_FA(b + 7)
    B200VQ_FA8(0) B200VQ_FA8(8) B200VQ_FA8(16) B200VQ_FA8(24) B200VQ_FA8(32) B200VQ_FA8(40) B200VQ_FA8(48)
    B200VQ_FA8(56)
#undef B200VQ_FA8
#undef B200VQ_FA
}
template <int CB>   // 32-column form (J = CB * 32 + i): halves the live accumulator values
__device__ __forceinline__ void flag_chunk32(const float (&v)[32], float thr, uint32_t one, uint32_t (&ia)[4]) {
#define B200VQ_FA(i) flag_accumulate<CB * 32 + (i)>(v[i], thr, one, ia[(i) & 3]);
#define B200VQ_FA8(b) B200VQ_FA(b) B200VQ_FA(b + 1) B200VQ_FA(b + 2) B200VQ_FA(b + 3) \
                      B200VQ_FA(b + 4) B200VQ_FA(b + 5) B200VQ_FA(b + 6) B200VQ_FA(b + 7)
    B200VQ_FA8(0) B200VQ_FA8(8) B200VQ_FA8(16) B200VQ_FA8(24)
#undef B200VQ_FA8
#undef B200VQ_FA
}
__device__ __forceinline__ float fmin3(float a, float b, float c) {
    float m;
    asm("min.f32 %0, %1, %2, %3;" : "=f"(m) : "f"(a), "f"(b), "f"(c));
    return m;
}
__device__ __forceinline__ float fset_le(float a, float b) {  // 1.0f if a <= b else 0.0f
    float f;
    asm("set.le.f32.f32 %0, %1, %2;" : "=f"(f) : "f"(a), "f"(b));
    return f;
}


// ------------------------------------------------------------- parameters ---
struct Params {
    const float* x;          // (n, ldx)
    int64_t n, ldx;
    const float* centroids;  // (M, 256, dsub) original fp32 centroids (exact re-evaluation)
    const float* cnorm;      // (M, 256)
    const float2* cmax;      // (M) error-bound constants of sub-quantizer m: 2E = |x_m| * .x + .y
    uint8_t* codes;          // (n, M)
    uint8_t* cand2;          // (n, M) second candidate of a pair awaiting exact evaluation (0 = settled)
    unsigned long long* fix_count;  // (M) rows appended per sub-quantizer (may exceed fix_cap)
    uint32_t* fix_list;             // (M, fix_cap) row indices needing the exact 256-way scan
    unsigned long long fix_cap;     // capacity per sub-quantizer
    unsigned long long* stats;      // optional: [0] += pairs re-ranked, [1] += candidates, [2] += fix-up entries
    int M, dsub;
    int KS;        // sub-quantizers resident per CTA
    int CH;        // floats per K chunk (= swizzle span / 4): 32, 16 or 8
    int nchunk;    // dsub / CH
    int nstages;   // A ring depth
    int groups;    // M / KS
    int64_t ntiles;          // ceil(n / 128)
    int64_t tiles_per_item;  // row tiles per work item
    int64_t nitems;          // groups * ceil(ntiles / tiles_per_item)
    int debug;               // B200VQ_TC_DEBUG (profiling experiments only; 0 in production)
};

__device__ __forceinline__ uint32_t swizzle_mask(int CH) { return (uint32_t)(CH >> 2) - 1u; }  // 7, 3, 1
__device__ __forceinline__ uint32_t layout_type_of(int CH) { return CH == 32 ? 2u : (CH == 16 ? 4u : 6u); }

// --------------------------------------------------------- prepare kernel ---
// grid M, block 256: thread j prepares centroid j of sub-quantizer m.
//   bprep  (M*256, dsub)  = -2 c           (exact scaling)
//   bextra (M*256, 8)     = {hi, mid, lo, 0...} TF32-truncation split of ||c||^2
//   cmax   (M)            = {2a, 2b}: the screen's error bound is 2E = |x_m| 2a + 2b with c = max_j |c_mj|,
//                            a = c (k1 + 2 k2), b = k2 c^2, k1 = 2^-9 * 1.02, k2 = 3.8147e-6
//   ones   (128, 8)       = {1,1,1,0,0,0,0,0} per row; counter zeroed.
__global__ void prepare_kernel(const float* __restrict__ centroids, const float* __restrict__ cnorm, int dsub,
                               float* __restrict__ bprep, float* __restrict__ bextra,
                               float2* __restrict__ cmax, float* __restrict__ ones,
                               unsigned long long* __restrict__ fix_count) {  // (M) counters
    const int m = blockIdx.x, j = threadIdx.x;
    const size_t row = (size_t)m * KSUB + j;
    const float* c = centroids + row * dsub;
    float* b = bprep + row * dsub;
    for (int i = 0; i < dsub; ++i) b[i] = -2.0f * c[i];
    const float cn = cnorm[row];
    // tcgen05 kind::tf32 TRUNCATES both operands (measured: tools/tf32_probe.cu), so every product
    // x_i c_i shrinks by a factor in (1 - 2^-9, 1]: centre that at (1 - 2^-10) by scaling the norm
    // term the same way. The screen then ranks (1 - 2^-10) * d with half the worst-case error.
    const float cns = cn * 0.9990234375f;
    const float hi = __uint_as_float(__float_as_uint(cns) & 0xFFFFE000u);
    const float r1 = cns - hi;
    const float mid = __uint_as_float(__float_as_uint(r1) & 0xFFFFE000u);
    const float r2 = r1 - mid;
    const float lo = __uint_as_float(__float_as_uint(r2) & 0xFFFFE000u);
    float* e = bextra + row * 8;
    e[0] = hi;
    e[1] = mid;
    e[2] = lo;
    e[3] = e[4] = e[5] = e[6] = e[7] = 0.f;
    __shared__ float red[KSUB / 32];
    float v = sqrtf(fmaxf(cn, 0.f)) * 1.0000005f;
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(FULL_MASK, v, o));
    if ((j & 31) == 0) red[j >> 5] = v;
    __syncthreads();
    if (j == 0) {
        float mx = red[0];
        for (int w = 1; w < KSUB / 32; ++w) mx = fmaxf(mx, red[w]);
        // |err| <= E = k1 |x||c| + k2 (2 |x||c| + c^2): TF32 truncation after centring (2^-10 |x||c| on
        // x.c, doubled by the -2 factor, widened by 2 %) plus slack for accumulation and the norm split
        const float k1 = 0.001953125f * 1.02f, k2 = 3.8147e-6f;
        cmax[m] = make_float2(2.f * mx * (k1 + 2.f * k2) * 1.000001f, 2.f * k2 * mx * mx * 1.000001f);
    }
    if (m == 0) {
        if (j < TILE_ROWS) {
            float* o = ones + j * 8;
            o[0] = o[1] = o[2] = 1.f;
            o[3] = o[4] = o[5] = o[6] = o[7] = 0.f;
        }
    }
    if (j == 0) fix_count[m] = 0ull;
}

// Exact distance of the SIMT kernel (pq.cu: pq_assign_simt_kernel): sequential fp32 FMA chain
// over the sub-vector, (xn + cn) - 2 acc, clamped at 0. x is read from global memory.
__device__ __forceinline__ float exact_dist(const float* __restrict__ xr, const float* __restrict__ c, int dsub,
                                            float xn, float cn) {
    float acc = 0.f;
    for (int i = 0; i < dsub; ++i) acc = __fmaf_rn(xr[i], c[i], acc);
    float d = __fmaf_rn(-2.f, acc, __fadd_rn(xn, cn));
    return fmaxf(d, 0.f);
}

// Exact 256-way scan by one thread (used when the fix-up list is full; practically never).
__device__ __noinline__ int exact_scan_thread(const float* xr, const float* cm, const float* cnm, int dsub,
                                              float xn) {
    float best = __int_as_float(0x7f800000);
    int bk = 0;
    for (int j = 0; j < KSUB; ++j) {
        float d = exact_dist(xr, cm + (size_t)j * dsub, dsub, xn, cnm[j]);
        if (d < best) {
            best = d;
            bk = j;
        }
    }
    return bk;
}

// ------------------------------------------------------ pair fix-up kernel ---
// Scans cand2 (16 pairs per thread and step) and compacts the marked (row, m) into a shared
// queue; then one thread per queued pair:
//   second > code : two candidates -> both evaluated with the SIMT kernel's exact arithmetic
//                   (operands from global memory / L2, 8 independent 128-bit loads per operand
//                   in flight before the sequential FMA chains), lower index wins ties
//   otherwise     : (255, 255) marker -> appended to the per-sub-quantizer list of fixup_kernel
constexpr int PF_THREADS = 256;
constexpr int PF_QUEUE = PF_THREADS * 16;
__global__ void __launch_bounds__(PF_THREADS) pair_fixup_kernel(Params p) {
    __shared__ uint32_t queue[PF_QUEUE];   // index of the pair inside the block's 4096-pair window
    __shared__ int qn;
    const int64_t total = p.n * (int64_t)p.M;
    const int64_t total16 = (total + 15) >> 4;
    const uint4* c16 = reinterpret_cast<const uint4*>(p.cand2);
    for (int64_t g0 = (int64_t)blockIdx.x * PF_THREADS; g0 < total16; g0 += (int64_t)gridDim.x * PF_THREADS) {
        if (threadIdx.x == 0) qn = 0;
        __syncthreads();
        const int64_t g = g0 + threadIdx.x;
        if (g < total16) {
            const uint4 w = __ldg(c16 + g);   // the scratch region is padded to a multiple of 16 bytes
            if ((w.x | w.y | w.z | w.w) != 0u) {
                const uint32_t wd[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
                for (int k = 0; k < 16; ++k)
                    if ((wd[k >> 2] >> (8 * (k & 3))) & 0xFF) queue[atomicAdd(&qn, 1)] = threadIdx.x * 16 + k;
            }
        }
        __syncthreads();
        const int nq = qn;
        for (int e = threadIdx.x; e < nq; e += PF_THREADS) {
            const int64_t idx = g0 * 16 + queue[e];
            if (idx >= total) continue;
            const int64_t row = idx / p.M;
            const int m = (int)(idx - row * p.M);
            const int code = p.codes[idx], second = p.cand2[idx];
            const float* xr = p.x + row * p.ldx + (int64_t)m * p.dsub;
            const float* cm = p.centroids + (size_t)m * KSUB * p.dsub;
            const float* cnm = p.cnorm + (size_t)m * KSUB;
            if (second > code) {
                float xn = 0.f, acca = 0.f, accb = 0.f;
                const float4* x4 = reinterpret_cast<const float4*>(xr);     // dsub % 8 == 0, 16-byte rows
                const float4* a4 = reinterpret_cast<const float4*>(cm + (size_t)code * p.dsub);
                const float4* b4 = reinterpret_cast<const float4*>(cm + (size_t)second * p.dsub);
                const int units = p.dsub >> 2;
                for (int u0 = 0; u0 < units; u0 += 8) {
                    float4 xv[8], av[8], bv[8];
#pragma unroll
                    for (int u = 0; u < 8; ++u)
                        if (u0 + u < units) {
                            xv[u] = x4[u0 + u];
                            av[u] = __ldg(a4 + u0 + u);
                            bv[u] = __ldg(b4 + u0 + u);
                        }
#pragma unroll
                    for (int u = 0; u < 8; ++u)
                        if (u0 + u < units) {
                            xn = __fmaf_rn(xv[u].x, xv[u].x, xn);
                            xn = __fmaf_rn(xv[u].y, xv[u].y, xn);
                            xn = __fmaf_rn(xv[u].z, xv[u].z, xn);
                            xn = __fmaf_rn(xv[u].w, xv[u].w, xn);
                            acca = __fmaf_rn(xv[u].x, av[u].x, acca);
                            acca = __fmaf_rn(xv[u].y, av[u].y, acca);
                            acca = __fmaf_rn(xv[u].z, av[u].z, acca);
                            acca = __fmaf_rn(xv[u].w, av[u].w, acca);
                            accb = __fmaf_rn(xv[u].x, bv[u].x, accb);
                            accb = __fmaf_rn(xv[u].y, bv[u].y, accb);
                            accb = __fmaf_rn(xv[u].z, bv[u].z, accb);
                            accb = __fmaf_rn(xv[u].w, bv[u].w, accb);
                        }
                }
                const float da = fmaxf(__fmaf_rn(-2.f, acca, __fadd_rn(xn, cnm[code])), 0.f);
                const float db = fmaxf(__fmaf_rn(-2.f, accb, __fadd_rn(xn, cnm[second])), 0.f);
                if (db < da) p.codes[idx] = (uint8_t)second;
            } else {
                const unsigned long long slot = atomicAdd(p.fix_count + m, 1ull);
                if (slot < p.fix_cap) {
                    p.fix_list[(size_t)m * p.fix_cap + slot] = (uint32_t)row;
                } else {
                    float xn = 0.f;
                    for (int i = 0; i < p.dsub; ++i) xn = __fmaf_rn(xr[i], xr[i], xn);
                    p.codes[idx] = (uint8_t)exact_scan_thread(xr, cm, cnm, p.dsub, xn);
                }
            }
        }
        __syncthreads();
    }
}

// ---------------------------------------------------------- fix-up kernel ---
// grid (chunks, M): CTA (c, m) owns entries [c, c + 1) * FIX_CHUNK of sub-quantizer m's list and exits
// at once when the list is shorter (the lists are very uneven: the error bound scales with |x_m||c|,
// so high-variance sub-spaces collect most entries; a fixed number of CTAs per m left most of the
// device idle behind the longest list). The CTA stages the 256 centroids of sub-quantizer m in shared
// memory dimension-major ([i][j]: lane j reads consecutive words, no bank conflicts) and its
// warps walk their share: exact scan of all 256 centroids with the arithmetic
// and tie-breaking (lowest index) of pq_assign_simt_kernel; lane l owns centroids l, l+32, ...
constexpr int FIX_WARPS = 8;
constexpr int FIX_ROWS = 4;   // list entries scored together by one warp (centroid loads are shared; 8 rows need 194
                              // registers and halve the occupancy: measured slower)
constexpr int FIX_CHUNK = 256; // list entries per CTA
__global__ void __launch_bounds__(FIX_WARPS * 32) fixup_kernel(Params p) {
    extern __shared__ __align__(16) float fsm[];
    const int dsub = p.dsub, m = blockIdx.y;
    float* cs = fsm;                         // [dsub][256]
    float* cn = cs + (size_t)dsub * KSUB;    // [256]
    float* xs_all = cn + KSUB;               // [FIX_WARPS][FIX_ROWS][dsub]
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    unsigned long long cnt = p.fix_count[m];
    if (p.stats && blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(p.stats + 2, cnt);
    if (cnt > p.fix_cap) cnt = p.fix_cap;
    const unsigned long long c_begin = (unsigned long long)blockIdx.x * FIX_CHUNK;
    if (c_begin >= cnt) return;
    const unsigned long long c_end = c_begin + FIX_CHUNK < cnt ? c_begin + FIX_CHUNK : cnt;
    const float* cm = p.centroids + (size_t)m * KSUB * dsub;
    // thread j stages centroid j: 128-bit global loads, conflict-free transposed stores
    for (int j = threadIdx.x; j < KSUB; j += blockDim.x) {
        const float4* c4 = reinterpret_cast<const float4*>(cm + (size_t)j * dsub);
        for (int i4 = 0; i4 < (dsub >> 2); ++i4) {
            const float4 c = __ldg(c4 + i4);
            cs[(4 * i4 + 0) * KSUB + j] = c.x;
            cs[(4 * i4 + 1) * KSUB + j] = c.y;
            cs[(4 * i4 + 2) * KSUB + j] = c.z;
            cs[(4 * i4 + 3) * KSUB + j] = c.w;
        }
        cn[j] = p.cnorm[(size_t)m * KSUB + j];
    }
    __syncthreads();
    float* xs = xs_all + (size_t)w * FIX_ROWS * dsub;
    const uint32_t* list = p.fix_list + (size_t)m * p.fix_cap;
    for (unsigned long long e0 = c_begin + (unsigned long long)w * FIX_ROWS; e0 < c_end;
         e0 += (unsigned long long)FIX_WARPS * FIX_ROWS) {
        int64_t rows[FIX_ROWS];
        __syncwarp();
#pragma unroll
        for (int t = 0; t < FIX_ROWS; ++t) {
            const unsigned long long e = e0 + t < c_end ? e0 + t : e0;   // pad with the first entry
            rows[t] = (int64_t)list[e];
            const float* xr = p.x + rows[t] * p.ldx + (int64_t)m * dsub;
            for (int i = lane; i < dsub; i += 32) xs[t * dsub + i] = xr[i];
        }
        __syncwarp();
        float xn[FIX_ROWS], acc[FIX_ROWS][8];
#pragma unroll
        for (int t = 0; t < FIX_ROWS; ++t) {
            xn[t] = 0.f;
#pragma unroll
            for (int k = 0; k < 8; ++k) acc[t][k] = 0.f;
        }
        for (int i0 = 0; i0 < dsub; i0 += 4) {
            // four dimensions of every row with one broadcast 128-bit load each
            float4 xq[FIX_ROWS];
#pragma unroll
            for (int t = 0; t < FIX_ROWS; ++t) xq[t] = *reinterpret_cast<const float4*>(xs + t * dsub + i0);
#pragma unroll
            for (int ii = 0; ii < 4; ++ii) {
                const float* ci = cs + (i0 + ii) * KSUB + lane;
                float cv[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) cv[k] = ci[32 * k];
#pragma unroll
                for (int t = 0; t < FIX_ROWS; ++t) {
                    const float xv = ii == 0 ? xq[t].x : (ii == 1 ? xq[t].y : (ii == 2 ? xq[t].z : xq[t].w));
                    xn[t] = __fmaf_rn(xv, xv, xn[t]);
#pragma unroll
                    for (int k = 0; k < 8; ++k) acc[t][k] = __fmaf_rn(xv, cv[k], acc[t][k]);
                }
            }
        }
#pragma unroll
        for (int t = 0; t < FIX_ROWS; ++t) {
            float best = __int_as_float(0x7f800000);
            int bk = 0x7fffffff;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const int j = lane + 32 * k;
                const float d = fmaxf(__fmaf_rn(-2.f, acc[t][k], __fadd_rn(xn[t], cn[j])), 0.f);
                if (d < best) {
                    best = d;
                    bk = j;
                }
            }
            for (int o = 16; o > 0; o >>= 1) {
                float ob = __shfl_xor_sync(FULL_MASK, best, o);
                int ok = __shfl_xor_sync(FULL_MASK, bk, o);
                if (ob < best || (ob == best && ok < bk)) {
                    best = ob;
                    bk = ok;
                }
            }
            if (bk == 0x7fffffff) bk = 0;  // all distances NaN: the SIMT kernel keeps index 0
            if (lane == 0 && e0 + t < c_end) p.codes[rows[t] * p.M + m] = (uint8_t)bk;
        }
    }
}

// ------------------------------------------------------------ main kernel ---
struct SmemPlan {
    uint32_t b_off, bx_off, ax_off, a_off, bar_off, total;
    uint32_t a_chunk_bytes, b_chunk_bytes, a_stage_bytes;
};
__host__ __device__ inline SmemPlan make_plan(int KS, int CH, int nchunk, int nstages) {
    SmemPlan s;
    s.a_chunk_bytes = TILE_ROWS * CH * 4;
    s.b_chunk_bytes = KSUB * CH * 4;
    s.a_stage_bytes = (uint32_t)KS * nchunk * s.a_chunk_bytes;
    s.b_off = 0;
    s.bx_off = s.b_off + (uint32_t)KS * nchunk * s.b_chunk_bytes;
    s.ax_off = s.bx_off + (uint32_t)KS * 8192;
    s.a_off = s.ax_off + 4096;
    s.bar_off = s.a_off + (uint32_t)nstages * s.a_stage_bytes;
    s.total = s.bar_off + 256 + 4096;   // barriers + epilogue exchange buffers
    return s;
}

// barrier slots (8 bytes each) inside the bar region
enum { BAR_B_FULL = 0, BAR_B_EMPTY = 1, BAR_AX_FULL = 2, BAR_A_FULL = 3, BAR_A_EMPTY = 7, BAR_T_FULL = 11,
       BAR_T_EMPTY = 13, BAR_TMEM_PTR = 15 };

// DBG = true compiles the ablation switches (Params::debug, set from B200VQ_TC_DEBUG) in; the
// production instantiation has none of them.
template <bool DBG>
__global__ void __launch_bounds__(THREADS, 1)
    assign_tc_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_b,
                     const __grid_constant__ CUtensorMap map_bx, const __grid_constant__ CUtensorMap map_ax,
                     const Params p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    // dynamic shared memory is only guaranteed 16-byte aligned: round up to 1024 for the swizzle atoms
    const uint32_t raw = smem_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;
    uint8_t* gbase = smem_raw + (base - raw);
    const SmemPlan sp = make_plan(p.KS, p.CH, p.nchunk, p.nstages);
    const uint32_t bar0 = base + sp.bar_off;
    auto bar = [&](int slot) { return bar0 + 8u * (uint32_t)slot; };
    volatile uint32_t* tmem_ptr_smem = reinterpret_cast<volatile uint32_t*>(gbase + sp.bar_off + 8 * BAR_TMEM_PTR);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int KS = p.KS, CH = p.CH, nchunk = p.nchunk, NST = p.nstages;
    const int dbg = DBG ? p.debug : 0;

    if (warp == SVC_WARP && lane == 0) {
        tma_prefetch_desc(&map_x);
        tma_prefetch_desc(&map_b);
        tma_prefetch_desc(&map_bx);
        tma_prefetch_desc(&map_ax);
        mbar_init(bar(BAR_B_FULL), 1);
        mbar_init(bar(BAR_B_EMPTY), 1);  // MMA commit + 4 epilogue warps (they read B for the exact path)
        mbar_init(bar(BAR_AX_FULL), 1);
        for (int i = 0; i < MAX_STAGES; ++i) {
            mbar_init(bar(BAR_A_FULL + i), 1);
            mbar_init(bar(BAR_A_EMPTY + i), 1 + EPI_WARPS);  // MMA commit + every epilogue warp
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(bar(BAR_T_FULL + i), 1);
            mbar_init(bar(BAR_T_EMPTY + i), 8);
        }
        fence_barrier_init();
    }
    if (warp == SVC_WARP + 1) tmem_alloc(bar(BAR_TMEM_PTR), 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;

    if (warp == SVC_WARP) {
        // ===================================================== TMA producer ===
        if (lane == 0) {
            mbar_expect_tx(bar(BAR_AX_FULL), 4096);
            tma_load_2d(base + sp.ax_off, &map_ax, bar(BAR_AX_FULL), 0, 0);
            uint32_t st = 0, a_phase = 0, b_phase = 0;
            for (int64_t item = blockIdx.x; item < p.nitems; item += gridDim.x) {
                const int g = (int)(item % p.groups);
                const int64_t rc = item / p.groups;
                const int64_t t0 = rc * p.tiles_per_item;
                int64_t t1 = t0 + p.tiles_per_item;
                if (t1 > p.ntiles) t1 = p.ntiles;
                mbar_wait_backoff(bar(BAR_B_EMPTY), b_phase ^ 1);
                mbar_expect_tx(bar(BAR_B_FULL), (uint32_t)KS * (nchunk * sp.b_chunk_bytes + 8192));
                for (int s = 0; s < KS; ++s) {
                    const int m = g * KS + s;
                    for (int c = 0; c < nchunk; ++c)
                        tma_load_2d(base + sp.b_off + (uint32_t)(s * nchunk + c) * sp.b_chunk_bytes, &map_b,
                                    bar(BAR_B_FULL), c * CH, m * KSUB);
                    tma_load_2d(base + sp.bx_off + (uint32_t)s * 8192, &map_bx, bar(BAR_B_FULL), 0, m * KSUB);
                }
                b_phase ^= 1;
                for (int64_t t = t0; t < t1; ++t) {
                    mbar_wait_backoff(bar(BAR_A_EMPTY + st), a_phase ^ 1);
                    mbar_expect_tx(bar(BAR_A_FULL + st), sp.a_stage_bytes);
                    for (int s = 0; s < KS; ++s)
                        for (int c = 0; c < nchunk; ++c)
                            tma_load_2d(base + sp.a_off + st * sp.a_stage_bytes +
                                            (uint32_t)(s * nchunk + c) * sp.a_chunk_bytes,
                                        &map_x, bar(BAR_A_FULL + st), (g * KS + s) * p.dsub + c * CH,
                                        (dbg & 16) ? 0 : (int)(t * TILE_ROWS));
                    if (++st == (uint32_t)NST) {
                        st = 0;
                        a_phase ^= 1;
                    }
                }
            }
        }
    } else if (warp == SVC_WARP + 1) {
        // ======================================================= MMA issuer ===
        if (lane == 0) {
            const uint32_t lt = layout_type_of(CH);
            const uint32_t sbo = 32u * (uint32_t)CH;
            const int kper = CH / 8;
            uint32_t st = 0, a_phase = 0, b_phase = 0, acc = 0, t_phase = 0;
            long long w_a = 0, w_t = 0, w_b = 0;
            const long long t_begin = clock64();
            mbar_wait_backoff(bar(BAR_AX_FULL), 0);
            const uint64_t ax_desc = make_kmajor_desc(base + sp.ax_off, 256, 6);
            for (int64_t item = blockIdx.x; item < p.nitems; item += gridDim.x) {
                const int64_t rc = item / p.groups;
                const int64_t t0 = rc * p.tiles_per_item;
                int64_t t1 = t0 + p.tiles_per_item;
                if (t1 > p.ntiles) t1 = p.ntiles;
                { const long long c0 = clock64(); mbar_wait_backoff(bar(BAR_B_FULL), b_phase); w_b += clock64() - c0; }
                b_phase ^= 1;
                for (int64_t t = t0; t < t1; ++t) {
                    { const long long c0 = clock64(); mbar_wait_backoff(bar(BAR_A_FULL + st), a_phase); w_a += clock64() - c0; }
                    for (int s = 0; s < KS; ++s) {
                        { const long long c0 = clock64(); mbar_wait_backoff(bar(BAR_T_EMPTY + acc), t_phase ^ 1); w_t += clock64() - c0; }
                        tc_fence_after();
                        const uint32_t d = tmem_base + acc * 256u;
                        uint32_t accum = 0;
                        for (int c = 0; c < nchunk; ++c) {
                            const uint32_t a_addr = base + sp.a_off + st * sp.a_stage_bytes +
                                                    (uint32_t)(s * nchunk + c) * sp.a_chunk_bytes;
                            const uint32_t b_addr = base + sp.b_off + (uint32_t)(s * nchunk + c) * sp.b_chunk_bytes;
                            for (int kk = 0; kk < kper; ++kk) {
                                umma_tf32(d, make_kmajor_desc(a_addr + kk * 32, sbo, lt),
                                          make_kmajor_desc(b_addr + kk * 32, sbo, lt), IDESC_TF32_128x256, accum);
                                accum = 1;
                            }
                        }
                        umma_tf32(d, ax_desc, make_kmajor_desc(base + sp.bx_off + (uint32_t)s * 8192, 256, 6),
                                  IDESC_TF32_128x256, 1);
                        umma_commit(bar(BAR_T_FULL + acc));
                        if (++acc == 2) {
                            acc = 0;
                            t_phase ^= 1;
                        }
                    }
                    umma_commit(bar(BAR_A_EMPTY + st));
                    if (++st == (uint32_t)NST) {
                        st = 0;
                        a_phase ^= 1;
                    }
                }
                umma_commit(bar(BAR_B_EMPTY));
            }
            if (p.stats && (dbg & 32)) {
                atomicAdd(p.stats + 3, (unsigned long long)w_a);
                atomicAdd(p.stats + 4, (unsigned long long)w_t);
                atomicAdd(p.stats + 5, (unsigned long long)w_b);
                atomicAdd(p.stats + 6, (unsigned long long)(clock64() - t_begin));
            }
        }
    } else {
        // ========================================================= epilogue ===
        const int q = warp & 3;                 // TMEM lane quarter this warp may read
        const int r = q * 32 + lane;            // row inside the tile
        const uint32_t lane_addr = tmem_base + ((uint32_t)(q * 32) << 16);
        const uint32_t pitch = (uint32_t)CH * 4u, swz = swizzle_mask(CH);
        const int units = CH / 4;
        // 16 epilogue warps: accumulator stage acc (units alternate), TMEM lane quarter q, column half hf.
        // The two warps of a (acc, q) pair exchange partial results through shared memory.
        const uint32_t eidx = (uint32_t)warp >> 2;
        const uint32_t acc = eidx & 1u, hf = eidx >> 1;
        float* xch = reinterpret_cast<float*>(gbase + sp.bar_off + 256) + (acc * 4 + q) * 128;  // [4][32]
        const uint32_t pair_bar = 1u + acc * 4u + (uint32_t)q;
        const uint32_t one = p.nstages > 0 ? 1u : 0u;   // runtime 1 (see flag_accumulate)
        uint32_t xoff[8];   // swizzled byte offsets of this lane's row inside one A chunk tile
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            uint32_t off = (uint32_t)r * pitch + (uint32_t)u * 16u;
            xoff[u] = off ^ (((off >> 7) & swz) << 4);
        }
        uint32_t st = 0, a_phase = 0, t_phase = 0, unit = 0;
        // 32-bit loop state (n < 2^32 rows: tile and item counts fit): the 64-bit form spilled four loop variables
        // to local memory, whose reloads sat on the critical path of every slice
        const uint32_t nitems32 = (uint32_t)p.nitems, ntiles32 = (uint32_t)p.ntiles, tpi32 = (uint32_t)p.tiles_per_item;
        for (uint32_t item = blockIdx.x; item < nitems32; item += gridDim.x) {
            const int g = (int)(item % (uint32_t)p.groups);
            const uint32_t rc = item / (uint32_t)p.groups;
            const uint32_t t0 = rc * tpi32;
            uint32_t t1 = t0 + tpi32;
            if (t1 > ntiles32) t1 = ntiles32;
            for (uint32_t t = t0; t < t1; ++t) {
                const uint32_t row32 = t * (uint32_t)TILE_ROWS + (uint32_t)r;
                const int64_t row = (int64_t)row32;
                const bool live = row < p.n;
                mbar_wait(bar(BAR_A_FULL + st), a_phase);
                // units alternate between the two accumulator stages: this warp takes every second one
                const int s_first = (int)((unit & 1u) ^ acc);
                unit += (uint32_t)KS;
                for (int s = s_first; s < KS; s += 2) {
                    const int m = g * KS + s;
                    const uint8_t* atile = gbase + sp.a_off + st * sp.a_stage_bytes +
                                           (uint32_t)(s * nchunk) * sp.a_chunk_bytes;
                    // ||x_m||^2 with the SIMT kernel's sequential FMA chain, from the swizzled A tile
                    // (upper-half warp only: the lower-half warp of the pair does the bookkeeping)
                    float xn = 0.f;
                    if (hf == 1 && !(dbg & 4)) {
                        for (int c = 0; c < nchunk; ++c) {
#pragma unroll
                            for (int u = 0; u < 8; ++u) {
                                if (u < units) {
                                    const float4 w = *reinterpret_cast<const float4*>(
                                        atile + (uint32_t)c * sp.a_chunk_bytes + xoff[u]);
                                    xn = __fmaf_rn(w.x, w.x, xn);
                                    xn = __fmaf_rn(w.y, w.y, xn);
                                    xn = __fmaf_rn(w.z, w.z, xn);
                                    xn = __fmaf_rn(w.w, w.w, xn);
                                }
                            }
                        }
                    }
                    mbar_wait(bar(BAR_T_FULL + acc), t_phase);
                    tc_fence_after();
                    const uint32_t taddr = lane_addr + acc * 256u + hf * 128u;
                    float v[32];
                    // pass 1: minimum of this warp's 128 columns (four 32-column TMEM loads: 64-column loads kept
                    // 64 values live next to the loop state and ptxas spilled that state to local memory)
                    float mn = __int_as_float(0x7f800000);
#pragma unroll
                    for (int cb = 0; cb < 4; ++cb) {
                        if (dbg & 8) break;
                        tmem_ld32(taddr + cb * 32, v);
                        float t[10];
#pragma unroll
                        for (int k = 0; k < 10; ++k) t[k] = fmin3(v[3 * k], v[3 * k + 1], v[3 * k + 2]);
                        const float u0 = fmin3(t[0], t[1], t[2]), u1 = fmin3(t[3], t[4], t[5]), u2 = fmin3(t[6], t[7], t[8]);
                        const float u3 = fmin3(t[9], v[30], v[31]);
                        mn = fmin3(fmin3(u0, u1, u2), u3, mn);
                    }
                    xch[hf * 32 + lane] = mn;
                    if (hf == 1) {
                        // 2E = |x_m| e2.x + e2.y (constants of prepare_kernel); only an upper bound of |x_m| is
                        // needed: 1.00001 covers the approximate square root (2^-22)
                        const float2 e2 = p.cmax[m];
                        float xl;
                        asm("sqrt.approx.f32 %0, %1;" : "=f"(xl) : "f"(xn));
                        xch[64 + lane] = __fmaf_rn(xl * 1.00001f, e2.x, e2.y);
                    }
                    asm volatile("bar.sync %0, 64;" ::"r"(pair_bar) : "memory");
                    const float thr = fminf(xch[lane], xch[32 + lane]) + xch[64 + lane];
                    // pass 2: one packed integer accumulator per candidate flag:
                    //   bits [0,8) count, [8,17) sum of j, [17,32) sum of j^2  (j relative to this half;
                    //   the upper fields are exact whenever the count is <= 2, which is all that reads them)
                    uint32_t ia[4] = {0, 0, 0, 0};
                    if (dbg & 1) { ia[0] = 1; } else {
                    tmem_ld32(taddr, v);
                    flag_chunk32<0>(v, thr, one, ia);
                    tmem_ld32(taddr + 32, v);
                    flag_chunk32<1>(v, thr, one, ia);
                    tmem_ld32(taddr + 64, v);
                    flag_chunk32<2>(v, thr, one, ia);
                    tmem_ld32(taddr + 96, v);
                    flag_chunk32<3>(v, thr, one, ia);
                    }
                    const uint32_t isum = (ia[0] + ia[1]) + (ia[2] + ia[3]);
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar(BAR_T_EMPTY + acc));
                    t_phase ^= 1;
                    // combine the halves: the upper-half warp hands its accumulator over and moves on
                    if (hf == 1) xch[96 + lane] = __uint_as_float(isum);
                    asm volatile("bar.sync %0, 64;" ::"r"(pair_bar) : "memory");
                    if (hf == 1) continue;
                    const uint32_t plo = isum, phi = __float_as_uint(xch[96 + lane]);
                    // Branch-free decode: with the upper half's moments moved to row-wide column indices
                    // (J = 128 + j) count, sum J and sum J^2 add up, and with D = cnt S2 - S1^2 (0 for one
                    // candidate, (J1 - J2)^2 for two) the candidates are (S1 -+ sqrt D) / cnt.
                    // 1 candidate: it is the exact argmin. 2: provisional code = lower index, the other goes
                    // to cand2 and pair_fixup_kernel evaluates both exactly. Otherwise (>= 3, or 0 for NaN
                    // rows) the (255, 255) marker sends the pair to the exact 256-way scan. Plain stores
                    // only: nothing here waits on memory.
                    const uint32_t cnt_hi = phi & 0xFFu, s1_hi = (phi >> 8) & 0x1FFu;
                    const uint32_t ucnt = (plo & 0xFFu) + cnt_hi;
                    const uint32_t S1 = ((plo >> 8) & 0x1FFu) + s1_hi + 128u * cnt_hi;
                    const uint32_t S2 = (plo >> 17) + (phi >> 17) + 256u * s1_hi + 16384u * cnt_hi;
                    const bool ok = (ucnt - 1u) < 2u || (DBG && (dbg & 2));
                    float df;
                    asm("sqrt.approx.f32 %0, %1;" : "=f"(df) : "f"((float)(int)(ucnt * S2 - S1 * S1)));
                    const uint32_t dd = (uint32_t)(int)(df + 0.5f), sh = (ucnt - 1u) & 1u;
                    const uint32_t lo_j = (S1 - dd) >> sh, hi_j = (S1 + dd) >> sh;
                    const int cnt = (int)ucnt;
                    const int code = ok ? (int)(lo_j & 0xFFu) : 255;
                    const int second = ok ? (hi_j != lo_j ? (int)(hi_j & 0xFFu) : 0) : 255;
                    const bool multi = live && second != 0;
                    if (live) {
                        p.codes[row * p.M + m] = (uint8_t)code;
                        if (second) p.cand2[row * p.M + m] = (uint8_t)second;
                    }
                    if (p.stats) {
                        const unsigned mm = __ballot_sync(FULL_MASK, multi);
                        if (lane == 0 && mm) atomicAdd(p.stats, (unsigned long long)__popc(mm));
                        unsigned long long cc = multi ? (unsigned long long)(cnt > 0 ? cnt : KSUB) : 0ull;
                        for (int o = 16; o > 0; o >>= 1) cc += __shfl_xor_sync(FULL_MASK, cc, o);
                        if (lane == 0 && cc) atomicAdd(p.stats + 1, cc);
                    }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(bar(BAR_A_EMPTY + st));
                if (++st == (uint32_t)NST) {
                    st = 0;
                    a_phase ^= 1;
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == SVC_WARP + 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

}  // namespace tc
}  // namespace b200vq

using namespace b200vq;
using namespace b200vq::tc;

// ------------------------------------------------------------- host side ---

struct TcShape {
    int KS, CH, nchunk, nstages;
    SmemPlan sp;
    bool ok;
};

static TcShape plan_shape(int M, int dsub) {
    TcShape t{};
    t.ok = false;
    if (dsub % 8 != 0 || dsub < 8 || dsub > 128) return t;
    t.CH = (dsub % 32 == 0) ? 32 : (dsub % 16 == 0 ? 16 : 8);
    t.nchunk = dsub / t.CH;
    int ks = 64 / dsub;
    if (ks < 1) ks = 1;
    while (ks > 1 && M % ks != 0) --ks;
    t.KS = ks;
    for (int nst = MAX_STAGES; nst >= 1; --nst) {
        SmemPlan sp = make_plan(t.KS, t.CH, t.nchunk, nst);
        if (sp.total + 1024 <= SMEM_BUDGET) {
            t.nstages = nst;
            t.sp = sp;
            t.ok = true;
            break;
        }
    }
    return t;
}

// scratch layout (all offsets 256-byte aligned)
struct ScratchPlan {
    size_t bprep, bextra, ones, cmax, counter, list, cand2, total;
    unsigned long long cap;
};
static ScratchPlan plan_scratch(int64_t n, int M, int dsub) {
    auto al = [](size_t v) { return (v + 255) & ~(size_t)255; };
    ScratchPlan s;
    s.bprep = 0;
    s.bextra = al(s.bprep + (size_t)M * KSUB * dsub * 4);
    s.ones = al(s.bextra + (size_t)M * KSUB * 8 * 4);
    s.cmax = al(s.ones + TILE_ROWS * 8 * 4);
    s.counter = al(s.cmax + (size_t)M * 8);
    s.list = al(s.counter + (size_t)M * 8);
    unsigned long long cap = (unsigned long long)(n / 8);   // per sub-quantizer; overflow falls back inline
    if (cap < 1024) cap = 1024;
    s.cap = cap;
    s.cand2 = al(s.list + (size_t)M * cap * 4);
    s.total = al(s.cand2 + (((size_t)n * M + 15) & ~(size_t)15));
    return s;
}

extern "C" size_t b200vq_pq_assign_tc_scratch_bytes2(int64_t n, int M, int dsub) {
    if (n < 0 || M <= 0 || dsub <= 0) return 0;
    return plan_scratch(n, M, dsub).total;
}

// Kept for ABI v1 callers that do not know dsub: sized for the largest supported dsub.
extern "C" size_t b200vq_pq_assign_tc_scratch_bytes(int64_t n, int M) {
    return b200vq_pq_assign_tc_scratch_bytes2(n, M, 128);
}

extern "C" int b200vq_pq_assign_tc(const float* x, int64_t n, int64_t ldx, const float* centroids,
                                   const float* cnorm, int M, int ksub, int dsub, uint8_t* codes,
                                   void* scratch, size_t scratch_bytes, unsigned long long* stats,
                                   void* stream) {
    VQ_CHECK_ARG(n >= 0 && M > 0 && dsub > 0, "pq_assign_tc: bad shape");
    VQ_CHECK_ARG(ksub == KSUB, "pq_assign_tc: needs ksub == 256 (got %d)", ksub);
    TcShape ts = plan_shape(M, dsub);
    if (!ts.ok) {
        set_error("pq_assign_tc: dsub=%d unsupported (needs dsub %% 8 == 0 and 8 <= dsub <= 128)", dsub);
        return B200VQ_ERR_UNSUPPORTED;
    }
    if (n == 0) return B200VQ_OK;
    VQ_CHECK_ARG(x && centroids && cnorm && codes, "pq_assign_tc: null pointer");
    VQ_CHECK_ARG(ldx >= (int64_t)M * dsub && ldx % 4 == 0 && ((uintptr_t)x & 15) == 0,
                 "pq_assign_tc: x must be 16-byte aligned with ldx %% 4 == 0 and ldx >= D");
    VQ_CHECK_ARG(n < (1ll << 32), "pq_assign_tc: n must be below 2^32 rows per call");
    ScratchPlan sc = plan_scratch(n, M, dsub);
    VQ_CHECK_ARG(scratch && scratch_bytes >= sc.total && ((uintptr_t)scratch & 255) == 0,
                 "pq_assign_tc: scratch too small or not 256-byte aligned (%zu < %zu)", scratch_bytes, sc.total);
    cudaStream_t st = as_stream(stream);
    uint8_t* sb = reinterpret_cast<uint8_t*>(scratch);
    float* bprep = reinterpret_cast<float*>(sb + sc.bprep);
    float* bextra = reinterpret_cast<float*>(sb + sc.bextra);
    float* ones = reinterpret_cast<float*>(sb + sc.ones);
    float2* cmax = reinterpret_cast<float2*>(sb + sc.cmax);
    unsigned long long* counter = reinterpret_cast<unsigned long long*>(sb + sc.counter);
    uint32_t* list = reinterpret_cast<uint32_t*>(sb + sc.list);
    uint8_t* cand2 = sb + sc.cand2;
    VQ_CUDA(cudaMemsetAsync(cand2, 0, (((size_t)n * M + 15) & ~(size_t)15), st));

    prepare_kernel<<<M, KSUB, 0, st>>>(centroids, cnorm, dsub, bprep, bextra, cmax, ones, counter);
    VQ_LAUNCH_CHECK();

    const CUtensorMapSwizzle swz = ts.CH == 32 ? CU_TENSOR_MAP_SWIZZLE_128B
                                   : ts.CH == 16 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B;
    CUtensorMap map_x, map_b, map_bx, map_ax;
    int rc;
    if ((rc = encode_2d(&map_x, x, (uint64_t)n, (uint64_t)M * dsub, (uint64_t)ldx, TILE_ROWS, ts.CH, swz))) return rc;
    if ((rc = encode_2d(&map_b, bprep, (uint64_t)M * KSUB, dsub, dsub, KSUB, ts.CH, swz))) return rc;
    if ((rc = encode_2d(&map_bx, bextra, (uint64_t)M * KSUB, 8, 8, KSUB, 8, CU_TENSOR_MAP_SWIZZLE_32B))) return rc;
    if ((rc = encode_2d(&map_ax, ones, TILE_ROWS, 8, 8, TILE_ROWS, 8, CU_TENSOR_MAP_SWIZZLE_32B))) return rc;

    Params p{};
    p.x = x;
    p.n = n;
    p.ldx = ldx;
    p.centroids = centroids;
    p.cnorm = cnorm;
    p.cmax = cmax;
    p.codes = codes;
    p.cand2 = cand2;
    p.fix_count = counter;
    p.fix_list = list;
    p.fix_cap = sc.cap;
    p.stats = stats;
    p.M = M;
    p.dsub = dsub;
    p.KS = ts.KS;
    p.CH = ts.CH;
    p.nchunk = ts.nchunk;
    p.nstages = ts.nstages;
    p.groups = M / ts.KS;
    p.ntiles = (n + TILE_ROWS - 1) / TILE_ROWS;
    const int sms = device_sm_count();
    // ~4 work items per CTA when the problem allows, at most 64 row tiles per item
    int64_t tpi = (p.ntiles * p.groups) / ((int64_t)sms * 4);
    if (tpi < 1) tpi = 1;
    if (tpi > 64) tpi = 64;
    p.tiles_per_item = tpi;
    p.nitems = (int64_t)p.groups * ((p.ntiles + tpi - 1) / tpi);
    p.debug = 0;
#ifdef B200VQ_DEBUG_BUILD   // tools/ A/B builds only: the debug bits change results, so release builds never read them
    {
        const char* dbg = getenv("B200VQ_TC_DEBUG");
        p.debug = dbg ? atoi(dbg) : 0;
    }
#endif
    int grid = (int)(p.nitems < sms ? p.nitems : sms);
    const size_t smem = ts.sp.total + 1024;
#ifdef B200VQ_DEBUG_BUILD
    auto kern = p.debug ? assign_tc_kernel<true> : assign_tc_kernel<false>;
#else
    auto kern = assign_tc_kernel<false>;
#endif
    VQ_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<grid, THREADS, smem, st>>>(map_x, map_b, map_bx, map_ax, p);
    VQ_LAUNCH_CHECK();
    {
        int64_t pg = ((p.n * (int64_t)M + 15) / 16 + PF_THREADS - 1) / PF_THREADS;
        if (pg > (int64_t)sms * 8) pg = (int64_t)sms * 8;
        pair_fixup_kernel<<<(int)pg, PF_THREADS, 0, st>>>(p);
        VQ_LAUNCH_CHECK();
    }
    {
        const size_t fsmem = ((size_t)dsub * KSUB + KSUB + (size_t)FIX_WARPS * FIX_ROWS * dsub) * sizeof(float);
        VQ_CUDA(cudaFuncSetAttribute(fixup_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fsmem));
        const unsigned long long chunks = (sc.cap + FIX_CHUNK - 1) / FIX_CHUNK;
        dim3 fgrid((unsigned)chunks, (unsigned)M);
        fixup_kernel<<<fgrid, FIX_WARPS * 32, fsmem, st>>>(p);
    }
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}
