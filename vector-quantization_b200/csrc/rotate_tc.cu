// fp32-accurate rotation GEMM on the sm_100a tensor cores:  Y (n x N) = X (n x K) . W^T,  W (N x K).
//
// Replaces the sgemm behind faiss OPQMatrix.apply / reverse_transform as driven by the reference at
// src/haag_vq/methods/optimized_product_quantization.py:30-34 (codes = pq.compute_codes(opq.apply(X)),
// decode = reverse_transform(pq.decode(codes))) and the per-segment rotations of the SAQ engine
// (external/saq/include/saq/quantizer.h:66-72, rotator.h:43-48). At config C2 the rotation is
// 2 D^2 = 4.7 MFLOP per vector against 0.8 MFLOP for the assignment that follows it, so it decides the
// OPQ encode rate.
//
// tcgen05 kind::tf32 keeps 11 significant bits per operand, so one product pass is only a screen.
// Here every operand is split into two TF32-representable parts, a = a_h + a_l with
// a_h = round-to-nearest(a) (11 bits) and a_l = round-to-nearest(a - a_h) (the subtraction is exact),
// |a - a_h - a_l| <= 2^-22 |a|, and three product passes accumulate in fp32 in TMEM:
//      x.w ~= x_l.w_h + x_h.w_l + x_h.w_h           (the dropped x_l.w_l is <= 2^-22 |x||w|)
// which leaves an error of a few 2^-22 per product. The tensor core's own accumulation is NOT
// round-to-nearest (measured: a 576-step chain loses ~1e-5 relative, about 2^-25.6 per step, one
// sided), so the TMEM accumulator only ever holds ONE K chunk (12 accumulation steps): the epilogue
// warps drain it after every chunk and keep the running sum in registers with ordinary fp32 adds.
// Result: the accuracy class of an fp32 sgemm (tests/test_gpu_rotate_tc.py measures both against fp64).
//
// W is split once by rotate_split_kernel (it is reused for every batch). X is split inside the GEMM:
// TMA lands a raw 128 x 32 tile, four converter warps rewrite it in place as x_h and write x_l to a
// second tile (same swizzled addresses, so the layout never has to be decoded), fence to the async
// proxy and hand the stage to the MMA thread. Persistent CTAs, one per SM; per (128-row, 256-column)
// unit and 32-float K chunk 12 tcgen05.mma (128 x 256 x 8) into one of two TMEM accumulators while
// eight epilogue warps (TMEM lane quarter x column half, 128 running sums per thread) drain the
// other. Warp groups: 0 = {TMA producer, MMA issuer (owns TMEM), 2 idle}, 1 = converters,
// 2-3 = epilogue; setmaxnreg moves registers from groups 0-1 (64) to the epilogue (184).
#include "tc_common.cuh"

namespace b200vq {
namespace rot {
using namespace tc;

constexpr int TILE_ROWS = 128;
constexpr int TILE_N = 256;
constexpr int CH = 32;                       // K chunk: 32 floats = one 128-byte swizzle row
constexpr int NST = 2;
constexpr int THREADS = 512;
constexpr uint32_t A_BYTES = TILE_ROWS * CH * 4;      // 16 KB
constexpr uint32_t B_BYTES = TILE_N * CH * 4;         // 32 KB
constexpr uint32_t OFF_XH = 0, OFF_XL = A_BYTES, OFF_WH = 2 * A_BYTES, OFF_WL = 2 * A_BYTES + B_BYTES;
constexpr uint32_t STAGE_BYTES = 2 * A_BYTES + 2 * B_BYTES;   // 96 KB
constexpr uint32_t TX_BYTES = A_BYTES + 2 * B_BYTES;          // what TMA delivers per stage
constexpr uint32_t BAR_OFF = NST * STAGE_BYTES;
constexpr uint32_t SMEM_BYTES = BAR_OFF + 256 + 1024;          // + alignment slack
enum { B_FULL = 0, B_CONV = NST, B_EMPTY = 2 * NST, B_TFULL = 3 * NST, B_TEMPTY = 3 * NST + 2, B_TMEM = 3 * NST + 4 };

struct Params {
    float* y;
    int64_t n, ldy;
    int N, nchunk, ntn;       // output columns, K / 32, column tiles
    int64_t ntr;              // row tiles
    const float* x_sub;       // (K) subtracted from every row of X before the split, or null
    const float* col_bias;    // (N) added to every output row, or null
    const float* row_scale;   // (n) output row i is multiplied by row_scale[i] (before the bias), or null
    int transpose_out;        // 1: element (row, col) is stored at y[col * ldy + row]
};

// round-to-nearest (ties away) to 11 significant bits; the result is exactly representable in TF32
__device__ __forceinline__ float rn_tf32(float a) {
    return __uint_as_float((__float_as_uint(a) + 0x1000u) & 0xFFFFE000u);
}

// W (N x K, row stride ldw) -> wh, wl (N x K, dense): the two TF32 parts of every element.
__global__ void rotate_split_kernel(const float* __restrict__ w, int64_t ldw, int64_t N, int64_t K,
                                    float* __restrict__ wh, float* __restrict__ wl) {
    const int64_t total = N * K;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / K, c = i - r * K;
        const float a = w[r * ldw + c];
        const float h = rn_tf32(a);
        wh[i] = h;
        wl[i] = rn_tf32(a - h);
    }
}

__global__ void __launch_bounds__(THREADS, 1)
rotate_tc_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_wh,
                 const __grid_constant__ CUtensorMap map_wl, const Params p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;
    uint8_t* gbase = smem_raw + (base - raw);
    const uint32_t bar0 = base + BAR_OFF;
    auto bar = [&](int slot) { return bar0 + 8u * (uint32_t)slot; };
    volatile uint32_t* tmem_ptr_smem = reinterpret_cast<volatile uint32_t*>(gbase + BAR_OFF + 8 * B_TMEM);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t nunits = p.ntr * p.ntn;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&map_x);
        tma_prefetch_desc(&map_wh);
        tma_prefetch_desc(&map_wl);
        for (int i = 0; i < NST; ++i) {
            mbar_init(bar(B_FULL + i), 1);
            mbar_init(bar(B_CONV + i), 4);     // one arrival per converter warp
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

    // setmaxnreg sits inside each role's branch: ptxas budgets the code it dominates
    if (warp < 4) {
      asm volatile("setmaxnreg.dec.sync.aligned.u32 64;");
      if (warp == 0) {
        // ===================================================== TMA producer ===
        if (lane == 0) {
            uint32_t st = 0, phase = 0;
            for (int64_t unit = blockIdx.x; unit < nunits; unit += gridDim.x) {
                const int64_t t = unit / p.ntn;
                const int u = (int)(unit - t * p.ntn);
                const int row0 = (int)(t * TILE_ROWS);
                for (int c = 0; c < p.nchunk; ++c) {
                    mbar_wait_backoff(bar(B_EMPTY + st), phase ^ 1, 256);
                    mbar_expect_tx(bar(B_FULL + st), TX_BYTES);
                    const uint32_t sb = base + st * STAGE_BYTES;
                    tma_load_2d(sb + OFF_XH, &map_x, bar(B_FULL + st), c * CH, row0);
                    tma_load_2d(sb + OFF_WH, &map_wh, bar(B_FULL + st), c * CH, u * TILE_N);
                    tma_load_2d(sb + OFF_WL, &map_wl, bar(B_FULL + st), c * CH, u * TILE_N);
                    if (++st == NST) { st = 0; phase ^= 1; }
                }
            }
        }
      } else if (warp == 1) {
        // ======================================================= MMA issuer ===
        if (lane == 0) {
            uint32_t st = 0, phase = 0, acc = 0, t_phase = 0;
            for (int64_t unit = blockIdx.x; unit < nunits; unit += gridDim.x) {
                for (int c = 0; c < p.nchunk; ++c) {
                    mbar_wait_backoff(bar(B_TEMPTY + acc), t_phase ^ 1, 32);
                    mbar_wait_backoff(bar(B_CONV + st), phase, 32);
                    tc_fence_after();
                    const uint32_t d = tmem_base + acc * 256u;
                    const uint32_t sb = base + st * STAGE_BYTES;
#pragma unroll
                    for (int kk = 0; kk < CH / 8; ++kk) {
                        const uint64_t xh = make_kmajor_desc(sb + OFF_XH + kk * 32, 1024, 2);
                        const uint64_t xl = make_kmajor_desc(sb + OFF_XL + kk * 32, 1024, 2);
                        const uint64_t wh = make_kmajor_desc(sb + OFF_WH + kk * 32, 1024, 2);
                        const uint64_t wl = make_kmajor_desc(sb + OFF_WL + kk * 32, 1024, 2);
                        umma_tf32(d, xl, wh, IDESC_TF32_128x256, kk ? 1u : 0u);   // small terms first
                        umma_tf32(d, xh, wl, IDESC_TF32_128x256, 1);
                        umma_tf32(d, xh, wh, IDESC_TF32_128x256, 1);
                    }
                    umma_commit(bar(B_EMPTY + st));
                    umma_commit(bar(B_TFULL + acc));      // one K chunk per accumulator: drained by the epilogue
                    if (++st == NST) { st = 0; phase ^= 1; }
                    if (++acc == 2) { acc = 0; t_phase ^= 1; }
                }
            }
        }
      }
    } else if (warp < 8) {
        // ======================================================= converters ===
        asm volatile("setmaxnreg.dec.sync.aligned.u32 64;");
        const int tid = threadIdx.x - 128;     // 0..127
        // float4 slot i * 128 + tid of the SWIZZLE_128B tile is row i * 16 + tid / 8, physical 16-byte chunk
        // tid % 8, i.e. LOGICAL chunk (tid % 8) ^ (row % 8) = (tid & 7) ^ ((tid >> 3) & 7) for every i: a thread
        // always sees the same four columns of the K chunk, so x_sub costs one float4 per chunk
        const int lcol = 4 * ((tid & 7) ^ ((tid >> 3) & 7));
        uint32_t st = 0, phase = 0;
        for (int64_t unit = blockIdx.x; unit < nunits; unit += gridDim.x) {
            for (int c = 0; c < p.nchunk; ++c) {
                float4 sub = make_float4(0.f, 0.f, 0.f, 0.f);
                if (p.x_sub) sub = __ldg(reinterpret_cast<const float4*>(p.x_sub + c * CH + lcol));
                mbar_wait(bar(B_FULL + st), phase);
                float4* xh = reinterpret_cast<float4*>(gbase + st * STAGE_BYTES + OFF_XH);
                float4* xl = reinterpret_cast<float4*>(gbase + st * STAGE_BYTES + OFF_XL);
                float4 v[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    v[i] = xh[i * 128 + tid];
                    v[i].x = __fsub_rn(v[i].x, sub.x); v[i].y = __fsub_rn(v[i].y, sub.y);
                    v[i].z = __fsub_rn(v[i].z, sub.z); v[i].w = __fsub_rn(v[i].w, sub.w);
                }
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    float4 h, l;
                    h.x = rn_tf32(v[i].x); l.x = rn_tf32(v[i].x - h.x);
                    h.y = rn_tf32(v[i].y); l.y = rn_tf32(v[i].y - h.y);
                    h.z = rn_tf32(v[i].z); l.z = rn_tf32(v[i].z - h.z);
                    h.w = rn_tf32(v[i].w); l.w = rn_tf32(v[i].w - h.w);
                    xh[i * 128 + tid] = h;
                    xl[i * 128 + tid] = l;
                }
                fence_proxy_async();            // generic-proxy stores -> visible to tcgen05.mma
                __syncwarp();
                if (lane == 0) mbar_arrive(bar(B_CONV + st));
                if (++st == NST) { st = 0; phase ^= 1; }
            }
        }
    } else {
        // ========================================================= epilogue ===
        asm volatile("setmaxnreg.inc.sync.aligned.u32 184;");
        const int q4 = warp & 3;                              // TMEM lane quarter this warp may read
        const int half = (warp - 8) >> 2;                     // column half of the 256-wide tile
        const uint32_t lane_addr = tmem_base + ((uint32_t)(q4 * 32) << 16) + (uint32_t)half * 128u;
        uint32_t acc = 0, t_phase = 0;
        for (int64_t unit = blockIdx.x; unit < nunits; unit += gridDim.x) {
            const int64_t t = unit / p.ntn;
            const int u = (int)(unit - t * p.ntn);
            const int64_t row = t * TILE_ROWS + q4 * 32 + lane;
            float sum[128];
#pragma unroll
            for (int j = 0; j < 128; ++j) sum[j] = 0.f;
            for (int c = 0; c < p.nchunk; ++c) {
                mbar_wait(bar(B_TFULL + acc), t_phase);
                tc_fence_after();
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    float v[32];
                    tmem_ld32(lane_addr + acc * 256u + g * 32, v);
#pragma unroll
                    for (int j = 0; j < 32; ++j) sum[g * 32 + j] = __fadd_rn(sum[g * 32 + j], v[j]);
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(bar(B_TEMPTY + acc));
                if (++acc == 2) { acc = 0; t_phase ^= 1; }
            }
            const int col0 = u * TILE_N + half * 128;
            const int ncol = p.N - col0;                  // valid columns from col0 on (may exceed 128)
            if (row < p.n && (p.row_scale || p.col_bias)) {
                const float rs = p.row_scale ? __ldg(p.row_scale + row) : 1.f;
#pragma unroll
                for (int j = 0; j < 128; ++j) {
                    float v = p.row_scale ? __fmul_rn(sum[j], rs) : sum[j];
                    if (p.col_bias && j < ncol) v = __fadd_rn(v, __ldg(p.col_bias + col0 + j));
                    sum[j] = v;
                }
            }
            if (p.transpose_out) {
                // lanes = 32 consecutive rows: every column is one coalesced 128-byte store
                if (row < p.n) {
#pragma unroll
                    for (int j = 0; j < 128; ++j)
                        if (j < ncol) p.y[(int64_t)(col0 + j) * p.ldy + row] = sum[j];
                }
            } else if (row < p.n) {
                float* yr = p.y + row * p.ldy + col0;
#pragma unroll
                for (int j = 0; j < 128; j += 4) {
                    if (j + 3 < ncol) {
                        *reinterpret_cast<float4*>(yr + j) = make_float4(sum[j], sum[j + 1], sum[j + 2], sum[j + 3]);
                    } else {
#pragma unroll
                        for (int e = 0; e < 4; ++e)
                            if (j + e < ncol) yr[j + e] = sum[j + e];
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

}  // namespace rot
}  // namespace b200vq

using namespace b200vq;
using namespace b200vq::rot;

// Shapes the tensor-core rotation covers; anything else is the caller's library GEMM.
extern "C" int b200vq_rotate_tc_applies(int64_t n, int64_t ldx, int64_t N, int64_t K, int64_t ldy) {
    return n > 0 && N > 0 && K >= CH && K % CH == 0 && ldx % 4 == 0 && ldy % 4 == 0 && n < (1ll << 31) &&
           N < (1 << 30);
}

extern "C" size_t b200vq_rotate_tc_scratch_bytes(int64_t N, int64_t K) {
    return (N > 0 && K > 0) ? (size_t)2 * (size_t)N * (size_t)K * sizeof(float) : 0;
}

// Split W (N x K) into its two TF32 parts inside `wsplit` (once per matrix).
extern "C" int b200vq_rotate_tc_prepare(const float* w, int64_t ldw, int64_t N, int64_t K, void* wsplit,
                                        size_t wsplit_bytes, void* stream) {
    VQ_CHECK_ARG(w && wsplit && N > 0 && K > 0 && ldw >= K, "rotate_tc_prepare: bad arguments");
    VQ_CHECK_ARG(wsplit_bytes >= (size_t)2 * N * K * sizeof(float), "rotate_tc_prepare: scratch too small");
    VQ_CHECK_ARG((reinterpret_cast<uintptr_t>(wsplit) & 15) == 0, "rotate_tc_prepare: scratch must be 16-byte aligned");
    float* wh = reinterpret_cast<float*>(wsplit);
    float* wl = wh + (size_t)N * K;
    const int64_t total = N * K;
    int64_t grid = (total + 255) / 256;
    if (grid > 148 * 16) grid = 148 * 16;
    rotate_split_kernel<<<(unsigned)grid, 256, 0, as_stream(stream)>>>(w, ldw, N, K, wh, wl);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

// Y (n x N, row stride ldy) = X (n x K, row stride ldx) . W^T with W prepared by b200vq_rotate_tc_prepare.
extern "C" int b200vq_rotate_tc(const float* x, int64_t n, int64_t ldx, const void* wsplit, int64_t N, int64_t K,
                                float* y, int64_t ldy, void* stream) {
    return b200vq_rotate_tc_ex(x, n, ldx, wsplit, N, K, nullptr, nullptr, nullptr, 0, y, ldy, stream);
}

// Y = row_scale[i] * ((X - x_sub) . W^T) + col_bias, optionally stored transposed (y[col * ldy + row]).
extern "C" int b200vq_rotate_tc_ex(const float* x, int64_t n, int64_t ldx, const void* wsplit, int64_t N, int64_t K,
                                   const float* x_sub, const float* row_scale, const float* col_bias,
                                   int transpose_out, float* y, int64_t ldy, void* stream) {
    VQ_CHECK_ARG(n >= 0 && N > 0 && K > 0 && ldx >= K && ldy >= (transpose_out ? n : N), "rotate_tc: bad shape");
    if (n == 0) return B200VQ_OK;
    VQ_CHECK_ARG(x && wsplit && y, "rotate_tc: null pointer");
    VQ_CHECK_ARG(b200vq_rotate_tc_applies(n, ldx, N, K, transpose_out ? 4 : ldy),
                 "rotate_tc: shape not covered (K %% 32, ld %% 4)");
    VQ_CHECK_ARG(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(wsplit)) & 15) == 0 &&
                     (reinterpret_cast<uintptr_t>(y) & (transpose_out ? 3 : 15)) == 0 &&
                     (reinterpret_cast<uintptr_t>(x_sub) & 15) == 0,
                 "rotate_tc: pointers must be 16-byte aligned");
    const float* wh = reinterpret_cast<const float*>(wsplit);
    const float* wl = wh + (size_t)N * K;
    CUtensorMap map_x, map_wh, map_wl;
    int rc;
    if ((rc = encode_2d(&map_x, x, (uint64_t)n, (uint64_t)K, (uint64_t)ldx, TILE_ROWS, CH, CU_TENSOR_MAP_SWIZZLE_128B)))
        return rc;
    if ((rc = encode_2d(&map_wh, wh, (uint64_t)N, (uint64_t)K, (uint64_t)K, TILE_N, CH, CU_TENSOR_MAP_SWIZZLE_128B)))
        return rc;
    if ((rc = encode_2d(&map_wl, wl, (uint64_t)N, (uint64_t)K, (uint64_t)K, TILE_N, CH, CU_TENSOR_MAP_SWIZZLE_128B)))
        return rc;
    Params p;
    p.y = y; p.n = n; p.ldy = ldy; p.N = (int)N; p.nchunk = (int)(K / CH);
    p.ntn = (int)((N + TILE_N - 1) / TILE_N);
    p.ntr = (n + TILE_ROWS - 1) / TILE_ROWS;
    p.x_sub = x_sub; p.col_bias = col_bias; p.row_scale = row_scale; p.transpose_out = transpose_out ? 1 : 0;
    const int64_t units = p.ntr * p.ntn;
    const int sms = device_sm_count();
    VQ_CUDA(cudaFuncSetAttribute(rotate_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES));
    rotate_tc_kernel<<<(unsigned)(units < sms ? units : sms), THREADS, SMEM_BYTES, as_stream(stream)>>>(map_x, map_wh,
                                                                                                     map_wl, p);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}
