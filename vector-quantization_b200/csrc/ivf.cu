// IVF search over a cluster-sorted matrix of reconstructions.
//
// Replaces the per-query Python probe loop of the reference's IVF shell
// (src/haag_vq/methods/search/ivf_quantized_index.py:98-146: for every query, decompress the probed clusters, exact
// distance to the reconstructions, top-k) with the same result: the exact top-k, in (distance, id) order, over the
// reconstructions of the clusters each query probes. The caller (haag_vq.methods.search.IvfQuantizedIndex) keeps the
// reconstructions R = decompress(codes_c) + centroid_c of all rows resident in HBM, sorted by cluster (CSR offsets
// row_off), with their squared norms and global ids; the probe lists come from b200vq_flat_topk over the centroids.
//
// A query-major scan would re-read every probed row once per query (nprobe / K of the database per query). The scan
// here is CLUSTER-major: the (query, probe) pairs are inverted into one query list per cluster (histogram, scan, fill:
// three small kernels, no host synchronisation), and CTA (cluster c, tile t) scores up to 64 of the queries probing c
// against all rows of c with the register-tiled exact kernel of b200vq_flat_topk (scan.cu: sequential fp32 FMA chain
// over d, key = max(0, (|q|^2 + |x|^2) - 2 q.x) or -q.x), so a row is read once per 64 probing queries. Every
// (query, probe) pair owns one partial list of k keys (value, global id); merge_partials_kernel merges the nprobe
// lists of a query. HBM-resident layout, no collective: the database shards by cluster across GPUs.
#include "common.cuh"
#include "topk.cuh"

namespace b200vq {
namespace ivf {

constexpr int Q_TILE = 64;    // queries per CTA tile
constexpr int N_TILE = 128;   // rows per step
constexpr int K_STEP = 16;    // reduction step
constexpr int THREADS = 256;

__global__ void count_kernel(const uint32_t* __restrict__ probe, int64_t total, int K, uint32_t* __restrict__ cnt) {
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const uint32_t c = probe[e];
        if (c < (uint32_t)K) atomicAdd(cnt + c, 1u);
    }
}
// exclusive scan of cnt[0..K) into off[0..K] (one CTA), cursor = copy of off
__global__ void __launch_bounds__(1024) offsets_kernel(const uint32_t* __restrict__ cnt, int K, uint32_t* __restrict__ off,
                                                       uint32_t* __restrict__ cursor) {
    __shared__ uint32_t wsum[32];
    __shared__ uint32_t carry;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < K; base += 1024) {
        const int i = base + threadIdx.x;
        const uint32_t v = i < K ? cnt[i] : 0u;
        uint32_t s = v;
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(FULL_MASK, s, o);
            if (lane >= o) s += t;
        }
        if (lane == 31) wsum[w] = s;
        __syncthreads();
        if (w == 0) {
            uint32_t t = wsum[lane];
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t u = __shfl_up_sync(FULL_MASK, t, o);
                if (lane >= o) t += u;
            }
            wsum[lane] = t;
        }
        __syncthreads();
        const uint32_t excl = carry + (w ? wsum[w - 1] : 0u) + s - v;
        if (i < K) {
            off[i] = excl;
            cursor[i] = excl;
        }
        __syncthreads();
        if (threadIdx.x == 1023) carry = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) off[K] = carry;
}
__global__ void fill_kernel(const uint32_t* __restrict__ probe, int64_t total, int K, uint32_t* __restrict__ cursor,
                            uint32_t* __restrict__ qlist) {
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const uint32_t c = probe[e];
        if (c < (uint32_t)K) qlist[atomicAdd(cursor + c, 1u)] = (uint32_t)e;
    }
}

// CTA (c, t): queries qlist[off[c] + 64 t ..) against rows [row_off[c], row_off[c + 1]).
template <int CAP>
__global__ void __launch_bounds__(THREADS)
scan_kernel(const float* __restrict__ R, int64_t ldr, int D, const float* __restrict__ rnorm,
            const uint32_t* __restrict__ gid, const int64_t* __restrict__ row_off, const float* __restrict__ queries,
            int64_t ldq, const uint32_t* __restrict__ qoff, const uint32_t* __restrict__ qlist, int nprobe, int metric,
            int k, uint64_t* __restrict__ part, const float* __restrict__ after_vals = nullptr,
            const uint32_t* __restrict__ after_ids = nullptr) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int c = blockIdx.x;
    const uint32_t e0 = qoff[c] + (uint32_t)blockIdx.y * Q_TILE, e1 = qoff[c + 1];
    if (e0 >= e1) return;
    const int nqt = (int)min((uint32_t)Q_TILE, e1 - e0);
    float* Qs = reinterpret_cast<float*>(smem_raw);            // [K_STEP][Q_TILE+4]
    float* Xs = Qs + K_STEP * (Q_TILE + 4);                    // [K_STEP][N_TILE+4]
    float* ds = Xs + K_STEP * (N_TILE + 4);                    // [Q_TILE][N_TILE+1]
    float* qn = ds + Q_TILE * (N_TILE + 1);                    // [Q_TILE]
    float* s_thr = qn + Q_TILE;                                // [Q_TILE]
    int* s_cnt = reinterpret_cast<int*>(s_thr + Q_TILE);       // [Q_TILE]
    uint32_t* s_ent = reinterpret_cast<uint32_t*>(s_cnt + Q_TILE);   // [Q_TILE] (query, probe) entry = q * nprobe + j
    uint64_t* lists = reinterpret_cast<uint64_t*>(
        (reinterpret_cast<uintptr_t>(s_ent + Q_TILE) + 15) & ~(uintptr_t)15);  // [Q_TILE][CAP]

    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int ty = tid >> 4, tx = tid & 15;
    const int64_t r0 = row_off[c], r1 = row_off[c + 1];
    if (tid < Q_TILE) s_ent[tid] = tid < nqt ? qlist[e0 + tid] : 0u;
    __syncthreads();
    for (int ql = w; ql < Q_TILE; ql += THREADS / 32) {
        float acc = 0.f;
        if (ql < nqt) {
            const float* p = queries + (int64_t)(s_ent[ql] / (uint32_t)nprobe) * ldq;
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

    for (int64_t t0 = r0; t0 < r1; t0 += N_TILE) {
        float acc[4][8];
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 8; ++b) acc[a][b] = 0.f;
        for (int k0 = 0; k0 < D; k0 += K_STEP) {
            __syncthreads();
            for (int idx = tid; idx < Q_TILE * K_STEP; idx += THREADS) {
                const int r = idx / K_STEP, kk = idx - r * K_STEP;
                float v = 0.f;
                if (r < nqt && k0 + kk < D) v = queries[(int64_t)(s_ent[r] / (uint32_t)nprobe) * ldq + k0 + kk];
                Qs[kk * (Q_TILE + 4) + r] = v;
            }
            for (int idx = tid; idx < N_TILE * K_STEP; idx += THREADS) {
                const int r = idx / K_STEP, kk = idx - r * K_STEP;
                const int64_t row = t0 + r;
                float v = 0.f;
                if (row < r1 && k0 + kk < D) v = R[row * ldr + k0 + kk];
                Xs[kk * (N_TILE + 4) + r] = v;
            }
            __syncthreads();
#pragma unroll
            for (int kk = 0; kk < K_STEP; ++kk) {
                const float4 qa = *reinterpret_cast<const float4*>(Qs + kk * (Q_TILE + 4) + ty * 4);
                const float4 xa = *reinterpret_cast<const float4*>(Xs + kk * (N_TILE + 4) + tx * 4);
                const float4 xb = *reinterpret_cast<const float4*>(Xs + kk * (N_TILE + 4) + 64 + tx * 4);
                const float qv[4] = {qa.x, qa.y, qa.z, qa.w};
                const float xv[8] = {xa.x, xa.y, xa.z, xa.w, xb.x, xb.y, xb.z, xb.w};
#pragma unroll
                for (int a = 0; a < 4; ++a)
#pragma unroll
                    for (int b = 0; b < 8; ++b) acc[a][b] = __fmaf_rn(qv[a], xv[b], acc[a][b]);
            }
        }
#pragma unroll
        for (int a = 0; a < 4; ++a) {
            const int ql = ty * 4 + a;
#pragma unroll
            for (int b = 0; b < 8; ++b) {
                const int rl = (b < 4) ? (tx * 4 + b) : (64 + tx * 4 + (b - 4));
                const int64_t row = t0 + rl;
                float v;
                if (metric == B200VQ_METRIC_L2) {
                    const float xn = (row < r1) ? rnorm[row] : 0.f;
                    v = fmaxf(__fmaf_rn(-2.f, acc[a][b], __fadd_rn(qn[ql], xn)), 0.f);
                } else {
                    v = -acc[a][b];
                }
                ds[ql * (N_TILE + 1) + rl] = v;
            }
        }
        __syncthreads();
        // per-query running top-k: warp w owns queries w*8 .. w*8+7; ids are the rows' global ids
        for (int qq = 0; qq < Q_TILE / (THREADS / 32); ++qq) {
            const int ql = w * (Q_TILE / (THREADS / 32)) + qq;
            if (ql >= nqt) break;
            WarpTopK<CAP> tk;
            tk.buf = lists + (size_t)ql * CAP;
            tk.k = k;
            tk.count = s_cnt[ql];
            tk.thr = s_thr[ql];
            uint64_t floor_key = 0;      // continuation (b200vq_ivf_search_after): only keys after the previous page's last
            if (after_vals) {
                const uint32_t qg = s_ent[ql] / (uint32_t)nprobe;
                const float av = after_vals[qg];
                floor_key = make_key(metric == B200VQ_METRIC_IP ? -av : av, after_ids[qg]);
            }
            for (int cc = 0; cc < N_TILE / 32; ++cc) {
                const int rl = cc * 32 + lane;
                const int64_t row = t0 + rl;
                const float dv = ds[ql * (N_TILE + 1) + rl];
                const uint32_t id = row < r1 ? gid[row] : 0u;
                bool ok = row < r1;
                if (after_vals) ok = ok && make_key(dv, id) > floor_key;
                tk.offer(ok, dv, id);
            }
            if (lane == 0) {
                s_cnt[ql] = tk.count;
                s_thr[ql] = tk.thr;
            }
            __syncwarp();
        }
    }
    __syncthreads();
    for (int qq = 0; qq < Q_TILE / (THREADS / 32); ++qq) {
        const int ql = w * (Q_TILE / (THREADS / 32)) + qq;
        if (ql >= nqt) break;
        WarpTopK<CAP> tk;
        tk.buf = lists + (size_t)ql * CAP;
        tk.k = k;
        tk.count = s_cnt[ql];
        tk.thr = s_thr[ql];
        tk.finish();
        uint64_t* dst = part + (size_t)s_ent[ql] * k;
        for (int j = lane; j < k; j += 32) dst[j] = (j < tk.count) ? tk.buf[j] : KEY_INF;
    }
}

// Row-wise arg-min of a score matrix (coarse k-means assignment: scores = |c|^2 - 2 x.c from the rotation GEMM with
// a column bias). One warp per row, 128-bit loads; the lowest index wins ties.
__global__ void __launch_bounds__(256) row_argmin_kernel(const float* __restrict__ y, int64_t n, int64_t ld, int K,
                                                         uint32_t* __restrict__ idx, float* __restrict__ val) {
    const int lane = threadIdx.x & 31;
    const int64_t wid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const bool vec = (ld % 4 == 0) && ((reinterpret_cast<uintptr_t>(y) & 15) == 0) && (K % 4 == 0);
    for (int64_t row = wid; row < n; row += nw) {
        const float* p = y + row * ld;
        float best = __int_as_float(0x7f800000);
        uint32_t bi = 0xffffffffu;
        if (vec) {
            for (int j = 4 * lane; j < K; j += 128) {
                const float4 v = ld_stream_f4(reinterpret_cast<const float4*>(p + j));
                if (v.x < best) { best = v.x; bi = j; }
                if (v.y < best) { best = v.y; bi = j + 1; }
                if (v.z < best) { best = v.z; bi = j + 2; }
                if (v.w < best) { best = v.w; bi = j + 3; }
            }
        } else {
            for (int j = lane; j < K; j += 32) {
                const float v = p[j];
                if (v < best) { best = v; bi = j; }
            }
        }
        for (int o = 16; o > 0; o >>= 1) {
            const float ob = __shfl_xor_sync(FULL_MASK, best, o);
            const uint32_t oi = __shfl_xor_sync(FULL_MASK, bi, o);
            if (ob < best || (ob == best && oi < bi)) { best = ob; bi = oi; }
        }
        if (lane == 0) {
            idx[row] = bi == 0xffffffffu ? 0u : bi;    // all-NaN row: cluster 0
            if (val) val[row] = best;
        }
    }
}

struct Plan {
    size_t off_cnt, off_off, off_cursor, off_qlist, off_part, total;
};
static Plan plan(int nq, int nprobe, int K, int k) {
    auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
    Plan p;
    size_t o = 0;
    p.off_cnt = o;     o = up(o + (size_t)K * 4);
    p.off_off = o;     o = up(o + (size_t)(K + 1) * 4);
    p.off_cursor = o;  o = up(o + (size_t)K * 4);
    p.off_qlist = o;   o = up(o + (size_t)nq * nprobe * 4);
    p.off_part = o;    o = up(o + (size_t)nq * nprobe * k * 8);
    p.total = o;
    return p;
}

template <int CAP>
static int launch_scan(const float* R, int64_t ldr, int D, const float* rnorm, const uint32_t* gid, const int64_t* row_off,
                       int K, const float* queries, int nq, int64_t ldq, const uint32_t* qoff, const uint32_t* qlist,
                       int nprobe, int metric, int k, uint64_t* part, cudaStream_t st, const float* after_vals,
                       const uint32_t* after_ids) {
    const size_t smem = (size_t)(K_STEP * (Q_TILE + 4) + K_STEP * (N_TILE + 4) + Q_TILE * (N_TILE + 1) + 4 * Q_TILE) * 4 +
                        16 + (size_t)Q_TILE * CAP * sizeof(uint64_t);
    VQ_CUDA(cudaFuncSetAttribute(scan_kernel<CAP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    // a cluster is probed by at most nq queries: ceil(nq / 64) tiles, the surplus CTAs exit at once
    dim3 grid((unsigned)K, (unsigned)((nq + Q_TILE - 1) / Q_TILE));
    scan_kernel<CAP><<<grid, THREADS, smem, st>>>(R, ldr, D, rnorm, gid, row_off, queries, ldq, qoff, qlist, nprobe, metric,
                                                  k, part, after_vals, after_ids);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

}  // namespace ivf
}  // namespace b200vq

using namespace b200vq;

extern "C" size_t b200vq_ivf_search_scratch_bytes(int nq, int nprobe, int K, int k) {
    if (nq <= 0 || nprobe <= 0 || K <= 0 || k <= 0) return 0;
    return ivf::plan(nq, nprobe, K, k).total;
}

static int ivf_search_impl(const float* recon, int64_t ldr, int D, const float* rnorm, const uint32_t* gid,
                           const int64_t* row_off, int K, const float* queries, int nq, int64_t ldq,
                           const uint32_t* probe, int nprobe, int metric, int k, const float* after_vals,
                           const uint32_t* after_ids, float* out_vals, uint32_t* out_ids,
                           void* scratch, size_t scratch_bytes, void* stream) {
    VQ_CHECK_ARG(recon && gid && row_off && queries && probe && out_vals && out_ids, "ivf_search: null pointer");
    VQ_CHECK_ARG(D > 0 && ldr >= D && ldq >= D && nq >= 0 && K > 0 && K <= (1 << 20) && nprobe > 0,
                 "ivf_search: bad shape");
    VQ_CHECK_ARG((int64_t)nq * nprobe < (1ll << 31), "ivf_search: nq * nprobe must be below 2^31");
    VQ_CHECK_ARG(k > 0 && k <= 128, "ivf_search: k must be in [1,128], got %d", k);
    VQ_CHECK_ARG(metric == B200VQ_METRIC_L2 || metric == B200VQ_METRIC_IP, "ivf_search: bad metric");
    VQ_CHECK_ARG(metric == B200VQ_METRIC_IP || rnorm, "ivf_search: L2 needs rnorm");
    if (nq == 0) return B200VQ_OK;
    const ivf::Plan p = ivf::plan(nq, nprobe, K, k);
    VQ_CHECK_ARG(scratch && scratch_bytes >= p.total && ((uintptr_t)scratch & 255) == 0,
                 "ivf_search: scratch too small or not 256-byte aligned (%zu < %zu)", scratch_bytes, p.total);
    cudaStream_t st = as_stream(stream);
    uint8_t* sb = reinterpret_cast<uint8_t*>(scratch);
    uint32_t* cnt = reinterpret_cast<uint32_t*>(sb + p.off_cnt);
    uint32_t* off = reinterpret_cast<uint32_t*>(sb + p.off_off);
    uint32_t* cursor = reinterpret_cast<uint32_t*>(sb + p.off_cursor);
    uint32_t* qlist = reinterpret_cast<uint32_t*>(sb + p.off_qlist);
    uint64_t* part = reinterpret_cast<uint64_t*>(sb + p.off_part);
    const int64_t total = (int64_t)nq * nprobe;
    VQ_CUDA(cudaMemsetAsync(cnt, 0, (size_t)K * 4, st));
    VQ_CUDA(cudaMemsetAsync(part, 0xFF, (size_t)total * k * 8, st));   // KEY_INF: pairs with an invalid / empty cluster
    int g = (int)((total + 255) / 256);
    const int gcap = device_sm_count() * 8;
    if (g > gcap) g = gcap;
    ivf::count_kernel<<<g, 256, 0, st>>>(probe, total, K, cnt);
    VQ_LAUNCH_CHECK();
    ivf::offsets_kernel<<<1, 1024, 0, st>>>(cnt, K, off, cursor);
    VQ_LAUNCH_CHECK();
    ivf::fill_kernel<<<g, 256, 0, st>>>(probe, total, K, cursor, qlist);
    VQ_LAUNCH_CHECK();
    int rc = (k <= 32) ? ivf::launch_scan<64>(recon, ldr, D, rnorm, gid, row_off, K, queries, nq, ldq, off, qlist, nprobe,
                                              metric, k, part, st, after_vals, after_ids)
                       : ivf::launch_scan<256>(recon, ldr, D, rnorm, gid, row_off, K, queries, nq, ldq, off, qlist, nprobe,
                                               metric, k, part, st, after_vals, after_ids);
    if (rc != B200VQ_OK) return rc;
    merge_partials_kernel<2048><<<nq, 256, 0, st>>>(part, nprobe, k, metric == B200VQ_METRIC_IP, out_vals, out_ids);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" int b200vq_ivf_search(const float* recon, int64_t ldr, int D, const float* rnorm, const uint32_t* gid,
                                 const int64_t* row_off, int K, const float* queries, int nq, int64_t ldq,
                                 const uint32_t* probe, int nprobe, int metric, int k, float* out_vals, uint32_t* out_ids,
                                 void* scratch, size_t scratch_bytes, void* stream) {
    return ivf_search_impl(recon, ldr, D, rnorm, gid, row_off, K, queries, nq, ldq, probe, nprobe, metric, k, nullptr,
                           nullptr, out_vals, out_ids, scratch, scratch_bytes, stream);
}

// Continuation: the best k probed rows AFTER (after_vals[q], after_ids[q]) in (distance, id) order - the next page of
// a search whose previous page ended there (any k = pages of at most 128). Same scratch as b200vq_ivf_search.
extern "C" int b200vq_ivf_search_after(const float* recon, int64_t ldr, int D, const float* rnorm, const uint32_t* gid,
                                       const int64_t* row_off, int K, const float* queries, int nq, int64_t ldq,
                                       const uint32_t* probe, int nprobe, int metric, int k, const float* after_vals,
                                       const uint32_t* after_ids, float* out_vals, uint32_t* out_ids, void* scratch,
                                       size_t scratch_bytes, void* stream) {
    VQ_CHECK_ARG(after_vals && after_ids, "ivf_search_after: null pointer");
    return ivf_search_impl(recon, ldr, D, rnorm, gid, row_off, K, queries, nq, ldq, probe, nprobe, metric, k, after_vals,
                           after_ids, out_vals, out_ids, scratch, scratch_bytes, stream);
}

extern "C" int b200vq_row_argmin(const float* y, int64_t n, int64_t ld, int K, uint32_t* idx, float* val, void* stream) {
    VQ_CHECK_ARG(n >= 0 && K > 0 && ld >= K, "row_argmin: bad shape");
    if (n == 0) return B200VQ_OK;
    VQ_CHECK_ARG(y && idx, "row_argmin: null pointer");
    int64_t g = (n * 32 + 255) / 256;
    const int64_t cap = (int64_t)device_sm_count() * 16;
    if (g > cap) g = cap;
    ivf::row_argmin_kernel<<<(int)g, 256, 0, as_stream(stream)>>>(y, n, ld, K, idx, val);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}
