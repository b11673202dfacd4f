// Data-fit per-dimension scalar codebooks on the device: 1-D k-means (Lloyd) over a sorted column with prefix sums.
//
// Replaces the SAQ engine's CPU builder as called by the reference's per-dimension methods:
//   external/saq/src/preprocessing/codebook_builder.cpp:341-420  build_codebook_lloyd (one level per call here)
//   external/saq/src/preprocessing/codebook_builder.cpp:146-188  init_kmeanspp (LloydOpts default init, seed 0)
//   external/saq/src/preprocessing/codebook_builder.cpp:64-128   lloyd_k (assign by upper_bound of the float
//       mid-point, centroid = prefix-sum mean, one empty-cell repair per iteration, budget 2k)
//   consumed by src/haag_vq/methods/rank_aware_quantization.py:189-214 (codebook="lloyd") and by
//   external/saq/src/quantization_plan.cpp:21-51 (per-dimension MSE table of the greedy allocator).
//
// One CTA per dimension. The column arrives sorted with its fp64 prefix sums (sum, sum of squares); the draws
// are std::mt19937_64 through libstdc++'s uniform_int_distribution / uniform_real_distribution, restated
// here (oracle/codebook.py holds the same restatement, pinned on the reference's compiled builder). The
// k-means++ seeding keeps the squared distance to the nearest chosen centre of every point in a global
// scratch row and fuses "update with the new centre" with "sum for the next draw" into one pass; the sums run
// over per-thread contiguous chunks (the reference adds left to right: the totals agree to fp64 rounding,
// so a draw can differ only when its target falls within ~1e-16 relative of a prefix boundary).
#include "common.cuh"

namespace b200vq {
namespace cbk {

constexpr int THREADS = 256;
constexpr int MAXK = 8192;      // 13 bits: the engine's kMaxQuantBit (the greedy allocator's MSE table goes that far)

struct Mt64 {
    unsigned long long mt[312];
    int idx;
};
__device__ void mt_seed(Mt64& g, unsigned long long seed) {
    g.mt[0] = seed;
    for (int i = 1; i < 312; ++i) g.mt[i] = 6364136223846793005ull * (g.mt[i - 1] ^ (g.mt[i - 1] >> 62)) + (unsigned long long)i;
    g.idx = 312;
}
__device__ unsigned long long mt_next(Mt64& g) {
    if (g.idx >= 312) {
        for (int i = 0; i < 312; ++i) {
            const unsigned long long x = (g.mt[i] & 0xFFFFFFFF80000000ull) | (g.mt[(i + 1) % 312] & 0x7FFFFFFFull);
            g.mt[i] = g.mt[(i + 156) % 312] ^ (x >> 1) ^ ((x & 1ull) ? 0xB5026F5AA96619E9ull : 0ull);
        }
        g.idx = 0;
    }
    unsigned long long x = g.mt[g.idx++];
    x ^= (x >> 29) & 0x5555555555555555ull;
    x ^= (x << 17) & 0x71D67FFFEDA60000ull;
    x ^= (x << 37) & 0xFFF7EEE000000000ull;
    x ^= x >> 43;
    return x;
}
// std::uniform_int_distribution<size_t>(0, hi) on a 64-bit generator (libstdc++ _S_nd: Lemire's method)
__device__ unsigned long long uniform_int(Mt64& g, unsigned long long hi) {
    const unsigned long long range = hi + 1ull;
    unsigned long long u = mt_next(g);
    unsigned long long low = u * range, high = __umul64hi(u, range);
    if (low < range) {
        const unsigned long long threshold = (0ull - range) % range;
        while (low < threshold) {
            u = mt_next(g);
            low = u * range;
            high = __umul64hi(u, range);
        }
    }
    return high;
}
// std::uniform_real_distribution<double>(0, b): generate_canonical<double, 53> * b
__device__ double uniform_real0(Mt64& g, double b) {
    double r = __ull2double_rn(mt_next(g)) / 18446744073709551616.0;
    if (r >= 1.0) r = 0.99999999999999988898;    // nextafter(1, 0)
    return __dmul_rn(r, b);
}

__device__ __forceinline__ double seg_mean(const double* ps, int64_t a, int64_t b) {      // [a, b] inclusive
    return __ddiv_rn(__dsub_rn(ps[b + 1], ps[a]), (double)(b - a + 1));
}
__device__ __forceinline__ double seg_sse(const double* ps, const double* psq, int64_t a, int64_t b) {
    const double c = (double)(b - a + 1), s = __dsub_rn(ps[b + 1], ps[a]);
    return __dsub_rn(__dsub_rn(psq[b + 1], psq[a]), __ddiv_rn(__dmul_rn(s, s), c));
}
__device__ void nudge(double* c, int k) {
    for (int i = 1; i < k; ++i)
        if (c[i] <= c[i - 1]) c[i] = __dadd_rn(c[i - 1], fmax(1e-9, __dmul_rn(fabs(c[i - 1]), 1e-6)));
}
// bitonic sort of k = 2^b doubles in shared memory by the whole CTA (ends with a barrier)
__device__ void sort_block(double* c, int k) {
    for (int size = 2; size <= k; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            __syncthreads();
            for (int t = threadIdx.x; t < (k >> 1); t += blockDim.x) {
                const int lo = 2 * t - (t & (stride - 1)), hi = lo + stride;
                const bool up = (lo & size) == 0;
                const double a = c[lo], b = c[hi];
                if ((a > b) == up) { c[lo] = b; c[hi] = a; }
            }
        }
    }
    __syncthreads();
}
__device__ void sort_small(double* c, int k) {      // insertion sort of a nearly sorted array, one thread
    for (int i = 1; i < k; ++i) {
        const double v = c[i];
        int j = i - 1;
        while (j >= 0 && c[j] > v) { c[j + 1] = c[j]; --j; }
        c[j + 1] = v;
    }
}
__device__ __forceinline__ int64_t upper_bound_f(const float* s, int64_t n, float v) {   // first i with s[i] > v
    int64_t lo = 0, hi = n;
    while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if (!(v < s[mid])) lo = mid + 1; else hi = mid;
    }
    return lo;
}

__global__ void __launch_bounds__(THREADS)
codebook_lloyd_kernel(const float* __restrict__ cols, int64_t ld, const double* __restrict__ ps_all,
                      const double* __restrict__ psq_all, int64_t n, const int* __restrict__ dim_bits, int fixed_bits,
                      unsigned long long seed, int max_iters, float tol_f, double* __restrict__ d2_all,
                      float* __restrict__ centroids, int* __restrict__ sizes, float* __restrict__ costs, int kmax) {
    extern __shared__ __align__(16) unsigned char cb_smem[];
    __shared__ Mt64 rng;
    __shared__ double part[THREADS];
    __shared__ double sh_d[4];
    __shared__ int64_t sh_i[4];
    const int dim = blockIdx.x, tid = threadIdx.x;
    const int bits = dim_bits ? dim_bits[dim] : fixed_bits;
    const float* s = cols + (int64_t)dim * ld;
    const double* ps = ps_all + (int64_t)dim * (n + 1);
    const double* psq = psq_all + (int64_t)dim * (n + 1);
    double* d2 = d2_all + (int64_t)dim * n;
    float* out = centroids + (int64_t)dim * kmax;
    const int k = 1 << bits;
    double* c = reinterpret_cast<double*>(cb_smem);                         // [kmax]
    int64_t* bnd = reinterpret_cast<int64_t*>(c + kmax);                    // [kmax + 1]
    if (bits == 0) {
        if (tid == 0) {
            out[0] = (float)seg_mean(ps, 0, n - 1);
            sizes[dim] = 1;
            costs[dim] = (float)__ddiv_rn(seg_sse(ps, psq, 0, n - 1), (double)n);
        }
        return;
    }
    // number of distinct values (block count of s[i] != s[i-1])
    {
        int64_t cnt = 0;
        for (int64_t i = tid + 1; i < n; i += THREADS) cnt += s[i] != s[i - 1];
        part[tid] = (double)cnt;
        __syncthreads();
        if (tid == 0) {
            double t = 1.0;
            for (int i = 0; i < THREADS; ++i) t += part[i];
            sh_i[0] = (int64_t)t;
        }
        __syncthreads();
    }
    const int64_t ndist = sh_i[0];
    if ((int64_t)k >= ndist) {          // every distinct value is its own centroid (not padded, like the reference)
        if (tid == 0) {
            int m = 0;
            for (int64_t i = 0; i < n; ++i)
                if (i == 0 || s[i] != s[i - 1]) out[m++] = s[i];
            sizes[dim] = m;
            costs[dim] = 0.f;
        }
        return;
    }
    // ---------------------------------------------------------------- k-means++ seeding
    const int64_t chunk = (n + THREADS - 1) / THREADS;
    const int64_t i0 = (int64_t)tid * chunk, i1 = (i0 + chunk < n) ? i0 + chunk : n;
    if (tid == 0) {
        mt_seed(rng, seed);
        c[0] = (double)s[uniform_int(rng, (unsigned long long)(n - 1))];
    }
    __syncthreads();
    {
        const double c0 = c[0];
        double acc = 0.0;
        for (int64_t i = i0; i < i1; ++i) {
            const double diff = __dsub_rn((double)s[i], c0);
            const double dd = __dmul_rn(diff, diff);
            d2[i] = dd;
            acc = __dadd_rn(acc, dd);
        }
        part[tid] = acc;
    }
    __syncthreads();
    for (int m = 1; m < k; ++m) {
        if (tid == 0) {
            double total = 0.0;
            for (int i = 0; i < THREADS; ++i) total = __dadd_rn(total, part[i]);
            if (total <= 0.0) {
                sh_i[1] = (int64_t)uniform_int(rng, (unsigned long long)(n - 1));
                sh_i[2] = -1;
            } else {
                const double target = uniform_real0(rng, total);
                double acc = 0.0;
                int t = 0;
                for (; t < THREADS - 1; ++t) {           // chunk whose running total reaches the target
                    const double nx = __dadd_rn(acc, part[t]);
                    if (nx >= target) break;
                    acc = nx;
                }
                sh_i[2] = t;
                sh_d[0] = acc;
                sh_d[1] = target;
                sh_i[1] = n - 1;
            }
        }
        __syncthreads();
        if (sh_i[2] == tid) {                            // the owning thread walks its chunk
            double acc = sh_d[0];
            const double target = sh_d[1];
            int64_t chosen = n - 1;
            for (int64_t i = i0; i < i1; ++i) {
                acc = __dadd_rn(acc, d2[i]);
                if (acc >= target) { chosen = i; break; }
            }
            sh_i[1] = chosen;
        }
        __syncthreads();
        const double cv = (double)s[sh_i[1]];
        if (tid == 0) c[m] = cv;
        double acc = 0.0;
        for (int64_t i = i0; i < i1; ++i) {              // update with the new centre + sum for the next draw
            const double diff = __dsub_rn((double)s[i], cv);
            const double dd = __dmul_rn(diff, diff);
            double cur = d2[i];
            if (dd < cur) { cur = dd; d2[i] = dd; }
            acc = __dadd_rn(acc, cur);
        }
        __syncthreads();
        part[tid] = acc;
        __syncthreads();
    }
    sort_block(c, k);
    if (tid == 0) nudge(c, k);
    __syncthreads();
    // ---------------------------------------------------------------- Lloyd iterations
    const double tol = (double)tol_f;
    int repairs = 0;
    auto assign = [&]() {
        for (int i = tid; i <= k; i += THREADS) {
            if (i == 0) bnd[0] = 0;
            else if (i == k) bnd[k] = n;
            else bnd[i] = upper_bound_f(s, n, (float)__dmul_rn(0.5, __dadd_rn(c[i - 1], c[i])));
        }
        __syncthreads();
        if (tid == 0)
            for (int i = 1; i < k; ++i)
                if (bnd[i] < bnd[i - 1]) bnd[i] = bnd[i - 1];
        __syncthreads();
    };
    for (int iter = 0; iter < max_iters; ++iter) {
        assign();
        double move = 0.0;
        for (int i = tid; i < k; i += THREADS) {
            const int64_t a = bnd[i], b = bnd[i + 1];
            const double nc = (b > a) ? seg_mean(ps, a, b - 1) : c[i];
            move = fmax(move, fabs(__dsub_rn(nc, c[i])));
            c[i] = nc;
        }
        part[tid] = move;
        __syncthreads();
        if (tid == 0) {
            double maxmove = 0.0;
            for (int i = 0; i < THREADS; ++i) maxmove = fmax(maxmove, part[i]);
            int state = maxmove < tol ? 1 : 0;           // 1 = converged, 2 = repaired (go on)
            if (repairs < 2 * k) {
                int empty = k;
                for (int i = 0; i < k; ++i)
                    if (bnd[i + 1] <= bnd[i]) { empty = i; break; }
                if (empty < k) {
                    int best = k;
                    double bsse = -1.0;
                    for (int j = 0; j < k; ++j) {
                        const int64_t a = bnd[j], b = bnd[j + 1];
                        if (b > a + 1) {
                            const double e = seg_sse(ps, psq, a, b - 1);
                            if (e > bsse) { bsse = e; best = j; }
                        }
                    }
                    if (best < k) {
                        const int64_t a = bnd[best], b = bnd[best + 1], mid = a + (b - a) / 2;
                        c[best] = seg_mean(ps, a, mid - 1);
                        c[empty] = seg_mean(ps, mid, b - 1);
                        sort_small(c, k);
                        nudge(c, k);
                        ++repairs;
                        state = 2;
                    }
                }
            }
            sh_i[3] = state;
        }
        __syncthreads();
        const int state = (int)sh_i[3];
        if (state == 2) continue;                         // reassign with the repaired centroids
        if (state == 1) break;
    }
    assign();
    if (tid == 0) {
        // the reference adds the cells' SSE in index order
        double sse = 0.0;
        for (int i = 0; i < k; ++i)
            if (bnd[i + 1] > bnd[i]) sse = __dadd_rn(sse, seg_sse(ps, psq, bnd[i], bnd[i + 1] - 1));
        float* o = out;
        int m = 0;
        // float(c) sorted, duplicates dropped, padded with the last value up to 2^bits
        for (int i = 0; i < k; ++i) {
            const float v = (float)c[i];
            if (m == 0 || v != o[m - 1]) o[m++] = v;     // c is non-decreasing, so is its fp32 image
        }
        const float last = o[m - 1];
        while (m < k) o[m++] = last;
        sizes[dim] = k;
        costs[dim] = (float)__ddiv_rn(sse, (double)n);
    }
}

}  // namespace cbk
}  // namespace b200vq

using namespace b200vq;

extern "C" size_t b200vq_codebook_lloyd_scratch_bytes(int64_t n, int D) {
    return (n > 0 && D > 0) ? (size_t)n * (size_t)D * sizeof(double) : 0;
}

extern "C" int b200vq_codebook_lloyd(const float* sorted_cols, int64_t ld, const double* ps, const double* psq, int64_t n,
                                     int D, const int* dim_bits, int fixed_bits, int max_bits, unsigned long long seed,
                                     int max_iters, float tol, float* centroids, int* sizes, float* costs, void* scratch,
                                     size_t scratch_bytes, void* stream) {
    VQ_CHECK_ARG(sorted_cols && ps && psq && centroids && sizes && costs && n >= 1 && D >= 1 && ld >= n && max_iters >= 0,
                 "codebook_lloyd: bad args");
    VQ_CHECK_ARG(max_bits >= 0 && max_bits <= 13 && (dim_bits || (fixed_bits >= 0 && fixed_bits <= max_bits)),
                 "codebook_lloyd: bits must be in [0, max_bits], max_bits <= 13");
    const int kmax = 1 << max_bits;
    const size_t smem = (size_t)kmax * 8 + (size_t)(kmax + 1) * 8;
    VQ_CUDA(cudaFuncSetAttribute(cbk::codebook_lloyd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    VQ_CHECK_ARG(scratch && scratch_bytes >= b200vq_codebook_lloyd_scratch_bytes(n, D),
                 "codebook_lloyd: scratch too small (%zu < %zu)", scratch_bytes, b200vq_codebook_lloyd_scratch_bytes(n, D));
    cbk::codebook_lloyd_kernel<<<D, cbk::THREADS, smem, as_stream(stream)>>>(
        sorted_cols, ld, ps, psq, n, dim_bits, fixed_bits, seed, max_iters, tol, reinterpret_cast<double*>(scratch), centroids,
        sizes, costs, kmax);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}
