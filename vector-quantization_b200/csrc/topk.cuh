// Block-level helpers shared by the scan kernels: bitonic sort of composite
// (value,id) keys in shared memory, partial-list emission and the cross-chunk
// merge kernel.
#pragma once
#include "common.cuh"

namespace b200vq {

#ifdef __CUDACC__

__host__ __device__ __forceinline__ int next_pow2(int v) {
    int p = 1;
    while (p < v) p <<= 1;
    return p;
}

// Sort keys[0..n2) ascending with the whole CTA. n2 must be a power of two.
// Ends with a __syncthreads().
// A warp's 32 consecutive compare-exchange slots t touch one aligned block of 64 keys whenever stride <= 32, and the
// slot -> thread map is the same in every stage: consecutive stages with stride <= 32 only need a warp barrier
// (10 CTA barriers instead of 45 for 512 keys; the CTA barrier per stage was 65 % of the screened scan's finish).
__device__ __forceinline__ void block_bitonic_sort(uint64_t* keys, int n2) {
    bool wide_prev = true;      // the keys were written by arbitrary threads
    for (int size = 2; size <= n2; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            const bool wide = stride >= 64;
            if (wide || wide_prev) __syncthreads();
            else __syncwarp();
            wide_prev = wide;
            for (int t = threadIdx.x; t < (n2 >> 1); t += blockDim.x) {
                int lo = 2 * t - (t & (stride - 1));
                int hi = lo + stride;
                bool up = ((lo & size) == 0);
                uint64_t a = keys[lo], b = keys[hi];
                if ((a > b) == up) {
                    keys[lo] = b;
                    keys[hi] = a;
                }
            }
        }
    }
    __syncthreads();
}

// k-th smallest (1 <= k <= n) of n order-preserving integer images of floats (float_to_ordered) in shared memory,
// found by a radix select: four passes over 8-bit digits, most significant first (histogram with shared atomics,
// warp 0 locates the digit). Whole CTA, any multiple of 32 threads; `hist` = 256 words and `sel` = 2 words of shared
// memory. Duplicates count with their multiplicity. A sort is only needed where the ORDER of the k keys matters.
__device__ __forceinline__ uint32_t block_kth_smallest(const uint32_t* keys, int n, uint32_t k, uint32_t* hist,
                                                       uint32_t* sel) {
    const int lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        sel[0] = 0u;
        sel[1] = k;
    }
    for (int pass = 0; pass < 4; ++pass) {
        const int shift = 24 - 8 * pass;
        __syncthreads();
        for (int i = threadIdx.x; i < 256; i += blockDim.x) hist[i] = 0u;
        __syncthreads();
        const uint32_t prefix = sel[0], kk = sel[1];
        for (int i = threadIdx.x; i < n; i += blockDim.x) {
            const uint32_t key = keys[i];
            if (pass == 0 || ((key ^ prefix) >> (shift + 8)) == 0u) atomicAdd(&hist[(key >> shift) & 255u], 1u);
        }
        __syncthreads();
        if (threadIdx.x < 32) {
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
                        sel[0] = prefix | ((uint32_t)(lane * 8 + j) << shift);
                        sel[1] = kk - run;
                    }
                    run += c[j];
                }
            }
        }
    }
    __syncthreads();
    return sel[0];
}

// Emit the CTA's best k: gather every warp's sorted list into `mbuf`
// (n2 = pow2 >= SCAN_WARPS*k), sort, write k keys to dst.
template <int CAP>
__device__ __forceinline__ void block_emit_topk(WarpTopK<CAP>& tk, uint64_t* mbuf, int n2, int k,
                                                uint64_t* dst) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    tk.finish();
    for (int i = threadIdx.x; i < n2; i += blockDim.x) mbuf[i] = KEY_INF;
    __syncthreads();
    for (int i = lane; i < tk.count; i += 32) mbuf[w * k + i] = tk.buf[i];
    block_bitonic_sort(mbuf, n2);
    for (int j = threadIdx.x; j < k; j += blockDim.x) dst[j] = mbuf[j];
}

// Partial results layout: part[(q * nparts + p) * k + j], u64 keys ascending,
// KEY_INF padded.
//
// Merge kernel: one CTA per query, streams nparts*k keys through a 2*TILE-key
// shared buffer, keeping the best k after every sort. `negate` un-does the
// sign flip applied to inner-product scores.
template <int TILE>
__global__ void merge_partials_kernel(const uint64_t* __restrict__ part, int nparts, int k,
                                      int negate, float* __restrict__ out_vals,
                                      uint32_t* __restrict__ out_ids, const uint32_t* __restrict__ gate = nullptr,
                                      uint32_t gate_cap = 0) {
    __shared__ uint64_t keys[TILE];
    const int q = blockIdx.x;
    if (gate && gate[q] <= gate_cap) return;   // gated launch: see adc_scan_rot_kernel
    const int total = nparts * k;
    const uint64_t* src = part + (size_t)q * total;
    const int kpad = k;  // the best k live in keys[0..k)
    for (int i = threadIdx.x; i < TILE; i += blockDim.x) keys[i] = KEY_INF;
    __syncthreads();
    const int step = TILE - kpad;
    for (int base = 0; base < total; base += step) {
        for (int i = threadIdx.x; i < step; i += blockDim.x) {
            int s = base + i;
            keys[kpad + i] = (s < total) ? src[s] : KEY_INF;
        }
        block_bitonic_sort(keys, TILE);
    }
    for (int j = threadIdx.x; j < k; j += blockDim.x) {
        uint64_t key = keys[j];
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

#endif

}  // namespace b200vq
