// b200vq — shared device/host helpers for the sm_100a kernels.
//
// Everything here is header-only and included by every .cu in csrc/.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/b200vq.h"

namespace b200vq {

// ---------------------------------------------------------------- errors ---
// Thread-local last-error string, returned by b200vq_last_error().
char* last_error_buf();
void set_error(const char* fmt, ...);

#define VQ_CHECK_ARG(cond, ...)                 \
    do {                                        \
        if (!(cond)) {                          \
            ::b200vq::set_error(__VA_ARGS__);   \
            return B200VQ_ERR_INVALID;          \
        }                                       \
    } while (0)

#define VQ_CUDA(expr)                                                          \
    do {                                                                       \
        cudaError_t _e = (expr);                                               \
        if (_e != cudaSuccess) {                                               \
            ::b200vq::set_error("%s failed: %s (%s:%d)", #expr,                \
                                cudaGetErrorString(_e), __FILE__, __LINE__);   \
            return B200VQ_ERR_CUDA;                                            \
        }                                                                      \
    } while (0)

#define VQ_LAUNCH_CHECK()                                                      \
    do {                                                                       \
        cudaError_t _e = cudaGetLastError();                                   \
        if (_e != cudaSuccess) {                                               \
            ::b200vq::set_error("kernel launch failed: %s (%s:%d)",            \
                                cudaGetErrorString(_e), __FILE__, __LINE__);   \
            return B200VQ_ERR_CUDA;                                            \
        }                                                                      \
    } while (0)

static inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

int device_sm_count();

// scan.cu: exact SIMT flat search restricted to the queries with gate[q] != 0 (scratch sized by
// b200vq_flat_topk_scratch_bytes); the gated fallback of the tensor-core search in flat_tc.cu
int flat_topk_gated(const float* x, int64_t n, int64_t ldx, int D, const float* xnorm, const float* row_scale,
                    const float* queries, int nq, int64_t ldq, int metric, int k, float* out_vals, uint32_t* out_ids,
                    void* scratch, const uint32_t* gate, cudaStream_t st);

// rabitq.cu: SIMT 1-bit RaBitQ scan restricted to the queries with gate[q] != 0; the gated fallback of
// the tensor-core scan in rabitq_tc.cu
int rabitq1_scan_gated(const uint8_t* codes, int64_t n, int D, const uint32_t* planes, const float* qfac, int nq,
                       int qb, int metric, int k, float* out_vals, uint32_t* out_ids, void* scratch,
                       const uint32_t* gate, cudaStream_t st);

// --------------------------------------------------------- device helpers ---
#ifdef __CUDACC__

constexpr unsigned FULL_MASK = 0xffffffffu;

// Monotone map float -> uint32 so that a < b  <=>  key(a) < key(b)  (no NaNs).
__device__ __forceinline__ uint32_t float_to_ordered(float f) {
    uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ordered_to_float(uint32_t k) {
    uint32_t u = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
    return __uint_as_float(u);
}
// Composite 64-bit key: ascending by (value, id).
__device__ __forceinline__ uint64_t make_key(float v, uint32_t id) {
    return (static_cast<uint64_t>(float_to_ordered(v)) << 32) | id;
}
__device__ __forceinline__ float key_value(uint64_t k) {
    return ordered_to_float(static_cast<uint32_t>(k >> 32));
}
__device__ __forceinline__ uint32_t key_id(uint64_t k) { return static_cast<uint32_t>(k); }

constexpr uint64_t KEY_INF = 0xffffffffffffffffull;

// 128-bit streaming load that does not pollute L1 (codes are read once per pass).
__device__ __forceinline__ uint4 ld_stream_u4(const uint4* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}
__device__ __forceinline__ float ld_stream_f32(const float* p) {
    float r;
    asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(r) : "l"(p));
    return r;
}
__device__ __forceinline__ float4 ld_stream_f4(const float4* p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(p));
    return r;
}

// ---------------------------------------------------------------------------
// Warp-level running top-k over a stream of (value, id) candidates.
//
// Each warp owns CAP u64 slots of shared memory. Lanes offer a candidate when
// value <= thr (warp-uniform float threshold). Accepted candidates are appended
// with a ballot prefix; when the buffer could overflow on the next offer it is
// bitonic-sorted in shared memory and truncated to the best k, and thr becomes
// the k-th best value. Ordering is total: ascending (value, id).
// ---------------------------------------------------------------------------
template <int CAP>
struct WarpTopK {
    static_assert((CAP & (CAP - 1)) == 0 && CAP >= 64, "CAP must be a power of two >= 64");
    uint64_t* buf;  // CAP entries, warp-private shared memory
    int count;      // warp-uniform
    int k;          // warp-uniform, k <= CAP/2
    float thr;      // warp-uniform

    __device__ __forceinline__ void init(uint64_t* smem, int k_) {
        buf = smem;
        k = k_;
        count = 0;
        thr = __int_as_float(0x7f800000);  // +inf
    }

    // Sort buf[0..CAP) ascending (unused slots are KEY_INF), keep the best k.
    __device__ __noinline__ void compact() {
        const int lane = threadIdx.x & 31;
        for (int i = count + lane; i < CAP; i += 32) buf[i] = KEY_INF;
        __syncwarp();
        for (int size = 2; size <= CAP; size <<= 1) {
            for (int stride = size >> 1; stride > 0; stride >>= 1) {
                for (int t = lane; t < CAP / 2; t += 32) {
                    int lo = 2 * t - (t & (stride - 1));
                    int hi = lo + stride;
                    bool up = ((lo & size) == 0);
                    uint64_t a = buf[lo], b = buf[hi];
                    if ((a > b) == up) {
                        buf[lo] = b;
                        buf[hi] = a;
                    }
                }
                __syncwarp();
            }
        }
        if (count > k) count = k;
        if (count == k) thr = key_value(buf[k - 1]);
        __syncwarp();
    }

    // All 32 lanes must call. `valid` lanes offer (v, id).
    __device__ __forceinline__ void offer(bool valid, float v, uint32_t id) {
        bool hit = valid && (v <= thr);
        unsigned m = __ballot_sync(FULL_MASK, hit);
        if (m == 0) return;
        if (count + 32 > CAP) {
            compact();
            hit = valid && (v <= thr);
            m = __ballot_sync(FULL_MASK, hit);
            if (m == 0) return;
        }
        const int lane = threadIdx.x & 31;
        if (hit) buf[count + __popc(m & ((1u << lane) - 1u))] = make_key(v, id);
        count += __popc(m);
        __syncwarp();
    }

    // Final: sorted best-k in buf[0..min(count,k)).
    __device__ __forceinline__ void finish() { compact(); }
};

#endif  // __CUDACC__

}  // namespace b200vq
