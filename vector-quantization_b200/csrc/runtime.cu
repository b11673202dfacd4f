// Library runtime: error string, device queries, synthetic-row generator and
// the squared-error reduction used by the on-device MSE metric
// (src/haag_vq/benchmarks/exact_search.py:104-122, metrics/distortion.py:4-7).
#include <stdarg.h>

#include "common.cuh"

namespace b200vq {

char* last_error_buf() {
    static thread_local char buf[512] = {0};
    return buf;
}

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(last_error_buf(), 512, fmt, ap);
    va_end(ap);
}

int device_sm_count() {
    static thread_local int cached_dev = -1;
    static thread_local int cached_sms = 148;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (dev != cached_dev) {
        int sms = 0;
        if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && sms > 0)
            cached_sms = sms;
        cached_dev = dev;
    }
    return cached_sms;
}

// ----------------------------------------------------------- Philox ---
__device__ __forceinline__ void philox_round(uint32_t& c0, uint32_t& c1, uint32_t& c2, uint32_t& c3,
                                             uint32_t k0, uint32_t k1) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
    uint32_t hi0 = __umulhi(M0, c0), lo0 = M0 * c0;
    uint32_t hi1 = __umulhi(M1, c2), lo1 = M1 * c2;
    uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0;
    c1 = n1;
    c2 = n2;
    c3 = n3;
}

__device__ __forceinline__ float4 philox_normal4(uint64_t seed, uint64_t row, uint32_t col4) {
    uint32_t c0 = (uint32_t)row, c1 = (uint32_t)(row >> 32), c2 = col4, c3 = 0x5eedu;
    uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        philox_round(c0, c1, c2, c3, k0, k1);
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    const float inv = 2.3283064365386963e-10f;  // 2^-32
    float u0 = ((float)c0 + 0.5f) * inv, u1 = ((float)c1 + 0.5f) * inv;
    float u2 = ((float)c2 + 0.5f) * inv, u3 = ((float)c3 + 0.5f) * inv;
    u0 = fminf(fmaxf(u0, 1e-10f), 1.f);
    u2 = fminf(fmaxf(u2, 1e-10f), 1.f);
    float r0 = sqrtf(-2.f * __logf(u0)), r1 = sqrtf(-2.f * __logf(u2));
    float s0, c0f, s1, c1f;
    __sincosf(6.283185307179586f * u1, &s0, &c0f);
    __sincosf(6.283185307179586f * u3, &s1, &c1f);
    return make_float4(r0 * c0f, r0 * s0, r1 * c1f, r1 * s1);
}

// One warp per row: generate, optionally L2-normalise, store.
__global__ void synth_rows_kernel(float* __restrict__ out, int64_t n, int64_t ldo, int D,
                                  int64_t row0, uint64_t seed, const float* __restrict__ col_scale,
                                  int normalize) {
    const int lane = threadIdx.x & 31;
    const int64_t wid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int D4 = (D + 3) >> 2;
    for (int64_t r = wid; r < n; r += nw) {
        float ss = 0.f;
        float* o = out + r * ldo;
        for (int c4 = lane; c4 < D4; c4 += 32) {
            float4 g = philox_normal4(seed, (uint64_t)(row0 + r), (uint32_t)c4);
            float v[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                int c = c4 * 4 + j;
                if (c < D) {
                    float t = col_scale ? v[j] * col_scale[c] : v[j];
                    ss = __fmaf_rn(t, t, ss);
                    o[c] = t;
                }
            }
        }
        if (normalize) {
            for (int of = 16; of > 0; of >>= 1) ss += __shfl_xor_sync(FULL_MASK, ss, of);
            float inv = rsqrtf(fmaxf(ss, 1e-30f));
            __syncwarp();
            for (int c = lane; c < D; c += 32) o[c] *= inv;
        }
    }
}

// One warp per row (no 64-bit index arithmetic per element): fp32 differences, fp64 squares and sums.
__global__ void __launch_bounds__(256) sq_err_sum_kernel(const float* __restrict__ a, int64_t lda,
                                                         const float* __restrict__ b, int64_t ldb, int64_t n, int D,
                                                         double* __restrict__ out) {
    const int64_t wid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    double acc = 0.0;
    for (int64_t row = wid; row < n; row += nw) {
        const float* ar = a + row * lda;
        const float* br = b + row * ldb;
#pragma unroll 4
        for (int c = threadIdx.x & 31; c < D; c += 32) {
            const float d = ar[c] - br[c];
            acc += (double)d * (double)d;
        }
    }
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(FULL_MASK, acc, o);
    __shared__ double part[32];
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) part[w] = acc;
    __syncthreads();
    if (w == 0) {
        double v = (lane < (blockDim.x >> 5)) ? part[lane] : 0.0;
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL_MASK, v, o);
        if (lane == 0) atomicAdd(out, v);
    }
}

}  // namespace b200vq

using namespace b200vq;

extern "C" const char* b200vq_last_error(void) { return last_error_buf(); }

extern "C" int b200vq_abi_version(void) { return 1; }

extern "C" int b200vq_device_cc(void) {
    int dev = 0, major = 0, minor = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) {
        set_error("cudaGetDevice failed (no CUDA device?)");
        return B200VQ_ERR_CUDA;
    }
    if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev) != cudaSuccess) {
        set_error("cudaDeviceGetAttribute failed");
        return B200VQ_ERR_CUDA;
    }
    return major * 10 + minor;
}

extern "C" int b200vq_synth_rows(float* out, int64_t n, int64_t ldo, int D, int64_t row0,
                                 uint64_t seed, const float* col_scale, int normalize,
                                 void* stream) {
    VQ_CHECK_ARG(out && n >= 0 && D > 0 && ldo >= D, "synth_rows: bad args");
    if (n == 0) return B200VQ_OK;
    int64_t g = (n * 32 + 255) / 256;
    int64_t cap = (int64_t)device_sm_count() * 16;
    if (g > cap) g = cap;
    synth_rows_kernel<<<(int)g, 256, 0, as_stream(stream)>>>(out, n, ldo, D, row0, seed, col_scale,
                                                             normalize);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" int b200vq_sq_err_sum(const float* a, int64_t lda, const float* b, int64_t ldb, int64_t n,
                                 int D, double* out, void* stream) {
    VQ_CHECK_ARG((a || n == 0) && (b || n == 0) && (out || n == 0) && n >= 0 && D > 0 && lda >= D && ldb >= D, "sq_err_sum: bad args");
    if (n == 0) return B200VQ_OK;
    int64_t g = (n * 32 + 255) / 256;
    int64_t cap = (int64_t)device_sm_count() * 8;
    if (g > cap) g = cap;
    sq_err_sum_kernel<<<(int)g, 256, 0, as_stream(stream)>>>(a, lda, b, ldb, n, D, out);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}
