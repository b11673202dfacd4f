// Locally-adaptive vector quantization (LVQ, single level) for sm_100a.
//
// Replaces the numpy implementation of src/haag_vq/methods/lvq_quantization.py:44-141
// (registry method "lvq", benchmarks/method_registry_saq.py:49-52) with the same bytes:
//   R = x - mu (fp32);  lo = min_d R, hi = max_d R;  delta = (hi - lo) / (2^B - 1), or the smallest
//   normal float when hi == lo;  idx_d = clip(rint((R_d - lo) / delta), 0, 2^B - 1)  (np.round = half to even)
//   row = [ idx bit-packed B bits each, MSB first | lo fp32 | delta fp32 ],  ceil(D B / 8) + 8 bytes
//   decode:  (lo + float(idx) * delta) + mu   (separate fp32 multiply and adds, as numpy evaluates it)
// All arithmetic uses the round-to-nearest intrinsics (no FMA contraction, IEEE division), so codes
// and reconstructions are bit-identical to the reference.
//
// One warp per row, HBM-bound: a lane owns 8 consecutive dimensions = exactly B output bytes, so no
// two lanes share a byte; the row is read twice (min/max, then quantise), the second time from L1/L2.
// Algorithmic bytes per vector: 4 D + ceil(D B / 8) + 8 (encode), the same for decode.
#include "common.cuh"

namespace b200vq {

__device__ __forceinline__ float lvq_load_f32(const uint8_t* p) {
    const uint32_t u = (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
    return __uint_as_float(u);
}
__device__ __forceinline__ void lvq_store_f32(uint8_t* p, float v) {
    const uint32_t u = __float_as_uint(v);
    p[0] = (uint8_t)u;
    p[1] = (uint8_t)(u >> 8);
    p[2] = (uint8_t)(u >> 16);
    p[3] = (uint8_t)(u >> 24);
}

// R for the 8 dimensions starting at j (dimensions >= D are reported through `cnt`)
__device__ __forceinline__ int lvq_load8(const float* __restrict__ xr, const float* __restrict__ mu, int j, int D,
                                         bool vec, float (&r)[8]) {
    const int cnt = D - j < 8 ? D - j : 8;
    if (vec && cnt == 8) {
        const float4 a = __ldg(reinterpret_cast<const float4*>(xr + j)), b = __ldg(reinterpret_cast<const float4*>(xr + j + 4));
        const float4 c = __ldg(reinterpret_cast<const float4*>(mu + j)), d = __ldg(reinterpret_cast<const float4*>(mu + j + 4));
        r[0] = __fsub_rn(a.x, c.x); r[1] = __fsub_rn(a.y, c.y); r[2] = __fsub_rn(a.z, c.z); r[3] = __fsub_rn(a.w, c.w);
        r[4] = __fsub_rn(b.x, d.x); r[5] = __fsub_rn(b.y, d.y); r[6] = __fsub_rn(b.z, d.z); r[7] = __fsub_rn(b.w, d.w);
    } else {
#pragma unroll
        for (int k = 0; k < 8; ++k) r[k] = k < cnt ? __fsub_rn(xr[j + k], mu[j + k]) : 0.f;
    }
    return cnt;
}

__global__ void __launch_bounds__(256) lvq_encode_kernel(const float* __restrict__ x, int64_t n, int64_t ldx, int D,
                                                         const float* __restrict__ mu, int B,
                                                         uint8_t* __restrict__ codes) {
    const int lane = threadIdx.x & 31;
    const int64_t wid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int index_bytes = (D * B + 7) >> 3, code_size = index_bytes + 8;
    const int levels = (1 << B) - 1;
    const bool vec = (ldx % 4 == 0) && ((reinterpret_cast<uintptr_t>(x) & 15) == 0) &&
                     ((reinterpret_cast<uintptr_t>(mu) & 15) == 0);
    for (int64_t row = wid; row < n; row += nw) {
        const float* xr = x + row * ldx;
        uint8_t* code = codes + row * (int64_t)code_size;
        float lo = __int_as_float(0x7f800000), hi = __int_as_float(0xff800000);
        for (int j = 8 * lane; j < D; j += 256) {
            float r[8];
            const int cnt = lvq_load8(xr, mu, j, D, vec, r);
#pragma unroll
            for (int k = 0; k < 8; ++k)
                if (k < cnt) {
                    lo = fminf(lo, r[k]);
                    hi = fmaxf(hi, r[k]);
                }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo = fminf(lo, __shfl_xor_sync(FULL_MASK, lo, o));
            hi = fmaxf(hi, __shfl_xor_sync(FULL_MASK, hi, o));
        }
        const float span = __fsub_rn(hi, lo);
        const float delta = span == 0.0f ? 1.17549435e-38f : __fdiv_rn(span, (float)levels);
        for (int j = 8 * lane; j < D; j += 256) {
            float r[8];
            const int cnt = lvq_load8(xr, mu, j, D, vec, r);
            unsigned long long acc = 0ull;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                int q = 0;
                if (k < cnt) {
                    q = __float2int_rn(__fdiv_rn(__fsub_rn(r[k], lo), delta));   // rint: half to even, as np.round
                    q = q < 0 ? 0 : (q > levels ? levels : q);
                }
                acc = (acc << B) | (unsigned long long)q;
            }
            // 8 B bits, most significant first; a ragged last group keeps ceil(cnt B / 8) bytes
            const int nb = (cnt * B + 7) >> 3;
            uint8_t* o = code + (j >> 3) * B;
            for (int b = 0; b < nb; ++b) o[b] = (uint8_t)(acc >> (8 * (B - 1 - b)));
        }
        if (lane == 0) {
            lvq_store_f32(code + index_bytes, lo);
            lvq_store_f32(code + index_bytes + 4, delta);
        }
    }
}

__global__ void __launch_bounds__(256) lvq_decode_kernel(const uint8_t* __restrict__ codes, int64_t n, int D,
                                                         const float* __restrict__ mu, int B,
                                                         float* __restrict__ out, int64_t ldo) {
    const int lane = threadIdx.x & 31;
    const int64_t wid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int index_bytes = (D * B + 7) >> 3, code_size = index_bytes + 8;
    const unsigned levels = (1u << B) - 1u;
    const bool vec = (ldo % 4 == 0) && ((reinterpret_cast<uintptr_t>(out) & 15) == 0) &&
                     ((reinterpret_cast<uintptr_t>(mu) & 15) == 0);
    for (int64_t row = wid; row < n; row += nw) {
        const uint8_t* code = codes + row * (int64_t)code_size;
        const float lo = lvq_load_f32(code + index_bytes), delta = lvq_load_f32(code + index_bytes + 4);
        float* o = out + row * ldo;
        for (int j = 8 * lane; j < D; j += 256) {
            const int cnt = D - j < 8 ? D - j : 8;
            const int nb = (cnt * B + 7) >> 3;
            const uint8_t* c = code + (j >> 3) * B;
            unsigned long long acc = 0ull;
            for (int b = 0; b < B; ++b) acc = (acc << 8) | (unsigned long long)(b < nb ? __ldg(c + b) : 0);
            float r[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const unsigned q = (unsigned)(acc >> (B * (7 - k))) & levels;
                r[k] = __fadd_rn(lo, __fmul_rn((float)q, delta));
            }
            if (vec && cnt == 8) {
                const float4 m0 = __ldg(reinterpret_cast<const float4*>(mu + j)), m1 = __ldg(reinterpret_cast<const float4*>(mu + j + 4));
                *reinterpret_cast<float4*>(o + j) = make_float4(__fadd_rn(r[0], m0.x), __fadd_rn(r[1], m0.y),
                                                                __fadd_rn(r[2], m0.z), __fadd_rn(r[3], m0.w));
                *reinterpret_cast<float4*>(o + j + 4) = make_float4(__fadd_rn(r[4], m1.x), __fadd_rn(r[5], m1.y),
                                                                    __fadd_rn(r[6], m1.z), __fadd_rn(r[7], m1.w));
            } else {
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    if (k < cnt) o[j + k] = __fadd_rn(r[k], mu[j + k]);
            }
        }
    }
}

// Column means in numpy's order: X.mean(axis=0) of a C-contiguous float32 matrix adds row after row
// into an fp32 accumulator per column, then divides by n in fp32. One thread per column; 16 rows of
// loads in flight before their (sequential) adds.
__global__ void __launch_bounds__(128) col_mean_seq_kernel(const float* __restrict__ x, int64_t n, int64_t ldx, int D,
                                                           float* __restrict__ mean) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= D) return;
    float acc = 0.f;
    int64_t i = 0;
    for (; i + 16 <= n; i += 16) {
        float v[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) v[k] = __ldg(x + (i + k) * ldx + j);
#pragma unroll
        for (int k = 0; k < 16; ++k) acc = __fadd_rn(acc, v[k]);
    }
    for (; i < n; ++i) acc = __fadd_rn(acc, __ldg(x + i * ldx + j));
    mean[j] = __fdiv_rn(acc, (float)n);
}

}  // namespace b200vq

using namespace b200vq;

static int lvq_grid(int64_t n) {
    int64_t g = (n * 32 + 255) / 256;
    const int64_t cap = (int64_t)device_sm_count() * 16;
    if (g > cap) g = cap;
    return (int)(g < 1 ? 1 : g);
}

extern "C" int b200vq_lvq_encode(const float* x, int64_t n, int64_t ldx, int D, const float* mu, int num_bits,
                                 uint8_t* codes, void* stream) {
    VQ_CHECK_ARG(n >= 0 && D > 0 && ldx >= D, "lvq_encode: bad shape");
    VQ_CHECK_ARG(num_bits >= 1 && num_bits <= 8, "lvq_encode: num_bits must be in [1, 8], got %d", num_bits);
    if (n == 0) return B200VQ_OK;
    VQ_CHECK_ARG(x && mu && codes, "lvq_encode: null pointer");
    lvq_encode_kernel<<<lvq_grid(n), 256, 0, as_stream(stream)>>>(x, n, ldx, D, mu, num_bits, codes);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" int b200vq_lvq_decode(const uint8_t* codes, int64_t n, int D, const float* mu, int num_bits,
                                 float* out, int64_t ldo, void* stream) {
    VQ_CHECK_ARG(n >= 0 && D > 0 && ldo >= D, "lvq_decode: bad shape");
    VQ_CHECK_ARG(num_bits >= 1 && num_bits <= 8, "lvq_decode: num_bits must be in [1, 8], got %d", num_bits);
    if (n == 0) return B200VQ_OK;
    VQ_CHECK_ARG(codes && mu && out, "lvq_decode: null pointer");
    lvq_decode_kernel<<<lvq_grid(n), 256, 0, as_stream(stream)>>>(codes, n, D, mu, num_bits, out, ldo);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" int b200vq_col_mean_seq(const float* x, int64_t n, int64_t ldx, int D, float* mean, void* stream) {
    VQ_CHECK_ARG(n > 0 && D > 0 && ldx >= D && x && mean, "col_mean_seq: bad args");
    col_mean_seq_kernel<<<(D + 127) / 128, 128, 0, as_stream(stream)>>>(x, n, ldx, D, mean);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}
