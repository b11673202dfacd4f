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

// clip(rint(a / delta), 0, levels) with the reference's IEEE division and half-to-even rounding (the exact form).
__device__ __forceinline__ unsigned lvq_quant_exact(float a, float delta, int levels) {
    const int q = __float2int_rn(__fdiv_rn(a, delta));
    return (unsigned)(q < 0 ? 0 : (q > levels ? levels : q));
}
// The same eight indices without the division in the common case. t = a * (1 / delta) is within 1.8e-7 t of the true
// quotient, and so is the rounded quotient, so both round to the same integer unless a half-integer lies within
// 6e-7 t <= 1.6e-4 of t (levels <= 255); t + 1.5 * 2^23 rounds t to the nearest even integer and leaves it in the low
// mantissa bits, which IMAD shifts into the packed word (the exponent bits are a constant, subtracted once). 0 <= a <=
// span gives 0 <= rint <= levels, so nothing is clipped here. A thread whose group has a value within 2e-4 of a
// half-integer, or a NaN, redoes the group with the exact form (about 0.3 % of the groups).
// hi = indices 0..3, lo = indices 4..7, each MSB first in 4 B bits.
__device__ __forceinline__ void lvq_quant8(const float (&v)[8], float lo_v, float delta, float inv, int B, int levels,
                                           uint32_t& hi, uint32_t& lo) {
    const uint32_t mul = 1u << B;
    uint32_t w[2] = {0u, 0u};
    bool risky = false;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const float t = __fmul_rn(__fsub_rn(v[k], lo_v), inv);
        const float tr = __fadd_rn(t, 12582912.f);
        const float d = __fsub_rn(t, __fsub_rn(tr, 12582912.f));
        risky |= !(fabsf(d) <= 0.4998f);
        w[k >> 2] = w[k >> 2] * mul + __float_as_uint(tr);
    }
    const uint32_t bias = 0x4B400000u * (1u + mul + mul * mul + mul * mul * mul);   // mod 2^32, like the sums
    hi = w[0] - bias;
    lo = w[1] - bias;
    if (risky) {
        hi = lo = 0u;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            hi = (hi << B) | lvq_quant_exact(__fsub_rn(v[k], lo_v), delta, levels);
            lo = (lo << B) | lvq_quant_exact(__fsub_rn(v[4 + k], lo_v), delta, levels);
        }
    }
}
// B bytes of one group, most significant first
__device__ __forceinline__ void lvq_store_group(uint8_t* o, uint32_t hi, uint32_t lo, int B, bool wide) {
    if (wide && B == 8) {
        *reinterpret_cast<uint2*>(o) = make_uint2(__byte_perm(hi, 0, 0x0123), __byte_perm(lo, 0, 0x0123));
    } else if (wide && B == 4) {
        *reinterpret_cast<uint32_t*>(o) = __byte_perm((hi << 16) | lo, 0, 0x0123);
    } else if (wide && B == 2) {
        *reinterpret_cast<uint16_t*>(o) = (uint16_t)__byte_perm((hi << 8) | lo, 0, 0x3201);
    } else {
        const unsigned long long acc = ((unsigned long long)hi << (4 * B)) | lo;
        for (int b = 0; b < B; ++b) o[b] = (uint8_t)(acc >> (8 * (B - 1 - b)));
    }
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
    // a lane's B bytes go out in one store when the row pitch keeps them aligned (B = 8, 4, 2)
    const bool wide = (B == 8 || B == 4 || B == 2) && code_size % B == 0 &&
                      (reinterpret_cast<uintptr_t>(codes) % B) == 0;
    for (int64_t row = wid; row < n; row += nw) {
        const float* xr = x + row * ldx;
        uint8_t* code = codes + row * (int64_t)code_size;
        float lo = __int_as_float(0x7f800000), hi = __int_as_float(0xff800000);
#pragma unroll 4
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
        const float inv = __fdiv_rn(1.0f, delta);
        // second read of the row: L1 / L2
#pragma unroll 2
        for (int j = 8 * lane; j < D; j += 256) {
            float r[8];
            const int cnt = lvq_load8(xr, mu, j, D, vec, r);
            uint8_t* o = code + (j >> 3) * B;
            if (cnt == 8) {
                uint32_t h32, l32;
                lvq_quant8(r, lo, delta, inv, B, levels, h32, l32);
                lvq_store_group(o, h32, l32, B, wide);
            } else {
                // ragged last group: ceil(cnt B / 8) bytes
                unsigned long long acc = 0ull;
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    acc = (acc << B) | (unsigned long long)(k < cnt ? lvq_quant_exact(__fsub_rn(r[k], lo), delta, levels) : 0u);
                const int nb = (cnt * B + 7) >> 3;
                for (int b = 0; b < nb; ++b) o[b] = (uint8_t)(acc >> (8 * (B - 1 - b)));
            }
        }
        if (lane == 0) {
            lvq_store_f32(code + index_bytes, lo);
            lvq_store_f32(code + index_bytes + 4, delta);
        }
    }
}

// Four indices (MSB first, 4 B bits) by the scheme of lvq_quant8.
__device__ __forceinline__ uint32_t lvq_quant4(const float4& v, float lo_v, float delta, float inv, int B, int levels) {
    const uint32_t mul = 1u << B;
    const float e[4] = {v.x, v.y, v.z, v.w};
    uint32_t w = 0u;
    bool risky = false;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float t = __fmul_rn(__fsub_rn(e[k], lo_v), inv);
        const float tr = __fadd_rn(t, 12582912.f);
        const float d = __fsub_rn(t, __fsub_rn(tr, 12582912.f));
        risky |= !(fabsf(d) <= 0.4998f);
        w = w * mul + __float_as_uint(tr);
    }
    w -= 0x4B400000u * (1u + mul + mul * mul + mul * mul * mul);
    if (risky) {
        w = 0u;
#pragma unroll
        for (int k = 0; k < 4; ++k) w = (w << B) | lvq_quant_exact(__fsub_rn(e[k], lo_v), delta, levels);
    }
    return w;
}

// Aligned rows, D % 4 == 0, D <= 128 NK, even B: the row lives in registers. Lane l owns the 4-dimension groups
// 32 k + l (every 128-bit load instruction of the warp is one contiguous 512-byte run; a lane owning 8 consecutive
// dimensions would touch every 32-byte sector twice), so x is read from HBM once with the whole row in flight and the
// second pass costs no loads. A group is 4 B bits = B / 2 whole bytes, stored by its lane (contiguous over the warp).
// Same arithmetic as lvq_encode_kernel.
template <int NK>
__global__ void __launch_bounds__(128, NK <= 8 ? 6 : (NK <= 16 ? 4 : (NK <= 24 ? 3 : 2))) lvq_encode_reg_kernel(const float* __restrict__ x, int64_t n, int64_t ldx, int D,
                                                             const float* __restrict__ mu, int B,
                                                             uint8_t* __restrict__ codes) {
    const int lane = threadIdx.x & 31;
    const int64_t wid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int index_bytes = (D * B) >> 3, code_size = index_bytes + 8;
    const int levels = (1 << B) - 1;
    const int ngroups = D >> 2, gb = B >> 1;                      // bytes per group
    const bool wide = (code_size % gb) == 0 && (reinterpret_cast<uintptr_t>(codes) % gb) == 0;
    extern __shared__ __align__(16) float4 lvq_mu_s[];            // the mean, staged once per CTA
    for (int i = threadIdx.x; i < ngroups; i += blockDim.x) lvq_mu_s[i] = __ldg(reinterpret_cast<const float4*>(mu) + i);
    __syncthreads();
    for (int64_t row = wid; row < n; row += nw) {
        const float4* xr = reinterpret_cast<const float4*>(x + row * ldx) + lane;
        const float4* mr = lvq_mu_s + lane;
        uint8_t* code = codes + row * (int64_t)code_size;
        float4 r[NK];
#pragma unroll
        for (int k = 0; k < NK; ++k)
            if (32 * k + lane < ngroups) r[k] = ld_stream_f4(xr + 32 * k);
        asm volatile("" ::: "memory");     // every load of the row is issued before the first use of one
        float lo = __int_as_float(0x7f800000), hi = __int_as_float(0xff800000);
#pragma unroll
        for (int k = 0; k < NK; ++k)
            if (32 * k + lane < ngroups) {
                const float4 m = mr[32 * k];
                r[k] = make_float4(__fsub_rn(r[k].x, m.x), __fsub_rn(r[k].y, m.y), __fsub_rn(r[k].z, m.z),
                                   __fsub_rn(r[k].w, m.w));
                lo = fminf(fminf(fminf(lo, r[k].x), fminf(r[k].y, r[k].z)), r[k].w);
                hi = fmaxf(fmaxf(fmaxf(hi, r[k].x), fmaxf(r[k].y, r[k].z)), r[k].w);
            }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo = fminf(lo, __shfl_xor_sync(FULL_MASK, lo, o));
            hi = fmaxf(hi, __shfl_xor_sync(FULL_MASK, hi, o));
        }
        const float span = __fsub_rn(hi, lo);
        const float delta = span == 0.0f ? 1.17549435e-38f : __fdiv_rn(span, (float)levels);
        const float inv = __fdiv_rn(1.0f, delta);
#pragma unroll
        for (int k = 0; k < NK; ++k)
            if (32 * k + lane < ngroups) {
                const uint32_t w = lvq_quant4(r[k], lo, delta, inv, B, levels);
                uint8_t* o = code + (32 * k + lane) * gb;
                if (wide && B == 8) *reinterpret_cast<uint32_t*>(o) = __byte_perm(w, 0, 0x0123);
                else if (wide && B == 4) *reinterpret_cast<uint16_t*>(o) = (uint16_t)__byte_perm(w, 0, 0x3201);
                else
                    for (int b = 0; b < gb; ++b) o[b] = (uint8_t)(w >> (8 * (gb - 1 - b)));
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
    const bool wide = (B == 8 || B == 4 || B == 2) && code_size % B == 0 &&
                      (reinterpret_cast<uintptr_t>(codes) % B) == 0;
    for (int64_t row = wid; row < n; row += nw) {
        const uint8_t* code = codes + row * (int64_t)code_size;
        const float lo = lvq_load_f32(code + index_bytes), delta = lvq_load_f32(code + index_bytes + 4);
        float* o = out + row * ldo;
        for (int j = 8 * lane; j < D; j += 256) {
            const int cnt = D - j < 8 ? D - j : 8;
            const int nb = (cnt * B + 7) >> 3;
            const uint8_t* c = code + (j >> 3) * B;
            unsigned long long acc = 0ull;
            if (wide && cnt == 8) {
                if (B == 8) {
                    const uint2 w = __ldg(reinterpret_cast<const uint2*>(c));
                    acc = ((unsigned long long)__byte_perm(w.x, 0, 0x0123) << 32) | __byte_perm(w.y, 0, 0x0123);
                } else if (B == 4) {
                    acc = __byte_perm(__ldg(reinterpret_cast<const uint32_t*>(c)), 0, 0x0123);
                } else {
                    acc = __byte_perm((uint32_t)__ldg(reinterpret_cast<const uint16_t*>(c)), 0, 0x3201) & 0xFFFFu;
                }
            } else {
                for (int b = 0; b < B; ++b) acc = (acc << 8) | (unsigned long long)(b < nb ? __ldg(c + b) : 0);
            }
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

// Two passes over the row with the same float4-granular ownership and few registers (the register-resident form
// runs 12 warps per SM at D = 3072): the second read of the row comes out of L1 / L2.
__global__ void __launch_bounds__(256) lvq_encode2p_kernel(const float* __restrict__ x, int64_t n, int64_t ldx, int D,
                                                           const float* __restrict__ mu, int B,
                                                           uint8_t* __restrict__ codes) {
    const int lane = threadIdx.x & 31;
    const int64_t wid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int index_bytes = (D * B) >> 3, code_size = index_bytes + 8;
    const int levels = (1 << B) - 1;
    const int ngroups = D >> 2, gb = B >> 1;
    const bool wide = (code_size % gb) == 0 && (reinterpret_cast<uintptr_t>(codes) % gb) == 0;
    extern __shared__ __align__(16) float4 lvq_mu_s[];
    for (int i = threadIdx.x; i < ngroups; i += blockDim.x) lvq_mu_s[i] = __ldg(reinterpret_cast<const float4*>(mu) + i);
    __syncthreads();
    for (int64_t row = wid; row < n; row += nw) {
        const float4* xr = reinterpret_cast<const float4*>(x + row * ldx);
        uint8_t* code = codes + row * (int64_t)code_size;
        float lo = __int_as_float(0x7f800000), hi = __int_as_float(0xff800000);
#pragma unroll 8
        for (int g = lane; g < ngroups; g += 32) {
            const float4 v = __ldg(xr + g), m = lvq_mu_s[g];
            const float a = __fsub_rn(v.x, m.x), b = __fsub_rn(v.y, m.y), c = __fsub_rn(v.z, m.z), d = __fsub_rn(v.w, m.w);
            lo = fminf(fminf(fminf(lo, a), fminf(b, c)), d);
            hi = fmaxf(fmaxf(fmaxf(hi, a), fmaxf(b, c)), d);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo = fminf(lo, __shfl_xor_sync(FULL_MASK, lo, o));
            hi = fmaxf(hi, __shfl_xor_sync(FULL_MASK, hi, o));
        }
        const float span = __fsub_rn(hi, lo);
        const float delta = span == 0.0f ? 1.17549435e-38f : __fdiv_rn(span, (float)levels);
        const float inv = __fdiv_rn(1.0f, delta);
#pragma unroll 4
        for (int g = lane; g < ngroups; g += 32) {
            const float4 v = __ldg(xr + g), m = lvq_mu_s[g];
            const float4 r = make_float4(__fsub_rn(v.x, m.x), __fsub_rn(v.y, m.y), __fsub_rn(v.z, m.z), __fsub_rn(v.w, m.w));
            const uint32_t w = lvq_quant4(r, lo, delta, inv, B, levels);
            uint8_t* o = code + g * gb;
            if (wide && B == 8) *reinterpret_cast<uint32_t*>(o) = __byte_perm(w, 0, 0x0123);
            else if (wide && B == 4) *reinterpret_cast<uint16_t*>(o) = (uint16_t)__byte_perm(w, 0, 0x3201);
            else
                for (int b = 0; b < gb; ++b) o[b] = (uint8_t)(w >> (8 * (gb - 1 - b)));
        }
        if (lane == 0) {
            lvq_store_f32(code + index_bytes, lo);
            lvq_store_f32(code + index_bytes + 4, delta);
        }
    }
}

// Decode with float4-granular lane ownership (B = 8, 4, 2; D % 4 == 0): a lane owns four consecutive dimensions =
// B/2 code bytes, so a warp's loads are 32 B (B) consecutive bytes and its stores 512 consecutive bytes (the 8-dimension
// form stored two float4 per lane at a 32-byte lane pitch: half of every 32-byte sector per store instruction). Two
// rows per warp step keep twice the loads in flight. Same arithmetic: fl(lo + fl(q delta)) + mu.
template <int B>
__global__ void __launch_bounds__(256) lvq_decode4_kernel(const uint8_t* __restrict__ codes, int64_t n, int D,
                                                          const float* __restrict__ mu, float* __restrict__ out,
                                                          int64_t ldo) {
    const int lane = threadIdx.x & 31;
    const int64_t wid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int index_bytes = (D * B) >> 3, code_size = index_bytes + 8;
    auto field4 = [&](const uint8_t* code, int j, unsigned (&q)[4]) {
        if (B == 8) {
            const uint32_t w = __ldg(reinterpret_cast<const uint32_t*>(code + j));
            q[0] = w & 0xFFu; q[1] = (w >> 8) & 0xFFu; q[2] = (w >> 16) & 0xFFu; q[3] = w >> 24;
        } else if (B == 4) {
            const uint32_t h = __ldg(reinterpret_cast<const uint16_t*>(code + (j >> 1)));
            q[0] = (h >> 4) & 15u; q[1] = h & 15u; q[2] = (h >> 12) & 15u; q[3] = (h >> 8) & 15u;
        } else {
            const uint32_t b = __ldg(code + (j >> 2));
            q[0] = (b >> 6) & 3u; q[1] = (b >> 4) & 3u; q[2] = (b >> 2) & 3u; q[3] = b & 3u;
        }
    };
    for (int64_t row = 2 * wid; row < n; row += 2 * nw) {
        const bool two = row + 1 < n;
        const uint8_t* c0 = codes + row * (int64_t)code_size;
        const uint8_t* c1 = two ? c0 + code_size : c0;
        const float lo0 = lvq_load_f32(c0 + index_bytes), de0 = lvq_load_f32(c0 + index_bytes + 4);
        const float lo1 = lvq_load_f32(c1 + index_bytes), de1 = lvq_load_f32(c1 + index_bytes + 4);
        float* o0 = out + row * ldo;
        float* o1 = o0 + ldo;
#pragma unroll 2
        for (int j = 4 * lane; j < D; j += 128) {
            unsigned q0[4], q1[4];
            field4(c0, j, q0);
            field4(c1, j, q1);
            const float4 m = __ldg(reinterpret_cast<const float4*>(mu + j));
            const float4 r0 = make_float4(__fadd_rn(__fadd_rn(lo0, __fmul_rn((float)q0[0], de0)), m.x),
                                          __fadd_rn(__fadd_rn(lo0, __fmul_rn((float)q0[1], de0)), m.y),
                                          __fadd_rn(__fadd_rn(lo0, __fmul_rn((float)q0[2], de0)), m.z),
                                          __fadd_rn(__fadd_rn(lo0, __fmul_rn((float)q0[3], de0)), m.w));
            *reinterpret_cast<float4*>(o0 + j) = r0;
            if (two) {
                const float4 r1 = make_float4(__fadd_rn(__fadd_rn(lo1, __fmul_rn((float)q1[0], de1)), m.x),
                                              __fadd_rn(__fadd_rn(lo1, __fmul_rn((float)q1[1], de1)), m.y),
                                              __fadd_rn(__fadd_rn(lo1, __fmul_rn((float)q1[2], de1)), m.z),
                                              __fadd_rn(__fadd_rn(lo1, __fmul_rn((float)q1[3], de1)), m.w));
                *reinterpret_cast<float4*>(o1 + j) = r1;
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
    const bool reg_ok = D % 4 == 0 && D <= 4096 && num_bits % 2 == 0 && ldx % 4 == 0 && ((uintptr_t)x & 15) == 0 &&
                        ((uintptr_t)mu & 15) == 0;
    if (reg_ok && D > 1024 && D <= 12288) {
        // long rows: two passes with few registers beat the register-resident row (12 warps per SM at D = 3072):
        // 1.06 -> 0.86 ms at 262 144 x 3072 x 8 bits. About 48 KB of rows per SM in flight (4 CTAs at D = 3072:
        // 3 -> 0.99 ms, 4 -> 0.86, 5 -> 0.86, 6 -> 1.11, 8 -> 0.91; the second pass wants its row still in L1 / L2)
        int64_t g = (n * 32 + 255) / 256;
        int per_sm = 49152 / (D * 4);
        per_sm = per_sm < 3 ? 3 : (per_sm > 8 ? 8 : per_sm);
        const int64_t cap = (int64_t)device_sm_count() * per_sm;
        if (g > cap) g = cap;
        lvq_encode2p_kernel<<<(int)g, 256, (size_t)D * sizeof(float), as_stream(stream)>>>(x, n, ldx, D, mu, num_bits, codes);
        VQ_LAUNCH_CHECK();
        return B200VQ_OK;
    }
    if (reg_ok) {
        int64_t g = (n * 32 + 127) / 128;
        const int64_t cap = (int64_t)device_sm_count() * 32;
        if (g > cap) g = cap;
        const int nk = (D + 127) / 128;
        auto k = nk <= 8 ? lvq_encode_reg_kernel<8> : nk <= 16 ? lvq_encode_reg_kernel<16>
                 : nk <= 24 ? lvq_encode_reg_kernel<24> : lvq_encode_reg_kernel<32>;
        k<<<(int)g, 128, (size_t)D * sizeof(float), as_stream(stream)>>>(x, n, ldx, D, mu, num_bits, codes);
        VQ_LAUNCH_CHECK();
        return B200VQ_OK;
    }
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
    const int code_size = ((D * num_bits + 7) >> 3) + 8;
    const bool fast = (num_bits == 8 || num_bits == 4 || num_bits == 2) && D % 4 == 0 && ldo % 4 == 0 &&
                      ((uintptr_t)out & 15) == 0 && ((uintptr_t)mu & 15) == 0 &&
                      (num_bits == 2 || (((uintptr_t)codes % (num_bits / 2)) == 0 && code_size % (num_bits / 2) == 0));
    if (fast) {
        const int g = lvq_grid((n + 1) / 2);
        if (num_bits == 8) lvq_decode4_kernel<8><<<g, 256, 0, as_stream(stream)>>>(codes, n, D, mu, out, ldo);
        else if (num_bits == 4) lvq_decode4_kernel<4><<<g, 256, 0, as_stream(stream)>>>(codes, n, D, mu, out, ldo);
        else lvq_decode4_kernel<2><<<g, 256, 0, as_stream(stream)>>>(codes, n, D, mu, out, ldo);
        VQ_LAUNCH_CHECK();
        return B200VQ_OK;
    }
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
