// Scalar-quantiser kernels: per-dimension min/max, uniform 4/8/16-bit
// encode/decode, and the segmented variable-bit (SAQ CLI contract) bit-string
// encode/decode.
//
// Reference: src/haag_vq/methods/scalar_quantization.py:37-90 (numpy, fp32).
//   scaled = (X - min) / (max - min + 1e-8)          each op rounded to fp32
//   q      = np.round(scaled * (L - 1)).astype(u8)   rint = half-to-even
//   4-bit  : (q[:,0::2] << 4) | q[:,1::2], odd D zero-padded
//   decode : q / (L - 1) * (max - min + 1e-8) + min
// The intrinsics below (__fsub_rn, __fdiv_rn, ...) forbid FMA contraction so
// the results are bit-identical to numpy's. Divergence (documented in
// DESIGN.md): out-of-range inputs saturate instead of wrapping.
#include "common.cuh"

namespace b200vq {

__device__ __forceinline__ void atomic_min_f32(float* addr, float v) {
    if (v >= 0.f) atomicMin(reinterpret_cast<int*>(addr), __float_as_int(v));
    else atomicMax(reinterpret_cast<unsigned int*>(addr), __float_as_uint(v));
}
__device__ __forceinline__ void atomic_max_f32(float* addr, float v) {
    if (v >= 0.f) atomicMax(reinterpret_cast<int*>(addr), __float_as_int(v));
    else atomicMin(reinterpret_cast<unsigned int*>(addr), __float_as_uint(v));
}

// block (32, 8): 32 consecutive columns x 8 row lanes; grid (ceil(D/32), row CTAs)
__global__ void sq_minmax_kernel(const float* __restrict__ x, int64_t n, int64_t ldx, int D,
                                 float* __restrict__ vmin, float* __restrict__ vmax) {
    __shared__ float smin[8][33], smax[8][33];
    const int col = blockIdx.x * 32 + threadIdx.x;
    float mn = __int_as_float(0x7f800000), mx = __int_as_float(0xff800000);
    if (col < D) {
        for (int64_t r = (int64_t)blockIdx.y * 8 + threadIdx.y; r < n; r += (int64_t)gridDim.y * 8) {
            float v = x[r * ldx + col];
            mn = fminf(mn, v);
            mx = fmaxf(mx, v);
        }
    }
    smin[threadIdx.y][threadIdx.x] = mn;
    smax[threadIdx.y][threadIdx.x] = mx;
    __syncthreads();
    if (threadIdx.y == 0 && col < D) {
        for (int j = 1; j < 8; ++j) {
            mn = fminf(mn, smin[j][threadIdx.x]);
            mx = fmaxf(mx, smax[j][threadIdx.x]);
        }
        if (mn <= mx) {  // skip when no row was seen
            atomic_min_f32(vmin + col, mn);
            atomic_max_f32(vmax + col, mx);
        }
    }
}

// Aligned rows: four columns per thread, rows strided over grid.y, four 128-bit loads in flight.
__global__ void __launch_bounds__(128)
sq_minmax_vec_kernel(const float* __restrict__ x, int64_t n, int64_t ldx, int D, float* __restrict__ vmin,
                     float* __restrict__ vmax) {
    const int c4 = blockIdx.x * blockDim.x + threadIdx.x;
    if (c4 * 4 >= D) return;
    const float inf = __int_as_float(0x7f800000);
    float4 mn = make_float4(inf, inf, inf, inf), mx = make_float4(-inf, -inf, -inf, -inf);
    auto acc = [&](const float4& v) {
        mn.x = fminf(mn.x, v.x); mn.y = fminf(mn.y, v.y); mn.z = fminf(mn.z, v.z); mn.w = fminf(mn.w, v.w);
        mx.x = fmaxf(mx.x, v.x); mx.y = fmaxf(mx.y, v.y); mx.z = fmaxf(mx.z, v.z); mx.w = fmaxf(mx.w, v.w);
    };
    const int64_t step = gridDim.y;
    int64_t row = blockIdx.y;
    for (; row + 3 * step < n; row += 4 * step) {
        float4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) v[u] = ld_stream_f4(reinterpret_cast<const float4*>(x + (row + u * step) * ldx) + c4);
#pragma unroll
        for (int u = 0; u < 4; ++u) acc(v[u]);
    }
    for (; row < n; row += step) acc(ld_stream_f4(reinterpret_cast<const float4*>(x + row * ldx) + c4));
    float* pm = vmin + c4 * 4;
    float* px = vmax + c4 * 4;
    if (mn.x <= mx.x) { atomic_min_f32(pm + 0, mn.x); atomic_max_f32(px + 0, mx.x); }
    if (mn.y <= mx.y) { atomic_min_f32(pm + 1, mn.y); atomic_max_f32(px + 1, mx.y); }
    if (mn.z <= mx.z) { atomic_min_f32(pm + 2, mn.z); atomic_max_f32(px + 2, mx.z); }
    if (mn.w <= mx.w) { atomic_min_f32(pm + 3, mn.w); atomic_max_f32(px + 3, mx.w); }
}

__device__ __forceinline__ float sq_range(float mn, float mx) {
    return __fadd_rn(__fsub_rn(mx, mn), 1e-8f);
}
// q in [0, lm1]
__device__ __forceinline__ uint32_t sq_quant(float x, float mn, float range, float lm1) {
    float scaled = __fdiv_rn(__fsub_rn(x, mn), range);
    float q = rintf(__fmul_rn(scaled, lm1));
    q = fminf(fmaxf(q, 0.f), lm1);  // NaN -> 0 via fmaxf
    return (uint32_t)q;
}
__device__ __forceinline__ float sq_dequant(uint32_t q, float mn, float range, float lm1) {
    float scaled = __fdiv_rn((float)q, lm1);
    return __fadd_rn(__fmul_rn(scaled, range), mn);
}

// Generic scalar kernels (any D / alignment). One thread per output element.
template <int BITS>
__global__ void sq_encode_kernel(const float* __restrict__ x, int64_t n, int64_t ldx, int D,
                                 const float* __restrict__ vmin, const float* __restrict__ vmax,
                                 void* __restrict__ codes) {
    const float lm1 = (float)((1 << BITS) - 1);
    if (BITS == 4) {
        const int W = (D + 1) >> 1;
        const int64_t total = n * (int64_t)W;
        uint8_t* out = reinterpret_cast<uint8_t*>(codes);
        for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total;
             e += (int64_t)gridDim.x * blockDim.x) {
            int64_t row = e / W;
            int j = (int)(e - row * W);
            int d0 = 2 * j, d1 = 2 * j + 1;
            uint32_t hi = sq_quant(x[row * ldx + d0], vmin[d0], sq_range(vmin[d0], vmax[d0]), lm1);
            uint32_t lo = 0;
            if (d1 < D) lo = sq_quant(x[row * ldx + d1], vmin[d1], sq_range(vmin[d1], vmax[d1]), lm1);
            out[e] = (uint8_t)((hi << 4) | lo);
        }
    } else {
        const int64_t total = n * (int64_t)D;
        for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total;
             e += (int64_t)gridDim.x * blockDim.x) {
            int64_t row = e / D;
            int d = (int)(e - row * D);
            uint32_t q = sq_quant(x[row * ldx + d], vmin[d], sq_range(vmin[d], vmax[d]), lm1);
            if (BITS == 8) reinterpret_cast<uint8_t*>(codes)[e] = (uint8_t)q;
            else reinterpret_cast<uint16_t*>(codes)[e] = (uint16_t)q;
        }
    }
}

template <int BITS>
__global__ void sq_decode_kernel(const void* __restrict__ codes, int64_t n, int D,
                                 const float* __restrict__ vmin, const float* __restrict__ vmax,
                                 float* __restrict__ out, int64_t ldo) {
    const float lm1 = (float)((1 << BITS) - 1);
    const int64_t total = n * (int64_t)D;
    const int W = (D + 1) >> 1;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total;
         e += (int64_t)gridDim.x * blockDim.x) {
        int64_t row = e / D;
        int d = (int)(e - row * D);
        uint32_t q;
        if (BITS == 4) {
            uint8_t b = reinterpret_cast<const uint8_t*>(codes)[row * W + (d >> 1)];
            q = (d & 1) ? (b & 0xf) : (b >> 4);
        } else if (BITS == 8) {
            q = reinterpret_cast<const uint8_t*>(codes)[e];
        } else {
            q = reinterpret_cast<const uint16_t*>(codes)[e];
        }
        out[row * ldo + d] = sq_dequant(q, vmin[d], sq_range(vmin[d], vmax[d]), lm1);
    }
}

// ------------------------------------------------------- vector kernels ---
// Division by a per-column constant, correctly rounded like __fdiv_rn but without the per-element
// reciprocal: with r = RN(1/b) precomputed, q0 = RN(a r), rem = a - q0 b (exact, one FMA),
// q1 = RN(q0 + rem r) is the IEEE quotient (Markstein's correction step) whenever every quantity
// is a normal number. Where it is not, the quantiser's result does not depend on the low bits:
// |q0| >= 2^100 saturates after the clamp either way (q0 is returned so that inf stays inf instead
// of turning into NaN), a quotient below 2^-100 rounds to level 0 either way, NaN stays NaN.
// Columns whose divisor is not a normal number in [2^-100, 2^100] take __fdiv_rn.
// tests/test_gpu_sq.py checks the codes against IEEE division on 2^26 random (a, b) pairs per width.
struct DivConst {
    float b, r;
    bool fast;
};
__device__ __forceinline__ DivConst make_div(float b) {
    DivConst c;
    c.b = b;
    c.r = __frcp_rn(b);
    c.fast = b >= 7.8886091e-31f && b <= 1.2676506e30f;   // [2^-100, 2^100]; false for NaN
    return c;
}
__device__ __forceinline__ float div_const(float a, const DivConst& c) {
    if (!c.fast) return __fdiv_rn(a, c.b);
    const float q0 = __fmul_rn(a, c.r);
    const float rem = __fmaf_rn(-q0, c.b, a);
    const float q1 = __fmaf_rn(rem, c.r, q0);
    return fabsf(q0) < 1.2676506e30f ? q1 : q0;
}
__device__ __forceinline__ uint32_t sq_quant_c(float x, float mn, const DivConst& c, float lm1) {
    float q = rintf(__fmul_rn(div_const(__fsub_rn(x, mn), c), lm1));
    q = fminf(fmaxf(q, 0.f), lm1);  // NaN -> 0 via fmaxf
    return (uint32_t)q;
}

// Four consecutive columns per thread (their constants live in registers), rows strided over
// grid.y and unrolled by four so that four 128-bit loads are in flight per thread.
// BITS = 4: 2 output bytes per row and thread, 8: 4 bytes, 16: 8 bytes.
template <int BITS>
__global__ void __launch_bounds__(128)
sq_encode_vec_kernel(const float* __restrict__ x, int64_t n, int64_t ldx, int D, const float* __restrict__ vmin,
                     const float* __restrict__ vmax, uint8_t* __restrict__ codes) {
    const int c4 = blockIdx.x * blockDim.x + threadIdx.x;
    if (c4 * 4 >= D) return;
    constexpr float lm1 = (float)((1 << BITS) - 1);
    const float4 mn = *reinterpret_cast<const float4*>(vmin + c4 * 4);
    const float4 mx = *reinterpret_cast<const float4*>(vmax + c4 * 4);
    const DivConst d0 = make_div(sq_range(mn.x, mx.x)), d1 = make_div(sq_range(mn.y, mx.y)),
                   d2 = make_div(sq_range(mn.z, mx.z)), d3 = make_div(sq_range(mn.w, mx.w));
    const int64_t row_bytes = BITS == 4 ? D / 2 : (BITS == 8 ? D : 2 * (int64_t)D);
    auto put = [&](int64_t row, const float4& v) {
        const uint32_t q0 = sq_quant_c(v.x, mn.x, d0, lm1), q1 = sq_quant_c(v.y, mn.y, d1, lm1),
                       q2 = sq_quant_c(v.z, mn.z, d2, lm1), q3 = sq_quant_c(v.w, mn.w, d3, lm1);
        uint8_t* o = codes + row * row_bytes;
        if (BITS == 4) {
            reinterpret_cast<uint16_t*>(o)[c4] = (uint16_t)(((q0 << 4) | q1) | (((q2 << 4) | q3) << 8));
        } else if (BITS == 8) {
            reinterpret_cast<uint32_t*>(o)[c4] = q0 | (q1 << 8) | (q2 << 16) | (q3 << 24);
        } else {
            reinterpret_cast<uint2*>(o)[c4] = make_uint2(q0 | (q1 << 16), q2 | (q3 << 16));
        }
    };
    const int64_t step = gridDim.y;
    int64_t row = blockIdx.y;
    for (; row + 3 * step < n; row += 4 * step) {
        float4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) v[u] = ld_stream_f4(reinterpret_cast<const float4*>(x + (row + u * step) * ldx) + c4);
#pragma unroll
        for (int u = 0; u < 4; ++u) put(row + u * step, v[u]);
    }
    for (; row < n; row += step) put(row, ld_stream_f4(reinterpret_cast<const float4*>(x + row * ldx) + c4));
}

template <int BITS>
__global__ void __launch_bounds__(128)
sq_decode_vec_kernel(const uint8_t* __restrict__ codes, int64_t n, int D, const float* __restrict__ vmin,
                     const float* __restrict__ vmax, float* __restrict__ out, int64_t ldo) {
    const int c4 = blockIdx.x * blockDim.x + threadIdx.x;
    if (c4 * 4 >= D) return;
    constexpr float lm1 = (float)((1 << BITS) - 1);
    const float4 mn = *reinterpret_cast<const float4*>(vmin + c4 * 4);
    const float4 mx = *reinterpret_cast<const float4*>(vmax + c4 * 4);
    const float r0 = sq_range(mn.x, mx.x), r1 = sq_range(mn.y, mx.y), r2 = sq_range(mn.z, mx.z),
                r3 = sq_range(mn.w, mx.w);
    const DivConst dl = make_div(lm1);   // q / (L - 1): exact-division constant, every q checked by the tests
    const int64_t row_bytes = BITS == 4 ? D / 2 : (BITS == 8 ? D : 2 * (int64_t)D);
    auto get = [&](int64_t row, uint32_t (&q)[4]) {
        const uint8_t* o = codes + row * row_bytes;
        if (BITS == 4) {
            const uint32_t w = __ldg(reinterpret_cast<const uint16_t*>(o) + c4);
            q[0] = (w >> 4) & 0xf; q[1] = w & 0xf; q[2] = (w >> 12) & 0xf; q[3] = (w >> 8) & 0xf;
        } else if (BITS == 8) {
            const uint32_t w = __ldg(reinterpret_cast<const uint32_t*>(o) + c4);
            q[0] = w & 0xff; q[1] = (w >> 8) & 0xff; q[2] = (w >> 16) & 0xff; q[3] = w >> 24;
        } else {
            const uint2 w = __ldg(reinterpret_cast<const uint2*>(o) + c4);
            q[0] = w.x & 0xffff; q[1] = w.x >> 16; q[2] = w.y & 0xffff; q[3] = w.y >> 16;
        }
    };
    auto deq = [&](uint32_t q, float m, float range) {
        return __fadd_rn(__fmul_rn(div_const((float)q, dl), range), m);
    };
    const int64_t step = gridDim.y;
    int64_t row = blockIdx.y;
    for (; row + 3 * step < n; row += 4 * step) {
        uint32_t q[4][4];
#pragma unroll
        for (int u = 0; u < 4; ++u) get(row + u * step, q[u]);
#pragma unroll
        for (int u = 0; u < 4; ++u)
            reinterpret_cast<float4*>(out + (row + u * step) * ldo)[c4] =
                make_float4(deq(q[u][0], mn.x, r0), deq(q[u][1], mn.y, r1), deq(q[u][2], mn.z, r2), deq(q[u][3], mn.w, r3));
    }
    for (; row < n; row += step) {
        uint32_t q[4];
        get(row, q);
        reinterpret_cast<float4*>(out + row * ldo)[c4] =
            make_float4(deq(q[0], mn.x, r0), deq(q[1], mn.y, r1), deq(q[2], mn.z, r2), deq(q[3], mn.w, r3));
    }
}

// ----------------------------------------------- segmented variable-bit ---
// Codes are an MSB-first bit string: dimension d occupies bits
// [bit_offset[d], bit_offset[d]+dim_bits[d]) counted from the MSB of byte 0.
// Gather formulation: one thread builds one output byte from the (at most 8)
// dimensions overlapping it, so there are no atomics and no write conflicts.
__device__ __forceinline__ int first_dim_at_bit(const int* __restrict__ bit_offset, int D, int bit) {
    // largest d with bit_offset[d] <= bit  (bit_offset is non-decreasing, [D] = total)
    int lo = 0, hi = D;
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (bit_offset[mid] <= bit) lo = mid;
        else hi = mid;
    }
    return lo;
}

__global__ void saq_encode_kernel(const float* __restrict__ x, int64_t n, int64_t ldx, int D,
                                  const float* __restrict__ vmin, const float* __restrict__ vmax,
                                  const int* __restrict__ dim_bits,
                                  const int* __restrict__ bit_offset, int code_bytes,
                                  uint8_t* __restrict__ codes) {
    const int total_bits = bit_offset[D];
    const int64_t total = n * (int64_t)code_bytes;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total;
         e += (int64_t)gridDim.x * blockDim.x) {
        int64_t row = e / code_bytes;
        int byte = (int)(e - row * code_bytes);
        int b0 = byte * 8, b1 = b0 + 8;
        uint32_t acc = 0;
        if (b0 < total_bits) {
            int d = first_dim_at_bit(bit_offset, D, b0);
            for (; d < D && bit_offset[d] < b1; ++d) {
                int nb = dim_bits[d];
                if (nb == 0) continue;
                int off = bit_offset[d];
                if (off + nb <= b0) continue;
                float lm1 = (float)((1 << nb) - 1);
                uint32_t q = sq_quant(x[row * ldx + d], vmin[d], sq_range(vmin[d], vmax[d]), lm1);
                // place q so that its MSB sits at absolute bit `off` (MSB-first)
                int shift = b1 - (off + nb);  // >0: left shift inside the byte; <0: spill right
                uint32_t piece = (shift >= 0) ? (q << shift) : (q >> (-shift));
                acc |= piece & 0xffu;
            }
        }
        codes[e] = (uint8_t)acc;
    }
}

// Aligned input rows: four consecutive dimensions per thread with their constants in registers, rows strided over
// grid.y, four 128-bit loads in flight (the structure of sq_encode_vec_kernel). A thread's four fields are one
// run of W <= 32 bits at the row-independent bit offset o. The thread OWNS every output byte whose first bit lies
// in [o, o + W): bytes that lie wholly inside the run come straight from the run (for the usual plans - widths
// from {0, 2, 4, 6, 8}, segment lengths a multiple of 4 - every run is whole bytes and nothing else happens);
// an owned byte that extends past the run takes the missing bits from the following dimensions, quantised here
// with the same arithmetic (any widths, any segment boundaries; rare, and only that one byte per thread).
__global__ void __launch_bounds__(128)
saq_encode_vec_kernel(const float* __restrict__ x, int64_t n, int64_t ldx, int D, const float* __restrict__ vmin,
                      const float* __restrict__ vmax, const int* __restrict__ dim_bits,
                      const int* __restrict__ bit_offset, int code_bytes, uint8_t* __restrict__ codes, int word_ok) {
    const int c4 = blockIdx.x * blockDim.x + threadIdx.x;
    if (c4 * 4 >= D) return;
    const float4 mn = *reinterpret_cast<const float4*>(vmin + c4 * 4);
    const float4 mx = *reinterpret_cast<const float4*>(vmax + c4 * 4);
    const int4 nb = *reinterpret_cast<const int4*>(dim_bits + c4 * 4);
    const DivConst d0 = make_div(sq_range(mn.x, mx.x)), d1 = make_div(sq_range(mn.y, mx.y)),
                   d2 = make_div(sq_range(mn.z, mx.z)), d3 = make_div(sq_range(mn.w, mx.w));
    const float l0 = (float)((1 << nb.x) - 1), l1 = (float)((1 << nb.y) - 1), l2 = (float)((1 << nb.z) - 1),
                l3 = (float)((1 << nb.w) - 1);
    const int o = bit_offset[c4 * 4], W = nb.x + nb.y + nb.z + nb.w, end = o + W;
    const int total_bits = bit_offset[D];
    const int fo = (o + 7) >> 3;                              // first owned byte
    const int nfull = W > 0 && (end >> 3) > fo ? (end >> 3) - fo : 0;     // owned bytes wholly inside the run
    const bool tail = W > 0 && (end & 7) != 0 && (end >> 3) >= fo;        // owned byte end/8 extends past the run
    const bool pad = total_bits == 0 && c4 == 0;              // empty plan: the single pad byte
    const bool word = word_ok && nfull == 4 && !tail && (fo & 3) == 0;
    const bool half_ok = word_ok != 0;                        // rows start on a 4-byte boundary   // 8-bit runs on a word boundary: one 32-bit store
    auto put = [&](int64_t row, const float4& v) {
        uint32_t V = 0;
        if (nb.x) V = sq_quant_c(v.x, mn.x, d0, l0);
        if (nb.y) V = (V << nb.y) | sq_quant_c(v.y, mn.y, d1, l1);
        if (nb.z) V = (V << nb.z) | sq_quant_c(v.z, mn.z, d2, l2);
        if (nb.w) V = (V << nb.w) | sq_quant_c(v.w, mn.w, d3, l3);
        uint8_t* out = codes + row * (int64_t)code_bytes;
        if (word) {
            *reinterpret_cast<uint32_t*>(out + fo) = __byte_perm(V, 0, 0x0123);
            return;
        }
        if (nfull) {
            // the run's whole bytes, most significant first, as the widest naturally aligned stores
            const uint32_t B = V >> (end - 8 * (fo + nfull));          // nfull bytes, big-endian in the low bits
            if (half_ok && nfull == 2 && !(fo & 1)) {
                *reinterpret_cast<uint16_t*>(out + fo) = (uint16_t)__byte_perm(B, 0, 0x4401);
            } else if (half_ok && nfull == 3) {
                if (fo & 1) {
                    out[fo] = (uint8_t)(B >> 16);
                    *reinterpret_cast<uint16_t*>(out + fo + 1) = (uint16_t)__byte_perm(B, 0, 0x4401);
                } else {
                    *reinterpret_cast<uint16_t*>(out + fo) = (uint16_t)__byte_perm(B, 0, 0x4412);
                    out[fo + 2] = (uint8_t)B;
                }
            } else {
#pragma unroll 4
                for (int j = 0; j < nfull; ++j) out[fo + j] = (uint8_t)(B >> (8 * (nfull - 1 - j)));
            }
        }
        if (tail) {
            const int b = end >> 3, b1 = 8 * b + 8;
            uint32_t acc = (V << (b1 - end)) & 0xffu;
            for (int d = c4 * 4 + 4; d < D && bit_offset[d] < b1; ++d) {
                const int w = dim_bits[d];
                if (w == 0) continue;
                const uint32_t q = sq_quant(x[row * ldx + d], vmin[d], sq_range(vmin[d], vmax[d]), (float)((1 << w) - 1));
                const int shift = b1 - (bit_offset[d] + w);
                acc |= ((shift >= 0) ? (q << shift) : (q >> (-shift))) & 0xffu;
            }
            out[b] = (uint8_t)acc;
        }
        if (pad) out[0] = 0;
    };
    const int64_t step = gridDim.y;
    int64_t row = blockIdx.y;
    // The usual plans (one even width w for the thread's four dimensions, run on a byte boundary): the run is
    // exactly w / 2 bytes and the row loop is straight-line code; zero-width groups read nothing at all.
    const bool uniform = nb.x == nb.y && nb.y == nb.z && nb.z == nb.w && d0.fast && d1.fast && d2.fast && d3.fast &&
                         (o & 7) == 0 && !pad && half_ok;
    if (W == 0 && !pad) return;
    if (uniform && (nb.x == 8 || nb.x == 6 || nb.x == 4 || nb.x == 2) && (nb.x != 8 || (fo & 3) == 0) &&
        (nb.x != 4 || (fo & 1) == 0)) {
        const int w = nb.x;
        auto fast = [&](int64_t r, const float4& v) {
            const uint32_t V = (((((sq_quant_c(v.x, mn.x, d0, l0) << w) | sq_quant_c(v.y, mn.y, d1, l1)) << w) |
                                 sq_quant_c(v.z, mn.z, d2, l2)) << w) | sq_quant_c(v.w, mn.w, d3, l3);
            uint8_t* out = codes + r * (int64_t)code_bytes + fo;
            if (w == 8) {
                *reinterpret_cast<uint32_t*>(out) = __byte_perm(V, 0, 0x0123);
            } else if (w == 4) {
                *reinterpret_cast<uint16_t*>(out) = (uint16_t)__byte_perm(V, 0, 0x4401);
            } else if (w == 2) {
                out[0] = (uint8_t)V;
            } else if (fo & 1) {          // w == 6: three bytes
                out[0] = (uint8_t)(V >> 16);
                *reinterpret_cast<uint16_t*>(out + 1) = (uint16_t)__byte_perm(V, 0, 0x4401);
            } else {
                *reinterpret_cast<uint16_t*>(out) = (uint16_t)__byte_perm(V, 0, 0x4412);
                out[2] = (uint8_t)V;
            }
        };
        for (; row + 7 * step < n; row += 8 * step) {
            float4 v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u)
                v[u] = ld_stream_f4(reinterpret_cast<const float4*>(x + (row + u * step) * ldx) + c4);
#pragma unroll
            for (int u = 0; u < 8; ++u) fast(row + u * step, v[u]);
        }
        for (; row < n; row += step) fast(row, ld_stream_f4(reinterpret_cast<const float4*>(x + row * ldx) + c4));
        return;
    }
    for (; row + 3 * step < n; row += 4 * step) {
        float4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) v[u] = ld_stream_f4(reinterpret_cast<const float4*>(x + (row + u * step) * ldx) + c4);
#pragma unroll
        for (int u = 0; u < 4; ++u) put(row + u * step, v[u]);
    }
    for (; row < n; row += step) put(row, ld_stream_f4(reinterpret_cast<const float4*>(x + row * ldx) + c4));
}

__global__ void saq_decode_kernel(const uint8_t* __restrict__ codes, int64_t n, int D,
                                  const float* __restrict__ vmin, const float* __restrict__ vmax,
                                  const int* __restrict__ dim_bits,
                                  const int* __restrict__ bit_offset, int code_bytes,
                                  float* __restrict__ out, int64_t ldo) {
    const int64_t total = n * (int64_t)D;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total;
         e += (int64_t)gridDim.x * blockDim.x) {
        int64_t row = e / D;
        int d = (int)(e - row * D);
        int nb = dim_bits[d];
        float v;
        if (nb == 0) {
            v = vmin[d];  // 0-bit dimensions decode to the fill value the host put in vmin
        } else {
            int off = bit_offset[d];
            const uint8_t* c = codes + row * (int64_t)code_bytes;
            int byte = off >> 3;
            uint32_t w = ((uint32_t)c[byte] << 8) | (byte + 1 < code_bytes ? c[byte + 1] : 0u);
            int shift = 16 - ((off & 7) + nb);
            uint32_t q = (w >> shift) & ((1u << nb) - 1u);
            v = sq_dequant(q, vmin[d], sq_range(vmin[d], vmax[d]), (float)((1 << nb) - 1));
        }
        out[row * ldo + d] = v;
    }
}

// Aligned output rows: four consecutive dimensions per thread (their widths, bit offsets and affine
// constants live in registers), rows strided over grid.y, two rows in flight. The code bytes of a
// warp's 128 dimensions are adjacent, so the byte loads coalesce in L1.
__global__ void __launch_bounds__(128)
saq_decode_vec_kernel(const uint8_t* __restrict__ codes, int64_t n, int D, const float* __restrict__ vmin,
                      const float* __restrict__ vmax, const int* __restrict__ dim_bits,
                      const int* __restrict__ bit_offset, int code_bytes, float* __restrict__ out, int64_t ldo) {
    const int c4 = blockIdx.x * blockDim.x + threadIdx.x;
    if (c4 * 4 >= D) return;
    int nb[4], byte[4], shift[4];
    float mn[4], range[4];
    DivConst dl[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int d = c4 * 4 + j;
        nb[j] = dim_bits[d];
        const int off = bit_offset[d];
        byte[j] = off >> 3;
        shift[j] = 16 - ((off & 7) + nb[j]);
        mn[j] = vmin[d];
        range[j] = sq_range(mn[j], vmax[d]);
        dl[j] = make_div((float)((1 << nb[j]) - 1));
    }
    // Word path (code rows 4-byte aligned, every width <= 8): the four fields span at most 32 bits, so
    // two aligned 32-bit loads, byte-swapped into one big-endian 64-bit window, hold all of them.
    bool words = (code_bytes & 3) == 0 && (reinterpret_cast<uintptr_t>(codes) & 3) == 0;
    const int w0 = bit_offset[c4 * 4] >> 5;
    int wsh[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        words = words && nb[j] <= 8;
        wsh[j] = 64 - ((bit_offset[c4 * 4 + j] - 32 * w0) + nb[j]);
    }
    const bool first = w0 < (code_bytes >> 2), second = w0 + 1 < (code_bytes >> 2);   // trailing 0-bit dims sit at the end
    auto dequant = [&](int j, uint32_t q) {
        // 0-bit dimensions decode to the fill value the host put in vmin
        return nb[j] ? __fadd_rn(__fmul_rn(div_const((float)q, dl[j]), range[j]), mn[j]) : mn[j];
    };
    const int64_t step = gridDim.y;
    int64_t row = blockIdx.y;
    if (words) {
        auto fetch = [&](int64_t r) {
            const uint32_t* cw = reinterpret_cast<const uint32_t*>(codes + r * (int64_t)code_bytes) + w0;
            const uint32_t a = first ? __ldg(cw) : 0u, b2 = second ? __ldg(cw + 1) : 0u;
            return ((uint64_t)__byte_perm(a, 0, 0x0123) << 32) | __byte_perm(b2, 0, 0x0123);
        };
        auto emit = [&](int64_t r, uint64_t win) {
            float v[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) v[j] = dequant(j, (uint32_t)(win >> (wsh[j] & 63)) & ((1u << nb[j]) - 1u));
            reinterpret_cast<float4*>(out + r * ldo)[c4] = make_float4(v[0], v[1], v[2], v[3]);
        };
        for (; row + 7 * step < n; row += 8 * step) {
            uint64_t win[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) win[u] = fetch(row + u * step);
#pragma unroll
            for (int u = 0; u < 8; ++u) emit(row + u * step, win[u]);
        }
        for (; row < n; row += step) emit(row, fetch(row));
        return;
    }
    auto fetch = [&](int64_t r, uint32_t (&w)[4]) {
        const uint8_t* c = codes + r * (int64_t)code_bytes;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            w[j] = 0;
            if (nb[j]) w[j] = ((uint32_t)__ldg(c + byte[j]) << 8) | (byte[j] + 1 < code_bytes ? (uint32_t)__ldg(c + byte[j] + 1) : 0u);
        }
    };
    auto emit = [&](int64_t r, const uint32_t (&w)[4]) {
        float v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) v[j] = dequant(j, (w[j] >> shift[j]) & ((1u << nb[j]) - 1u));
        reinterpret_cast<float4*>(out + r * ldo)[c4] = make_float4(v[0], v[1], v[2], v[3]);
    };
    for (; row + step < n; row += 2 * step) {
        uint32_t wa[4], wb[4];
        fetch(row, wa);
        fetch(row + step, wb);
        emit(row, wa);
        emit(row + step, wb);
    }
    for (; row < n; row += step) {
        uint32_t wa[4];
        fetch(row, wa);
        emit(row, wa);
    }
}

static inline int sq_grid(int64_t work, int block) {
    int64_t g = (work + block - 1) / block;
    int64_t cap = (int64_t)device_sm_count() * 16;
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    return (int)g;
}

}  // namespace b200vq

using namespace b200vq;

extern "C" int b200vq_sq_minmax(const float* x, int64_t n, int64_t ldx, int D, float* vmin,
                                float* vmax, void* stream) {
    VQ_CHECK_ARG((x || n == 0) && vmin && vmax && n >= 0 && D > 0 && ldx >= D, "sq_minmax: bad args");
    if (n == 0) return B200VQ_OK;
    if ((D % 4 == 0) && (ldx % 4 == 0) && (((uintptr_t)x & 15) == 0) && n >= 64) {
        int gx = (D / 4 + 127) / 128;
        int64_t gy = ((int64_t)device_sm_count() * 16 + gx - 1) / gx;
        if (gy > n) gy = n;
        if (gy > 65535) gy = 65535;
        sq_minmax_vec_kernel<<<dim3(gx, (unsigned)gy), 128, 0, as_stream(stream)>>>(x, n, ldx, D, vmin, vmax);
        VQ_LAUNCH_CHECK();
        return B200VQ_OK;
    }
    int gx = (D + 31) / 32;
    int64_t gy = (n + 7) / 8;
    int64_t cap = ((int64_t)device_sm_count() * 8 + gx - 1) / gx;
    if (cap < 1) cap = 1;
    if (gy > cap) gy = cap;
    if (gy > 65535) gy = 65535;
    sq_minmax_kernel<<<dim3(gx, (unsigned)gy), dim3(32, 8), 0, as_stream(stream)>>>(x, n, ldx, D,
                                                                                    vmin, vmax);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" int b200vq_sq_encode(const float* x, int64_t n, int64_t ldx, int D, const float* vmin,
                                const float* vmax, int bits, void* codes, void* stream) {
    VQ_CHECK_ARG((x || n == 0) && vmin && vmax && (codes || n == 0) && n >= 0 && D > 0 && ldx >= D, "sq_encode: bad args");
    VQ_CHECK_ARG(bits == 4 || bits == 8 || bits == 16, "sq_encode: bits must be 4, 8 or 16, got %d",
                 bits);
    if (n == 0) return B200VQ_OK;
    cudaStream_t st = as_stream(stream);
    const bool vec = (D % 4 == 0) && (ldx % 4 == 0) && (((uintptr_t)x & 15) == 0) && (((uintptr_t)vmin & 15) == 0) &&
                     (((uintptr_t)vmax & 15) == 0) && (((uintptr_t)codes & 15) == 0);
    if (vec) {
        int gx = (D / 4 + 127) / 128;
        int64_t gy = ((int64_t)device_sm_count() * 16 + gx - 1) / gx;
        if (gy > n) gy = n;
        if (gy > 65535) gy = 65535;
        const dim3 grid(gx, (unsigned)gy);
        uint8_t* c8 = reinterpret_cast<uint8_t*>(codes);
        if (bits == 4) sq_encode_vec_kernel<4><<<grid, 128, 0, st>>>(x, n, ldx, D, vmin, vmax, c8);
        else if (bits == 8) sq_encode_vec_kernel<8><<<grid, 128, 0, st>>>(x, n, ldx, D, vmin, vmax, c8);
        else sq_encode_vec_kernel<16><<<grid, 128, 0, st>>>(x, n, ldx, D, vmin, vmax, c8);
    } else if (bits == 8) {
        sq_encode_kernel<8><<<sq_grid(n * (int64_t)D, 256), 256, 0, st>>>(x, n, ldx, D, vmin, vmax, codes);
    } else if (bits == 4) {
        sq_encode_kernel<4><<<sq_grid(n * (int64_t)((D + 1) / 2), 256), 256, 0, st>>>(
            x, n, ldx, D, vmin, vmax, codes);
    } else {
        sq_encode_kernel<16><<<sq_grid(n * (int64_t)D, 256), 256, 0, st>>>(x, n, ldx, D, vmin, vmax,
                                                                           codes);
    }
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" int b200vq_sq_decode(const void* codes, int64_t n, int D, const float* vmin,
                                const float* vmax, int bits, float* out, int64_t ldo,
                                void* stream) {
    VQ_CHECK_ARG((codes || n == 0) && vmin && vmax && (out || n == 0) && n >= 0 && D > 0 && ldo >= D, "sq_decode: bad args");
    VQ_CHECK_ARG(bits == 4 || bits == 8 || bits == 16, "sq_decode: bits must be 4, 8 or 16, got %d",
                 bits);
    if (n == 0) return B200VQ_OK;
    cudaStream_t st = as_stream(stream);
    int g = sq_grid(n * (int64_t)D, 256);
    const bool vec = (D % 4 == 0) && (ldo % 4 == 0) && (((uintptr_t)out & 15) == 0) && (((uintptr_t)vmin & 15) == 0) &&
                     (((uintptr_t)vmax & 15) == 0) && (((uintptr_t)codes & 15) == 0);
    if (vec) {
        int gx = (D / 4 + 127) / 128;
        int64_t gy = ((int64_t)device_sm_count() * 16 + gx - 1) / gx;
        if (gy > n) gy = n;
        if (gy > 65535) gy = 65535;
        const dim3 grid(gx, (unsigned)gy);
        const uint8_t* c8 = reinterpret_cast<const uint8_t*>(codes);
        if (bits == 4) sq_decode_vec_kernel<4><<<grid, 128, 0, st>>>(c8, n, D, vmin, vmax, out, ldo);
        else if (bits == 8) sq_decode_vec_kernel<8><<<grid, 128, 0, st>>>(c8, n, D, vmin, vmax, out, ldo);
        else sq_decode_vec_kernel<16><<<grid, 128, 0, st>>>(c8, n, D, vmin, vmax, out, ldo);
    } else if (bits == 8) {
        sq_decode_kernel<8><<<g, 256, 0, st>>>(codes, n, D, vmin, vmax, out, ldo);
    } else if (bits == 4) {
        sq_decode_kernel<4><<<g, 256, 0, st>>>(codes, n, D, vmin, vmax, out, ldo);
    } else {
        sq_decode_kernel<16><<<g, 256, 0, st>>>(codes, n, D, vmin, vmax, out, ldo);
    }
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" int b200vq_saq_encode(const float* x, int64_t n, int64_t ldx, int D, const float* vmin,
                                 const float* vmax, const int* dim_bits, const int* bit_offset,
                                 int code_bytes, uint8_t* codes, void* stream) {
    VQ_CHECK_ARG((x || n == 0) && vmin && vmax && dim_bits && bit_offset && (codes || n == 0) && n >= 0 && D > 0 &&
                     ldx >= D && code_bytes > 0,
                 "saq_encode: bad args");
    if (n == 0) return B200VQ_OK;
    if ((D % 4 == 0) && (ldx % 4 == 0) && (((uintptr_t)x | (uintptr_t)vmin | (uintptr_t)vmax | (uintptr_t)dim_bits) & 15) == 0 &&
        true) {
        const int word_ok = (code_bytes % 4 == 0) && (((uintptr_t)codes & 3) == 0);
        int gx = (D / 4 + 127) / 128;
        int64_t gy = ((int64_t)device_sm_count() * 16 + gx - 1) / gx;
        if (gy > n) gy = n;
        if (gy > 65535) gy = 65535;
        saq_encode_vec_kernel<<<dim3(gx, (unsigned)gy), 128, 0, as_stream(stream)>>>(
            x, n, ldx, D, vmin, vmax, dim_bits, bit_offset, code_bytes, codes, word_ok);
        VQ_LAUNCH_CHECK();
        return B200VQ_OK;
    }
    saq_encode_kernel<<<sq_grid(n * (int64_t)code_bytes, 256), 256, 0, as_stream(stream)>>>(
        x, n, ldx, D, vmin, vmax, dim_bits, bit_offset, code_bytes, codes);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" int b200vq_saq_decode(const uint8_t* codes, int64_t n, int D, const float* vmin,
                                 const float* vmax, const int* dim_bits, const int* bit_offset,
                                 int code_bytes, float* out, int64_t ldo, void* stream) {
    VQ_CHECK_ARG((codes || n == 0) && vmin && vmax && dim_bits && bit_offset && (out || n == 0) && n >= 0 && D > 0 &&
                     ldo >= D && code_bytes > 0,
                 "saq_decode: bad args");
    if (n == 0) return B200VQ_OK;
    if ((D % 4 == 0) && (ldo % 4 == 0) && (((uintptr_t)out & 15) == 0)) {
        int gx = (D / 4 + 127) / 128;
        int64_t gy = ((int64_t)device_sm_count() * 16 + gx - 1) / gx;
        if (gy > n) gy = n;
        if (gy > 65535) gy = 65535;
        saq_decode_vec_kernel<<<dim3(gx, (unsigned)gy), 128, 0, as_stream(stream)>>>(
            codes, n, D, vmin, vmax, dim_bits, bit_offset, code_bytes, out, ldo);
        VQ_LAUNCH_CHECK();
        return B200VQ_OK;
    }
    saq_decode_kernel<<<sq_grid(n * (int64_t)D, 256), 256, 0, as_stream(stream)>>>(
        codes, n, D, vmin, vmax, dim_bits, bit_offset, code_bytes, out, ldo);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}
