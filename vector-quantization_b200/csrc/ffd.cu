// First-fit-decreasing byte packing of per-dimension codes (sm_100a).
//
// Replaces the numpy loops of src/haag_vq/methods/ffd_packing.py:84-117 (ffd_encode / ffd_decode,
// used by RankAwareQuantizer(packing="ffd"), methods/rank_aware_quantization.py:4,106-109): every
// dimension's b_d-bit code sits wholly inside one byte at a fixed shift, so packing is a per-byte
// OR of <= 8 shifted fields and unpacking one shift-and-mask per dimension. Integer work: bit-exact.
// The layout (which byte, which shift) is host state computed once per quantizer (ffd_layout).
// HBM-bound: D bytes in, n_bytes out per vector (pack) and the reverse (unpack).
#include "common.cuh"

namespace b200vq {

// byte_dims (n_bytes, 8) int32: dimensions packed into each byte (-1 = none); dim_shift (D) uint8.
__global__ void __launch_bounds__(256) ffd_pack_kernel(const uint8_t* __restrict__ codes, int64_t n, int D,
                                                       const int* __restrict__ byte_dims,
                                                       const uint8_t* __restrict__ dim_shift, int n_bytes,
                                                       uint8_t* __restrict__ out) {
    // a CTA walks rows; its threads own output bytes, the byte's field list stays in registers
    for (int b0 = threadIdx.x; b0 < n_bytes; b0 += blockDim.x) {
        int dims[8], sh[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            dims[k] = byte_dims[b0 * 8 + k];
            sh[k] = dims[k] >= 0 ? dim_shift[dims[k]] : 0;
        }
        for (int64_t row = blockIdx.x; row < n; row += gridDim.x) {
            const uint8_t* c = codes + row * D;
            uint32_t v = 0;
#pragma unroll
            for (int k = 0; k < 8; ++k)
                if (dims[k] >= 0) v |= (uint32_t)c[dims[k]] << sh[k];
            out[row * n_bytes + b0] = (uint8_t)v;   // fields wider than their slot wrap like numpy's uint8 shift
        }
    }
}

// dim_byte (D) int32 (-1 for zero-width dimensions), dim_shift (D), dim_mask (D) uint8.
__global__ void __launch_bounds__(256) ffd_unpack_kernel(const uint8_t* __restrict__ packed, int64_t n, int n_bytes,
                                                         const int* __restrict__ dim_byte,
                                                         const uint8_t* __restrict__ dim_shift,
                                                         const uint8_t* __restrict__ dim_mask, int D,
                                                         uint8_t* __restrict__ codes) {
    for (int d = threadIdx.x; d < D; d += blockDim.x) {
        const int bi = dim_byte[d];
        const int sh = dim_shift[d], mk = dim_mask[d];
        for (int64_t row = blockIdx.x; row < n; row += gridDim.x)
            codes[row * D + d] = bi >= 0 ? (uint8_t)((packed[row * n_bytes + bi] >> sh) & mk) : (uint8_t)0;
    }
}

}  // namespace b200vq

using namespace b200vq;

static int ffd_grid(int64_t n) {
    const int64_t cap = (int64_t)device_sm_count() * 8;
    return (int)(n < cap ? (n < 1 ? 1 : n) : cap);
}

extern "C" int b200vq_ffd_pack(const uint8_t* codes, int64_t n, int D, const int* byte_dims,
                               const uint8_t* dim_shift, int n_bytes, uint8_t* out, void* stream) {
    VQ_CHECK_ARG(n >= 0 && D > 0 && n_bytes >= 0, "ffd_pack: bad shape");
    if (n == 0 || n_bytes == 0) return B200VQ_OK;
    VQ_CHECK_ARG(codes && byte_dims && dim_shift && out, "ffd_pack: null pointer");
    ffd_pack_kernel<<<ffd_grid(n), 256, 0, as_stream(stream)>>>(codes, n, D, byte_dims, dim_shift, n_bytes, out);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" int b200vq_ffd_unpack(const uint8_t* packed, int64_t n, int n_bytes, const int* dim_byte,
                                 const uint8_t* dim_shift, const uint8_t* dim_mask, int D, uint8_t* codes,
                                 void* stream) {
    VQ_CHECK_ARG(n >= 0 && D > 0 && n_bytes >= 0, "ffd_unpack: bad shape");
    if (n == 0) return B200VQ_OK;
    VQ_CHECK_ARG(dim_byte && dim_shift && dim_mask && codes && (packed || n_bytes == 0), "ffd_unpack: null pointer");
    ffd_unpack_kernel<<<ffd_grid(n), 256, 0, as_stream(stream)>>>(packed, n, n_bytes, dim_byte, dim_shift, dim_mask,
                                                                  D, codes);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}
