// Per-dimension codebook quantizer: nearest-level search + variable-width bit packing (sm_100a).
//
// Data path of RankAwareQuantizer (src/haag_vq/methods/rank_aware_quantization.py): `_quantize_dim`
// (:229-235: np.searchsorted of the column into the mid-points of its sorted codebook), the dense
// MSB-first bit string of `_compress_block` (:262-279, np.packbits) or the FFD byte layout (:252-260,
// ffd_packing.py:84-98), and the inverse gathers of `decompress` (:283-320). Both layouts are "a b_d-bit
// field at MSB-first bit position pos_d of the row": pos_d = cumulative widths (dense) or
// 8 byte_idx_d + bit_off_d (FFD, never crossing a byte), so one kernel pair serves both.
// The projected coordinates are fp64 (the reference projects in float64); comparisons are exact, the
// packed bytes are integer work: bit-identical for identical inputs.
// HBM-bound: 8 D bytes in + code_size out per vector (encode), the reverse for decode.
#include "common.cuh"

namespace b200vq {

// One warp per row: indices are OR-ed into a zeroed shared-memory copy of the row's bytes, then stored.
__global__ void __launch_bounds__(256) perdim_encode_kernel(const double* __restrict__ y, int64_t n, int64_t ldy, int D,
                                                            const double* __restrict__ bounds,
                                                            const int* __restrict__ bound_off,
                                                            const uint8_t* __restrict__ bits,
                                                            const int* __restrict__ bitpos, int code_size,
                                                            uint8_t* __restrict__ codes) {
    extern __shared__ uint32_t pd_smem[];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int words = (code_size + 1 + 3) >> 2;   // + 1: a field's second byte may be the (unused) byte after the row
    uint32_t* buf = pd_smem + (size_t)w * words;
    const int64_t wid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t row = wid; row < n; row += nw) {
        for (int i = lane; i < words; i += 32) buf[i] = 0u;
        __syncwarp();
        const double* yr = y + row * ldy;
        for (int d = lane; d < D; d += 32) {
            const int b = bits[d];
            if (b == 0) continue;
            const double v = yr[d];
            // np.searchsorted(boundaries, v) (side='left'): number of boundaries strictly below v
            const double* bd = bounds + bound_off[d];
            int lo = 0, hi = (1 << b) - 1;
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                if (bd[mid] < v) lo = mid + 1; else hi = mid;
            }
            const int pos = bitpos[d], byte = pos >> 3;
            const uint32_t v16 = (uint32_t)lo << (16 - b - (pos & 7));   // field inside a big-endian 16-bit window
            const uint32_t hi8 = v16 >> 8, lo8 = v16 & 0xFFu;
            if (hi8) atomicOr(buf + (byte >> 2), hi8 << (8 * (byte & 3)));
            if (lo8) atomicOr(buf + ((byte + 1) >> 2), lo8 << (8 * ((byte + 1) & 3)));
        }
        __syncwarp();
        uint8_t* o = codes + row * (int64_t)code_size;
        const uint8_t* src = reinterpret_cast<const uint8_t*>(buf);
        for (int i = lane; i < code_size; i += 32) o[i] = src[i];
        __syncwarp();
    }
}

__global__ void __launch_bounds__(256) perdim_decode_kernel(const uint8_t* __restrict__ codes, int64_t n, int code_size,
                                                            int D, const uint8_t* __restrict__ bits,
                                                            const int* __restrict__ bitpos,
                                                            const double* __restrict__ cb,
                                                            const int* __restrict__ cb_off, double* __restrict__ yhat,
                                                            int64_t ldy) {
    const int lane = threadIdx.x & 31;
    const int64_t wid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t row = wid; row < n; row += nw) {
        const uint8_t* c = codes + row * (int64_t)code_size;
        double* o = yhat + row * ldy;
        for (int d = lane; d < D; d += 32) {
            const int b = bits[d];
            double r = 0.0;   // zero-width dimensions reconstruct to 0 (the component mean)
            if (b) {
                const int pos = bitpos[d], byte = pos >> 3;
                const uint32_t v16 = ((uint32_t)c[byte] << 8) | (byte + 1 < code_size ? (uint32_t)c[byte + 1] : 0u);
                const uint32_t idx = (v16 >> (16 - b - (pos & 7))) & ((1u << b) - 1u);
                r = cb[cb_off[d] + idx];
            }
            o[d] = r;
        }
    }
}

}  // namespace b200vq

using namespace b200vq;

static int perdim_grid(int64_t n) {
    int64_t g = (n * 32 + 255) / 256;
    const int64_t cap = (int64_t)device_sm_count() * 8;
    if (g > cap) g = cap;
    return (int)(g < 1 ? 1 : g);
}

extern "C" int b200vq_perdim_encode(const double* y, int64_t n, int64_t ldy, int D, const double* bounds,
                                    const int* bound_off, const uint8_t* bits, const int* bitpos, int code_size,
                                    uint8_t* codes, void* stream) {
    VQ_CHECK_ARG(n >= 0 && D > 0 && ldy >= D && code_size >= 0, "perdim_encode: bad shape");
    if (n == 0 || code_size == 0) return B200VQ_OK;
    VQ_CHECK_ARG(y && bounds && bound_off && bits && bitpos && codes, "perdim_encode: null pointer");
    const size_t smem = (size_t)8 * ((code_size + 4) / 4) * 4;
    VQ_CHECK_ARG(smem <= 200 * 1024, "perdim_encode: code_size %d too large", code_size);
    VQ_CUDA(cudaFuncSetAttribute(perdim_encode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    perdim_encode_kernel<<<perdim_grid(n), 256, smem, as_stream(stream)>>>(y, n, ldy, D, bounds, bound_off, bits, bitpos,
                                                                         code_size, codes);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}

extern "C" int b200vq_perdim_decode(const uint8_t* codes, int64_t n, int code_size, int D, const uint8_t* bits,
                                    const int* bitpos, const double* cb, const int* cb_off, double* yhat, int64_t ldy,
                                    void* stream) {
    VQ_CHECK_ARG(n >= 0 && D > 0 && ldy >= D && code_size >= 0, "perdim_decode: bad shape");
    if (n == 0) return B200VQ_OK;
    VQ_CHECK_ARG(bits && bitpos && cb && cb_off && yhat && (codes || code_size == 0), "perdim_decode: null pointer");
    perdim_decode_kernel<<<perdim_grid(n), 256, 0, as_stream(stream)>>>(codes, n, code_size, D, bits, bitpos, cb, cb_off,
                                                                      yhat, ldy);
    VQ_LAUNCH_CHECK();
    return B200VQ_OK;
}
