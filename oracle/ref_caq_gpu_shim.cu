// C-linkage shim around the REFERENCE's CUDA CAQ encoder, so the tests can call it through ctypes.
// TEST INFRASTRUCTURE ONLY. This file holds no reference code: it includes the reference's public
// header where it lies under /root/reference (external/saq/include/saq/gpu/gpu_encoder.cuh) and is
// linked with the reference's own src/gpu/gpu_encoder.cu by oracle/Makefile into
// oracle/_ref/libsaq_ref_gpu.so (git-ignored, travels to the GPU box).
#include <cstdint>
#include <cuda_runtime.h>

#include "saq/gpu/gpu_encoder.cuh"

// All pointers are device pointers. vectors_rotated (n, d_seg) row-major, centroid_rot (1, d_seg);
// short_raw (n, d_seg/8) and long_raw (n, d_seg*(bits-1)/8) must be zero-filled by the caller.
extern "C" int ref_saq_gpu_caq_encode(const float* vectors_rotated, const float* centroid_rot,
                                      const uint32_t* cluster_ids, float* o_l2norm, float* fac_rescale,
                                      float* fac_error, float* ip_cent_oa, uint8_t* short_raw, uint8_t* long_raw,
                                      size_t d_seg, size_t n, size_t num_bits, int rd_lmt, float adj_eps,
                                      int sequential) {
    try {
        saq::gpu::launch_fused_caq_encode(vectors_rotated, centroid_rot, cluster_ids, o_l2norm, fac_rescale,
                                          fac_error, ip_cent_oa, short_raw, long_raw, d_seg, n, /*K=*/1, num_bits,
                                          (uint16_t)((1u << num_bits) - 1u), rd_lmt, adj_eps, /*caq_ori_qB=*/0,
                                          sequential, 0);
    } catch (...) {
        return -1;
    }
    return (int)cudaDeviceSynchronize();
}
