// Canonical fp32 arithmetic of the 1-bit RaBitQ distance estimate (faiss RaBitDistanceComputerQ::
// distance_to_code, as driven by src/haag_vq/methods/search/rabitq_index.py:62-70), shared by the SIMT
// scan (rabitq.cu) and the tensor-core scan (rabitq_tc.cu) so that both produce the same bits.
//   dot_f, sumq_f : <code bits, quantised query> and popc(code) as exact fp32 integers
//   f0            : ||o - c||^2  (minus ||o||^2 for IP), a = -2 * dp_multiplier (exact scaling)
//   c1, c2, c34, qc2, qn2 : per-query factors written by rabitq1_prepare_queries_kernel
#pragma once
#include "common.cuh"

namespace b200vq {

__device__ __forceinline__ float rabitq1_estimate(float dot_f, float sumq_f, float f0, float a, float c1, float c2,
                                                  float c34, float qc2, float qn2, int metric) {
    float t = __fmul_rn(c1, dot_f);
    t = __fmaf_rn(c2, sumq_f, t);
    t = __fsub_rn(t, c34);
    const float pre = __fmaf_rn(a, t, __fadd_rn(f0, qc2));
    return metric == B200VQ_METRIC_L2 ? pre : __fmul_rn(0.5f, __fsub_rn(pre, qn2));  // IP: key = -ip
}

}  // namespace b200vq
