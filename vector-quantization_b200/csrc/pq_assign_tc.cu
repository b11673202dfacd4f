// placeholder until the tcgen05 kernel lands (replaced in a later commit)
#include "common.cuh"
using namespace b200vq;
extern "C" size_t b200vq_pq_assign_tc_scratch_bytes(int64_t n, int M) {
    if (n <= 0 || M <= 0) return 0;
    return (size_t)n * M * 4;
}
extern "C" int b200vq_pq_assign_tc(const float*, int64_t, int64_t, const float*, const float*, int, int,
                                   int, uint8_t*, void*, size_t, unsigned long long*, void*) {
    set_error("pq_assign_tc: not built yet");
    return B200VQ_ERR_UNSUPPORTED;
}
