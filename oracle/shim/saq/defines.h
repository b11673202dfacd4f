// Test-infrastructure shim (oracle/): stands in for the reference's saq/defines.h (which pulls Eigen, glog,
// gflags and fmt from the network) when compiling external/saq/src/preprocessing/codebook_builder.cpp in
// place. Only what that file touches: a row-major float matrix with rows() / cols() / operator().
#pragma once
#include <cstddef>
#include <vector>
namespace Eigen { using Index = long; }
namespace saq {
struct FloatRowMat {
    long r = 0, c = 0;
    const float* p = nullptr;
    long rows() const { return r; }
    long cols() const { return c; }
    float operator()(long i, long j) const { return p[i * c + j]; }
};
}  // namespace saq
