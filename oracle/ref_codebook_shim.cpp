// C-linkage shim (ours) over the REFERENCE's 1-D codebook builders, compiled together with
// external/saq/src/preprocessing/codebook_builder.cpp where it lies (oracle/Makefile, target ref_codebook).
// Test infrastructure only: pins oracle/codebook.py and the CUDA builder against the reference's own code.
#include <cstring>
#include <span>

#include "saq/preprocessing/codebook_builder.h"

extern "C" {
// init: 0 EqualMassQuantile, 1 UniformSpaced, 2 KMeansPlusPlus, 3 CubeRootDensity (the enum's order).
// centroids_out: (max_bits + 1) * (1 << max_bits) floats, level b at offset b << max_bits; sizes_out / costs_out: max_bits + 1.
int ref_codebook_lloyd(const float* values, long n, int max_bits, int init, int restarts, int max_iters, float tol,
                       unsigned long long seed, long sample_size, float* centroids_out, int* sizes_out, float* costs_out) {
    saq::LloydOpts o;
    o.max_bits = (size_t)max_bits;
    o.init = static_cast<saq::CodebookInit>(init);
    o.restarts = (size_t)restarts;
    o.max_iters = (size_t)max_iters;
    o.tol = tol;
    o.seed = seed;
    o.sample_size = (size_t)sample_size;
    saq::CodebookResult r = saq::build_codebook_lloyd(std::span<const float>(values, (size_t)n), o);
    for (int b = 0; b <= max_bits; ++b) {
        sizes_out[b] = (int)r.codebooks[b].centroids.size();
        costs_out[b] = r.costs[b];
        std::memcpy(centroids_out + ((size_t)b << max_bits), r.codebooks[b].centroids.data(),
                    r.codebooks[b].centroids.size() * sizeof(float));
    }
    return 0;
}
int ref_codebook_exact(const float* values, long n, int max_bits, float* centroids_out, int* sizes_out, float* costs_out) {
    saq::CodebookResult r = saq::build_codebook_exact(std::span<const float>(values, (size_t)n), (size_t)max_bits);
    for (int b = 0; b <= max_bits; ++b) {
        sizes_out[b] = (int)r.codebooks[b].centroids.size();
        costs_out[b] = r.costs[b];
        std::memcpy(centroids_out + ((size_t)b << max_bits), r.codebooks[b].centroids.data(),
                    r.codebooks[b].centroids.size() * sizeof(float));
    }
    return 0;
}
}
