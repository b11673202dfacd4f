/* CPU restatement (TEST INFRASTRUCTURE / timed CPU baseline ONLY) of faiss ProductQuantizer::compute_codes as the
 * reference drives it (src/haag_vq/methods/product_quantization.py:76-80): per sub-quantizer the nearest of ksub
 * centroids under d = |x|^2 + |c|^2 - 2 x.c (the decomposition of faiss' BLAS path), the first index winning ties.
 * faiss is a third-party dependency that is not vendored in the reference tree; oracle/pq.py is the numpy form of
 * the same arithmetic and tests/test_oracle.py checks this file against it. Rows are shared out over OpenMP threads,
 * the centroids are transposed once ([m][d][j]) so that the loop over the ksub centroids vectorises.
 * Build: oracle/Makefile -> _ref/liboracle_pq.so. */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <omp.h>

/* torchrun exports OMP_NUM_THREADS=1 to its workers: the CPU arm asks for its threads explicitly. */
void oracle_pq_set_threads(int n) { if (n > 0) omp_set_num_threads(n); }
int oracle_pq_max_threads(void) { return omp_get_max_threads(); }

int oracle_pq_assign_dec(const float* x, int64_t n, int D, const float* cent, int M, int ksub, uint8_t* codes) {
    const int dsub = D / M;
    if (dsub * M != D || ksub > 256 || ksub < 1) return -1;
    float* ct = (float*)malloc((size_t)M * dsub * ksub * sizeof(float));   /* [m][d][j] */
    float* cn = (float*)malloc((size_t)M * ksub * sizeof(float));
    if (!ct || !cn) { free(ct); free(cn); return -2; }
    for (int m = 0; m < M; ++m)
        for (int j = 0; j < ksub; ++j) {
            const float* c = cent + ((size_t)m * ksub + j) * dsub;
            float s = 0.f;
            for (int d = 0; d < dsub; ++d) {
                ct[((size_t)m * dsub + d) * ksub + j] = c[d];
                s += c[d] * c[d];
            }
            cn[(size_t)m * ksub + j] = s;
        }
#pragma omp parallel
    {
        float dot[256];
#pragma omp for schedule(static)
        for (int64_t i = 0; i < n; ++i) {
            const float* xi = x + i * (int64_t)D;
            for (int m = 0; m < M; ++m) {
                const float* xm = xi + m * dsub;
                const float* ctm = ct + (size_t)m * dsub * ksub;
                float xn = 0.f;
                for (int j = 0; j < ksub; ++j) dot[j] = 0.f;
                for (int d = 0; d < dsub; ++d) {
                    const float xv = xm[d];
                    const float* row = ctm + (size_t)d * ksub;
                    xn += xv * xv;
                    for (int j = 0; j < ksub; ++j) dot[j] += xv * row[j];
                }
                const float* cnm = cn + (size_t)m * ksub;
                float best = (xn + cnm[0]) - 2.f * dot[0];
                int bj = 0;
                for (int j = 1; j < ksub; ++j) {
                    const float dj = (xn + cnm[j]) - 2.f * dot[j];
                    if (dj < best) { best = dj; bj = j; }
                }
                codes[i * (int64_t)M + m] = (uint8_t)bj;
            }
        }
    }
    free(ct);
    free(cn);
    return 0;
}
