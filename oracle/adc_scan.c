/* CPU oracle (TEST INFRASTRUCTURE ONLY): C restatement of the reference's ADC
 * scan inner loop, bench/ffd_speed.cpp:120-126 ("uniform-width ADC"):
 *     score(i) = sum_d LUT[d][code_i[d]]
 * specialised to one byte per sub-quantizer (PQ, nbits = 8), plus the
 * nearest-centroid assignment loop of faiss ProductQuantizer::compute_codes
 * (call site src/haag_vq/methods/product_quantization.py:76-80) in its direct
 * (non-BLAS) form. Used by bench.py as the timed CPU baseline ("port") and by
 * the tests as a second opinion on oracle/pq.py. OpenMP over rows.
 */
#include <float.h>
#include <stdint.h>
#include <stddef.h>

/* scores[q*n + i] = sum_m lut[q][m][codes[i][m]] */
void oracle_adc_scores(const uint8_t* codes, int64_t n, int M, int ksub, const float* lut, int nq,
                       float* scores) {
    for (int q = 0; q < nq; ++q) {
        const float* L = lut + (size_t)q * M * ksub;
        float* out = scores + (size_t)q * n;
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < n; ++i) {
            const uint8_t* c = codes + (size_t)i * M;
            float a = 0.f;
            for (int m = 0; m < M; ++m) a += L[(size_t)m * ksub + c[m]];
            out[i] = a;
        }
    }
}

/* best-1 scan (keeps the timed loop free of an n-sized output): returns argmin id per query */
void oracle_adc_argmin(const uint8_t* codes, int64_t n, int M, int ksub, const float* lut, int nq,
                       int64_t* best_id, float* best_val) {
    for (int q = 0; q < nq; ++q) {
        const float* L = lut + (size_t)q * M * ksub;
        float gbest = FLT_MAX;
        int64_t gid = -1;
#pragma omp parallel
        {
            float lb = FLT_MAX;
            int64_t li = -1;
#pragma omp for schedule(static) nowait
            for (int64_t i = 0; i < n; ++i) {
                const uint8_t* c = codes + (size_t)i * M;
                float a = 0.f;
                for (int m = 0; m < M; ++m) a += L[(size_t)m * ksub + c[m]];
                if (a < lb) { lb = a; li = i; }
            }
#pragma omp critical
            { if (lb < gbest || (lb == gbest && li < gid)) { gbest = lb; gid = li; } }
        }
        best_id[q] = gid;
        best_val[q] = gbest;
    }
}

/* codes[i][m] = argmin_k ||x_im - c_mk||^2 (direct form, first index wins ties) */
void oracle_pq_assign(const float* x, int64_t n, int D, const float* centroids, int M, int ksub,
                      uint8_t* codes) {
    const int dsub = D / M;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i) {
        for (int m = 0; m < M; ++m) {
            const float* xv = x + (size_t)i * D + (size_t)m * dsub;
            float best = FLT_MAX;
            int bk = 0;
            for (int k = 0; k < ksub; ++k) {
                const float* c = centroids + ((size_t)m * ksub + k) * dsub;
                float a = 0.f;
                for (int j = 0; j < dsub; ++j) { float d = xv[j] - c[j]; a += d * d; }
                if (a < best) { best = a; bk = k; }
            }
            codes[(size_t)i * M + m] = (uint8_t)bk;
        }
    }
}

/* Top-k scan without the n-sized score matrix: queries are shared out over the threads, each keeps a sorted list of
 * its k best (score, id) pairs (ascending score, lower id first on ties; inner product callers pass negated tables).
 * The scores are those of oracle_adc_scores (same loop, same order). k <= 1024. */
int oracle_adc_topk(const uint8_t* codes, int64_t n, int M, int ksub, const float* lut, int nq, int k,
                    float* out_vals, int64_t* out_ids) {
    if (k < 1 || k > 1024) return -1;
#pragma omp parallel for schedule(dynamic, 1)
    for (int q = 0; q < nq; ++q) {
        const float* L = lut + (size_t)q * M * ksub;
        float bv[1024];
        int64_t bi[1024];
        int cnt = 0;
        for (int64_t i = 0; i < n; ++i) {
            const uint8_t* c = codes + (size_t)i * M;
            float a = 0.f;
            for (int m = 0; m < M; ++m) a += L[(size_t)m * ksub + c[m]];
            if (cnt == k && !(a < bv[k - 1])) continue;      /* ids ascend: an equal score never displaces */
            int p = cnt < k ? cnt : k - 1;
            while (p > 0 && bv[p - 1] > a) {
                bv[p] = bv[p - 1];
                bi[p] = bi[p - 1];
                --p;
            }
            bv[p] = a;
            bi[p] = i;
            if (cnt < k) ++cnt;
        }
        for (int j = 0; j < k; ++j) {
            out_vals[(size_t)q * k + j] = j < cnt ? bv[j] : FLT_MAX;
            out_ids[(size_t)q * k + j] = j < cnt ? bi[j] : -1;
        }
    }
    return 0;
}
