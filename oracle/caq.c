/* CPU oracle: CAQ segment encoder and the SAQ joint segmentation / bit-allocation DP.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle/__init__.py). Restates, in plain C with the same
 * float/double types and the same operation order:
 *   external/saq/include/saq/caq_encoder.h:170-218   CAQEncoder::encode
 *   external/saq/include/saq/caq_encoder.h:67-140    CAQEncoder::code_adjustment
 *   external/saq/include/saq/caq_encoder.h:220-261   CAQEncoder::encode_and_fac (+ rescale_vmx_to1, :29-40)
 *   external/saq/src/quantization_plan.cpp:144-255   SaqDataMaker::dynamic_programming
 * Built with -ffp-contract=off: every multiply and add rounds separately (the reference's build
 * flags are not recoverable; its Eigen reductions o.sum() / o.squaredNorm() are restated as
 * sequential float sums). PARITY UNPINNED against the CPU engine (it needs Eigen/glog, absent
 * here); pinned instead against the reference's own CUDA twin (src/gpu/gpu_encoder.cu,
 * caq_sequential=1) in tests/test_gpu_caq.py.
 */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* One row of one segment. o: len floats (rotated residual). Outputs the integer codes and the
 * three float factors the engine stores (o_l2norm, fac_rescale after rescale_vmx_to1, fac_error),
 * plus recon_mse (caq_encoder.h:233-246) and <o,oa>/(|o||oa|). */
static void caq_encode_row(const float* o, int len, int bits, int rd_lmt, float adj_eps, uint16_t* code,
                           float* out_l2norm, float* out_rescale, float* out_error, float* out_mse,
                           float* out_cos) {
    if (bits == 0) { /* caq_encoder.h:171-174: CaqCode() zero-initialised */
        float sq = 0.f;
        for (int j = 0; j < len; ++j) sq += o[j] * o[j];
        double o_l2sqr = sq;
        *out_l2norm = (float)sqrt(o_l2sqr);
        *out_rescale = 0.f; /* ip_o_oa == 0 */
        /* the CPU engine's expression is 0/0 here (a NaN no estimator reads: 0-bit segments skip the
         * factors); like the reference's CUDA twin (gpu_encoder.cu:66-78) the value stored is 0 */
        *out_error = 0.f;
        *out_mse = sq / (float)len;
        *out_cos = 0.f;
        for (int j = 0; j < len; ++j) code[j] = 0;
        return;
    }
    const int code_max = (1 << bits) - 1;
    float mx = o[0], mn = o[0], fsum = 0.f, fsq = 0.f;
    for (int j = 0; j < len; ++j) {
        if (o[j] > mx) mx = o[j];
        if (o[j] < mn) mn = o[j];
        fsum += o[j];
        fsq += o[j] * o[j];
    }
    double v_mx = (mx > -mn) ? mx : -mn;
    double v_mi = -v_mx;
    double delta = (v_mx - v_mi) / (code_max + 1);
    double vec_sum = fsum;

    double ip_o_code = 0;
    uint64_t code_l2sqr = 0;
    uint32_t code_sum = 0;
    const float fmi = (float)v_mi, fdelta = (float)delta;
    for (int j = 0; j < len; ++j) {
        float cf = 0.f;
        if (delta) {
            cf = floorf((o[j] - fmi) / fdelta);
            if (cf > (float)code_max) cf = (float)code_max;
        }
        ip_o_code += cf * o[j]; /* float product, double accumulate */
        int c = (int)cf;
        code[j] = (uint16_t)c;
        code_l2sqr += (uint64_t)(c * c);
        code_sum += (uint32_t)c;
    }
    double ip_o_oa = ip_o_code * delta + (v_mi + 0.5 * delta) * vec_sum;
    double oa_l2sqr = delta * delta * code_l2sqr + (delta * delta + 2 * delta * v_mi) * code_sum;
    oa_l2sqr += (0.25 * delta * delta + delta * v_mi + v_mi * v_mi) * len;

    if (rd_lmt && oa_l2sqr) {
        const double re_eps = adj_eps * oa_l2sqr;
        for (int adj_cnt = 1, round = 1; adj_cnt; round++) {
            adj_cnt = 0;
            for (int j = 0; j < len; j++) {
                int curr = 0;
                const float of = o[j];
                double oa = (code[j] + 0.5) * delta + v_mi;
                int c = code[j];
                const double tmp = oa_l2sqr - oa * oa;
                const double ip_delta = delta * of;
                while (c < code_max) {
                    double new_q = oa + delta;
                    double new_length = tmp + new_q * new_q;
                    double new_ip = ip_o_oa + ip_delta;
                    if ((ip_o_oa * ip_o_oa + re_eps) * new_length >= new_ip * new_ip * oa_l2sqr) break;
                    c++;
                    ip_o_oa = new_ip;
                    oa = new_q;
                    oa_l2sqr = new_length;
                    curr++;
                }
                while (c > 0) {
                    double new_q = oa - delta;
                    double new_length = tmp + new_q * new_q;
                    double new_ip = ip_o_oa - ip_delta;
                    if ((ip_o_oa * ip_o_oa + re_eps) * new_length >= new_ip * new_ip * oa_l2sqr) break;
                    c--;
                    ip_o_oa = new_ip;
                    oa = new_q;
                    oa_l2sqr = new_length;
                    curr++;
                }
                if (code[j] != c) {
                    code[j] = (uint16_t)c;
                    adj_cnt += curr;
                }
            }
            if (rd_lmt && round >= rd_lmt) break;
            double chk_l2 = 0, chk_ip = 0;
            for (int j = 0; j < len; ++j) {
                double q = (code[j] + 0.5) * delta + v_mi;
                chk_ip += q * o[j];
                chk_l2 += q * q;
            }
            oa_l2sqr = chk_l2;
            ip_o_oa = chk_ip;
        }
    }
    /* encode_and_fac */
    double o_l2sqr = fsq;
    double o_l2norm = sqrt(o_l2sqr);
    double fac_rescale = ip_o_oa ? o_l2sqr / ip_o_oa : 0;
    double fac_error =
        o_l2sqr * (double)1.9f * sqrt((((o_l2sqr * oa_l2sqr) / (ip_o_oa * ip_o_oa)) - 1) / (double)(len - 1));
    /* recon_mse uses CaqCode::get_oa(): code * delta + v_mi in float (no half step) */
    float diff = 0.f;
    for (int j = 0; j < len; ++j) {
        float oa = (float)code[j] * (float)delta + (float)v_mi;
        float d = o[j] - oa;
        diff += d * d;
    }
    *out_mse = diff / (float)len;
    *out_cos = (bits > 1) ? (float)(ip_o_oa / o_l2norm / sqrt(oa_l2sqr)) : 0.f;
    if (v_mx) fac_rescale *= v_mx; /* rescale_vmx_to1 */
    *out_l2norm = (float)o_l2norm;
    *out_rescale = (float)fac_rescale;
    *out_error = (float)fac_error;
}

/* o: (n, len) row-major with row stride ld. codes: (n, len) uint16. */
void oracle_caq_encode(const float* o, int64_t n, int len, int64_t ld, int bits, int rd_lmt, float adj_eps,
                       uint16_t* codes, float* l2norm, float* rescale, float* error, float* mse, float* cosv) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i)
        caq_encode_row(o + i * ld, len, bits, rd_lmt, adj_eps, codes + i * (int64_t)len, l2norm + i, rescale + i,
                       error + i, mse + i, cosv + i);
}

/* SaqDataMaker::dynamic_programming. variance: d_pad floats. Writes up to d_pad/64 (len, bits)
 * pairs; returns the number of segments, or -1 on allocation failure. */
int oracle_saq_plan_dp(const float* variance, int d_pad, float avg_bits, int* seg_len, int* seg_bits) {
    const int BLK = 64, MAXB = 13;
    const size_t nbf = 2 * sizeof(float) * 8;
    const size_t tot_bits = (size_t)(avg_bits * d_pad + nbf);
    const size_t i_end = (size_t)d_pad / BLK;
    size_t max_segs = avg_bits < 2 ? i_end : i_end / 2;
    if (max_segs < 1) max_segs = 1;
    const double LMT = DBL_MAX;
    const size_t W = tot_bits + 1, plane = (i_end + 1) * W;
    double* fv = (double*)malloc(sizeof(double) * (max_segs + 2) * plane);
    uint32_t* fb = (uint32_t*)calloc((max_segs + 2) * plane, sizeof(uint32_t));
    if (!fv || !fb) { free(fv); free(fb); return -1; }
    for (size_t k = 0; k < (max_segs + 2) * plane; ++k) fv[k] = LMT;
#define F(ns, i, b) ((ns) * plane + (i) * W + (b))
    size_t ans_ns = 0, ans_i = i_end, ans_b = tot_bits;
    fv[F(0, 0, 0)] = 0;
    for (size_t ns = 0; ns <= max_segs; ns++)
        for (size_t i = 0; i <= i_end; i++)
            for (size_t ub = 0; ub <= tot_bits; ++ub) {
                const double cur = fv[F(ns, i, ub)];
                if (!(cur < LMT)) continue;
                if (i == i_end) {
                    if (cur * (1.01) < fv[F(ans_ns, ans_i, ans_b)]) { ans_ns = ns; ans_i = i; ans_b = ub; }
                    continue;
                }
                if (ns == max_segs) continue;
                double var_sum = 0;
                for (size_t j = 1; (i + j) * BLK <= (size_t)d_pad; j++) {
                    float blk = 0.f; /* Eigen segment(...).sum() in float, then added to the double */
                    for (int t = 0; t < BLK; ++t) blk += variance[(i + j - 1) * BLK + t];
                    var_sum += blk;
                    for (size_t b = 1; b <= (size_t)MAXB; ++b) {
                        size_t bn = ub + b * j * BLK + nbf;
                        if (bn > tot_bits) break;
                        double v = var_sum / (1 << b);
                        size_t to = F(ns + 1, i + j, bn);
                        if (fv[to] > cur + v) { fv[to] = cur + v; fb[to] = (uint32_t)((i << 4) + b); }
                    }
                }
                size_t to0 = F(ns + 1, i_end, ub);
                if (cur + var_sum < fv[to0]) { fv[to0] = cur + var_sum; fb[to0] = (uint32_t)((i << 4) + 0); }
            }
    int nseg = 0;
    {
        size_t ns = ans_ns, i = ans_i, B = ans_b;
        while (i > 0) {
            uint32_t s = fb[F(ns, i, B)];
            size_t prev_i = s >> 4, bits = s & 0xf, len = (i - prev_i) * BLK;
            seg_len[nseg] = (int)len;
            seg_bits[nseg] = (int)bits;
            nseg++;
            ns--;
            i = prev_i;
            if (bits) B -= bits * len + nbf;
        }
    }
#undef F
    for (int a = 0, b = nseg - 1; a < b; ++a, --b) {
        int t = seg_len[a]; seg_len[a] = seg_len[b]; seg_len[b] = t;
        t = seg_bits[a]; seg_bits[a] = seg_bits[b]; seg_bits[b] = t;
    }
    free(fv);
    free(fb);
    return nseg;
}
