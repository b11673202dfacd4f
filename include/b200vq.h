/* b200vq.h — C ABI of the B200-native quantize-and-search hot path.
 *
 * This is the drop-in boundary below the reference's Python quantizer classes.
 * Every entry point replaces one arithmetic provider call the reference makes
 * (faiss via SWIG, numpy, or the `saq` pybind module); the reference file:line
 * each one stands in for is cited beside it (paths relative to the reference
 * repository root, Human-Augment-Analytics/vector-quantization).
 *
 * Conventions
 *   - All data pointers are DEVICE pointers unless the name ends in `_host`.
 *   - Matrices are row-major, C-contiguous. `n` counts rows (vectors).
 *   - `stream` is a cudaStream_t passed as void* (NULL = default stream).
 *     Calls are asynchronous with respect to the host; no hidden allocation,
 *     no hidden synchronisation. Scratch buffers are caller-owned; their
 *     required sizes come from the matching *_scratch_bytes() function.
 *   - Return value: B200VQ_OK (0) or a negative status; on error
 *     b200vq_last_error() returns a thread-local, NUL-terminated message.
 *   - Top-k results are ordered by (value ascending, id ascending) for L2 and
 *     by (value descending, id ascending) for inner product; ids are uint32
 *     row indices local to the scanned shard (callers add the shard offset).
 */
#ifndef B200VQ_H
#define B200VQ_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200VQ_OK 0
#define B200VQ_ERR_INVALID (-1)     /* bad argument / unsupported shape */
#define B200VQ_ERR_CUDA (-2)        /* a CUDA runtime call failed        */
#define B200VQ_ERR_UNSUPPORTED (-3) /* valid request the build cannot serve */

#define B200VQ_METRIC_L2 0
#define B200VQ_METRIC_IP 1

const char* b200vq_last_error(void);
/* ABI version: bumped whenever a signature below changes. */
int b200vq_abi_version(void);
/* Compute capability of the current device as major*10+minor (100 on B200), <0 on error. */
int b200vq_device_cc(void);

/* ------------------------------------------------------------------ PQ ---
 * centroids: (M, ksub, dsub) fp32, the layout of faiss ProductQuantizer.centroids
 * (reference: src/haag_vq/methods/product_quantization.py:70-73). D = M*dsub.
 * cnorm: (M, ksub) fp32 squared norms of the centroids (b200vq_pq_centroid_norms).
 */

/* ||c_{m,k}||^2 for every centroid. */
int b200vq_pq_centroid_norms(const float* centroids, int M, int ksub, int dsub,
                             float* cnorm, void* stream);

/* Nearest-centroid assignment, exact fp32 SIMT path (any dsub, ksub <= 256).
 * Replaces faiss ProductQuantizer.compute_codes as called at
 * src/haag_vq/methods/product_quantization.py:76-80 and
 * src/haag_vq/methods/optimized_product_quantization.py:30-31.
 * x: (n, ldx) fp32 with ldx >= D (row stride in floats); codes: (n, M) uint8.
 * Distance = (||x_m||^2 + ||c||^2) - 2 x_m.c in fp32, ties -> lowest index.
 * If sqdist != NULL it receives (n, M) fp32 winning distances. */
int b200vq_pq_assign_f32(const float* x, int64_t n, int64_t ldx, const float* centroids,
                         const float* cnorm, int M, int ksub, int dsub, uint8_t* codes,
                         float* sqdist, void* stream);

/* Tensor-core assignment: tcgen05 kind::tf32 screen (128 x 256 TMEM accumulators, TMA-staged
 * swizzled tiles, ||c||^2 folded in as an extra K slice) + exact fp32 re-evaluation of every
 * (row, sub-quantizer) whose screen leaves more than one candidate inside the TF32 error
 * margin. Same contract and bit-identical codes to b200vq_pq_assign_f32. Requires
 * ksub == 256, dsub % 8 == 0, 8 <= dsub <= 128, x 16-byte aligned, ldx % 4 == 0
 * (B200VQ_ERR_UNSUPPORTED / B200VQ_ERR_INVALID otherwise). scratch: 256-byte aligned,
 * b200vq_pq_assign_tc_scratch_bytes2(n, M, dsub) bytes (the two-argument form sizes for the
 * largest dsub). stats (optional, 3 x uint64 device): [0] += pairs re-evaluated exactly,
 * [1] += candidates they held, [2] += pairs sent to the exact 256-way fix-up scan. */
size_t b200vq_pq_assign_tc_scratch_bytes(int64_t n, int M);
size_t b200vq_pq_assign_tc_scratch_bytes2(int64_t n, int M, int dsub);
int b200vq_pq_assign_tc(const float* x, int64_t n, int64_t ldx, const float* centroids,
                        const float* cnorm, int M, int ksub, int dsub, uint8_t* codes,
                        void* scratch, size_t scratch_bytes, unsigned long long* stats,
                        void* stream);

/* Decode: out[i, m*dsub:(m+1)*dsub] = centroids[m, codes[i,m], :].
 * Replaces faiss ProductQuantizer.decode (product_quantization.py:82-86). */
int b200vq_pq_decode(const uint8_t* codes, int64_t n, const float* centroids, int M, int ksub,
                     int dsub, float* out, int64_t ldo, void* stream);

/* Sub-byte code layout of faiss ProductQuantizer for nbits != 8 (the reference passes any B to faiss,
 * product_quantization.py:67-73; faiss PQEncoderGeneric / PQDecoderGeneric): the M indices of a row as a
 * little-endian bit stream of ceil(M nbits / 8) bytes. codes (n, M) uint8 <-> packed (n, ceil(M nbits / 8)). */
int b200vq_pq_pack_bits(const uint8_t* codes, int64_t n, int M, int nbits, uint8_t* packed, void* stream);
int b200vq_pq_unpack_bits(const uint8_t* packed, int64_t n, int M, int nbits, uint8_t* codes, void* stream);

/* Decode selected rows: out[j] = decode(codes[ids[j]]) — the adapter's
 * reconstruct(ids) (src/haag_vq/benchmarks/quantizer_adapters.py:48-54). */
int b200vq_pq_decode_ids(const uint8_t* codes, const int64_t* ids, int64_t n_ids,
                         const float* centroids, int M, int ksub, int dsub, float* out,
                         int64_t ldo, void* stream);

/* One Lloyd accumulation pass for all M sub-quantizers (k-means inside
 * faiss ProductQuantizer.train, product_quantization.py:67-68):
 * sums (M,ksub,dsub) fp32 and counts (M,ksub) int32 must be zeroed by the
 * caller; codes are the assignments of x under the current centroids. */
int b200vq_pq_kmeans_accumulate(const float* x, int64_t n, int64_t ldx, const uint8_t* codes,
                                int M, int ksub, int dsub, float* sums, int* counts,
                                void* stream);
/* centroids = sums / counts; empty clusters are re-seeded by splitting the
 * currently largest cluster with the +-1/1024 perturbation faiss uses.
 * n_split (device int, optional) receives the number of re-seeded clusters. */
int b200vq_pq_kmeans_finalize(const float* sums, int* counts, int M, int ksub, int dsub,
                              float* centroids, int* n_split, void* stream);

/* Sum over rows and sub-quantizers of the squared distance to the assigned
 * centroid (the k-means objective / reconstruction SSE). out: 1 x double, +=. */
int b200vq_pq_sse(const float* x, int64_t n, int64_t ldx, const uint8_t* codes,
                  const float* centroids, int M, int ksub, int dsub, double* out,
                  void* stream);

/* ----------------------------------------------------------------- ADC ---
 * Per-query look-up tables: lut[q, m, k] = ||q_m - c_{m,k}||^2 (L2) or q_m.c_{m,k} (IP).
 * This is the code-domain equivalent of the reference's decompress-then-distance
 * search (src/haag_vq/methods/search/flat_quantized_index.py:57-68) and of the ADC
 * inner loop of bench/ffd_speed.cpp:87-127. */
int b200vq_adc_lut(const float* queries, int nq, int64_t ldq, const float* centroids, int M,
                   int ksub, int dsub, int metric, float* lut, void* stream);

/* Scan (n, M) uint8 codes against nq LUTs and keep the best k per query.
 * score(q, i) = sum_m lut[q, m, codes[i, m]] (+ bias[i] if bias != NULL).
 * out_vals (nq, k) fp32, out_ids (nq, k) uint32; rows with fewer than k hits
 * are padded with +inf/-inf and id 0xffffffff. k <= 128.
 * scratch: b200vq_adc_scan_scratch_bytes(n, nq, k). */
size_t b200vq_adc_scan_scratch_bytes(int64_t n, int nq, int k);
int b200vq_adc_scan_topk(const uint8_t* codes, int64_t n, int M, int ksub, const float* lut,
                         int nq, int metric, int k, float* out_vals, uint32_t* out_ids,
                         void* scratch, size_t scratch_bytes, void* stream);

/* Allocation-free variant of b200vq_adc_scan_topk for callers that cannot call cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid it). It has NO counterpart in the
 * reference. Same inputs, outputs and results as
 * b200vq_adc_scan_topk, but the partial top-k lists live in a file-scope `static __device__` array of
 * B200VQ_ADC_STATIC_SCRATCH_BYTES (8 MiB, held by every GPU that loads the library until exit) whose
 * device address the library resolves once per device with cudaGetSymbolAddress. The call performs no
 * cudaMalloc / cudaFree / allocating library routine. Requests whose partial lists would not fit
 * (b200vq_adc_scan_scratch_bytes(n, nq, k) > B200VQ_ADC_STATIC_SCRATCH_BYTES) are rejected with
 * B200VQ_ERR_INVALID. All calls on a device share the array: calls that overlap on the GPU (different
 * streams, no ordering) race; calls ordered on one stream do not. Nothing in it needs clearing between
 * calls (every slot a call reads is written by that call first). */
#define B200VQ_ADC_STATIC_SCRATCH_BYTES (8u << 20)
int b200vq_adc_scan_topk_static(const uint8_t* codes, int64_t n, int M, int ksub, const float* lut,
                                int nq, int metric, int k, float* out_vals, uint32_t* out_ids,
                                void* stream);

/* Same scan over the conflict-free "scan layout" produced by
 * b200vq_pq_codes_to_scan_layout: blocks of 32 rows x groups of 32 sub-quantizers
 * (1 KiB each, row block major). Inside a block, lane i = row i of the block owns 32
 * bytes, byte s = its code for sub-quantizer (i+s) mod 32 of the group (so at step s
 * the 32 lanes address 32 distinct table columns = 32 distinct banks); the bytes are
 * stored as two 16-byte halves, each half contiguous across the 32 lanes:
 * out[block*1024 + (s/16)*512 + i*16 + s%16]. Requires M % 32 == 0, M <= 96, or M == 16: the
 * half-group layout of BASELINE config 1 - blocks of 32 rows x 16 bytes (512 B), byte s of lane i =
 * sub-quantizer (i+s) mod 16, out[block*512 + i*16 + s]. out has ceil(n/32)*32*M bytes (rows past n
 * are padding). */
int b200vq_pq_codes_to_scan_layout(const uint8_t* codes, int64_t n, int M, uint8_t* out,
                                   void* stream);
int b200vq_adc_scan_topk_rot(const uint8_t* codes_rot, int64_t n, int M, int ksub,
                             const float* lut, int nq, int metric, int k, float* out_vals,
                             uint32_t* out_ids, void* scratch, size_t scratch_bytes,
                             void* stream);

/* Screened scan over the scan layout: the SAME result as b200vq_adc_scan_topk_rot (exact fp32
 * scores, (value, id) order), obtained faster for large n. The table entries of four queries are
 * quantised to 6 bits and packed into one 32-bit word, so one shared-memory lookup serves four
 * queries; the integer sum is a rigorous lower bound of the score and only rows that can still
 * reach the top-k (threshold = exact k-th best of a sample: n/8 rows up to 65 536, or n/16 rows up
 * to 8192 for shards of at most 262 144 rows while k n / sample <= 256) are re-evaluated exactly.
 * Queries whose survivor list overflows are re-done by the exact kernel. Falls back to the exact
 * scan for n < 65 536, nq < 4 or k > 64. M in {16, 32, 64, 96}.
 * Scratch: b200vq_adc_scan_screen_scratch_bytes. */
size_t b200vq_adc_scan_screen_scratch_bytes(int64_t n, int M, int nq, int k);
int b200vq_adc_scan_topk_screen(const uint8_t* codes_rot, int64_t n, int M, int ksub,
                                const float* lut, int nq, int metric, int k, float* out_vals,
                                uint32_t* out_ids, void* scratch, size_t scratch_bytes,
                                void* stream);

/* ---------------------------------------------------- exact flat search ---
 * Exact top-k of queries against raw fp32 rows (faiss IndexFlatL2/IndexFlatIP.search:
 * src/haag_vq/benchmarks/exact_search.py:44-51,72-77, search_bench.py:42-56).
 * L2: ||q||^2 + ||x||^2 - 2 q.x (fp32); IP: q.x. xnorm (n) may be NULL for IP.
 * row_scale (n, optional): x rows are multiplied by row_scale[i] before the
 * product (the 1/||x|| pre-scaling of build_scaled_ip_index). */
size_t b200vq_flat_topk_scratch_bytes(int64_t n, int nq, int k);
int b200vq_flat_topk(const float* x, int64_t n, int64_t ldx, int D, const float* xnorm,
                     const float* row_scale, const float* queries, int nq, int64_t ldq,
                     int metric, int k, float* out_vals, uint32_t* out_ids, void* scratch,
                     size_t scratch_bytes, void* stream);

/* Continuation of b200vq_flat_topk: the best k rows AFTER (after_vals[q], after_ids[q]) in its (key, id) order, i.e.
 * ranks r+1 .. r+k when given every query's rank-r entry. Chained calls lift the k <= 128 limit of one launch
 * (nprobe = 200 of the reference's IVF shell, ivf_quantized_index.py:87-90). Scratch as b200vq_flat_topk. */
int b200vq_flat_topk_after(const float* x, int64_t n, int64_t ldx, int D, const float* xnorm,
                           const float* row_scale, const float* queries, int nq, int64_t ldq, int metric,
                           int k, const float* after_vals, const uint32_t* after_ids, float* out_vals,
                           uint32_t* out_ids, void* scratch, size_t scratch_bytes, void* stream);

/* ------------------------------------------------------------ IVF search ---
 * Exact top-k over the reconstructions of the clusters each query probes: the result of the per-query loop of
 * src/haag_vq/methods/search/ivf_quantized_index.py:98-146 (decompress the probed clusters, exact distance, top-k).
 * recon (N, ldr): reconstructions decompress(codes_c) + centroid_c of all rows, sorted by cluster; rnorm (N) their
 * squared norms (L2); gid (N) the rows' global ids; row_off (K + 1) int64 CSR offsets of the clusters;
 * probe (nq, nprobe) uint32 cluster ids per query (ids >= K are skipped). Values as b200vq_flat_topk (squared L2,
 * or the inner product), ids = global ids, (inf, 0xffffffff) where fewer than k rows were probed. The scan is
 * cluster-major (csrc/ivf.cu): a row is read once per 64 probing queries. k <= 128. */
/* Row-wise arg-min (lowest index on ties) of an (n, K) score matrix: the assignment step of the coarse k-means
 * (faiss.Kmeans.train / km.index.search(X, 1), ivf_quantized_index.py:52-57) on scores |c|^2 - 2 x.c produced by
 * b200vq_rotate_tc_ex with a column bias. val (n, optional) receives the minima. */
int b200vq_row_argmin(const float* y, int64_t n, int64_t ld, int K, uint32_t* idx, float* val, void* stream);
size_t b200vq_ivf_search_scratch_bytes(int nq, int nprobe, int K, int k);
int b200vq_ivf_search(const float* recon, int64_t ldr, int D, const float* rnorm, const uint32_t* gid,
                      const int64_t* row_off, int K, const float* queries, int nq, int64_t ldq,
                      const uint32_t* probe, int nprobe, int metric, int k, float* out_vals, uint32_t* out_ids,
                      void* scratch, size_t scratch_bytes, void* stream);

/* Continuation of b200vq_ivf_search: the best k probed rows AFTER (after_vals[q], after_ids[q]) in (distance, id)
 * order, so that any k (the reference's loop takes any k, ivf_quantized_index.py:133-146) is served in pages of at
 * most 128. Scratch as b200vq_ivf_search. */
int b200vq_ivf_search_after(const float* recon, int64_t ldr, int D, const float* rnorm, const uint32_t* gid,
                            const int64_t* row_off, int K, const float* queries, int nq, int64_t ldq,
                            const uint32_t* probe, int nprobe, int metric, int k, const float* after_vals,
                            const uint32_t* after_ids, float* out_vals, uint32_t* out_ids, void* scratch,
                            size_t scratch_bytes, void* stream);

/* The same search on the tensor cores: the screen / candidate-pool machinery of b200vq_flat_topk_tc over the
 * cluster-sorted reconstructions, where a row only counts for the queries whose probe list holds its cluster
 * (rowcluster (N): cluster of every row; the probe lists become one K-bit mask per query). Same values and ids as
 * b200vq_ivf_search. Needs N >= 1024, D % 32 == 0, 16-byte aligned rows (else an error: use b200vq_ivf_search).
 * flags_out (nq): 1 for a query whose candidate lists overflowed - its output row is undefined and the caller redoes
 * it with b200vq_ivf_search. init_vals (nq, optional): an upper bound of every query's k-th best value (e.g. the exact
 * k-th best over its nearest cluster, from b200vq_ivf_search with nprobe = 1): the screen then starts with finite
 * thresholds instead of admitting every probed row until k of them have been seen. */
size_t b200vq_ivf_search_tc_scratch_bytes(int64_t n, int nq, int K, int k);
int b200vq_ivf_search_tc(const float* recon, int64_t n, int64_t ldr, int D, const float* rnorm,
                         const uint32_t* gid, const uint32_t* rowcluster, int K, const float* queries, int nq,
                         int64_t ldq, const uint32_t* probe, int nprobe, int metric, int k, float* out_vals,
                         uint32_t* out_ids, uint32_t* flags_out, const float* init_vals, void* scratch,
                         size_t scratch_bytes, void* stream);

/* Same search, same results (ids and values are those of b200vq_flat_topk's fp32 arithmetic), with the
 * tensor cores as a screen: TF32 tcgen05 products of 128 rows x 256 queries give a rigorous lower
 * bound of every exact key; rows are processed in geometric stages against the thresholds of the
 * exact top-k so far and only the survivors are scored exactly. xnorm (n) = squared norms of the
 * RAW rows is required (it also bounds the screen error). Falls back to b200vq_flat_topk for
 * n < 1024, D % 32 != 0, unaligned rows, or L2 with row_scale. */
size_t b200vq_flat_topk_tc_scratch_bytes(int64_t n, int nq, int k);
int b200vq_flat_topk_tc(const float* x, int64_t n, int64_t ldx, int D, const float* xnorm,
                        const float* row_scale, const float* queries, int nq, int64_t ldq,
                        int metric, int k, float* out_vals, uint32_t* out_ids, void* scratch,
                        size_t scratch_bytes, void* stream);

/* fp32-accurate rotation GEMM on the tensor cores: Y (n x N) = X (n x K) . W^T, W (N x K) row-major.
 * Replaces the sgemm of faiss OPQMatrix.apply / reverse_transform (optimized_product_quantization.py:30-34)
 * and the SAQ engine's segment rotations (external/saq/include/saq/quantizer.h:66-72). Three TF32
 * product passes over round-to-nearest hi/lo operand splits, fp32 accumulation: error of a few 2^-22
 * per product. _prepare splits W once into `wsplit` (rotate_tc_scratch_bytes(N, K) bytes, 16-byte
 * aligned, reusable for every batch); _applies tells whether a shape is covered (K % 32 == 0,
 * ldx % 4 == 0, ldy % 4 == 0); x, y 16-byte aligned. */
int b200vq_rotate_tc_applies(int64_t n, int64_t ldx, int64_t N, int64_t K, int64_t ldy);
size_t b200vq_rotate_tc_scratch_bytes(int64_t N, int64_t K);
int b200vq_rotate_tc_prepare(const float* w, int64_t ldw, int64_t N, int64_t K, void* wsplit,
                             size_t wsplit_bytes, void* stream);
int b200vq_rotate_tc(const float* x, int64_t n, int64_t ldx, const void* wsplit, int64_t N, int64_t K,
                     float* y, int64_t ldy, void* stream);

/* Same GEMM with the element-wise work around it fused in (the centring / scaling the reference does in
 * separate numpy / Eigen passes): Y = row_scale[i] * ((X - x_sub) . W^T) + col_bias.
 * x_sub (K, 16-byte aligned), row_scale (n), col_bias (N) are optional (NULL). transpose_out != 0 stores
 * element (row, col) at y[col * ldy + row] (ldy >= n): the layout b200vq_caq_encode_t consumes
 * (external/saq/include/saq/quantizer.h:66-72 rotates, caq_encoder.h walks the row). Used for the SAQ
 * engine's PCA projection + segment rotations (external/saq/src/ivf_index.cpp:186-243), its inverse
 * (ivf_index.cpp:296-374), and the rotations of ExtendedRaBitQuantizer
 * (src/haag_vq/methods/extended_rabitq.py:140-170,180-198) and RankAwareQuantizer
 * (rank_aware_quantization.py:195,218). */
int b200vq_rotate_tc_ex(const float* x, int64_t n, int64_t ldx, const void* wsplit, int64_t N, int64_t K,
                        const float* x_sub, const float* row_scale, const float* col_bias, int transpose_out,
                        float* y, int64_t ldy, void* stream);

/* Row squared norms (compute_exact_norms, exact_search.py:20-24, before the sqrt). */
int b200vq_row_sqnorms(const float* x, int64_t n, int64_t ldx, int D, float* out, void* stream);

/* Merge P sorted partial lists per query into the final top-k.
 * in_vals (nq, P, k) fp32 ascending-better after sign handling by `metric`,
 * in_ids (nq, P, k) int64 global ids (-1 = empty). Used for the per-shard
 * merge after the NCCL all-gather. out_ids int64. */
int b200vq_topk_merge(const float* in_vals, const int64_t* in_ids, int nq, int P, int k,
                      int metric, float* out_vals, int64_t* out_ids, void* stream);

/* ------------------------------------------------------------------ SQ ---
 * Reference: src/haag_vq/methods/scalar_quantization.py:37-90. */
/* Column-wise min / max; vmin/vmax (D) must be pre-filled with +inf/-inf
 * (so that several calls accumulate over row batches / shards). */
int b200vq_sq_minmax(const float* x, int64_t n, int64_t ldx, int D, float* vmin, float* vmax,
                     void* stream);
/* bits in {4, 8, 16}. codes: (n, D) u8 | (n, ceil(D/2)) u8 (even dim in the high
 * nibble) | (n, D) u16. q = rint((x-min)/(max-min+1e-8f) * (2^bits-1)), fp32,
 * each operation rounded separately like numpy; out-of-range values saturate. */
int b200vq_sq_encode(const float* x, int64_t n, int64_t ldx, int D, const float* vmin,
                     const float* vmax, int bits, void* codes, void* stream);
int b200vq_sq_decode(const void* codes, int64_t n, int D, const float* vmin, const float* vmax,
                     int bits, float* out, int64_t ldo, void* stream);

/* Segmented variable-bit scalar quantiser (the `SAQ(num_bits|total_bits,
 * allowed_bits, n_segments)` CLI contract, src/haag_vq/benchmarks/sweep.py:418-430).
 * dim_bits (D) int32 in [0, 8]: bits of every dimension (constant inside a
 * segment). bit_offset (D+1) int32: exclusive prefix sum of dim_bits. Codes are an
 * MSB-first bit string of bit_offset[D] bits per row, padded to code_bytes. */
int b200vq_saq_encode(const float* x, int64_t n, int64_t ldx, int D, const float* vmin,
                      const float* vmax, const int* dim_bits, const int* bit_offset,
                      int code_bytes, uint8_t* codes, void* stream);
int b200vq_saq_decode(const uint8_t* codes, int64_t n, int D, const float* vmin,
                      const float* vmax, const int* dim_bits, const int* bit_offset,
                      int code_bytes, float* out, int64_t ldo, void* stream);

/* ------------------------------------------------------ Extended RaBitQ ---
 * Reference: src/haag_vq/methods/extended_rabitq.py:125-198 (fp64 numpy).
 * s = ((x-c)/||x-c||) P sqrt(D) is produced by the caller (rotation GEMM);
 * these kernels do the per-row tail. `srot` (n, D) fp64 rotated+scaled rows,
 * `nrm` (n) fp64 residual norms, levels (2^B) fp64 sorted, bounds (2^B-1) fp64
 * mid-points. codes (n, ceil(D*B/8)+8) u8: MSB-first B-bit indices, then
 * float32 nrm, float32 t. */
int b200vq_erabitq_pack(const double* srot, const double* nrm, int64_t n, int D, int B,
                        const double* levels, const double* bounds, uint8_t* codes,
                        void* stream);
/* The same encode with the O(n D^2) rotation on the tensor cores instead of an fp64 GEMM
 * (extended_rabitq.py:134-149): nrm = |x - c| in fp64 (b200vq_resid_norms_f64; x fp32, or fp64 when x_is_f64),
 * y (n, ldy) fp32 = (x - c32) P32 from b200vq_rotate_tc_ex, then b200vq_erabitq_pack_tc takes level indices from
 * s = y sqrt(D) / nrm and recomputes, as exact fp64 dot products sum_k (x_k - c_k) PT[d][k] (PT = P^T, (D, D) fp64
 * row-major), every coordinate within eps = (err_rel nrm + err_abs) sqrt(D) / nrm of a level mid-point, so the
 * indices are those of the fp64 computation whenever |y - (x - c) P| <= err_rel |x - c| + err_abs. t uses the
 * fp32-accurate s (the exact one where recomputed). y == NULL computes every coordinate exactly (SIMT fp64, for
 * shapes the GEMM does not cover). stats (optional, device) += number of recomputed coordinates. */
int b200vq_resid_norms_f64(const void* x, int x_is_f64, int64_t n, int64_t ldx, int D, const double* c,
                           double* nrm, void* stream);
int b200vq_erabitq_pack_tc(const float* y, int64_t ldy, const void* x, int x_is_f64, int64_t ldx,
                           const double* c, const double* PT, const double* nrm, int64_t n, int D, int B,
                           const double* levels, const double* bounds, double err_rel, double err_abs,
                           uint8_t* codes, unsigned long long* stats, void* stream);

/* The same encoder, same codes and trailers, with the exact re-evaluations deferred to a second kernel that visits them
 * coordinate by coordinate (bucket per coordinate d: row d of P^T and the centre sit in shared memory and are read
 * once for all the rows flagged at d, instead of 12 KB of P^T per flagged coordinate); a third kernel writes the
 * rescale factors from the corrected sums. y (the fp32 rotated residuals) and scratch are required. */
size_t b200vq_erabitq_pack_tc2_scratch_bytes(int64_t n, int D);
int b200vq_erabitq_pack_tc2(const float* y, int64_t ldy, const void* x, int x_is_f64, int64_t ldx,
                            const double* c, const double* PT, const double* nrm, int64_t n, int D, int B,
                            const double* levels, const double* bounds, double err_rel, double err_abs,
                            uint8_t* codes, unsigned long long* stats, void* scratch, size_t scratch_bytes,
                            void* stream);
/* a_i = nrm_i t_i / sqrt(D) read from the code trailers: the row scale of the code-domain estimator. */
int b200vq_erabitq_rowscale(const uint8_t* codes, int64_t n, int D, int B, float* a, void* stream);
/* Unpack: shat (n, D) fp64 = levels[idx]/sqrt(D)*t ; nrm_out (n) fp64. The caller
 * applies P^T, the norm and the centroid. */
int b200vq_erabitq_unpack(const uint8_t* codes, int64_t n, int D, int B, const double* levels,
                          double* shat, double* nrm_out, void* stream);
/* Code-domain estimator scan: <s_hat, qrot> per (query, code) with a per-query
 * level LUT, score = ||q-c||^2 + nrm^2 - 2 nrm t/sqrt(D) <levels[idx], qrot>  (L2)
 * with qrot = (q-c) P and qconst = ||q-c||^2; inner product: score = qconst +
 * nrm t/sqrt(D) <levels[idx], qrot> with qrot = q P and qconst = q.c. */
/* Decode for the tensor-core form of the same estimator: x (n, D) fp32 = a_i * levels[idx], a_i = nrm t / sqrt(D),
 * xnorm (n) = ||x_i||^2. The estimator then equals the exact flat search of the rotated queries over x
 * (b200vq_flat_topk_tc with this xnorm): squared L2 for the L2 form, inner product for the IP form. */
int b200vq_erabitq_decode_f32(const uint8_t* codes, int64_t n, int D, int B, const float* levels,
                              float* x, int64_t ldx, float* xnorm, void* stream);
size_t b200vq_erabitq_scan_scratch_bytes(int64_t n, int nq, int k);
int b200vq_erabitq_scan_topk(const uint8_t* codes, int64_t n, int D, int B, const float* levels,
                             const float* qrot, const float* qconst, int nq, int metric, int k,
                             float* out_vals, uint32_t* out_ids, void* scratch,
                             size_t scratch_bytes, void* stream);

/* ------------------------------------- code-domain estimator scan (tcgen05) ---
 * Top-k of an estimator that is linear in per-dimension codes, scored straight from the packed codes (no
 * decoded matrix in HBM). Serves (1) the SAQ engine's accurate estimator, external/saq/include/saq/
 * caq_estimator.h:254-316 + saq_estimator.h:27-129,183-190 (GPU twin src/gpu/gpu_search.cu:37-99), and
 * (2) extended RaBitQ searched through FlatQuantizedIndex (src/haag_vq/methods/extended_rabitq.py:173-198 +
 * methods/search/flat_quantized_index.py:57-68), whose distance to the reconstruction is
 * qc + a^2 sum(l^2) - 2 a <levels[idx], q_rot>.
 *
 * A stored row is nseg segments; segment s has seg_len[s] dimensions (multiple of 32) of seg_bits[s]-bit codes
 * (1..13) in an MSB-first bit string at seg_codes[s] + row * seg_stride[s] (device pointers in HOST arrays;
 * 4-byte aligned, stride % 4 == 0). A code means seg_table[s][code] (device table of 2^bits floats, bits <= 8)
 * or, when seg_table[s] is NULL, code + seg_base[s]; the virtual row is x_i = concat_s seg_rowscale[s][i] *
 * value (seg_rowscale[s] device (n) or NULL = 1). queries (nq, ldq) fp32 hold the matching concatenated query
 * coordinates (sum of seg_len columns, 16-byte aligned, ldq % 4 == 0), qconst (nq, optional) is C_q,
 * row_bias (n) is B_i and row_norm2 (n) = |x_i|^2 from b200vq_est_row_norms (it bounds the screen error).
 *   L2 form: value = C_q + B_i - 2 <x_i, q>, ascending;   IP form: value = C_q + <x_i, q>, descending.
 * Arithmetic of the returned values: per segment one fp32 FMA chain in dimension order, then
 * acc = fma(rowscale, ip_s, acc) in segment order, value = fma(-2, acc, C_q + B_i) or C_q + acc. Rows are
 * screened on the tensor cores (TF32, rigorous bound) and only survivors are scored; overflowing queries are
 * redone exactly, so ids / values never depend on the screen. n < 16 384 is scored exactly without a screen.
 * SAQ engine: seg_base = -(2^b - 1)/2, rowscale = rescale * 2/2^b, B_i = sum_s |o_s|^2, C_q = |q|^2.
 * Extended RaBitQ: one segment, table = levels, rowscale = nrm t / sqrt(D), B_i = |x_i|^2, C_q = |q - c|^2.
 * b200vq_est_row_norms needs 8192 bytes of scratch. */
int b200vq_est_row_norms(int nseg, const int* seg_len, const int* seg_bits, const uint8_t* const* seg_codes,
                         const int64_t* seg_stride, const float* const* seg_rowscale,
                         const float* const* seg_table, const float* seg_base, int64_t n, float* norm2,
                         void* scratch, size_t scratch_bytes, void* stream);
size_t b200vq_est_scan_scratch_bytes(int64_t n, int nq, int k);
int b200vq_est_scan_topk(int nseg, const int* seg_len, const int* seg_bits, const uint8_t* const* seg_codes,
                         const int64_t* seg_stride, const float* const* seg_rowscale,
                         const float* const* seg_table, const float* seg_base, const float* row_norm2,
                         const float* row_bias, int64_t n, const float* queries, int nq, int64_t ldq,
                         const float* qconst, int metric, int k, float* out_vals, uint32_t* out_ids,
                         void* scratch, size_t scratch_bytes, void* stream);

/* ----------------------------------------------------------------- LVQ ---
 * Locally-adaptive vector quantization, single level: replaces the numpy code of
 * src/haag_vq/methods/lvq_quantization.py:44-141 (LVQQuantizer.compress / decompress; registry
 * method "lvq", benchmarks/method_registry_saq.py:49-52) with identical bytes. Row layout:
 * ceil(D * num_bits / 8) bytes of MSB-first packed indices, then lo and delta as fp32.
 * codes (n, ceil(D*num_bits/8) + 8) uint8; mu (D) fp32 global mean; 1 <= num_bits <= 8. */
int b200vq_lvq_encode(const float* x, int64_t n, int64_t ldx, int D, const float* mu, int num_bits,
                      uint8_t* codes, void* stream);
int b200vq_lvq_decode(const uint8_t* codes, int64_t n, int D, const float* mu, int num_bits,
                      float* out, int64_t ldo, void* stream);
/* Column means with numpy's summation order (X.mean(axis=0) of a C-contiguous fp32 matrix: rows
 * added one after another in fp32, then / n): lvq_quantization.py:44-46. mean (D) fp32. */
int b200vq_col_mean_seq(const float* x, int64_t n, int64_t ldx, int D, float* mean, void* stream);

/* --------------------------------------------------------- FFD packing ---
 * Byte-aligned packing of per-dimension codes: src/haag_vq/methods/ffd_packing.py:84-98 (ffd_encode)
 * and :101-117 (ffd_decode). The layout of ffd_layout (:25-81) is passed as small device tables:
 * byte_dims (n_bytes, 8) int32 = dimensions stored in each byte (-1 padded); dim_byte (D) int32 (-1 for
 * zero-width dimensions); dim_shift (D) uint8 = byte_cap - bit_off - width; dim_mask (D) uint8.
 * codes (n, D) uint8 (one code per dimension, < 2^width); packed (n, n_bytes) uint8. */
int b200vq_ffd_pack(const uint8_t* codes, int64_t n, int D, const int* byte_dims, const uint8_t* dim_shift,
                    int n_bytes, uint8_t* out, void* stream);
int b200vq_ffd_unpack(const uint8_t* packed, int64_t n, int n_bytes, const int* dim_byte,
                      const uint8_t* dim_shift, const uint8_t* dim_mask, int D, uint8_t* codes, void* stream);

/* ------------------------------------------ per-dimension codebooks ---
 * RankAwareQuantizer data path (src/haag_vq/methods/rank_aware_quantization.py:229-235 nearest level by
 * searchsorted into codebook mid-points; :252-279 FFD / dense MSB-first packing; :283-320 decode).
 * y / yhat (n, D) fp64 projected coordinates; bits (D) uint8 widths (0..8); bitpos (D) int32 MSB-first bit
 * position of each field in the row (dense: cumulative widths; FFD: 8 * byte_idx + bit_off);
 * bounds = concatenated mid-points (2^b - 1 per dimension) at bound_off (D); cb = concatenated levels
 * (2^b per dimension) at cb_off (D); codes (n, code_size) uint8. */
int b200vq_perdim_encode(const double* y, int64_t n, int64_t ldy, int D, const double* bounds,
                         const int* bound_off, const uint8_t* bits, const int* bitpos, int code_size,
                         uint8_t* codes, void* stream);
int b200vq_perdim_decode(const uint8_t* codes, int64_t n, int code_size, int D, const uint8_t* bits,
                         const int* bitpos, const double* cb, const int* cb_off, double* yhat, int64_t ldy,
                         void* stream);

/* The same data path with the O(n D^2) PCA projection on the tensor cores (rank_aware_quantization.py:246-250 and
 * :318-320 project in float64): y (n, ldy) fp32 = (x - mu32) V32 from b200vq_rotate_tc_ex, nrm (n) = |x - mu| from
 * b200vq_resid_norms_f64; every coordinate within err_rel * nrm + err_abs of one of its mid-points is recomputed as
 * the exact fp64 dot product sum_k (x_k - mu_k) VT[d][k] (VT = V^T, (D, D) fp64), so the indices are the fp64 ones.
 * stats (optional) += recomputed coordinates. _decode_f32 gathers fp32 levels (cb fp32) for the inverse GEMM. */
int b200vq_perdim_encode_tc(const float* y, int64_t ldy, const void* x, int x_is_f64, int64_t ldx,
                            const double* mu, const double* VT, const double* nrm, int64_t n, int D,
                            const double* bounds, const int* bound_off, const uint8_t* bits, const int* bitpos,
                            int code_size, double err_rel, double err_abs, uint8_t* codes,
                            unsigned long long* stats, void* stream);
int b200vq_perdim_decode_f32(const uint8_t* codes, int64_t n, int code_size, int D, const uint8_t* bits,
                             const int* bitpos, const float* cb, const int* cb_off, float* yhat, int64_t ldy,
                             void* stream);

/* Data-fit per-dimension codebooks: 1-D k-means (k-means++ seeding with std::mt19937_64(seed), Lloyd over the sorted
 * column with prefix sums, one empty-cell repair per iteration) = the SAQ engine's build_codebook_lloyd
 * (external/saq/src/preprocessing/codebook_builder.cpp:341-420 with :146-188 and :64-128; LloydOpts defaults
 * max_iters 50, tol 1e-6, seed 0), called per dimension by rank_aware_quantization.py:189-214 (codebook="lloyd") and
 * by quantization_plan.cpp:21-51 (MSE table of the greedy allocator). sorted_cols (D, ld) fp32: row d = the sorted
 * sample of dimension d (n values); ps / psq (D, n + 1) fp64: exclusive prefix sums of the values / their squares.
 * dim_bits (D, device, optional): bit-rate b_d in [0, max_bits] of every dimension (NULL: fixed_bits for all),
 * max_bits <= 13 (the engine's kMaxQuantBit). Outputs: centroids (D, 2^max_bits) fp32, the first sizes[d] entries
 * valid (2^b, or the distinct values when there are fewer than 2^b); costs (D) = MSE on the sample.
 * scratch: b200vq_codebook_lloyd_scratch_bytes(n, D). */
size_t b200vq_codebook_lloyd_scratch_bytes(int64_t n, int D);
int b200vq_codebook_lloyd(const float* sorted_cols, int64_t ld, const double* ps, const double* psq, int64_t n,
                          int D, const int* dim_bits, int fixed_bits, int max_bits, unsigned long long seed,
                          int max_iters, float tol, float* centroids, int* sizes, float* costs, void* scratch,
                          size_t scratch_bytes, void* stream);

/* ADC scan directly on the packed per-dimension layout (dense or FFD) with a top-k: the reference's micro-kernel
 * bench/ffd_speed.cpp:87-127 (score = sum_d LUT[d][code_d], LUT[d][c] = q_d cb_d[c]; :119-123 is the FFD-packed
 * variant), without decoding the rows to HBM. Layout arguments as b200vq_perdim_decode_f32 (cb fp32). queries
 * (nq, ldq) fp32 are the PROJECTED query coordinates; value = C_q + B_i - 2 <y^_i, q> (L2; row_bias = B_i = |y^_i|^2,
 * qconst = C_q = |q|^2; both optional) or C_q + <y^_i, q> (IP). One fp32 FMA chain per (row, query) in dimension
 * order. k <= 128; scratch: b200vq_perdim_scan_scratch_bytes(n, nq, k). */
size_t b200vq_perdim_scan_scratch_bytes(int64_t n, int nq, int k);
int b200vq_perdim_scan_topk(const uint8_t* codes, int64_t n, int code_size, int D, const uint8_t* bits,
                            const int* bitpos, const float* cb, const int* cb_off, const float* row_bias,
                            const float* queries, int nq, int64_t ldq, const float* qconst, int metric, int k,
                            float* out_vals, uint32_t* out_ids, void* scratch, size_t scratch_bytes, void* stream);

/* ------------------------------------------------------- 1-bit RaBitQ ---
 * faiss RaBitQuantizer / IndexRaBitQ semantics (src/haag_vq/methods/
 * rabit_quantization.py:20-29, methods/search/rabitq_index.py:42-70):
 * code = ceil(D/8) sign bytes of (x - centroid), LSB-first, + {float
 * or_minus_c_l2sqr, float dp_multiplier}. */
int b200vq_rabitq1_encode(const float* x, int64_t n, int64_t ldx, int D, const float* centroid,
                          int metric, uint8_t* codes, void* stream);
int b200vq_rabitq1_decode(const uint8_t* codes, int64_t n, int D, const float* centroid,
                          float* out, int64_t ldo, void* stream);
/* Query preparation: residual, qb-bit scalar quantisation, bit-plane transpose.
 * planes (nq, qb, W) uint32 with W = ceil(D/32); qfac (nq, 8) fp32:
 * {c1, c2, c34, qr_to_c_l2sqr, qr_norm_l2sqr, 0, 0, 0}. qb in [1, 8]. */
int b200vq_rabitq1_prepare_queries(const float* queries, int nq, int64_t ldq, int D,
                                   const float* centroid, int qb, uint32_t* planes, float* qfac,
                                   void* stream);
/* AND+popcount estimator scan with top-k. */
size_t b200vq_rabitq1_scan_scratch_bytes(int64_t n, int nq, int k);
int b200vq_rabitq1_scan_topk(const uint8_t* codes, int64_t n, int D, const uint32_t* planes,
                             const float* qfac, int nq, int qb, int metric, int k,
                             float* out_vals, uint32_t* out_ids, void* scratch,
                             size_t scratch_bytes, void* stream);
/* Continuation: the best k rows AFTER (after_vals[q], after_ids[q]) in (estimate, id) order; chained calls serve the
 * k > 128 of faiss IndexRaBitQ.search (src/haag_vq/methods/search/rabitq_index.py:63-75). Same scratch. */
int b200vq_rabitq1_scan_topk_after(const uint8_t* codes, int64_t n, int D, const uint32_t* planes,
                                   const float* qfac, int nq, int qb, int metric, int k,
                                   const float* after_vals, const uint32_t* after_ids, float* out_vals,
                                   uint32_t* out_ids, void* scratch, size_t scratch_bytes, void* stream);

/* Same scan, same results, with the integer part of the estimator on the tensor cores: the
 * (row, query) term sum_d bit_d * qv_d is an exact FP8 (E4M3) GEMM of the sign bits with the quantised
 * query values (tcgen05 kind::f8f6f4, fp32 accumulate), screened against the running top-k thresholds;
 * survivors get the canonical estimate. Falls back to b200vq_rabitq1_scan_topk for n < 32 768,
 * D % 64 != 0, qb > 4 or unaligned codes. */
size_t b200vq_rabitq1_scan_tc_scratch_bytes(int64_t n, int D, int nq, int k);
int b200vq_rabitq1_scan_topk_tc(const uint8_t* codes, int64_t n, int D, const uint32_t* planes,
                                const float* qfac, int nq, int qb, int metric, int k, float* out_vals,
                                uint32_t* out_ids, void* scratch, size_t scratch_bytes, void* stream);

/* ------------------------------------------------ SAQ engine (CAQ) path ---
 * The reference's C++ engine external/saq, reached through SaqEngineAdapter
 * (src/haag_vq/benchmarks/quantizer_adapters.py:113-135): IVF(n, D, K=1).fit / decompress.
 *
 * b200vq_saq_plan_dp: HOST function (no device work). Joint segmentation + bit allocation,
 * SaqDataMaker::dynamic_programming (external/saq/src/quantization_plan.cpp:144-255):
 * variance (d_pad) fp32 on the host, d_pad % 64 == 0; writes up to d_pad/64 (length, bits)
 * pairs and their count.
 *
 * b200vq_caq_encode_t: CAQEncoder::encode_and_fac for every row of one segment
 * (external/saq/include/saq/caq_encoder.h:170-261, code_adjustment :67-140; driven by
 * quantizer.h:84-102). oT is the TRANSPOSED rotated segment, element (j, row) at
 * oT[j*ld + row]; c_rot (len, may be NULL) is subtracted first (quantizer.h:66-72).
 * codesT (len, ldc) uint16 receives the integer codes, also transposed; l2norm / rescale /
 * error are the stored factors (o_l2norm, fac_rescale after rescale_vmx_to1, fac_error),
 * mse = recon_mse, cosv = norm_ip_o_oa. Codes are bit-identical to the CPU order of
 * operations (fp64 state, no contraction); rd_lmt = 6 and adj_eps = 1e-8 are the engine's
 * defaults (config.h:17-18).
 *
 * b200vq_caq_pack_t: codesT -> packed (n, len*bits/8) MSB-first bit string (the HBM-resident
 * form; the engine's short/long AVX-512 layouts, cluster_packer.h:62-142, are not kept).
 *
 * b200vq_caq_decode: per-segment dequantisation of IVF::decompress
 * (external/saq/src/ivf_index.cpp:296-338) for rows ids[0..m) (NULL = 0..m-1):
 * mode 0: out = ((code+0.5)*2/2^bits - 1) * l2norm/|.|  (before the P^T un-rotation);
 * mode 1: out = code as fp32 (operand of the estimator, caq_estimator.h:254-283). */
int b200vq_saq_plan_dp(const float* variance, int d_pad, float avg_bits, int* seg_len, int* seg_bits,
                       int* n_segments);
int b200vq_caq_encode_t(const float* oT, int64_t n, int64_t ld, int len, const float* c_rot, int bits,
                        int rd_lmt, float adj_eps, uint16_t* codesT, int64_t ldc, float* l2norm,
                        float* rescale, float* error, float* mse, float* cosv, void* stream);
/* Every segment of the plan in one launch (blockIdx.y = segment, longest first), so the latency-bound per-row
 * chains of different segments overlap: oT (d_pad, ld) holds all segments stacked (segment s = rows
 * seg_off[s] .. seg_off[s] + seg_len[s]), already centred; codesT (d_pad, ldc) likewise; the five factor outputs
 * are (nseg, n) row-major. seg_* are HOST arrays. Same codes and factors as one b200vq_caq_encode_t per segment. */
int b200vq_caq_encode_t_segments(const float* oT, int64_t n, int64_t ld, int nseg, const int* seg_off,
                                 const int* seg_len, const int* seg_bits, int rd_lmt, float adj_eps,
                                 uint16_t* codesT, int64_t ldc, float* l2norm, float* rescale, float* error,
                                 float* mse, float* cosv, void* stream);
int b200vq_caq_pack_t(const uint16_t* codesT, int64_t n, int64_t ldc, int len, int bits, uint8_t* packed,
                      void* stream);
int b200vq_caq_decode(const uint8_t* packed, int64_t n_total, const int64_t* ids, int64_t m, int len,
                      int bits, const float* l2norm, int mode, float* out, int64_t ldo, void* stream);

/* ------------------------------------------------------------ utilities ---
 * Deterministic synthetic rows for throughput runs (Philox-4x32-10 keyed by
 * (seed, global row, column/4); anisotropic N(0,1)*scale[col], optional row
 * L2-normalisation). Used so 53M x 1024 never crosses PCIe. */
int b200vq_synth_rows(float* out, int64_t n, int64_t ldo, int D, int64_t row0, uint64_t seed,
                      const float* col_scale, int normalize, void* stream);

/* Per-element squared error sum between two (n, D) matrices: out (1 double) +=. */
int b200vq_sq_err_sum(const float* a, int64_t lda, const float* b, int64_t ldb, int64_t n, int D,
                      double* out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200VQ_H */
