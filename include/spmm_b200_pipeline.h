/*
 * spmm_b200_pipeline.h -- C ABI of the sparse-attention pipeline.
 *
 * Replaces the plug-in interface of the reference's pipeline bench,
 *     benchmark_code/CPU/AMD/pipeline_code_bench/sddmm_kernel.h:9-31
 *         struct Matrix_Format { spmm(type, m, k, n, ia, ja, a, x, y, threads); sddmm(y, threads); }
 *         csr_to_format(row_ptr, col_ind, values, m, nnz, n)
 * as implemented by pipeline_code_bench/sddmm_taco_naive.cpp and driven by sddmm_bench.cpp:918-936:
 *
 *     K = Wk * x,  Q = Wq * x,  V = Wv * x        spmm('K'|'Q'|'V', ...)      sddmm_taco_naive.cpp:219-260
 *     y[p] = mask[p] * sum_d Q[i][d] * K[i][d]    sddmm(y)                    sddmm_taco_naive.cpp:98-140 (compute2)
 *     y_final = Mask(values y) * V                spmm('final', ...)          sddmm_bench.cpp:934-936
 *
 * Wk, Wq, Wv are CSR (m x k), x is dense k x n row-major, K, Q, V and y_final are dense m x n row-major
 * (mkl_dcsrmm with matdescra[3] = 'C': zero-based indices, row-major dense operands, alpha = 1, beta = 0).
 * Mask is the m x m CSR the format was built from; p runs over its nonzeros, i is the row of p.
 *
 * SDDMM modes.  The reference's compute2 reads both factors at the OUTPUT row (O_vals[pO2] * D_vals[pO2],
 * pO2 = i*n + d; the column index kA is computed and left unused), and the gold its harness checks against
 * does the same (sddmm_bench.cpp:268-281).  SPMM_B200_SDDMM_REFERENCE reproduces that bit for bit.
 * SPMM_B200_SDDMM_ATTENTION is the sampled product the names describe, y[p] = mask[p] * sum_d Q[i][d] * K[ja[p]][d];
 * its fast kernels sum in tree order (agreement with a left-to-right CPU loop: 1e-12 of |Q[i]|.|K[j]|),
 * SPMM_B200_SDDMM_ATTENTION_ORDERED keeps the loop's order at about half the speed.
 *
 * Everything runs on one B200; host pointers in and out unless a function says otherwise.  All functions
 * return SPMM_B200_OK or a negative SPMM_B200_E* code from spmm_b200.h.
 */
#ifndef SPMM_B200_PIPELINE_H
#define SPMM_B200_PIPELINE_H

#include "spmm_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct spmm_b200_pipeline spmm_b200_pipeline_t;

enum {
	SPMM_B200_SDDMM_REFERENCE = 0,  /* both factors at row i (sddmm_taco_naive.cpp:128)           */
	SPMM_B200_SDDMM_ATTENTION = 1,  /* Q at row i, K at row ja[p]; tree-ordered sum for n = 16, 32, 64, 128 */
	SPMM_B200_SDDMM_ATTENTION_ORDERED = 2  /* same, summed left to right like the CPU loop (bit-reproducible against it) */
};

/* csr_to_format(row_ptr, col_ind, values, m, nnz, n) of the pipeline bench (sddmm_taco_naive.cpp:263-277):
 * the format is built from the attention mask (m x m CSR); n is the dense width (NUM_COLS).  The mask is
 * copied to `device`; the caller keeps the host arrays. */
int spmm_b200_pipeline_create(spmm_b200_pipeline_t **p,
                              const int32_t *mask_row_ptr, const int32_t *mask_col_ind, const double *mask_values,
                              int64_t m, int64_t nnz, int n, int device);

/* Matrix_Format::spmm(type, m, k, n, ia, ja, a, x, y) (sddmm_kernel.h:19): y = A * x.
 *   type 'K', 'Q', 'V': A = (ia, ja, a) is a projection weight, m must equal the mask's m; the result also stays
 *       on the device under that name for sddmm / final.  The weight is uploaded on the first call with that
 *       type and reused for as long as the same (ia, ja, a, m, k, nnz) are passed; x is uploaded on every call.
 *   any other type (the harness passes 'final'): a plain y = A * x; nothing is kept.
 * y may be NULL for 'K'/'Q'/'V' (result kept on the device only). */
int spmm_b200_pipeline_spmm(spmm_b200_pipeline_t *p, char type, int64_t m, int64_t k, int n,
                            const int32_t *ia, const int32_t *ja, const double *a,
                            const double *x_host, double *y_host);

/* Matrix_Format::sddmm(y) (sddmm_kernel.h:20): y (nnz doubles, mask order) from the resident Q and K.
 * SPMM_B200_EINVAL when 'Q' or 'K' has not been computed.  y_host may be NULL (result stays on the device). */
int spmm_b200_pipeline_sddmm(spmm_b200_pipeline_t *p, double *y_host, int mode);

/* softmax(y, nnz) of the reference (sddmm_taco_naive.cpp:191-209; the call is commented out at :215 and
 * sddmm_bench.cpp:279): ONE softmax over all nnz values of the resident y.  rowwise != 0 normalises each mask
 * row on its own instead (what attention does; not in the reference). */
int spmm_b200_pipeline_softmax(spmm_b200_pipeline_t *p, double *y_host, int rowwise);

/* spmm('final', m, m, n, mask_ia, mask_ja, y, V, y_final) on the RESIDENT y and V (no upload):
 * y_final = Mask(values y) * V, m x n row-major. */
int spmm_b200_pipeline_final(spmm_b200_pipeline_t *p, double *y_final_host);

/* The whole layer with one upload and one download: the three projections of x_host (k x n), sddmm in `mode`,
 * optional softmax (0 none, 1 global, 2 row-wise), final product -> y_final_host (m x n).  The weights must
 * have been registered by earlier spmm('K'|'Q'|'V') calls or spmm_b200_pipeline_set_weight.  The three
 * projections run side by side.  With mode = SPMM_B200_SDDMM_ATTENTION, softmax = 2 and n = 32, 64 or 128 the
 * SDDMM, the softmax and the final product are ONE kernel (online softmax); y is then never materialised and
 * spmm_b200_pipeline_device_ptr(p, 'y') returns NULL afterwards.  SPMM_B200_ATTENTION_FUSED=0 forces the
 * step-by-step kernels. */
int spmm_b200_pipeline_set_weight(spmm_b200_pipeline_t *p, char type, int64_t m, int64_t k, int64_t nnz,
                                  const int32_t *ia, const int32_t *ja, const double *a);
/* Residency hazard: set_weight / spmm('K'|'Q'|'V') keep a weight on the device and take a later call with the SAME host
 * pointers and sizes for the same weight (after comparing a cheap fingerprint: sizes + ~1024 strided samples of the
 * arrays).  A caller that rewrites a weight in place should call spmm_b200_pipeline_invalidate(p, type) ('K', 'Q',
 * 'V' or '*' for all) before handing it over again -- the fingerprint catches a refill, not a single changed entry. */
int spmm_b200_pipeline_invalidate(spmm_b200_pipeline_t *p, char type);
/* Tuning knobs (SPMM_B200_SDDMM_GROUP, _SDDMM_STAGED, _SDDMM_TUNE, _TICKET_REV, _PIPELINE_FORK, _ATTENTION_FUSED) are
 * read once, at creation; this changes one afterwards ("SPMM_B200_" prefix optional). */
int spmm_b200_pipeline_set_option(spmm_b200_pipeline_t *p, const char *name, int value);
int spmm_b200_pipeline_attention(spmm_b200_pipeline_t *p, const double *x_host, double *y_final_host,
                                 int mode, int softmax);

/* Device-resident results of the last calls (NULL when absent): 'K', 'Q', 'V' (m x n), 'y' (nnz), 'f' (m x n). */
const double *spmm_b200_pipeline_device_ptr(const spmm_b200_pipeline_t *p, char which);
/* CUDA-event time of the last call's kernels, and kernels launched so far. */
double  spmm_b200_pipeline_last_kernel_ms(const spmm_b200_pipeline_t *p);
int64_t spmm_b200_pipeline_launch_count(const spmm_b200_pipeline_t *p);
void    spmm_b200_pipeline_destroy(spmm_b200_pipeline_t *p);

/* ---- the same step-by-step interface for the pipeline bench's "_f" build (ValueType = float, mkl_scsrmm +
 * compute2 on floats): fp32 SpMM handles and ordered fp32 SDDMM kernels (both attention modes keep the loop's
 * order); no softmax and no fused path in fp32 yet. */
typedef struct spmm_b200_pipeline_f32 spmm_b200_pipeline_f32_t;
int spmm_b200_pipeline_create_f32(spmm_b200_pipeline_f32_t **p,
                                  const int32_t *mask_row_ptr, const int32_t *mask_col_ind, const float *mask_values,
                                  int64_t m, int64_t nnz, int n, int device);
int spmm_b200_pipeline_spmm_f32(spmm_b200_pipeline_f32_t *p, char type, int64_t m, int64_t k, int n,
                                const int32_t *ia, const int32_t *ja, const float *a,
                                const float *x_host, float *y_host);
int spmm_b200_pipeline_sddmm_f32(spmm_b200_pipeline_f32_t *p, float *y_host, int mode);
/* softmax of y in place, fp32 (running sums in double): rowwise != 0 = per mask row, 0 = over all nonzeros (the
 * reference's commented-out host softmax(y)); y_host (optional) receives the nnz values */
int spmm_b200_pipeline_softmax_f32(spmm_b200_pipeline_f32_t *p, float *y_host, int rowwise);
int spmm_b200_pipeline_final_f32(spmm_b200_pipeline_f32_t *p, float *y_final_host);
int64_t spmm_b200_pipeline_launch_count_f32(const spmm_b200_pipeline_f32_t *p);
void    spmm_b200_pipeline_destroy_f32(spmm_b200_pipeline_f32_t *p);

#ifdef __cplusplus
}
#endif
#endif /* SPMM_B200_PIPELINE_H */
