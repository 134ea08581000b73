/*
 * mkl.h -- stand-in for the Intel MKL header, TEST INFRASTRUCTURE ONLY.
 *
 * The reference's pipeline TU (pipeline_code_bench/sddmm_taco_naive.cpp:7-10) includes <mkl.h> and calls
 * mkl_dcsrmm / mkl_scsrmm for its projections.  MKL is not in this image, so oracle/Makefile puts this
 * directory on the include path when it compiles that TU into oracle/_ref/libpipeline_ref.so; the two
 * functions are defined in oracle/ref_shim_pipeline.cpp from MKL's published contract
 * (C := alpha*op(A)*B + beta*C; matdescra[3] = 'C': zero-based indices, row-major dense operands).
 * The reference code that is being pinned through this build is its SDDMM (compute2), which is its own.
 */
#ifndef ORACLE_MKL_STUB_H
#define ORACLE_MKL_STUB_H
#ifdef __cplusplus
extern "C" {
#endif
#ifndef MKL_INT
#define MKL_INT int
#endif
void mkl_dcsrmm(const char *transa, const MKL_INT *m, const MKL_INT *n, const MKL_INT *k, const double *alpha,
                const char *matdescra, const double *val, const MKL_INT *indx, const MKL_INT *pntrb,
                const MKL_INT *pntre, const double *b, const MKL_INT *ldb, const double *beta, double *c,
                const MKL_INT *ldc);
void mkl_scsrmm(const char *transa, const MKL_INT *m, const MKL_INT *n, const MKL_INT *k, const float *alpha,
                const char *matdescra, const float *val, const MKL_INT *indx, const MKL_INT *pntrb,
                const MKL_INT *pntre, const float *b, const MKL_INT *ldb, const float *beta, float *c,
                const MKL_INT *ldc);
void mkl_set_dynamic(int flag);
void mkl_set_num_threads(int n);
int mkl_get_max_threads(void);
#ifdef __cplusplus
}
#endif
#endif
