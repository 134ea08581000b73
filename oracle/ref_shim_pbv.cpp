/*
 * ref_shim_pbv.cpp -- the reference's "perfect nnz balance with row carry" SpMV kernel (Custom_CSR_PBV) as plain C
 * symbols.  TEST INFRASTRUCTURE ONLY (see oracle/spmm_oracle.c header).  Holds no reference source: it #includes
 * benchmark_code/CPU/AMD/spmv_code_bench/spmv_kernel_csr.cpp from where it lies (-DCUSTOM_VECTOR_PERFECT_NNZ_BALANCE).
 *
 * That translation unit was written against the plug-in interface BEFORE this fork added the dense-column count
 * (spmv(x, y), 6-argument csr_to_format) and no longer matches the fork's spmv_kernel.h, so the
 * shim declares that earlier interface itself and blanks the header's include guard.  Output: oracle/_ref/libpbv_ref.so.
 *
 *   ref_pbv_partition -> CSRArrays ctor, spmv_kernel_csr.cpp:68-80 : thread_i_s / thread_i_e / thread_j_s / thread_j_e
 *   ref_pbv_spmv      -> compute_csr_vector_perfect_nnz_balance, :625-680
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <omp.h>

#include "macros/cpp_defines.h"
#include "spmv_bench_common.h"

#define SPMV_KERNELS_H            /* the fork's spmv_kernel.h becomes empty; the interface the TU implements: */
struct Matrix_Format
{
	char * format_name;
	INT_T m, n, nnz;
	double mem_footprint, csr_mem_footprint;
	virtual void spmv(ValueType * x, ValueType * y) = 0;
	virtual void statistics_start() = 0;
	virtual int statistics_print_data(char * buf, long buf_n) = 0;
	Matrix_Format(long m, long n, long nnz) : m(m), n(n), nnz(nnz)
	{
		csr_mem_footprint = nnz * (sizeof(ValueType) + sizeof(INT_T)) + (m+1) * sizeof(INT_T);
	}
};
struct Matrix_Format * csr_to_format(INT_T * row_ptr, INT_T * col_ind, ValueType * values, long m, long n, long nnz);
int prefetch_distance = 8;        /* extern in the TU, defined by the harness */

#include "spmv_kernel_csr.cpp"    /* reference TU, resolved through -I (AMD/spmv_code_bench) */

static CSRArrays * build(const int32_t *ia, const int32_t *ja, const ValueType *a, long m, long c, long nnz, int P)
{
	/* the format object frees its arrays: hand it private copies, as the harness does */
	INT_T *ia2 = (INT_T *) aligned_alloc(64, (((m + 1 + VECTOR_ELEM_NUM) * sizeof(INT_T) + 63) / 64) * 64);
	INT_T *ja2 = (INT_T *) aligned_alloc(64, (((nnz + VECTOR_ELEM_NUM) * sizeof(INT_T) + 63) / 64) * 64);
	ValueType *a2 = (ValueType *) aligned_alloc(64, (((nnz + VECTOR_ELEM_NUM) * sizeof(ValueType) + 63) / 64) * 64);
	memcpy(ia2, ia, (m + 1) * sizeof(INT_T));
	memcpy(ja2, ja, nnz * sizeof(INT_T));
	memcpy(a2, a, nnz * sizeof(ValueType));
	setenv("USE_PROCESSES", "0", 0);   /* the ctor does atoi(getenv("USE_PROCESSES")) (:61): mandatory, as in run.sh */
	omp_set_dynamic(0);
	omp_set_num_threads(P);       /* the worker count of the partition is omp_get_max_threads() (:43) */
	return (CSRArrays *) csr_to_format(ia2, ja2, a2, m, c, nnz);
}

extern "C" {

/* out[4t..4t+3] = thread_i_s, thread_i_e, thread_j_s, thread_j_e */
void ref_pbv_partition(const int32_t *ia, const int32_t *ja, const ValueType *a, long m, long c, long nnz, int P, int64_t *out)
{
	CSRArrays *csr = build(ia, ja, a, m, c, nnz, P);
	for (int t = 0; t < P; t++) {
		out[4 * t] = thread_i_s[t];
		out[4 * t + 1] = thread_i_e[t];
		out[4 * t + 2] = thread_j_s[t];
		out[4 * t + 3] = thread_j_e[t];
	}
	delete csr;
}

void ref_pbv_spmv(const int32_t *ia, const int32_t *ja, const ValueType *a, long m, long c, long nnz, int P,
                  ValueType *x, ValueType *y)
{
	CSRArrays *csr = build(ia, ja, a, m, c, nnz, P);
	csr->spmv(x, y);
	delete csr;
}

}
