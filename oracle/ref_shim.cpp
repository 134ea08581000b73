/*
 * ref_shim.cpp -- exposes the UNMODIFIED reference code as plain C symbols.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle/spmm_oracle.c header).  This file holds
 * no reference source: it #includes the reference translation unit from where
 * it lies under /root/reference (include paths come from oracle/Makefile) and
 * wraps it.  Output: oracle/_ref/libspmm_ref.so.
 *
 *   ref_spmm_csr           -> csr_to_format() + Matrix_Format::spmm()
 *                             (spmv_code_bench/spmm_kernel_csr.cpp:50-66,69-97)
 *   ref_balance_prefix_sums-> loop_partitioner_balance_prefix_sums  (lib/parallel_util.h:141-165)
 *   ref_balance_iterations -> loop_partitioner_balance_iterations   (lib/parallel_util.h:116-126)
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "spmm_kernel_csr.cpp"   /* reference TU, resolved through -I (AMD/spmv_code_bench) */

extern "C" {

/* y = A*x through the reference plug-in interface.  The format object takes
 * ownership of the CSR arrays (its dtor frees them), so hand it private copies
 * exactly as the harness does (spmv_bench.cpp:871-893). */
void ref_spmm_csr(const int32_t *ia, const int32_t *ja, const ValueType *a,
                  long m, long c, long nnz, ValueType *x, ValueType *y, int k)
{
	INT_T *ia2 = (INT_T *) aligned_alloc(64, (((m + 1 + VECTOR_ELEM_NUM) * sizeof(INT_T) + 63) / 64) * 64);
	INT_T *ja2 = (INT_T *) aligned_alloc(64, (((nnz + VECTOR_ELEM_NUM) * sizeof(INT_T) + 63) / 64) * 64);
	ValueType *a2 = (ValueType *) aligned_alloc(64, (((nnz + VECTOR_ELEM_NUM) * sizeof(ValueType) + 63) / 64) * 64);
	memcpy(ia2, ia, (m + 1) * sizeof(INT_T));
	memcpy(ja2, ja, nnz * sizeof(INT_T));
	memcpy(a2, a, nnz * sizeof(ValueType));
	struct Matrix_Format *MF = csr_to_format(ia2, ja2, a2, m, c, nnz, k);
	MF->spmm(x, y, k);
	delete (CSRArrays *) MF;
}

/* Zero-copy variant for timing loops: the caller keeps ownership. */
void ref_spmm_csr_nocopy(int32_t *ia, int32_t *ja, ValueType *a,
                         long m, long c, long nnz, ValueType *x, ValueType *y, int k)
{
	CSRArrays csr(m, c, nnz);
	csr.ia = ia; csr.ja = ja; csr.a = a;
	csr.spmm(x, y, k);
	csr.ia = NULL; csr.ja = NULL; csr.a = NULL;
}

void ref_balance_prefix_sums(int32_t *ia, long m, long nnz, long P, int64_t *out)
{
	for (long t = 0; t < P; t++) {
		long s, e;
		long total = nnz;               /* spmv_kernel_csr.cpp:44,112: the ctor's 'long nnz' */
		loop_partitioner_balance_prefix_sums(P, t, ia, m, total, &s, &e);
		out[t] = s;
		if (t == P - 1)
			out[P] = e;
	}
}

/* Per-worker [start, end) as the macro returns them: out[2t], out[2t+1]. */
void ref_balance_prefix_sums_pairs(int32_t *ia, long m, long nnz, long P, int64_t *out)
{
	for (long t = 0; t < P; t++) {
		long s, e;
		long total = nnz;
		loop_partitioner_balance_prefix_sums(P, t, ia, m, total, &s, &e);
		out[2 * t] = s;
		out[2 * t + 1] = e;
	}
}

void ref_balance_iterations(long start, long end, long P, int64_t *out)
{
	for (long t = 0; t < P; t++) {
		long s, e;
		loop_partitioner_balance_iterations(P, t, start, end, &s, &e);
		out[2 * t] = s;
		out[2 * t + 1] = e;
	}
}

int ref_num_threads(void) { return omp_get_max_threads(); }

}
