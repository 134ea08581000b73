/*
 * ref_shim_pipeline.cpp -- exposes the UNMODIFIED pipeline TU of the reference
 * (benchmark_code/CPU/AMD/pipeline_code_bench/sddmm_taco_naive.cpp) as plain C symbols.
 *
 * TEST INFRASTRUCTURE ONLY.  This file holds no reference source: it #includes the reference translation
 * unit from where it lies under /root/reference (include paths from oracle/Makefile) and wraps its
 * csr_to_format / Matrix_Format::spmm / Matrix_Format::sddmm / softmax.  The TU's only external dependency,
 * mkl_?csrmm, is defined below from MKL's published contract (oracle/mkl_stub/mkl.h).
 * Output: oracle/_ref/libpipeline_ref.so.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "sddmm_taco_naive.cpp"   /* reference TU, resolved through -I (AMD/pipeline_code_bench) */

#include "mkl_stub.cpp"             /* mkl_?csrmm stand-ins (oracle/mkl_stub) */

extern "C" {

/* The harness's sequence (sddmm_bench.cpp:918-936) through the reference's own objects:
 *   MF = csr_to_format(mask...);  MF->spmm('K'), ('Q'), ('V');  MF->sddmm(y);  [softmax];  MF->spmm('final', mask, y, V).
 * Every output array is the caller's. */
void ref_pipeline(int32_t *mask_ia, int32_t *mask_ja, ValueType *mask_a, long m, long mask_nnz, int n,
                  int32_t *k_ia, int32_t *k_ja, ValueType *k_a,
                  int32_t *q_ia, int32_t *q_ja, ValueType *q_a,
                  int32_t *v_ia, int32_t *v_ja, ValueType *v_a, long kdim,
                  ValueType *x, ValueType *K, ValueType *Q, ValueType *V, ValueType *y, ValueType *y_final,
                  int do_softmax)
{
	struct Matrix_Format *MF = csr_to_format(mask_ia, mask_ja, mask_a, m, mask_nnz, n);
	MF->spmm('K', m, kdim, n, k_ia, k_ja, k_a, x, K, 1);
	MF->spmm('Q', m, kdim, n, q_ia, q_ja, q_a, x, Q, 1);
	MF->spmm('V', m, kdim, n, v_ia, v_ja, v_a, x, V, 1);
	MF->sddmm(y, 1);
	if (do_softmax)
		softmax(y, (int) mask_nnz);
	MF->spmm('f', m, m, n, mask_ia, mask_ja, y, V, y_final, 1);
}

/* sddmm alone on caller-supplied Q and K */
void ref_sddmm(int32_t *mask_ia, int32_t *mask_ja, ValueType *mask_a, long m, long mask_nnz, int n,
               ValueType *Q, ValueType *K, ValueType *y)
{
	struct CSRTensors *MF = (struct CSRTensors *) csr_to_format(mask_ia, mask_ja, mask_a, m, mask_nnz, n);
	MF->K->vals = K;
	MF->Q->vals = Q;
	MF->sddmm(y, 1);
}

void ref_softmax(ValueType *v, int size)
{
	softmax(v, size);
}

}
