/*
 * mkl_stub.cpp -- definitions behind oracle/mkl_stub/mkl.h.  TEST INFRASTRUCTURE ONLY.
 * mkl_?csrmm from MKL's published contract (C := alpha*A*B + beta*C, 'C' indexing = zero-based, row-major dense
 * operands), rows summed left to right with one fused multiply-add per term (the rounding of oracle_spmm_csr);
 * the thread-control calls the pipeline harness makes are no-ops.
 */
#include <vector>
#include "mkl.h"
#include <stdlib.h>

static inline double fma_t(double a, double b, double c) { return __builtin_fma(a, b, c); }
static inline float fma_t(float a, float b, float c) { return __builtin_fmaf(a, b, c); }

template <typename T>
static void csrmm_contract(const char *transa, int m, int n, int k, T alpha, const char *matdescra,
                           const T *val, const int *indx, const int *pntrb, const int *pntre,
                           const T *b, int ldb, T beta, T *c, int ldc)
{
	(void) k;
	if (transa[0] != 'N' && transa[0] != 'n') abort();
	if (matdescra[0] != 'G' || matdescra[3] != 'C') abort();
	/* Row-major loop nest (nonzero outer, dense column inner): every A entry is read once and the B row it selects is
	 * walked contiguously, as a CSR x row-major-dense kernel would do -- not a strawman for the timings quoted next to
	 * it.  Per (row, column) the terms are still added left to right with one fused multiply-add each: the rounding of
	 * the SpMM oracle (oracle_spmm_csr). */
	#pragma omp parallel
	{
		std::vector<T> sum((size_t) (n > 0 ? n : 1));
		#pragma omp for
		for (int i = 0; i < m; i++) {
			for (int d = 0; d < n; d++) sum[d] = 0;
			for (int j = pntrb[i]; j < pntre[i]; j++) {
				const T v = val[j];
				const T *brow = b + (long) indx[j] * ldb;
				for (int d = 0; d < n; d++) sum[d] = fma_t(v, brow[d], sum[d]);
			}
			for (int d = 0; d < n; d++)
				c[(long) i * ldc + d] = (beta == 0) ? alpha * sum[d] : alpha * sum[d] + beta * c[(long) i * ldc + d];
		}
	}
}

extern "C" {

void mkl_dcsrmm(const char *transa, const int *m, const int *n, const int *k, const double *alpha,
                const char *matdescra, const double *val, const int *indx, const int *pntrb,
                const int *pntre, const double *b, const int *ldb, const double *beta, double *c, const int *ldc)
{
	csrmm_contract<double>(transa, *m, *n, *k, *alpha, matdescra, val, indx, pntrb, pntre, b, *ldb, *beta, c, *ldc);
}

void mkl_scsrmm(const char *transa, const int *m, const int *n, const int *k, const float *alpha,
                const char *matdescra, const float *val, const int *indx, const int *pntrb,
                const int *pntre, const float *b, const int *ldb, const float *beta, float *c, const int *ldc)
{
	csrmm_contract<float>(transa, *m, *n, *k, *alpha, matdescra, val, indx, pntrb, pntre, b, *ldb, *beta, c, *ldc);
}

void mkl_set_dynamic(int flag) { (void) flag; }
void mkl_set_num_threads(int n) { (void) n; }
int mkl_get_max_threads(void) { return 1; }

}
