/*
 * ref_shim_merge.cpp -- exposes the reference's CPU merge-path code as C symbols.
 *
 * TEST INFRASTRUCTURE ONLY.  Holds no reference source: #includes the reference
 * translation unit benchmark_code/CPU/ARM/spmv_code_bench/spmv_kernel_merge.cpp
 * (MergePathSearch :214-234, OmpMergeCsrmv :240-302) from /root/reference via
 * the include paths in oracle/Makefile.  Output: oracle/_ref/libmerge_ref.so.
 */
#include <stdint.h>
#include "spmv_kernel_merge.cpp"   /* reference TU (ARM tree) */

extern "C" {

void ref_merge_path_search(long diagonal, int32_t *row_end, long m, long nnz,
                           int64_t *row_out, int64_t *nz_out)
{
	CountingInputIterator nonzero_indices(0);
	int2 coord;
	MergePathSearch((INT_T) diagonal, row_end, nonzero_indices, (INT_T) m, (INT_T) nnz, coord);
	*row_out = coord.x;
	*nz_out = coord.y;
}

/* P+1 coordinates exactly as OmpMergeCsrmv derives them (:262-272). */
void ref_merge_partition(int32_t *ia, long m, long nnz, long P, int64_t *out)
{
	INT_T num_merge_items = m + nnz;
	INT_T items_per_thread = (num_merge_items + P - 1) / P;
	for (long t = 0; t <= P; t++) {
		int d = std::min((INT_T) (items_per_thread * t), num_merge_items);
		ref_merge_path_search(d, ia + 1, m, nnz, &out[2 * t], &out[2 * t + 1]);
	}
}

/* y = A*x (SpMV) by the reference's merge-path kernel with P threads. */
void ref_spmv_merge(int32_t *ia, int32_t *ja, double *a, long m, long nnz,
                    double *x, double *y, int P)
{
	OmpMergeCsrmv(P, ia + 1, ja, a, x, y, (INT_T) m, (INT_T) nnz);
}

}
