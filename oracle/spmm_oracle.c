/*
 * spmm_oracle.c -- CPU restatement of the reference's SpMM hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under spmm-research_b200/ (the product) may
 * link, import or execute this file; only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs use it, and there only as
 * the checker / the timed CPU baseline.
 *
 * Parity status: PINNED.  Every function here is checked in tests/ against
 *   (1) the golden vectors of SURVEY.md section 8(c) (generator FNV hashes,
 *       partition row starts, merge-path coordinates), and
 *   (2) the reference's own code compiled from /root/reference into
 *       oracle/_ref/libspmm_ref.so (see oracle/Makefile, oracle/ref_shim.cpp), and for the sparse-attention
 *       pipeline (N4) into oracle/_ref/libpipeline_ref.so / libpipeline_ref_f.so (oracle/ref_shim_pipeline.cpp;
 *       the TU's one external dependency, MKL's mkl_?csrmm, is served from its published contract by
 *       oracle/mkl_stub/ -- so SDDMM and softmax are pinned to reference bits, the projections to that contract).
 *
 * All "ref:" citations are relative to /root/reference.
 */
#include <stdint.h>
#include <stddef.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* ---------------------------------------------------------------------------
 * a3: naive CSR SpMM  ("Naive_CSR_CPU")
 * ref: benchmark_code/CPU/AMD/spmv_code_bench/spmm_kernel_csr.cpp:69-97
 *
 *   for n < k:  for i < m (omp for):  sum = 0
 *       for j in [ia[i], ia[i+1]):  sum += a[j] * x[n*c + ja[j]]
 *       y[i*k + n] = sum
 *
 * x = k contiguous vectors of length c (column-major, ld = c);
 * y = row-major m x k (ld = k), fully overwritten.  Left-to-right fp64 sum.
 * ------------------------------------------------------------------------- */
void oracle_spmm_csr(const int32_t *ia, const int32_t *ja, const double *a,
                     long m, long c, const double *x, double *y, int k)
{
	#pragma omp parallel
	{
		for (long n = 0; n < k; n++) {
			#pragma omp for
			for (long i = 0; i < m; i++) {
				double sum = 0;
				for (long j = ia[i]; j < ia[i + 1]; j++)
					sum += a[j] * x[n * c + ja[j]];
				y[i * (long) k + n] = sum;
			}
		}
	}
}

int oracle_num_threads(void)
{
#ifdef _OPENMP
	return omp_get_max_threads();
#else
	return 1;
#endif
}

/* ---------------------------------------------------------------------------
 * binary_search with the reference's "nearest value" semantics.
 * ref: lib/macros/macrolib.h:471-515 (__binary_search, default cmp/dist).
 *
 * Inclusive index range [lo, hi].  T < A[lo] -> lo;  T > A[hi] -> hi;
 * otherwise bisect keeping A[s] <= T <= A[e] (s moves only when T > A[mid]);
 * exact hit on s, then on e; else the nearer of the two, tie -> e.
 * ------------------------------------------------------------------------- */
static long nearest_search(const int32_t *A, long lo, long hi, long T)
{
	long s = lo, e = hi, mid;
	if (T < A[s])
		return s;
	if (T > A[e])
		return e;
	for (;;) {
		mid = (s + e) / 2;
		if (mid == s || mid == e)
			break;
		if (T > A[mid])
			s = mid;
		else
			e = mid;
	}
	if (T == A[s])
		return s;
	if (T == A[e])
		return e;
	long ds = T - A[s]; if (ds < 0) ds = -ds;
	long de = T - A[e]; if (de < 0) de = -de;
	return (ds < de) ? s : e;
}

/* ---------------------------------------------------------------------------
 * a5: nnz-balanced row split.
 * ref: lib/parallel_util.h:141-165 (loop_partitioner_balance_prefix_sums),
 *      call site spmv_kernel_csr.cpp:112 (Sums = ia, N = m, total_sum = nnz).
 *
 * out[t] = first row of worker t, out[P] = m.  Target T_t = (nnz*t)/P in long
 * integer division; the search range is ia[0..m-1] (N-1), never m.
 * Worker t's end is worker t+1's start (same target), last end forced to m.
 * ------------------------------------------------------------------------- */
void oracle_balance_prefix_sums(const int32_t *ia, long m, long nnz, long P, int64_t *out)
{
	out[0] = 0;
	for (long t = 1; t < P; t++) {
		long T = (nnz * t) / P;
		out[t] = (m > 0) ? nearest_search(ia, 0, m - 1, T) : 0;
	}
	out[P] = m;
}

/* ---------------------------------------------------------------------------
 * a5': equal-iterations split of [start, end), increment 1.
 * ref: lib/parallel_util.h:47-93 (loop_partitioner_balance_iterations_base).
 * out[2t], out[2t+1] = [local_start, local_end) of worker t.
 * ------------------------------------------------------------------------- */
void oracle_balance_iterations(long start, long end, long P, int64_t *out)
{
	long len = end - start;
	for (long t = 0; t < P; t++) {
		long s, e;
		if (1 > len) {                /* |incr| > len: worker 0 gets all */
			s = (t == 0) ? start : end;
			e = end;
		} else {
			long per = len / P, rem = len % P;
			if (rem != 0 && t < rem) {
				per += 1;
				rem = 0;
			}
			s = start + per * t + rem;
			e = s + per;
			if (t == P - 1)
				e = end;
		}
		out[2 * t] = s;
		out[2 * t + 1] = e;
	}
}

/* ---------------------------------------------------------------------------
 * a6: "perfect nnz balance with row carry" (Custom_CSR_PBV).
 * ref: benchmark_code/CPU/AMD/spmv_code_bench/spmv_kernel_csr.cpp:68-80 (set-up, CUSTOM_VECTOR_PERFECT_NNZ_BALANCE)
 *      and :625-680 (compute_csr_vector_perfect_nnz_balance).
 *
 * Set-up, per worker t of P: [j_s, j_e) = loop_partitioner_balance_iterations over [0, nnz);
 * i_s = LOWER boundary of binary_search(ia, 0, m, j_s) -- the range is 0..m here (ia has m+1 entries), and the
 * lower boundary is: the index hit exactly (s first, then e), else s with ia[s] < j_s < ia[s+1];
 * i_e[t] = i_s[t+1] + 1, last worker: m.   out[4t..4t+3] = i_s, i_e, j_s, j_e.
 * ------------------------------------------------------------------------- */
static long lower_boundary(const int32_t *A, long lo, long hi, long T)
{
	long s = lo, e = hi, mid;
	if (T < A[s])
		return -1;                   /* macrolib.h:481-486: s = -1 */
	if (T > A[e])
		return e;                    /* :487-492: s = e */
	for (;;) {
		mid = (s + e) / 2;
		if (mid == s || mid == e)
			break;
		if (T > A[mid])
			s = mid;
		else
			e = mid;
	}
	if (T == A[s])
		return s;
	if (T == A[e])
		return e;
	return s;
}

void oracle_nnz_carry_partition(const int32_t *ia, long m, long nnz, long P, int64_t *out)
{
	int64_t *js = (int64_t *) malloc(sizeof(int64_t) * 2 * (size_t) P);
	oracle_balance_iterations(0, nnz, P, js);
	for (long t = 0; t < P; t++) {
		out[4 * t + 2] = js[2 * t];
		out[4 * t + 3] = js[2 * t + 1];
		out[4 * t] = lower_boundary(ia, 0, m, js[2 * t]);
	}
	for (long t = 0; t < P; t++)
		out[4 * t + 1] = (t == P - 1) ? m : out[4 * (t + 1)] + 1;
	free(js);
}

/* SpMV over that partition, workers run one after the other (the reference runs them in parallel, separated by a
 * barrier after the zeroing of the shared rows; the arithmetic per worker and the order of the serial fix-up are the
 * same, so the result is the same bits).  ref: spmv_kernel_csr.cpp:625-680. */
void oracle_spmv_nnz_carry(const int32_t *ia, const int32_t *ja, const double *a, long m, long nnz, long P,
                           const double *x, double *y)
{
	int64_t *part = (int64_t *) malloc(sizeof(int64_t) * 4 * (size_t) P);
	double *v_e = (double *) malloc(sizeof(double) * (size_t) P);
	oracle_nnz_carry_partition(ia, m, nnz, P, part);
	for (long t = 0; t < P; t++)
		if (part[4 * t + 1] - 1 >= 0)
			y[part[4 * t + 1] - 1] = 0;
	for (long t = 0; t < P; t++) {
		long i_s = part[4 * t], i_e = part[4 * t + 1], j_s = part[4 * t + 2], j_e = part[4 * t + 3];
		long i, j = j_s;
		for (i = i_s; i < i_e - 1; i++) {
			long je = ia[i + 1];
			double sum = 0;
			for (; j < je; j++)
				sum += a[j] * x[ja[j]];
			y[i] = sum;
			j = je;
		}
		i = i_e - 1;
		j = ia[i];
		if (j_s > j)
			j = j_s;
		double sum = 0;
		for (; j < j_e; j++)
			sum += a[j] * x[ja[j]];
		v_e[t] = sum;
	}
	for (long t = 0; t < P; t++)
		if (part[4 * t + 1] - 1 < m)
			y[part[4 * t + 1] - 1] += v_e[t];
	free(part);
	free(v_e);
}

/* ---------------------------------------------------------------------------
 * a7: merge-path diagonal search.
 * ref: benchmark_code/CPU/ARM/spmv_code_bench/spmv_kernel_merge.cpp:214-234
 *      (MergePathSearch), list A = row_end_offsets = ia+1 (length m),
 *      list B = counting iterator 0..nnz-1.
 * ------------------------------------------------------------------------- */
void oracle_merge_path_search(long diagonal, const int32_t *row_end, long m, long nnz,
                              int64_t *row_out, int64_t *nz_out)
{
	long x_min = diagonal - nnz; if (x_min < 0) x_min = 0;
	long x_max = (diagonal < m) ? diagonal : m;
	while (x_min < x_max) {
		long p = (x_min + x_max) >> 1;
		if (row_end[p] <= diagonal - p - 1)
			x_min = p + 1;
		else
			x_max = p;
	}
	*row_out = (x_min < m) ? x_min : m;
	*nz_out = diagonal - x_min;
}

/* P+1 coordinates at diagonals min(items*t, m+nnz), items = ceil((m+nnz)/P).
 * ref: spmv_kernel_merge.cpp:262-272.  out[2t] = row, out[2t+1] = nnz index. */
void oracle_merge_partition(const int32_t *ia, long m, long nnz, long P, int64_t *out)
{
	long L = m + nnz;
	long items = (L + P - 1) / P;
	for (long t = 0; t <= P; t++) {
		long d = items * t; if (d > L) d = L;
		oracle_merge_path_search(d, ia + 1, m, nnz, &out[2 * t], &out[2 * t + 1]);
	}
}

/* ---------------------------------------------------------------------------
 * a7 generalised to k columns: merge-path SpMM with k-wide carries.
 * ref: spmv_kernel_merge.cpp:240-302 (OmpMergeCsrmv) is the k=1 case; the
 * k-wide generalisation keeps its structure (whole rows, partial last row,
 * serial carry fix-up) and the oracle's B/C layouts (a3).
 * ------------------------------------------------------------------------- */
void oracle_spmm_merge(const int32_t *ia, const int32_t *ja, const double *a,
                       long m, long c, const double *x, double *y, int k, int P)
{
	int64_t *coord = malloc(sizeof(int64_t) * 2 * (P + 1));
	int64_t *row_carry = malloc(sizeof(int64_t) * P);
	double *val_carry = malloc(sizeof(double) * (size_t) P * k);
	long nnz = ia[m];
	oracle_merge_partition(ia, m, nnz, P, coord);
	const int32_t *row_end = ia + 1;
	#pragma omp parallel for schedule(static)
	for (int t = 0; t < P; t++) {
		long r = coord[2 * t], j = coord[2 * t + 1];
		long r_end = coord[2 * t + 2], j_end = coord[2 * t + 3];
		double *acc = val_carry + (size_t) t * k;
		for (; r < r_end; r++) {
			for (int n = 0; n < k; n++) acc[n] = 0;
			for (; j < row_end[r]; j++)
				for (int n = 0; n < k; n++)
					acc[n] += a[j] * x[(long) n * c + ja[j]];
			for (int n = 0; n < k; n++) y[r * (long) k + n] = acc[n];
		}
		for (int n = 0; n < k; n++) acc[n] = 0;
		for (; j < j_end; j++)
			for (int n = 0; n < k; n++)
				acc[n] += a[j] * x[(long) n * c + ja[j]];
		row_carry[t] = r_end;
	}
	for (int t = 0; t < P - 1; t++)
		if (row_carry[t] < m)
			for (int n = 0; n < k; n++)
				y[row_carry[t] * (long) k + n] += val_carry[(size_t) t * k + n];
	free(coord); free(row_carry); free(val_carry);
}

/* ---------------------------------------------------------------------------
 * FNV-1a 64 over raw bytes: the hash SURVEY.md 8(c) pins the generator with.
 * ------------------------------------------------------------------------- */
uint64_t oracle_fnv1a64(const void *buf, size_t nbytes)
{
	const unsigned char *p = buf;
	uint64_t h = 0xcbf29ce484222325ULL;
	for (size_t i = 0; i < nbytes; i++) {
		h ^= p[i];
		h *= 0x100000001b3ULL;
	}
	return h;
}

/* ---------------------------------------------------------------------------
 * a8: COO -> CSR with per-row column sort.
 * ref: lib/storage_formats/csr/csr_gen.c:162-217 (coo_to_csr) and :82-156
 *      (csr_sort_columns); lib/sort/bucketsort/bucketsort_gen.c:163-205.
 *
 * Reference: histogram of the row ids, prefix sums -> row_ptr, every element
 * takes a free slot of its row (atomic decrement of the row's offset: the
 * slot order inside a row depends on the thread schedule), then each row is
 * sorted by column (quicksort on a permutation, or a stable bucket sort for
 * rows longer than n/5).  With transpose != 0 the roles of R and C (and of m
 * and n) are swapped.  The output is therefore defined up to the order of
 * duplicate (row, col) pairs; this restatement orders a row by (column,
 * position in the input), and with sort_columns == 0 by input position.
 * ------------------------------------------------------------------------- */
typedef struct { int32_t col; int64_t pos; } oracle_ent_t;

static int oracle_ent_cmp(const void *pa, const void *pb)
{
	const oracle_ent_t *a = pa, *b = pb;
	if (a->col != b->col)
		return (a->col < b->col) ? -1 : 1;
	return (a->pos < b->pos) ? -1 : (a->pos > b->pos);
}

void oracle_coo_to_csr(const int32_t *R, const int32_t *C, const double *V, long m, long n, long nnz,
                       int32_t *row_ptr, int32_t *col_idx, double *values, int sort_columns, int transpose)
{
	const int32_t *rows = transpose ? C : R;
	const int32_t *cols = transpose ? R : C;
	long nrows = transpose ? n : m;
	long *cursor = malloc(sizeof(long) * (nrows + 1));
	oracle_ent_t *ent = malloc(sizeof(oracle_ent_t) * (nnz > 0 ? nnz : 1));
	memset(row_ptr, 0, sizeof(int32_t) * (nrows + 1));
	for (long i = 0; i < nnz; i++)
		row_ptr[rows[i] + 1]++;
	for (long r = 0; r < nrows; r++)
		row_ptr[r + 1] += row_ptr[r];
	for (long r = 0; r <= nrows; r++)
		cursor[r] = row_ptr[r];
	for (long i = 0; i < nnz; i++) {                 /* stable scatter */
		long p = cursor[rows[i]]++;
		ent[p].col = sort_columns ? cols[i] : 0;
		ent[p].pos = i;
	}
	if (sort_columns)
		for (long r = 0; r < nrows; r++)
			qsort(ent + row_ptr[r], row_ptr[r + 1] - row_ptr[r], sizeof(oracle_ent_t), oracle_ent_cmp);
	for (long j = 0; j < nnz; j++) {
		col_idx[j] = cols[ent[j].pos];
		if (V != NULL)
			values[j] = V[ent[j].pos];
	}
	free(cursor);
	free(ent);
}

/* ---------------------------------------------------------------------------
 * a10 / N1: CheckAccuracy -- fp128 Kahan gold + the error metrics the harness prints
 * ref: benchmark_code/CPU/AMD/spmv_code_bench/spmv_bench.cpp:121-206 (gold :141-160,
 *      maxDiff :163-188, metrics :189-202) and lib/array_metrics.c (mae, max_ae, mse,
 *      mare->mape, smare->smape, lnQ_error, mlare, gmare: *_process / *_output).
 *
 *   gold[i*k+n]: sum in __float128 with Kahan compensation over val = a[j]*x[n*c+ja[j]]
 *                (csr_a_ref and x_ref are double*, so the product itself is rounded to fp64);
 *   stats[0] = max over entries with gold > eps of |gold - y| / |gold|   (eps = 1e-10, fp64 build)
 *   stats[1..8] = mae, max_ae, mse, mape, smape, lnQ_error, mlare, gmare over
 *                 a = (double) gold, f = y;  stats[9] = (stats[0] <= eps).
 * y_gold (may be NULL) receives (double) gold.
 * ------------------------------------------------------------------------- */
#include <math.h>
#include <float.h>
void oracle_check_accuracy(const int32_t *ia, const int32_t *ja, const double *a, long m, long c,
                           const double *x, const double *y, int k, double *stats, double *y_gold)
{
	const long N = m * (long) k;
	const __float128 eps = 1e-10;
	double *gold = (double *) malloc(sizeof(double) * (N > 0 ? N : 1));
	__float128 max_diff = 0;
	#pragma omp parallel
	{
		__float128 local_max = 0;
		#pragma omp for
		for (long i = 0; i < m; i++) {
			for (long n = 0; n < k; n++) {
				__float128 sum = 0, comp = 0, val, tmp;
				for (long j = ia[i]; j < ia[i + 1]; j++) {
					val = a[j] * x[n * c + ja[j]] - comp;   /* fp64 product (both operands are double*, :152), fp128 from there */
					tmp = sum + val;
					comp = (tmp - sum) - val;
					sum = tmp;
				}
				gold[i * k + n] = (double) sum;
				if (sum > eps) {
					__float128 d = sum - (__float128) y[i * k + n];
					if (d < 0) d = -d;
					d = d / (sum < 0 ? -sum : sum);
					if (d > local_max) local_max = d;
				}
			}
		}
		#pragma omp critical
		if (local_max > max_diff) max_diff = local_max;
	}
	double s_ae = 0, max_ae = 0, s_se = 0, s_are = 0, s_sare = 0, s_lnq = 0;
	for (long i = 0; i < N; i++) {
		const double g = gold[i], f = y[i];
		const double ae = fabs(g - f);
		s_ae += ae;
		if (ae > max_ae) max_ae = ae;
		s_se += (g - f) * (g - f);
		s_are += ae / fmax(fabs(g), DBL_EPSILON);
		s_sare += ae / fmax(fabs(g) + fabs(f), DBL_EPSILON);
		s_lnq += log10(fmax(fabs(f), DBL_EPSILON)) - log10(fmax(fabs(g), DBL_EPSILON));
	}
	stats[0] = (double) max_diff;
	stats[1] = s_ae / N;
	stats[2] = max_ae;
	stats[3] = s_se / N;
	stats[4] = 100.0 * (s_are / N);
	stats[5] = 100.0 * (s_sare / N);
	stats[6] = s_lnq / N;
	{
		long double e = stats[6];
		stats[7] = (double) log10l(fabsl(powl(10, e) - 1));
	}
	stats[8] = pow(10, stats[7]);
	stats[9] = (max_diff <= eps) ? 1.0 : 0.0;
	if (y_gold) memcpy(y_gold, gold, sizeof(double) * N);
	free(gold);
}

/* ---------------------------------------------------------------------------
 * N3: the same naive CSR SpMM in fp32 -- what spmm_kernel_csr.cpp:69-97 computes in an "_f" build
 * (ValueType = float, make.sh:98-108): float sum, float product, left to right.
 * ------------------------------------------------------------------------- */
void oracle_spmm_csr_f32(const int32_t *ia, const int32_t *ja, const float *a,
                         long m, long c, const float *x, float *y, int k)
{
	#pragma omp parallel
	{
		for (long n = 0; n < k; n++) {
			#pragma omp for
			for (long i = 0; i < m; i++) {
				float sum = 0;
				for (long j = ia[i]; j < ia[i + 1]; j++)
					sum += a[j] * x[n * c + ja[j]];
				y[i * (long) k + n] = sum;
			}
		}
	}
}

/* ---------------------------------------------------------------------------
 * N4: sparse-attention pipeline (benchmark_code/CPU/AMD/pipeline_code_bench).
 *
 * oracle_csrmm_rowmajor -- Matrix_Format::spmm of sddmm_taco_naive.cpp:219-260.  The reference calls
 *   mkl_dcsrmm('N', m, n, k, alpha = 1, "GLNC", a, ja, ia, ia+1, x, ldb = n, beta = 0, y, ldc = n).
 *   Intel MKL (oneMKL Sparse BLAS level 3, deprecated mkl_?csrmm) is a third-party dependency absent from
 *   /root/reference; its published contract: C := alpha*A*B + beta*C, and with matdescra[3] = 'C' zero-based
 *   indices and ROW-major B (k x n) and C (m x n).  The summation order inside MKL is unspecified; this
 *   restatement sums each row left to right like the reference's own gold (spmm_helper, sddmm_bench.cpp:
 *   141-162), so agreement with an MKL build is to rounding (1e-12 relative), not bit for bit.
 * ------------------------------------------------------------------------- */
void oracle_csrmm_rowmajor(const int32_t *ia, const int32_t *ja, const double *a,
                           long m, long n, const double *x, double *y)
{
	#pragma omp parallel for
	for (long i = 0; i < m; i++) {
		for (long d = 0; d < n; d++) {
			double sum = 0;
			for (long j = ia[i]; j < ia[i + 1]; j++)
				sum = __builtin_fma(a[j], x[(long) ja[j] * n + d], sum);   /* what gcc -O3 -mfma makes of `sum += a*x` in the scalar gather loop */
			y[i * n + d] = sum;
		}
	}
}

/* oracle_sddmm -- Matrix_Format::sddmm, sddmm_taco_naive.cpp:280-290 (compute_csr: y[] = 0, then compute2) and
 *   compute2, :98-140:
 *       for row i, for nonzero p of row i (column kA = ja[p]):
 *           for d < n:  y[p] += Q[i*n + d] * K[i*n + d]          <- BOTH factors at the output row (pO2, :128)
 *           y[p] *= mask[p]
 *   mode 0 restates exactly that.  mode 1 is the sampled product the variable names describe (pD2 = kA*n + d
 *   is computed at :124 and left unused): K is read at row ja[p].
 *   Rounding: gcc -O3 -mfma vectorises the reference's d loop as rounded products (vmulpd) followed by in-order
 *   adds, i.e. NO fused multiply-add anywhere in the sum (checked against oracle/_ref/libpipeline_ref.so for
 *   widths that are and are not multiples of the vector length); contraction is therefore switched off here. */
__attribute__((optimize("-ffp-contract=off")))
void oracle_sddmm(const int32_t *ia, const int32_t *ja, const double *mask, long m, long n,
                  const double *Q, const double *K, double *y, int mode)
{
	#pragma omp parallel for
	for (long i = 0; i < m; i++) {
		for (long p = ia[i]; p < ia[i + 1]; p++) {
			const double *krow = K + (mode ? (long) ja[p] : i) * n;
			double s = 0.0;
			for (long d = 0; d < n; d++)
				s += Q[i * n + d] * krow[d];
			y[p] = s * mask[p];
		}
	}
}

/* oracle_softmax -- softmax(input, size), sddmm_taco_naive.cpp:191-209 (its call sites are commented out,
 *   :215 and sddmm_bench.cpp:279): one softmax over the whole value array, sequential sums. */
#include <math.h>
void oracle_softmax(double *v, long size)
{
	if (size <= 0)
		return;
	double mx = v[0], sum = 0.0;
	for (long i = 1; i < size; i++)
		if (v[i] > mx)
			mx = v[i];
	for (long i = 0; i < size; i++) {
		v[i] = exp(v[i] - mx);
		sum += v[i];
	}
	for (long i = 0; i < size; i++)
		v[i] /= sum;
}

/* row-wise variant (not in the reference): the same three passes per mask row */
void oracle_softmax_rows(const int32_t *ia, long m, double *v)
{
	for (long i = 0; i < m; i++)
		oracle_softmax(v + ia[i], ia[i + 1] - ia[i]);
}

/* N4 in fp32 (the pipeline bench's "_f" build, ValueType = float: mkl_scsrmm + the same compute2 on floats).
 * Same restatements as above with float sums and float products. */
void oracle_csrmm_rowmajor_f32(const int32_t *ia, const int32_t *ja, const float *a,
                               long m, long n, const float *x, float *y)
{
	#pragma omp parallel for
	for (long i = 0; i < m; i++) {
		for (long d = 0; d < n; d++) {
			float sum = 0;
			for (long j = ia[i]; j < ia[i + 1]; j++)
				sum = __builtin_fmaf(a[j], x[(long) ja[j] * n + d], sum);
			y[i * n + d] = sum;
		}
	}
}

__attribute__((optimize("-ffp-contract=off")))
void oracle_sddmm_f32(const int32_t *ia, const int32_t *ja, const float *mask, long m, long n,
                      const float *Q, const float *K, float *y, int mode)
{
	#pragma omp parallel for
	for (long i = 0; i < m; i++) {
		for (long p = ia[i]; p < ia[i + 1]; p++) {
			const float *krow = K + (mode ? (long) ja[p] : i) * n;
			float s = 0.0f;
			for (long d = 0; d < n; d++)
				s += Q[i * n + d] * krow[d];
			y[p] = s * mask[p];
		}
	}
}
