// partition.cuh -- bit-exact device versions of the reference's work partitioners.
//
// These produce the same integers as the reference's host code, so that CTA work
// assignment and GPU shard boundaries are reproducible against it:
//   nnz-balanced row split : lib/parallel_util.h:141-165 + lib/macros/macrolib.h:471-515
//   equal-rows split       : lib/parallel_util.h:47-93
//   merge-path search      : benchmark_code/CPU/ARM/spmv_code_bench/spmv_kernel_merge.cpp:214-234
#pragma once
#include <stdint.h>

namespace sb200 {

// "Nearest value" search over the inclusive index range [lo, hi] of a
// non-decreasing array: below A[lo] -> lo, above A[hi] -> hi, exact hit -> the
// lowest matching index reachable by the bisection, else the closer neighbour
// with ties going to the upper one.
__host__ __device__ inline int64_t nearest_index(const int32_t* __restrict__ A, int64_t lo, int64_t hi, int64_t target)
{
	int64_t s = lo, e = hi;
	if (target < (int64_t) A[s]) return s;
	if (target > (int64_t) A[e]) return e;
	while (true) {
		int64_t mid = (s + e) / 2;
		if (mid == s || mid == e) break;
		if (target > (int64_t) A[mid]) s = mid; else e = mid;
	}
	int64_t vs = A[s], ve = A[e];
	if (target == vs) return s;
	if (target == ve) return e;
	int64_t ds = target - vs; if (ds < 0) ds = -ds;
	int64_t de = target - ve; if (de < 0) de = -de;
	return (ds < de) ? s : e;
}

// Row where worker t of P starts under the nnz-balanced split.
__host__ __device__ inline int64_t nnz_split_start(const int32_t* __restrict__ ia, int64_t m, int64_t nnz, int64_t P, int64_t t)
{
	if (t <= 0) return 0;
	if (t >= P) return m;
	if (m <= 0) return 0;
	int64_t target = (nnz * t) / P;
	return nearest_index(ia, 0, m - 1, target);
}

// Lower boundary of the reference's binary_search (lib/macros/macrolib.h:471-515) over the inclusive index range
// [lo, hi] of a non-decreasing array: the index s with A[s] < target < A[s+1], or -- on an exact hit -- the index
// the bisection ends on (the lowest i in (lo, hi] with A[i] == target, or lo itself when A[lo] == target); -1 when the
// target is below A[lo], hi when it is above A[hi].
__host__ __device__ inline int64_t lower_boundary(const int32_t* __restrict__ A, int64_t lo, int64_t hi, int64_t target)
{
	int64_t s = lo, e = hi;
	if (target < (int64_t) A[s]) return -1;
	if (target > (int64_t) A[e]) return e;
	while (true) {
		int64_t mid = (s + e) / 2;
		if (mid == s || mid == e) break;
		if (target > (int64_t) A[mid]) s = mid; else e = mid;
	}
	if (target == (int64_t) A[s]) return s;
	if (target == (int64_t) A[e]) return e;
	return s;
}

// [start, end) of worker t of P over the iteration space [start, end), step 1.
__host__ __device__ inline void equal_split(int64_t start, int64_t end, int64_t P, int64_t t, int64_t* s_out, int64_t* e_out)
{
	int64_t len = end - start;
	if (len < 1) {                     // worker 0 owns the (empty) range
		*s_out = (t == 0) ? start : end;
		*e_out = end;
		return;
	}
	int64_t per = len / P, rem = len % P;
	if (rem != 0 && t < rem) { per += 1; rem = 0; }
	int64_t s = start + per * t + rem;
	int64_t e = s + per;
	if (t == P - 1) e = end;
	*s_out = s;
	*e_out = e;
}

// Merge-path coordinate of `diagonal` between list A = row end offsets
// (row_end[i] = ia[i+1], length m) and list B = 0..nnz-1.
__host__ __device__ inline void merge_path_search(int64_t diagonal, const int32_t* __restrict__ row_end, int64_t m, int64_t nnz,
                                                  int64_t* row_out, int64_t* nz_out)
{
	int64_t x_min = diagonal - nnz; if (x_min < 0) x_min = 0;
	int64_t x_max = diagonal < m ? diagonal : m;
	while (x_min < x_max) {
		int64_t p = (x_min + x_max) >> 1;
		if ((int64_t) row_end[p] <= diagonal - p - 1) x_min = p + 1; else x_max = p;
	}
	*row_out = x_min < m ? x_min : m;
	*nz_out = diagonal - x_min;
}

// "Perfect nnz balance with row carry" (spmv_kernel_csr.cpp:68-80, CUSTOM_VECTOR_PERFECT_NNZ_BALANCE): worker t
// owns the nonzeros [j_s, j_e) of an equal split of [0, nnz); its first row i_s is the lower boundary of the search
// for j_s in ia[0..m] (range 0..m here, not m-1); its row range ends one past the next worker's first row (that row is
// shared: the worker's partial sum of it is the carry), the last worker's at m.  out4 = {i_s, i_e, j_s, j_e}.
__host__ __device__ inline void nnz_carry_split(const int32_t* __restrict__ ia, int64_t m, int64_t nnz, int64_t P, int64_t t, int64_t* out4)
{
	int64_t js, je;
	equal_split(0, nnz, P, t, &js, &je);
	out4[0] = lower_boundary(ia, 0, m, js);
	if (t == P - 1) {
		out4[1] = m;
	} else {
		int64_t js2, je2;
		equal_split(0, nnz, P, t + 1, &js2, &je2);
		out4[1] = lower_boundary(ia, 0, m, js2) + 1;
	}
	out4[2] = js;
	out4[3] = je;
}

__global__ void partition_nnz_carry_kernel(const int32_t* __restrict__ ia, int64_t m, int64_t nnz, int64_t P, int64_t* __restrict__ out)
{
	int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= P) return;
	nnz_carry_split(ia, m, nnz, P, t, out + 4 * t);
}

__global__ void partition_nnz_kernel(const int32_t* __restrict__ ia, int64_t m, int64_t nnz, int64_t P, int64_t* __restrict__ out)
{
	int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (t > P) return;
	out[t] = nnz_split_start(ia, m, nnz, P, t);
}

__global__ void partition_rows_kernel(int64_t start, int64_t end, int64_t P, int64_t* __restrict__ out)
{
	int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= P) return;
	equal_split(start, end, P, t, &out[2 * t], &out[2 * t + 1]);
}

// P+1 coordinates at diagonals min(t * ceil((m+nnz)/P), m+nnz).
__global__ void partition_merge_kernel(const int32_t* __restrict__ ia, int64_t m, int64_t nnz, int64_t P, int64_t* __restrict__ out)
{
	int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (t > P) return;
	int64_t L = m + nnz;
	int64_t items = (L + P - 1) / P;
	int64_t d = items * t; if (d > L) d = L;
	merge_path_search(d, ia + 1, m, nnz, &out[2 * t], &out[2 * t + 1]);
}

// int32 flavours used for CTA work lists inside the SpMM kernels.
__global__ void cta_rows_nnz_kernel(const int32_t* __restrict__ ia, int m, int64_t nnz, int P, int32_t* __restrict__ out)
{
	int t = blockIdx.x * blockDim.x + threadIdx.x;
	if (t > P) return;
	out[t] = (int32_t) nnz_split_start(ia, m, nnz, P, t);
}

// Tile t starts at diagonal t*items (fixed items per tile, last tile short).
__global__ void tile_merge_kernel(const int32_t* __restrict__ ia, int m, int64_t nnz, int64_t items, int ntiles, int2* __restrict__ out)
{
	int t = blockIdx.x * blockDim.x + threadIdx.x;
	if (t > ntiles) return;
	int64_t L = (int64_t) m + nnz;
	int64_t d = items * t; if (d > L) d = L;
	int64_t r, j;
	merge_path_search(d, ia + 1, m, nnz, &r, &j);
	out[t] = make_int2((int) r, (int) j);
}

} // namespace sb200
