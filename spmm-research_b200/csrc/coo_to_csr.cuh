// coo_to_csr.cuh -- COO -> CSR on the device.
//
// Reference: coo_to_csr() + csr_sort_columns()  (lib/storage_formats/csr/csr_gen.c:162-217,82-156)
// with bucketsort() (lib/sort/bucketsort/bucketsort_gen.c:163-205):
//   histogram of row ids -> prefix sums = row_ptr -> every element takes a slot of its row
//   (atomic, non-stable, thread-order dependent) -> col_idx/values scattered -> each row sorted
//   by column.
// The result is defined by the reference only up to the order of duplicate (row, col) pairs.
// Here the order inside a row is made total: by (column, position in the input), so the output
// is deterministic, equals the reference's whenever the input has no duplicates (every matrix
// the harness feeds it, spmv_bench.cpp:823), and with sort_columns = 0 is the STABLE order.
//
// Device plan:  (1) histogram with integer atomics, (2) exclusive scan (scan.cuh), (3) scatter
// of 64-bit keys (column << 32 | input position) to an arbitrary free slot of the row, (4) a
// per-row bitonic sort of the keys -- a warp per short row and a CTA per long row in shared
// memory, a CTA in global memory for rows beyond 4096 -- which erases the scatter order,
// (5) gather col_idx / values through the sorted keys.
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>
#include "scan.cuh"

namespace sb200 {

__global__ void coo_histogram_kernel(const int32_t* __restrict__ keys, int64_t nnz, int32_t* __restrict__ counts)
{
	for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < nnz; i += (int64_t) gridDim.x * blockDim.x)
		atomicAdd(counts + keys[i], 1);
}

// cursor[r] starts at row_ptr[r]; element i lands at cursor[row]++ (any order: step 4 sorts).
__global__ void coo_scatter_kernel(const int32_t* __restrict__ rows, const int32_t* __restrict__ cols, int64_t nnz,
                                   int32_t* __restrict__ cursor, unsigned long long* __restrict__ keys, int sort_columns)
{
	for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < nnz; i += (int64_t) gridDim.x * blockDim.x) {
		const int pos = atomicAdd(cursor + rows[i], 1);
		const unsigned long long hi = sort_columns ? (unsigned long long) (uint32_t) cols[i] : 0ull;
		keys[pos] = (hi << 32) | (unsigned long long) (uint32_t) i;
	}
}

// Bitonic network in its "all comparisons ascending" form (flip, then disperse), which
// tolerates a length that is not a power of two: a pair whose upper index is past the end is
// skipped (as if padded with +inf).  `nthreads` threads cooperate; SYNC separates sub-steps.
template <typename SyncFn>
__device__ __forceinline__ void bitonic_sort_u64(unsigned long long* v, int len, int tid, int nthreads, SyncFn sync)
{
	if (len < 2) return;
	int n2 = 1;
	while (n2 < len) n2 <<= 1;
	for (int k = 2; k <= n2; k <<= 1) {
		for (int i = tid; i < n2 / 2; i += nthreads) {              // flip: partner = i ^ (k-1) inside blocks of k
			const int blk = i / (k / 2), off = i % (k / 2);
			const int lo = blk * k + off, hi = blk * k + (k - 1 - off);
			if (hi < len) {
				const unsigned long long a = v[lo], b = v[hi];
				if (a > b) { v[lo] = b; v[hi] = a; }
			}
		}
		sync();
		for (int j = k >> 2; j > 0; j >>= 1) {                       // disperse: partner = i ^ j
			for (int i = tid; i < n2 / 2; i += nthreads) {
				const int lo = (i / j) * 2 * j + (i % j), hi = lo + j;
				if (hi < len) {
					const unsigned long long a = v[lo], b = v[hi];
					if (a > b) { v[lo] = b; v[hi] = a; }
				}
			}
			sync();
		}
	}
}

constexpr int COO_WARP_MAX = 128;     // keys a warp sorts in its shared-memory slice
constexpr int COO_CTA_MAX = 4096;     // keys a CTA sorts in shared memory

// One warp per row for rows of <= COO_WARP_MAX entries (others skipped).
__global__ void __launch_bounds__(256)
coo_sort_rows_warp_kernel(const int32_t* __restrict__ row_ptr, int m, unsigned long long* __restrict__ keys)
{
	__shared__ unsigned long long s[8][COO_WARP_MAX];
	const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
	const int row = blockIdx.x * 8 + w;
	if (row >= m) return;
	const int js = row_ptr[row], len = row_ptr[row + 1] - js;
	if (len < 2 || len > COO_WARP_MAX) return;
	for (int i = lane; i < len; i += 32) s[w][i] = keys[js + i];
	__syncwarp();
	bitonic_sort_u64(s[w], len, lane, 32, [] { __syncwarp(); });
	for (int i = lane; i < len; i += 32) keys[js + i] = s[w][i];
}

// One CTA per listed long row: shared memory up to COO_CTA_MAX keys, else in place in global memory.
__global__ void __launch_bounds__(256)
coo_sort_rows_cta_kernel(const int32_t* __restrict__ row_ptr, const int32_t* __restrict__ long_rows, unsigned long long* __restrict__ keys)
{
	__shared__ unsigned long long s[COO_CTA_MAX];
	const int row = long_rows[blockIdx.x];
	const int js = row_ptr[row], len = row_ptr[row + 1] - js;
	if (len <= COO_CTA_MAX) {
		for (int i = threadIdx.x; i < len; i += 256) s[i] = keys[js + i];
		__syncthreads();
		bitonic_sort_u64(s, len, threadIdx.x, 256, [] { __syncthreads(); });
		for (int i = threadIdx.x; i < len; i += 256) keys[js + i] = s[i];
	} else {
		bitonic_sort_u64(keys + js, len, threadIdx.x, 256, [] { __syncthreads(); });
	}
}

__global__ void coo_list_long_rows_kernel(const int32_t* __restrict__ row_ptr, int m, int32_t* __restrict__ long_rows, int32_t* __restrict__ n_long)
{
	const int row = blockIdx.x * blockDim.x + threadIdx.x;
	if (row >= m) return;
	if (row_ptr[row + 1] - row_ptr[row] > COO_WARP_MAX) long_rows[atomicAdd(n_long, 1)] = row;
}

__global__ void coo_gather_kernel(const unsigned long long* __restrict__ keys, int64_t nnz, const int32_t* __restrict__ cols,
                                  const double* __restrict__ vals, int32_t* __restrict__ col_idx, double* __restrict__ values)
{
	for (int64_t j = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; j < nnz; j += (int64_t) gridDim.x * blockDim.x) {
		const uint32_t src = (uint32_t) (keys[j] & 0xffffffffull);
		col_idx[j] = cols[src];
		if (vals) values[j] = vals[src];
	}
}

} // namespace sb200
