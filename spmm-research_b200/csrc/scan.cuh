// scan.cuh -- in-place exclusive prefix sum over int32 on the device (three short kernels).
// Used by the format conversions: CSR -> blocked CSR (spmm_bcsr.cuh) and COO -> CSR
// (coo_to_csr.cuh; the reference's scan_reduce_concurrent, lib/storage_formats/csr/csr_gen.c:186).
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>

namespace sb200 {

constexpr int SCAN_THREADS = 256;
constexpr int SCAN_PER_THREAD = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_PER_THREAD;

__device__ __forceinline__ int block_exclusive_scan(int v, int* s_warp, int* total)
{
	const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
	int inc = v;
	#pragma unroll
	for (int off = 1; off < 32; off <<= 1) {
		int t = __shfl_up_sync(0xffffffffu, inc, off);
		if (lane >= off) inc += t;
	}
	if (lane == 31) s_warp[wid] = inc;
	__syncthreads();
	if (wid == 0) {
		int w = lane < SCAN_THREADS / 32 ? s_warp[lane] : 0;
		int winc = w;
		#pragma unroll
		for (int off = 1; off < 32; off <<= 1) {
			int t = __shfl_up_sync(0xffffffffu, winc, off);
			if (lane >= off) winc += t;
		}
		if (lane < SCAN_THREADS / 32) s_warp[lane] = winc - w;
		if (lane == 31) s_warp[32] = winc;
	}
	__syncthreads();
	const int res = inc - v + s_warp[wid];
	if (total) *total = s_warp[32];
	__syncthreads();
	return res;
}

__global__ void __launch_bounds__(SCAN_THREADS) scan_tile_sums_kernel(const int32_t* __restrict__ data, int64_t n, int32_t* __restrict__ sums)
{
	__shared__ int s_warp[33];
	const int64_t base = (int64_t) blockIdx.x * SCAN_TILE + (int64_t) threadIdx.x * SCAN_PER_THREAD;
	int v = 0;
	#pragma unroll
	for (int i = 0; i < SCAN_PER_THREAD; i++)
		if (base + i < n) v += data[base + i];
	int total;
	block_exclusive_scan(v, s_warp, &total);
	if (threadIdx.x == 0) sums[blockIdx.x] = total;
}

// one block: exclusive scan of the tile sums
__global__ void __launch_bounds__(SCAN_THREADS) scan_sums_kernel(int32_t* __restrict__ sums, int ntiles)
{
	__shared__ int s_warp[33];
	int carry = 0;
	for (int b = 0; b < ntiles; b += SCAN_THREADS) {
		const int i = b + threadIdx.x;
		const int v = i < ntiles ? sums[i] : 0;
		int total;
		const int ex = block_exclusive_scan(v, s_warp, &total);
		if (i < ntiles) sums[i] = ex + carry;
		carry += total;
	}
}

__global__ void __launch_bounds__(SCAN_THREADS) scan_apply_kernel(int32_t* __restrict__ data, int64_t n, const int32_t* __restrict__ sums)
{
	__shared__ int s_warp[33];
	const int64_t base = (int64_t) blockIdx.x * SCAN_TILE + (int64_t) threadIdx.x * SCAN_PER_THREAD;
	int x[SCAN_PER_THREAD];
	int v = 0;
	#pragma unroll
	for (int i = 0; i < SCAN_PER_THREAD; i++) {
		x[i] = base + i < n ? data[base + i] : 0;
		v += x[i];
	}
	int run = block_exclusive_scan(v, s_warp, nullptr) + sums[blockIdx.x];
	#pragma unroll
	for (int i = 0; i < SCAN_PER_THREAD; i++) {
		if (base + i < n) data[base + i] = run;
		run += x[i];
	}
}

// data[0..n) <- exclusive prefix sums, in place.  tmp: at least ceil(n / SCAN_TILE) int32.
// Returns the number of kernels launched.
static inline int exclusive_scan_i32(int32_t* data, int64_t n, int32_t* tmp, cudaStream_t s)
{
	if (n <= 0) return 0;
	const int ntiles = (int) ((n + SCAN_TILE - 1) / SCAN_TILE);
	scan_tile_sums_kernel<<<ntiles, SCAN_THREADS, 0, s>>>(data, n, tmp);
	scan_sums_kernel<<<1, SCAN_THREADS, 0, s>>>(tmp, ntiles);
	scan_apply_kernel<<<ntiles, SCAN_THREADS, 0, s>>>(data, n, tmp);
	return 3;
}

static inline size_t scan_tmp_elems(int64_t n) { return (size_t) ((n + SCAN_TILE - 1) / SCAN_TILE) + 1; }

} // namespace sb200
