// spmm_merge.cuh -- CSR SpMM, merge-path variant (north_star kernel (2), skewed row lengths).
//
// The reference's merge-path code is an SpMV (k = 1):
//   benchmark_code/CPU/ARM/spmv_code_bench/spmv_kernel_merge.cpp:214-234 (MergePathSearch)
//   and :240-302 (OmpMergeCsrmv: whole rows, one partial row, serial carry fix-up).
// This file generalises it to k dense columns with k-wide carries and two levels of
// decomposition:
//   level 1  the merge path (m row-end items + nnz nonzero items) is cut into tiles of
//            ITEMS = NG*IPG items, one CTA per tile; tile start coordinates come from
//            partition.cuh:tile_merge_kernel (same search as the reference, bit-exact);
//   level 2  inside a CTA the tile's row ends and (col, val) pairs are staged in shared
//            memory with coalesced loads, and each group of G lanes (a lane owns V of the
//            k = G*V columns, as in spmm_rowsplit.cuh) takes IPG consecutive items found by
//            the same search on the staged copy.
// A group stores the rows that END inside its range (C[row] = sum, the reference's
// "y[row] = running_total") and leaves the nonzeros after its last row end as a carry
// (row, k values).  Carries whose row ends inside the tile are added after a CTA barrier;
// the rest leave the CTA as the tile's carry and are added by spmm_merge_fixup_kernel.
// Both fix-ups walk runs of equal carry rows in tile/group order, so the result is
// deterministic (no floating-point atomics).
#pragma once
#include <stdint.h>
#include <limits.h>
#include <cuda_runtime.h>
#include "spmm_rowsplit.cuh"

namespace sb200 {

// A lane's V columns of one dense row, for double and float operands alike: 16-byte pieces where V allows
// (piece i of the group covers the contiguous bytes [i*G*16, (i+1)*G*16) of the row), scalar loads otherwise.
template <typename T, int V> struct FragT { T v[V]; };

template <typename T, int G, int V>
__device__ __forceinline__ FragT<T, V> load_row_t(const T* __restrict__ B, int col, int lig, int ldb)
{
	constexpr int E = 16 / sizeof(T);            // elements per 16-byte piece
	FragT<T, V> f;
	const T* p = B + (size_t) col * ldb;
	if constexpr (V % E == 0) {
		#pragma unroll
		for (int i = 0; i < V / E; i++) {
			const uint4 t = __ldg(reinterpret_cast<const uint4*>(p) + i * G + lig);
			memcpy(&f.v[i * E], &t, 16);
		}
	} else {
		#pragma unroll
		for (int v = 0; v < V; v++) f.v[v] = __ldg(p + lig * V + v);
	}
	return f;
}

template <typename T, int G, int V>
__device__ __forceinline__ void store_row_t(T* __restrict__ C, size_t row, int lig, const T (&acc)[V], int ldc)
{
	constexpr int E = 16 / sizeof(T);
	T* p = C + row * ldc;
	if constexpr (V % E == 0) {
		#pragma unroll
		for (int i = 0; i < V / E; i++) {
			uint4 t;
			memcpy(&t, &acc[i * E], 16);
			reinterpret_cast<uint4*>(p)[i * G + lig] = t;      // (also used for the shared-memory carries: plain store)
		}
	} else {
		#pragma unroll
		for (int v = 0; v < V; v++) p[lig * V + v] = acc[v];
	}
}

template <typename T> __device__ __forceinline__ T fma_t(T a, T b, T c);
template <> __device__ __forceinline__ double fma_t<double>(double a, double b, double c) { return fma(a, b, c); }
template <> __device__ __forceinline__ float fma_t<float>(float a, float b, float c) { return fmaf(a, b, c); }

#ifndef MERGE_MIN_CTAS
#define MERGE_MIN_CTAS 1      // 4 (64 registers, 4 CTAs per SM instead of 3) together with streaming loads measured 4-7 % slower on cfg3a / cfg3b (r2c)
#endif

struct MergePlan {
	int ntiles = 0;            // CTAs of the main kernel
	int items = 0;             // merge items per tile
	int2* d_tile = nullptr;    // [ntiles+1] (row, nnz) start coordinate of each tile
	void* d_carry = nullptr;   // [ntiles][k] values (double or float)
	int32_t* d_carry_row = nullptr; // [ntiles]
	int carry_k = 0;
};

template <int G, int V, int IPG, int THREADS, typename T = double>
struct MergeCfg {
	static constexpr int K = G * V;
	static constexpr int NG = THREADS / G;
	static constexpr int ITEMS = NG * IPG;
	static constexpr size_t SMEM = (size_t) ITEMS * (sizeof(T) + 2 * sizeof(int32_t))
	                               + (size_t) NG * K * sizeof(T) + (size_t) NG * sizeof(int32_t);
};

template <int G, int V, int IPG, int UNROLL, int THREADS, typename T = double>
__global__ void __launch_bounds__(THREADS, MERGE_MIN_CTAS)
spmm_merge_kernel(const int32_t* __restrict__ ia, const int32_t* __restrict__ ja, const T* __restrict__ a,
                  const T* __restrict__ B, T* __restrict__ C,
                  const int2* __restrict__ tile_coord, T* __restrict__ tile_carry, int32_t* __restrict__ tile_carry_row)
{
	using Cfg = MergeCfg<G, V, IPG, THREADS, T>;
	constexpr int K = Cfg::K, NG = Cfg::NG, ITEMS = Cfg::ITEMS;
	static_assert((ITEMS * sizeof(T)) % 16 == 0 && (NG * K * sizeof(T)) % 4 == 0, "shared-memory carve-up alignment");
	extern __shared__ __align__(16) unsigned char smem_raw[];
	T* s_a = reinterpret_cast<T*>(smem_raw);                           // [ITEMS]
	T* s_carry = s_a + ITEMS;                                          // [NG][K]
	int32_t* s_ja = reinterpret_cast<int32_t*>(s_carry + NG * K);      // [ITEMS]
	int32_t* s_end = s_ja + ITEMS;                                     // [ITEMS] row ends, tile-local rows
	int32_t* s_crow = s_end + ITEMS;                                   // [NG]

	const int tile = blockIdx.x;
	const int2 c0 = tile_coord[tile], c1 = tile_coord[tile + 1];
	const int r0 = c0.x, j0 = c0.y, r1 = c1.x, j1 = c1.y;
	const int nr = r1 - r0, nn = j1 - j0;                              // nr + nn <= ITEMS

	for (int i = threadIdx.x; i < nr; i += THREADS) s_end[i] = __ldg(ia + r0 + 1 + i);
	for (int i = threadIdx.x; i < nn; i += THREADS) {
		s_ja[i] = __ldg(ja + j0 + i);
		s_a[i] = __ldg(a + j0 + i);
	}
	__syncthreads();

	const int lig = threadIdx.x % G, gid = threadIdx.x / G;
	// level-2 search: local diagonals gid*IPG and (gid+1)*IPG over (s_end - j0, 0..nn-1)
	int gr[2], gj[2];
	#pragma unroll
	for (int e = 0; e < 2; e++) {
		int d = min((gid + e) * IPG, nr + nn);
		int lo = max(d - nn, 0), hi = min(d, nr);
		while (lo < hi) {
			int p = (lo + hi) >> 1;
			if (s_end[p] - j0 <= d - p - 1) lo = p + 1; else hi = p;
		}
		gr[e] = lo; gj[e] = d - lo;
	}
	int row = gr[0];                           // tile-local row being accumulated
	const int row_e = gr[1];
	const int je = gj[1];
	int next_end = (row < row_e) ? s_end[row] - j0 : INT_MAX;
	T acc[V];
	#pragma unroll
	for (int v = 0; v < V; v++) acc[v] = T(0);

	for (int j = gj[0]; j < je; j += UNROLL) {
		FragT<T, V> bb[UNROLL];
		T vv[UNROLL];
		#pragma unroll
		for (int u = 0; u < UNROLL; u++) {
			const bool live = j + u < je;
			const int jj = live ? j + u : je - 1;
			vv[u] = live ? s_a[jj] : T(0);
			bb[u] = load_row_t<T, G, V>(B, s_ja[jj], lig, G * V);
		}
		#pragma unroll
		for (int u = 0; u < UNROLL; u++) {
			if (j + u < je) {
				while (j + u >= next_end) {      // rows that ended before this nonzero
					store_row_t<T, G, V>(C, (size_t) (r0 + row), lig, acc, G * V);
					#pragma unroll
					for (int v = 0; v < V; v++) acc[v] = T(0);
					row++;
					next_end = (row < row_e) ? s_end[row] - j0 : INT_MAX;
				}
				#pragma unroll
				for (int v = 0; v < V; v++) acc[v] = fma_t<T>(vv[u], bb[u].v[v], acc[v]);
			}
		}
	}
	for (; row < row_e; row++) {                 // row ends after the group's last nonzero
		store_row_t<T, G, V>(C, (size_t) (r0 + row), lig, acc, G * V);
		#pragma unroll
		for (int v = 0; v < V; v++) acc[v] = T(0);
	}
	// what is left belongs to tile-local row row_e, which ends in a later group or tile
	store_row_t<T, G, V>(s_carry, (size_t) gid, lig, acc, G * V);
	if (lig == 0) s_crow[gid] = row_e;
	__syncthreads();

	for (int idx = threadIdx.x; idx < NG * K; idx += THREADS) {
		const int g = idx / K, n = idx % K;
		const int crow = s_crow[g];
		if (g > 0 && s_crow[g - 1] == crow) continue;        // not the head of its run
		T sum = T(0);
		for (int g2 = g; g2 < NG && s_crow[g2] == crow; g2++) sum += s_carry[g2 * K + n];
		if (crow < nr) C[(size_t) (r0 + crow) * K + n] += sum; // row ended inside this tile
		else tile_carry[(size_t) tile * K + n] = sum;          // crow == nr: leaves the tile
	}
	if (threadIdx.x == 0) tile_carry_row[tile] = r1;
}

// Adds the tile carries to the rows they belong to.  A run of tiles with the same carry row
// (a row longer than a tile) is summed in tile order by the thread group of its first tile.
template <int THREADS, typename T = double>
__global__ void __launch_bounds__(THREADS)
spmm_merge_fixup_kernel(const T* __restrict__ tile_carry, const int32_t* __restrict__ tile_carry_row,
                        int ntiles, int m, int k, T* __restrict__ C)
{
	const int64_t idx = (int64_t) blockIdx.x * THREADS + threadIdx.x;
	if (idx >= (int64_t) ntiles * k) return;
	const int t = (int) (idx / k), n = (int) (idx % k);
	const int row = tile_carry_row[t];
	if (row >= m) return;
	if (t > 0 && tile_carry_row[t - 1] == row) return;
	T sum = T(0);
	for (int t2 = t; t2 < ntiles && tile_carry_row[t2] == row; t2++) sum += tile_carry[(size_t) t2 * k + n];
	C[(size_t) row * k + n] += sum;
}

} // namespace sb200
