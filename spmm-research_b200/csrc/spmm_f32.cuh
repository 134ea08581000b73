// spmm_f32.cuh -- fp32 CSR SpMM kernels (the reference's "_f" executables, make.sh:98-108,
// ValueType = float).  Same algorithm and work mapping as spmm_rowsplit.cuh:
//     C[i*k + n] = sum_{j in [ia[i], ia[i+1])} a[j] * B[ja[j]][n]      (spmm_kernel_csr.cpp:69-97)
// CTA b owns rows [cta_rows[b], cta_rows[b+1]) of the reference's nnz-balanced split; a group of G lanes owns
// a row, a lane owns V = 4 or 8 of the k = G*V columns (one or two 16-byte loads), the G lanes read G
// consecutive (col, val) with one coalesced load each and pass them round with shuffles; the row sum runs
// left to right in fp32 with fused multiply-add.
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>

namespace sb200 {

template <int G, int V>
__device__ __forceinline__ void load_b_f32(const float* __restrict__ B, int col, int lig, int ldb, float (&f)[V])
{
	static_assert(V % 4 == 0, "V must be a multiple of 4");
	const float4* p = reinterpret_cast<const float4*>(B + (size_t) col * ldb);
	#pragma unroll
	for (int i = 0; i < V / 4; i++) {
		const float4 t = __ldg(p + i * G + lig);
		f[4 * i] = t.x; f[4 * i + 1] = t.y; f[4 * i + 2] = t.z; f[4 * i + 3] = t.w;
	}
}

// TICKET = false: CTA b owns the fixed row range cta_rows[b..b+1).  TICKET = true: one resident wave of persistent
// warps draws `rows_per_ticket`-row chunks from an atomic counter (cta_rows is unused) -- the scheduling of
// spmm_rowsplit_ticket_kernel, same per-row arithmetic.
template <int G, int V, int UNROLL, int THREADS, bool TICKET = false>
__global__ void __launch_bounds__(THREADS)
spmm_rowsplit_f32_kernel(const int32_t* __restrict__ ia, const int32_t* __restrict__ ja, const float* __restrict__ a,
                         const float* __restrict__ B, float* __restrict__ C,
                         const int32_t* __restrict__ cta_rows, int ldb, int ldc,
                         int m = 0, int rows_per_ticket = 0, int* __restrict__ ticket = nullptr)
{
	constexpr int NG = TICKET ? 32 / G : THREADS / G;      // row groups that advance together (warp / CTA)
	constexpr int U = UNROLL < G ? UNROLL : G;
	static_assert(G % U == 0, "UNROLL and G must nest");
	const int lig = threadIdx.x % G;
	const int gid = TICKET ? (threadIdx.x & 31) / G : threadIdx.x / G;
	for (;;) {
	int r0, r1;
	if (TICKET) {
		int t = 0;
		if ((threadIdx.x & 31) == 0) t = atomicAdd(ticket, 1);
		t = __shfl_sync(0xffffffffu, t, 0);
		const int64_t lo = (int64_t) t * rows_per_ticket;
		if (lo >= m) break;
		r0 = (int) lo;
		r1 = (int) min((int64_t) m, lo + rows_per_ticket);
	} else {
		r0 = cta_rows[blockIdx.x];
		r1 = cta_rows[blockIdx.x + 1];
	}
	for (int base = r0; base < r1; base += NG) {
		const int row = base + gid;
		const bool live = row < r1;
		int j = 0, je = 0;
		if (live) { j = __ldg(ia + row); je = __ldg(ia + row + 1); }
		float acc[V];
		#pragma unroll
		for (int v = 0; v < V; v++) acc[v] = 0.f;
		while (true) {
			const int rem = je - j;
			const int rem_warp = __reduce_max_sync(0xffffffffu, rem);
			if (rem_warp <= 0) break;
			int cj = 0;
			float ca = 0.f;
			if (lig < rem) { cj = __ldg(ja + j + lig); ca = __ldg(a + j + lig); }
			const int cnt = min(rem_warp, G);
			for (int u = 0; u < cnt; u += U) {
				int cc[U];
				float vv[U];
				float bb[U][V];
				#pragma unroll
				for (int i = 0; i < U; i++) {
					cc[i] = __shfl_sync(0xffffffffu, cj, u + i, G);
					vv[i] = __shfl_sync(0xffffffffu, ca, u + i, G);
				}
				#pragma unroll
				for (int i = 0; i < U; i++)
					if (u + i < rem) load_b_f32<G, V>(B, cc[i], lig, ldb, bb[i]);
				#pragma unroll
				for (int i = 0; i < U; i++) {
					if (u + i < rem) {
						#pragma unroll
						for (int v = 0; v < V; v++) acc[v] = fmaf(vv[i], bb[i][v], acc[v]);
					}
				}
			}
			j += G;
		}
		if (live) {
			float4* p = reinterpret_cast<float4*>(C + (size_t) row * ldc);
			#pragma unroll
			for (int i = 0; i < V / 4; i++) p[i * G + lig] = make_float4(acc[4 * i], acc[4 * i + 1], acc[4 * i + 2], acc[4 * i + 3]);
		}
	}
	if (!TICKET) break;
	}
}

// Any k (including 1, 2, 4): a warp per row, lanes own columns in chunks of 32; for k < 32 the idle lanes
// cost nothing the memory system notices.  A is re-read per chunk out of L1.
template <int THREADS>
__global__ void __launch_bounds__(THREADS)
spmm_rowsplit_generic_f32_kernel(const int32_t* __restrict__ ia, const int32_t* __restrict__ ja, const float* __restrict__ a,
                                 const float* __restrict__ B, float* __restrict__ C,
                                 const int32_t* __restrict__ cta_rows, int k)
{
	constexpr int NW = THREADS / 32;
	const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
	const int r0 = cta_rows[blockIdx.x], r1 = cta_rows[blockIdx.x + 1];
	for (int row = r0 + wid; row < r1; row += NW) {
		const int js = __ldg(ia + row), je = __ldg(ia + row + 1);
		for (int n0 = 0; n0 < k; n0 += 32) {
			const int n = n0 + lane;
			float acc = 0.f;
			if (n < k) {
				for (int j = js; j < je; j++)
					acc = fmaf(__ldg(a + j), __ldg(B + (size_t) __ldg(ja + j) * k + n), acc);
				C[(size_t) row * k + n] = acc;
			}
		}
	}
}

// oracle layout x[n*c + j] (k vectors of length c) -> row-major B[j*k + n]
__global__ void transpose_b_f32_kernel(const float* __restrict__ x, float* __restrict__ B, int64_t c, int k)
{
	__shared__ float tile[32][33];
	const int64_t j0 = (int64_t) blockIdx.x * 32;
	const int n0 = blockIdx.y * 32;
	for (int dy = threadIdx.y; dy < 32; dy += blockDim.y) {
		const int n = n0 + dy;
		const int64_t j = j0 + threadIdx.x;
		if (n < k && j < c) tile[dy][threadIdx.x] = x[(int64_t) n * c + j];
	}
	__syncthreads();
	for (int dy = threadIdx.y; dy < 32; dy += blockDim.y) {
		const int64_t j = j0 + dy;
		const int n = n0 + threadIdx.x;
		if (n < k && j < c) B[j * k + n] = tile[threadIdx.x][dy];
	}
}

// The same for `rows` consecutive rows of B out of a column-major operand whose columns are `pitch` elements apart
// (window upload of spmm_b200_run_f32): x[n*pitch + j] -> B[j*k + n], j < rows.
__global__ void transpose_b_rows_f32_kernel(const float* __restrict__ x, float* __restrict__ B, int64_t pitch, int64_t rows, int k)
{
	__shared__ float tile[32][33];
	const int64_t j0 = (int64_t) blockIdx.x * 32;
	const int n0 = blockIdx.y * 32;
	for (int dy = threadIdx.y; dy < 32; dy += blockDim.y) {
		const int n = n0 + dy;
		const int64_t j = j0 + threadIdx.x;
		if (n < k && j < rows) tile[dy][threadIdx.x] = x[(int64_t) n * pitch + j];
	}
	__syncthreads();
	for (int dy = threadIdx.y; dy < 32; dy += blockDim.y) {
		const int64_t j = j0 + dy;
		const int n = n0 + threadIdx.x;
		if (n < k && j < rows) B[j * k + n] = tile[threadIdx.x][dy];
	}
}

} // namespace sb200
