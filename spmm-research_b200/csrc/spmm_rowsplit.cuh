// spmm_rowsplit.cuh -- CSR SpMM, "row-split" variant (north_star kernel (2), warp-per-row family).
//
// Computes what the reference's compute_csr does
// (benchmark_code/CPU/AMD/spmv_code_bench/spmm_kernel_csr.cpp:69-97):
//     C[i*k + n] = sum_{j in [ia[i], ia[i+1])} a[j] * B[ja[j]][n]
// but with the loop nest turned inside out: A is streamed ONCE (not k times) and every
// gathered B row (k contiguous doubles, B kept row-major on the device) is consumed for
// all k columns at once.  The per-row sum runs left to right like the oracle's.
//
// Work mapping
//   * CTA b owns the contiguous rows [cta_rows[b], cta_rows[b+1]) of the nnz-balanced
//     split (partition.cuh: cta_rows_nnz_kernel, bit-exact with lib/parallel_util.h:141-165).
//   * Inside a CTA a group of G lanes owns one row at a time; a lane owns V of the k = G*V
//     dense columns, so one group-wide load of a B row is k*8 contiguous bytes
//     (k=32: 16 lanes x 16 B = 256 B = two full 128-B lines).
//   * (col, val) of a nonzero are read once per group as same-address (broadcast) loads
//     through L1; UNROLL nonzeros are in flight per group so that each warp keeps
//     UNROLL * 32/G independent B gathers outstanding (the kernel is bound by L2->SM
//     gather bandwidth, not by HBM).
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>

namespace sb200 {

template <int V> struct Frag { double v[V]; };

// Load this lane's V columns of B row `col`.  Columns are dealt to lanes in pairs:
// lane l holds columns {2l, 2l+1} (+ 2G for the second pair when V == 4), so every
// load instruction of the group covers one contiguous span.
template <int G, int V, int LD = 0>
__device__ __forceinline__ Frag<V> load_b(const double* __restrict__ B, int col, int lig, int ldb)
{
	Frag<V> f;
	// LD > 0: the row stride is a compile-time constant and c*LD < 2^32 (checked by the launcher), so the element
	// offset is ONE 32-bit multiply by a constant instead of a 64-bit multiply by a kernel parameter that the
	// compiler re-reads from constant memory under every predicate (SASS study in profiles/r1_rowsplit_cfg2_v3.md)
	const double* p = LD > 0 ? B + (unsigned) col * (unsigned) LD : B + (size_t) col * ldb;
	if constexpr (V == 1) {
		f.v[0] = __ldg(p + lig);
	} else {
		// V/2 16-byte loads per lane; load i of the group covers the contiguous span [i*G*16, (i+1)*G*16) bytes
		// of the row, so with G = 8 every load instruction of a group is exactly one 128-byte line
		static_assert(V % 2 == 0, "V must be 1 or even");
		#pragma unroll
		for (int i = 0; i < V / 2; i++) {
			const double2 t = __ldg(reinterpret_cast<const double2*>(p) + i * G + lig);
			f.v[2 * i] = t.x; f.v[2 * i + 1] = t.y;
		}
	}
	return f;
}

template <int G, int V>
__device__ __forceinline__ void store_c(double* __restrict__ C, size_t row, int lig, const double (&acc)[V], int ldc)
{
	double* p = C + row * ldc;
	if constexpr (V == 1) {
		p[lig] = acc[0];
	} else {
		#pragma unroll
		for (int i = 0; i < V / 2; i++)
			reinterpret_cast<double2*>(p)[i * G + lig] = make_double2(acc[2 * i], acc[2 * i + 1]);
	}
}

// Broadcast of a double inside a lane group of width G.
template <int G>
__device__ __forceinline__ double shfl_f64(double v, int src)
{
	int lo = __double2loint(v), hi = __double2hiint(v);
	lo = __shfl_sync(0xffffffffu, lo, src, G);
	hi = __shfl_sync(0xffffffffu, hi, src, G);
	return __hiloint2double(hi, lo);
}

// (col, val) staging: the G lanes of a group read G consecutive nonzeros of their row with ONE
// coalesced load each (the 32/G groups of a warp own consecutive rows, so the warp's loads are
// contiguous in ja / a) and hand them round with shuffles.  The previous version read every
// (col, val) as a same-address load per group; ncu (profiles/r1_rowsplit_cfg2.md) showed those
// broadcast loads were half of the L1 data-pipe wavefronts, the busiest unit of the kernel.
// The per-row sum still runs strictly left to right.
template <int G, int V, int UNROLL, int THREADS, int LD = 0>
__global__ void __launch_bounds__(THREADS)
spmm_rowsplit_kernel(const int32_t* __restrict__ ia, const int32_t* __restrict__ ja, const double* __restrict__ a,
                     const double* __restrict__ B, double* __restrict__ C,
                     const int32_t* __restrict__ cta_rows, int ldb, int ldc)
{
	// ldb / ldc = row strides of B and C in doubles (G*V for a whole operand; larger when the launch
	// works on a column slice of a wider B / C, see launch_spmm's L2-aware k slicing)
	constexpr int NG = THREADS / G;          // row groups per CTA
	static_assert(G % UNROLL == 0 || UNROLL % G == 0, "UNROLL and G must nest");
	constexpr int U = UNROLL < G ? UNROLL : G;
	const int lig = threadIdx.x % G;         // lane in group
	const int gid = threadIdx.x / G;
	const int r0 = cta_rows[blockIdx.x];
	const int r1 = cta_rows[blockIdx.x + 1];

	// all lanes of a warp run the same trip counts (the shuffles need the full warp)
	for (int base = r0; base < r1; base += NG) {
		const int row = base + gid;
		const bool live = row < r1;
		int j = 0, je = 0;
		if (live) { j = __ldg(ia + row); je = __ldg(ia + row + 1); }
		double acc[V];
		#pragma unroll
		for (int v = 0; v < V; v++) acc[v] = 0.0;

		while (true) {
			const int rem = je - j;                                   // nonzeros left in my row
			const int rem_warp = __reduce_max_sync(0xffffffffu, rem); // longest row of the warp
			if (rem_warp <= 0) break;
			int cj = 0;
			double ca = 0.0;
			if (lig < rem) { cj = __ldg(ja + j + lig); ca = __ldg(a + j + lig); }
			const int cnt = min(rem_warp, G);
			// at most G/U steps: unrolled so that the shuffle source lanes are immediates
			#pragma unroll
			for (int u = 0; u < G; u += U) {
				if (u >= cnt) break;                                  // warp-uniform
				int cc[U];
				double vv[U];
				Frag<V> bb[U];
				#pragma unroll
				for (int i = 0; i < U; i++) {
					cc[i] = __shfl_sync(0xffffffffu, cj, u + i, G);
					vv[i] = shfl_f64<G>(ca, u + i);
				}
				#pragma unroll
				for (int i = 0; i < U; i++)
					if (u + i < rem) bb[i] = load_b<G, V, LD>(B, cc[i], lig, ldb);
				#pragma unroll
				for (int i = 0; i < U; i++) {
					if (u + i < rem) {
						#pragma unroll
						for (int v = 0; v < V; v++) acc[v] = fma(vv[i], bb[i].v[v], acc[v]);
					}
				}
			}
			j += G;
		}
		if (live) store_c<G, V>(C, (size_t) row, lig, acc, ldc);
	}
}

// Experiment (SPMM_B200_TICKET=R): the same row loop on ONE resident wave of persistent warps that draw R-row
// chunks from a ticket counter instead of owning a fixed nnz-balanced range -- removes the tail of the last wave.
template <int G, int V, int UNROLL, int THREADS, int LD = 0>
__global__ void __launch_bounds__(THREADS)
spmm_rowsplit_ticket_kernel(const int32_t* __restrict__ ia, const int32_t* __restrict__ ja, const double* __restrict__ a,
                            const double* __restrict__ B, double* __restrict__ C,
                            int m, int rows_per_ticket, int* __restrict__ ticket, int ldb, int ldc)
{
	constexpr int NGW = 32 / G;              // row groups per warp
	constexpr int U = UNROLL < G ? UNROLL : G;
	const int lane = threadIdx.x & 31;
	const int lig = lane % G;
	const int gid = lane / G;
	for (;;) {
		int t = 0;
		if (lane == 0) t = atomicAdd(ticket, 1);
		t = __shfl_sync(0xffffffffu, t, 0);
		const int64_t r0l = (int64_t) t * rows_per_ticket;
		if (r0l >= m) break;
		const int r0 = (int) r0l;
		const int r1 = (int) min((int64_t) m, r0l + rows_per_ticket);
		for (int base = r0; base < r1; base += NGW) {
			const int row = base + gid;
			const bool live = row < r1;
			int j = 0, je = 0;
			if (live) { j = __ldg(ia + row); je = __ldg(ia + row + 1); }
			double acc[V];
			#pragma unroll
			for (int v = 0; v < V; v++) acc[v] = 0.0;
			while (true) {
				const int rem = je - j;
				const int rem_warp = __reduce_max_sync(0xffffffffu, rem);
				if (rem_warp <= 0) break;
				int cj = 0;
				double ca = 0.0;
				if (lig < rem) { cj = __ldg(ja + j + lig); ca = __ldg(a + j + lig); }
				const int cnt = min(rem_warp, G);
				#pragma unroll
				for (int u = 0; u < G; u += U) {
					if (u >= cnt) break;
					int cc[U];
					double vv[U];
					Frag<V> bb[U];
					#pragma unroll
					for (int i = 0; i < U; i++) {
						cc[i] = __shfl_sync(0xffffffffu, cj, u + i, G);
						vv[i] = shfl_f64<G>(ca, u + i);
					}
					#pragma unroll
					for (int i = 0; i < U; i++)
						if (u + i < rem) bb[i] = load_b<G, V, LD>(B, cc[i], lig, ldb);
					#pragma unroll
					for (int i = 0; i < U; i++) {
						if (u + i < rem) {
							#pragma unroll
							for (int v = 0; v < V; v++) acc[v] = fma(vv[i], bb[i].v[v], acc[v]);
						}
					}
				}
				j += G;
			}
			if (live) store_c<G, V>(C, (size_t) row, lig, acc, ldc);
		}
	}
}

// Any k: one warp per row, lanes own columns, column chunks of 32 (A re-read per chunk
// out of L1/L2).  Used for k values without a specialised instantiation.
template <int THREADS>
__global__ void __launch_bounds__(THREADS)
spmm_rowsplit_generic_kernel(const int32_t* __restrict__ ia, const int32_t* __restrict__ ja, const double* __restrict__ a,
                             const double* __restrict__ B, double* __restrict__ C,
                             const int32_t* __restrict__ cta_rows, int k)
{
	constexpr int NW = THREADS / 32;
	const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
	const int r0 = cta_rows[blockIdx.x], r1 = cta_rows[blockIdx.x + 1];
	for (int row = r0 + wid; row < r1; row += NW) {
		const int js = __ldg(ia + row), je = __ldg(ia + row + 1);
		for (int n0 = 0; n0 < k; n0 += 32) {
			const int n = n0 + lane;
			double acc = 0.0;
			if (n < k) {
				for (int j = js; j < je; j++)
					acc = fma(__ldg(a + j), __ldg(B + (size_t) __ldg(ja + j) * k + n), acc);
				C[(size_t) row * k + n] = acc;
			}
		}
	}
}

// Small k (1, 2, 4): the SpMV-like regime.  G lanes share a row and stride over its
// nonzeros (coalesced reads of ja/a), each lane keeps K partial sums, then a shuffle
// tree combines them.  Summation order differs from the oracle's (tolerance, not bits).
template <int K, int G, int THREADS>
__global__ void __launch_bounds__(THREADS)
spmm_smallk_kernel(const int32_t* __restrict__ ia, const int32_t* __restrict__ ja, const double* __restrict__ a,
                   const double* __restrict__ B, double* __restrict__ C,
                   const int32_t* __restrict__ cta_rows)
{
	constexpr int NG = THREADS / G;
	const int lig = threadIdx.x % G, gid = threadIdx.x / G;
	const int r0 = cta_rows[blockIdx.x], r1 = cta_rows[blockIdx.x + 1];
	// all groups of a warp must run the same number of iterations for the shuffles
	const int nrows = r1 - r0;
	const int iters = (nrows + NG - 1) / NG;
	for (int it = 0; it < iters; it++) {
		const int row = r0 + it * NG + gid;
		const bool live = row < r1;
		double acc[K];
		#pragma unroll
		for (int n = 0; n < K; n++) acc[n] = 0.0;
		if (live) {
			const int js = __ldg(ia + row), je = __ldg(ia + row + 1);
			for (int j = js + lig; j < je; j += G) {
				const double v = __ldg(a + j);
				const double* p = B + (size_t) __ldg(ja + j) * K;
				if constexpr (K == 1) {
					acc[0] = fma(v, __ldg(p), acc[0]);
				} else if constexpr (K == 2) {
					double2 t = __ldg(reinterpret_cast<const double2*>(p));
					acc[0] = fma(v, t.x, acc[0]); acc[1] = fma(v, t.y, acc[1]);
				} else {
					double2 t = __ldg(reinterpret_cast<const double2*>(p));
					double2 u = __ldg(reinterpret_cast<const double2*>(p) + 1);
					acc[0] = fma(v, t.x, acc[0]); acc[1] = fma(v, t.y, acc[1]);
					acc[2] = fma(v, u.x, acc[2]); acc[3] = fma(v, u.y, acc[3]);
				}
			}
		}
		#pragma unroll
		for (int off = G / 2; off > 0; off >>= 1) {
			#pragma unroll
			for (int n = 0; n < K; n++) acc[n] += __shfl_down_sync(0xffffffffu, acc[n], off, G);
		}
		if (live && lig == 0) {
			#pragma unroll
			for (int n = 0; n < K; n++) C[(size_t) row * K + n] = acc[n];
		}
	}
}

// B upload helper: oracle layout x[n*c + j] (k vectors of length c) -> row-major B[j*k + n].
__global__ void transpose_b_kernel(const double* __restrict__ x, double* __restrict__ B, int64_t c, int k)
{
	__shared__ double tile[32][33];
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

} // namespace sb200

namespace sb200 {

// ------------------------------------------------------------------------------------------------
// Row-split on a row-PADDED copy of A ("vec" variant).
//
// ncu on spmm_rowsplit_kernel (profiles/r1_rowsplit_cfg2_v2.md) shows 13.9 warp instructions and ~3.5 L1
// data-pipe wavefronts per nonzero (2 for the 256-byte B row, the rest for handing (col, val) round a lane
// group with shuffles), issue slots 67 % and the L1 data pipe 78 % busy -- while a pure gather of the same B
// rows reaches 21.6 TB/s against the kernel's 14.6 (scripts/gather_probe.cu).  The cure is fewer
// instructions and wavefronts per nonzero, not more loads in flight:
//   * every row of A is padded to a multiple of 4 entries (col = the row's last column, val = 0), so a
//     lane reads 4 column indices with ONE 16-byte load and 4 values with TWO, as same-address loads of
//     its lane group (one wavefront per group and load, no shuffles, no per-step warp votes);
//   * lane groups are independent: no shuffle means no warp-synchronous trip counts;
//   * 4 B-row gathers (V doubles each) are in flight per lane and step.
// The row sum still runs left to right; padded entries are skipped by a predicate, never multiplied.
// ------------------------------------------------------------------------------------------------
template <int G, int V, int THREADS>
__global__ void __launch_bounds__(THREADS)
spmm_rowsplit_vec_kernel(const int32_t* __restrict__ ia, const int32_t* __restrict__ iap,
                         const int32_t* __restrict__ jap, const double* __restrict__ ap,
                         const double* __restrict__ B, double* __restrict__ C,
                         const int32_t* __restrict__ cta_rows, int ldb, int ldc)
{
	constexpr int NG = THREADS / G;
	const int lig = threadIdx.x % G;
	const int gid = threadIdx.x / G;
	const int r0 = cta_rows[blockIdx.x];
	const int r1 = cta_rows[blockIdx.x + 1];
	for (int row = r0 + gid; row < r1; row += NG) {
		const int len = __ldg(ia + row + 1) - __ldg(ia + row);
		const int p = __ldg(iap + row);
		const int4* cp = reinterpret_cast<const int4*>(jap + p);
		const double2* vp = reinterpret_cast<const double2*>(ap + p);
		double acc[V];
		#pragma unroll
		for (int v = 0; v < V; v++) acc[v] = 0.0;
		for (int u = 0; u < len; u += 4) {
			const int4 c4 = __ldg(cp + (u >> 2));
			const double2 v01 = __ldg(vp + (u >> 1));
			const double2 v23 = __ldg(vp + (u >> 1) + 1);
			const Frag<V> b0 = load_b<G, V>(B, c4.x, lig, ldb);
			const Frag<V> b1 = load_b<G, V>(B, c4.y, lig, ldb);
			const Frag<V> b2 = load_b<G, V>(B, c4.z, lig, ldb);
			const Frag<V> b3 = load_b<G, V>(B, c4.w, lig, ldb);
			const int rem = len - u;
			#pragma unroll
			for (int v = 0; v < V; v++) acc[v] = fma(v01.x, b0.v[v], acc[v]);
			if (rem > 1) {
				#pragma unroll
				for (int v = 0; v < V; v++) acc[v] = fma(v01.y, b1.v[v], acc[v]);
			}
			if (rem > 2) {
				#pragma unroll
				for (int v = 0; v < V; v++) acc[v] = fma(v23.x, b2.v[v], acc[v]);
			}
			if (rem > 3) {
				#pragma unroll
				for (int v = 0; v < V; v++) acc[v] = fma(v23.y, b3.v[v], acc[v]);
			}
		}
		store_c<G, V>(C, (size_t) row, lig, acc, ldc);
	}
}

// padded row lengths: out[i] = (len_i + 3) & ~3 for i < m, out[m] = 0 (then exclusive scan -> iap)
__global__ void pad_len_kernel(const int32_t* __restrict__ ia, int m, int32_t* __restrict__ out)
{
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < m) out[i] = (ia[i + 1] - ia[i] + 3) & ~3;
	else if (i == m) out[i] = 0;
}

// one thread per padded slot of a row group of 4: copies (col, val), pads with (last col of the row, 0)
__global__ void pad_fill_kernel(const int32_t* __restrict__ ia, const int32_t* __restrict__ iap,
                                const int32_t* __restrict__ ja, const double* __restrict__ a, int m,
                                int32_t* __restrict__ jap, double* __restrict__ ap)
{
	const int lane = threadIdx.x & 31;
	const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;      // a warp per row
	if (row >= m) return;
	const int s = ia[row], len = ia[row + 1] - s, p = iap[row], plen = iap[row + 1] - p;
	const int last = len > 0 ? ja[s + len - 1] : 0;
	for (int u = lane; u < plen; u += 32) {
		jap[p + u] = u < len ? ja[s + u] : last;
		ap[p + u] = u < len ? a[s + u] : 0.0;
	}
}

} // namespace sb200
