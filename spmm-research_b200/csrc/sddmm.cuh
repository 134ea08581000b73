// sddmm.cuh -- SDDMM and softmax kernels of the sparse-attention pipeline (SURVEY 8(f) N4;
// benchmark_code/CPU/AMD/pipeline_code_bench/sddmm_taco_naive.cpp).
//
//   reference mode (compute2, sddmm_taco_naive.cpp:98-140):
//       y[p] = ( sum_{d=0..n-1} Q[i*n+d] * K[i*n+d] ) * mask[p]           i = row of nonzero p
//     the sum starts at 0 and runs left to right.  gcc -O3 -mfma compiles the reference's loop `B[pA2] += O*D`
//     as rounded products followed by in-order adds (no fused multiply-add), so the chain here is
//     __dadd_rn(s, __dmul_rn(q, k)) -- bit for bit what the compiled reference returns.
//   attention mode:  y[p] = ( sum_d Q[i*n+d] * K[ja[p]*n+d] ) * mask[p], same order, same rounding.
//
// Both are HBM/L2 streams with a latency chain per row or nonzero: a warp per mask row, no tensor-core shape.
#pragma once
#include <stdint.h>
#include <float.h>
#include <cuda_runtime.h>

namespace sb200 {

// Reference mode.  Every lane of the warp walks the same Q and K row (one broadcast 16-byte load per two
// columns), so all 32 hold the row's dot product without a shuffle; then the lanes stride over the row's
// nonzeros.  Algorithmic bytes: 16 n per row + 16 per nonzero.
template <int THREADS>
__global__ void __launch_bounds__(THREADS)
sddmm_rowdot_kernel(const int32_t* __restrict__ ia, const double* __restrict__ mask_a,
                    const double* __restrict__ Q, const double* __restrict__ K,
                    double* __restrict__ y, int64_t m, int n)
{
	constexpr int NW = THREADS / 32;
	const int lane = threadIdx.x & 31;
	for (int64_t row = (int64_t) blockIdx.x * NW + (threadIdx.x >> 5); row < m; row += (int64_t) gridDim.x * NW) {
		const int js = __ldg(ia + row), je = __ldg(ia + row + 1);
		if (js == je) continue;
		const double* q = Q + (size_t) row * n;
		const double* kk = K + (size_t) row * n;
		double s = 0.0;
		if ((n & 1) == 0) {
			const double2* q2 = reinterpret_cast<const double2*>(q);
			const double2* k2 = reinterpret_cast<const double2*>(kk);
			#pragma unroll 4
			for (int d = 0; d < n / 2; d++) {
				const double2 a = __ldg(q2 + d), b = __ldg(k2 + d);
				s = __dadd_rn(s, __dmul_rn(a.x, b.x));
				s = __dadd_rn(s, __dmul_rn(a.y, b.y));
			}
		} else {
			for (int d = 0; d < n; d++) s = __dadd_rn(s, __dmul_rn(__ldg(q + d), __ldg(kk + d)));
		}
		for (int p = js + lane; p < je; p += 32) y[p] = s * __ldg(mask_a + p);
	}
}

// Attention mode.  A warp per mask row; a lane owns a nonzero and walks K[ja[p]] against the row's Q (the Q
// loads are warp-wide broadcasts that stay in L1).  Algorithmic bytes: 8 n per row (Q) + (8 n + 20) per nonzero.
template <int THREADS>
__global__ void __launch_bounds__(THREADS)
sddmm_attention_kernel(const int32_t* __restrict__ ia, const int32_t* __restrict__ ja, const double* __restrict__ mask_a,
                       const double* __restrict__ Q, const double* __restrict__ K,
                       double* __restrict__ y, int64_t m, int n)
{
	constexpr int NW = THREADS / 32;
	const int lane = threadIdx.x & 31;
	for (int64_t row = (int64_t) blockIdx.x * NW + (threadIdx.x >> 5); row < m; row += (int64_t) gridDim.x * NW) {
		const int js = __ldg(ia + row), je = __ldg(ia + row + 1);
		const double* q = Q + (size_t) row * n;
		for (int p = js + lane; p < je; p += 32) {
			const double* kk = K + (size_t) __ldg(ja + p) * n;
			double s = 0.0;
			if ((n & 1) == 0) {
				const double2* q2 = reinterpret_cast<const double2*>(q);
				const double2* k2 = reinterpret_cast<const double2*>(kk);
				#pragma unroll 4
				for (int d = 0; d < n / 2; d++) {
					const double2 a = __ldg(q2 + d), b = __ldg(k2 + d);
					s = __dadd_rn(s, __dmul_rn(a.x, b.x));
					s = __dadd_rn(s, __dmul_rn(a.y, b.y));
				}
			} else {
				for (int d = 0; d < n; d++) s = __dadd_rn(s, __dmul_rn(__ldg(q + d), __ldg(kk + d)));
			}
			y[p] = s * __ldg(mask_a + p);
		}
	}
}

// fp32 twins of the two ordered kernels above (the pipeline bench's "_f" build, ValueType = float): float
// products and in-order float adds, a warp per mask row.
template <int THREADS>
__global__ void __launch_bounds__(THREADS)
sddmm_rowdot_f32_kernel(const int32_t* __restrict__ ia, const float* __restrict__ mask_a,
                        const float* __restrict__ Q, const float* __restrict__ K,
                        float* __restrict__ y, int64_t m, int n)
{
	constexpr int NW = THREADS / 32;
	const int lane = threadIdx.x & 31;
	for (int64_t row = (int64_t) blockIdx.x * NW + (threadIdx.x >> 5); row < m; row += (int64_t) gridDim.x * NW) {
		const int js = __ldg(ia + row), je = __ldg(ia + row + 1);
		if (js == je) continue;
		const float* q = Q + (size_t) row * n;
		const float* kk = K + (size_t) row * n;
		float s = 0.f;
		for (int d = 0; d < n; d++) s = __fadd_rn(s, __fmul_rn(__ldg(q + d), __ldg(kk + d)));
		for (int p = js + lane; p < je; p += 32) y[p] = __fmul_rn(s, __ldg(mask_a + p));
	}
}

template <int THREADS>
__global__ void __launch_bounds__(THREADS)
sddmm_attention_f32_kernel(const int32_t* __restrict__ ia, const int32_t* __restrict__ ja, const float* __restrict__ mask_a,
                           const float* __restrict__ Q, const float* __restrict__ K,
                           float* __restrict__ y, int64_t m, int n)
{
	constexpr int NW = THREADS / 32;
	const int lane = threadIdx.x & 31;
	for (int64_t row = (int64_t) blockIdx.x * NW + (threadIdx.x >> 5); row < m; row += (int64_t) gridDim.x * NW) {
		const int js = __ldg(ia + row), je = __ldg(ia + row + 1);
		const float* q = Q + (size_t) row * n;
		for (int p = js + lane; p < je; p += 32) {
			const float* kk = K + (size_t) __ldg(ja + p) * n;
			float s = 0.f;
			for (int d = 0; d < n; d++) s = __fadd_rn(s, __fmul_rn(__ldg(q + d), __ldg(kk + d)));
			y[p] = __fmul_rn(s, __ldg(mask_a + p));
		}
	}
}

// Attention mode, fast path (n = G*V, V a multiple of 4).  Same work mapping as the row-split SpMM: a group of G
// lanes reads one K row per load instruction (V/2 16-byte chunks per lane, chunk i at i*G + lane-in-group), with
// U rows in flight.  A lane multiplies its V columns into a partial for each of the group's G nonzeros, then the
// G x G partials are reduced with a transposing butterfly (G-1 shuffle-adds per lane for G results instead of
// G log G), which leaves nonzero t's sum on lane t -- the y store is coalesced.  The sum order is a tree, not the
// CPU loop's left-to-right order: results agree with the ordered kernels to rounding (tests: 1e-12, norm-wise).
template <int G, int V, int U, int THREADS, int MINB = 1>
__global__ void __launch_bounds__(THREADS, MINB)
sddmm_attention_group_kernel(const int32_t* __restrict__ ia, const int32_t* __restrict__ ja, const double* __restrict__ mask_a,
                             const double* __restrict__ Q, const double* __restrict__ K,
                             double* __restrict__ y, int64_t m, int* __restrict__ ticket, int rev)
{
	static_assert(V % 4 == 0 && G % U == 0 && 32 % G == 0, "shape");
	constexpr int N = G * V;
	const int lane = threadIdx.x & 31;
	const int lig = lane % G;
	// persistent warps draw rows from a ticket counter, last row first (with causal-shaped masks the late rows are
	// the long ones): one resident wave, no tail of short CTAs
	for (;;) {
		int t = 0;
		if (lane == 0) t = atomicAdd(ticket, 1);
		t = __shfl_sync(0xffffffffu, t, 0);
		if (t >= m) break;
		const int64_t row = rev ? m - 1 - t : t;
		const int js = __ldg(ia + row), je = __ldg(ia + row + 1);
		if (js == je) continue;
		double q[V];
		{
			const double2* qp = reinterpret_cast<const double2*>(Q + (size_t) row * N);
			#pragma unroll
			for (int i = 0; i < V / 2; i++) {
				const double2 t = __ldg(qp + i * G + lig);
				q[2 * i] = t.x; q[2 * i + 1] = t.y;
			}
		}
		for (int p0 = js; p0 < je; p0 += 32) {
			const int p = p0 + lane;
			const int col = p < je ? __ldg(ja + p) : -1;
			double part[G];
			#pragma unroll
			for (int t0 = 0; t0 < G; t0 += U) {
				double2 kv[U][V / 2];
				int cc[U];
				#pragma unroll
				for (int u = 0; u < U; u++) cc[u] = __shfl_sync(0xffffffffu, col, t0 + u, G);
				#pragma unroll
				for (int u = 0; u < U; u++) {
					if (cc[u] >= 0) {
						const double2* kp = reinterpret_cast<const double2*>(K + (size_t) cc[u] * N);
						#pragma unroll
						for (int i = 0; i < V / 2; i++) kv[u][i] = __ldg(kp + i * G + lig);
					}
				}
				#pragma unroll
				for (int u = 0; u < U; u++) {
					double s = 0.0;
					if (cc[u] >= 0) {
						#pragma unroll
						for (int i = 0; i < V / 2; i++) {
							s = fma(q[2 * i], kv[u][i].x, s);
							s = fma(q[2 * i + 1], kv[u][i].y, s);
						}
					}
					part[t0 + u] = s;
				}
			}
			#pragma unroll
			for (int off = G / 2; off >= 1; off >>= 1) {
				const bool up = (lig & off) != 0;
				#pragma unroll
				for (int i = 0; i < off; i++) {
					const double send = up ? part[i] : part[i + off];
					const double keep = up ? part[i + off] : part[i];
					part[i] = keep + __shfl_xor_sync(0xffffffffu, send, off, G);
				}
			}
			if (p < je) y[p] = part[0] * __ldg(mask_a + p);
		}
	}
}

// SDDMM -> row-wise softmax -> SpMM in ONE kernel (attention mode; y never leaves the SM).  A warp owns a mask
// row i and walks its nonzeros 32 at a time:
//   scores   s_p = mask[p] * Q[i].K[ja[p]]            as in sddmm_attention_group_kernel (lane t ends up with nonzero t's)
//   softmax  online (running maximum mrun and sum l; the accumulators are rescaled by exp(mrun_old - mrun_new))
//   product  acc += exp(s_p - mrun) * V[ja[p]]        group g of the warp gathers the V rows of ITS G nonzeros,
//                                                    lanes own V columns each (the row-split SpMM's mapping)
// and at the end the 32/G group partials are added and divided by l.  Bytes per nonzero: 2 * 8n (K and V rows)
// + 12, against 2 * 8n + 12 + 3 * 12 + 48 for the three separate kernels; two launches and the y round trip go.
// Rounding: tree-ordered dot products, online softmax -- agrees with the step-by-step path to 1e-12 (tested).
template <int G, int V, int U, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB)
attention_fused_kernel(const int32_t* __restrict__ ia, const int32_t* __restrict__ ja, const double* __restrict__ mask_a,
                       const double* __restrict__ Q, const double* __restrict__ K, const double* __restrict__ Vm,
                       double* __restrict__ out, int64_t m, int* __restrict__ ticket, int rev)
{
	static_assert(V % 4 == 0 && G % U == 0 && 32 % G == 0, "shape");
	constexpr int N = G * V;
	const int lane = threadIdx.x & 31;
	const int lig = lane % G;
	for (;;) {                                     // persistent warps, rows by ticket, last row first
		int t = 0;
		if (lane == 0) t = atomicAdd(ticket, 1);
		t = __shfl_sync(0xffffffffu, t, 0);
		if (t >= m) break;
		const int64_t row = rev ? m - 1 - t : t;
		const int js = __ldg(ia + row), je = __ldg(ia + row + 1);
		double2* op = reinterpret_cast<double2*>(out + (size_t) row * N);
		if (js == je) {
			if (lane < G) {
				#pragma unroll
				for (int i = 0; i < V / 2; i++) op[i * G + lig] = make_double2(0.0, 0.0);
			}
			continue;
		}
		double q[V], acc[V];
		{
			const double2* qp = reinterpret_cast<const double2*>(Q + (size_t) row * N);
			#pragma unroll
			for (int i = 0; i < V / 2; i++) {
				const double2 t = __ldg(qp + i * G + lig);
				q[2 * i] = t.x; q[2 * i + 1] = t.y;
				acc[2 * i] = 0.0; acc[2 * i + 1] = 0.0;
			}
		}
		double mrun = -DBL_MAX, l = 0.0;
		for (int p0 = js; p0 < je; p0 += 32) {
			const int p = p0 + lane;
			const int col = p < je ? __ldg(ja + p) : -1;
			// ---- scores
			double part[G];
			#pragma unroll
			for (int t0 = 0; t0 < G; t0 += U) {
				double2 kv[U][V / 2];
				int cc[U];
				#pragma unroll
				for (int u = 0; u < U; u++) cc[u] = __shfl_sync(0xffffffffu, col, t0 + u, G);
				#pragma unroll
				for (int u = 0; u < U; u++) {
					if (cc[u] >= 0) {
						const double2* kp = reinterpret_cast<const double2*>(K + (size_t) cc[u] * N);
						#pragma unroll
						for (int i = 0; i < V / 2; i++) kv[u][i] = __ldg(kp + i * G + lig);
					}
				}
				#pragma unroll
				for (int u = 0; u < U; u++) {
					double s = 0.0;
					if (cc[u] >= 0) {
						#pragma unroll
						for (int i = 0; i < V / 2; i++) {
							s = fma(q[2 * i], kv[u][i].x, s);
							s = fma(q[2 * i + 1], kv[u][i].y, s);
						}
					}
					part[t0 + u] = s;
				}
			}
			#pragma unroll
			for (int off = G / 2; off >= 1; off >>= 1) {
				const bool up = (lig & off) != 0;
				#pragma unroll
				for (int i = 0; i < off; i++) {
					const double send = up ? part[i] : part[i + off];
					const double keep = up ? part[i + off] : part[i];
					part[i] = keep + __shfl_xor_sync(0xffffffffu, send, off, G);
				}
			}
			const double sc = p < je ? part[0] * __ldg(mask_a + p) : -DBL_MAX;
			// ---- online softmax
			double bmax = sc;
			#pragma unroll
			for (int o = 16; o > 0; o >>= 1) bmax = fmax(bmax, __shfl_xor_sync(0xffffffffu, bmax, o));
			const double mnew = fmax(mrun, bmax);
			const double scale = exp(mrun - mnew);          // first batch: exp(-huge) = 0, acc and l are 0 anyway
			const double pe = p < je ? exp(sc - mnew) : 0.0;
			double bsum = pe;
			#pragma unroll
			for (int o = 16; o > 0; o >>= 1) bsum += __shfl_xor_sync(0xffffffffu, bsum, o);
			l = l * scale + bsum;
			mrun = mnew;
			#pragma unroll
			for (int v = 0; v < V; v++) acc[v] *= scale;
			// ---- weighted V rows
			#pragma unroll
			for (int t0 = 0; t0 < G; t0 += U) {
				double2 vv[U][V / 2];
				int cc[U];
				double ww[U];
				#pragma unroll
				for (int u = 0; u < U; u++) {
					cc[u] = __shfl_sync(0xffffffffu, col, t0 + u, G);
					ww[u] = __shfl_sync(0xffffffffu, pe, t0 + u, G);
				}
				#pragma unroll
				for (int u = 0; u < U; u++) {
					if (cc[u] >= 0) {
						const double2* vp = reinterpret_cast<const double2*>(Vm + (size_t) cc[u] * N);
						#pragma unroll
						for (int i = 0; i < V / 2; i++) vv[u][i] = __ldg(vp + i * G + lig);
					}
				}
				#pragma unroll
				for (int u = 0; u < U; u++) {
					if (cc[u] >= 0) {
						#pragma unroll
						for (int i = 0; i < V / 2; i++) {
							acc[2 * i] = fma(ww[u], vv[u][i].x, acc[2 * i]);
							acc[2 * i + 1] = fma(ww[u], vv[u][i].y, acc[2 * i + 1]);
						}
					}
				}
			}
		}
		// the 32 / G groups hold partial rows: add them, normalise, store from group 0
		#pragma unroll
		for (int o = G; o < 32; o <<= 1) {
			#pragma unroll
			for (int v = 0; v < V; v++) acc[v] += __shfl_xor_sync(0xffffffffu, acc[v], o);
		}
		if (lane < G) {
			const double inv = 1.0 / l;
			#pragma unroll
			for (int i = 0; i < V / 2; i++) op[i * G + lig] = make_double2(acc[2 * i] * inv, acc[2 * i + 1] * inv);
		}
	}
}

// Attention mode for n % 32 == 0: the simple kernel above makes every load instruction touch 32 different
// 128-byte lines (one K row per lane), which the L1 tag stage serves at ~16 B/clk/SM (3.9 TB/s measured).
// Here the warp stages the K rows of 32 nonzeros through shared memory in 256-byte column chunks: two rows per
// coalesced load instruction (lanes 0-15 one row, 16-31 the next), rows padded by 16 bytes so that the
// per-lane 16-byte reads that follow are conflict-free.  Each lane still walks ITS nonzero's row left to
// right, so the sum order -- and every bit of y -- is unchanged.
constexpr int SDDMM_CH = 32;                       // doubles per staged chunk (256 bytes)
constexpr int SDDMM_LD = SDDMM_CH + 2;             // padded row length in shared memory

template <int THREADS>
__global__ void __launch_bounds__(THREADS)
sddmm_attention_staged_kernel(const int32_t* __restrict__ ia, const int32_t* __restrict__ ja, const double* __restrict__ mask_a,
                              const double* __restrict__ Q, const double* __restrict__ K,
                              double* __restrict__ y, int64_t m, int n)
{
	constexpr int NW = THREADS / 32;
	extern __shared__ __align__(16) double sddmm_smem[];
	const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
	double* tile = sddmm_smem + (size_t) wid * (32 * SDDMM_LD + n);     // 32 padded K-row chunks
	double* qrow = tile + 32 * SDDMM_LD;                                 // the mask row's Q row
	const int half = lane >> 4, l16 = lane & 15;
	for (int64_t row = (int64_t) blockIdx.x * NW + wid; row < m; row += (int64_t) gridDim.x * NW) {
		const int js = __ldg(ia + row), je = __ldg(ia + row + 1);
		if (js == je) continue;
		__syncwarp();
		for (int d = lane * 2; d < n; d += 64)
			*reinterpret_cast<double2*>(qrow + d) = __ldg(reinterpret_cast<const double2*>(Q + (size_t) row * n + d));
		for (int p0 = js; p0 < je; p0 += 32) {
			const int p = p0 + lane;
			const int cnt = min(32, je - p0);
			const int col = p < je ? __ldg(ja + p) : 0;
			double s = 0.0;
			for (int c0 = 0; c0 < n; c0 += SDDMM_CH) {
				__syncwarp();
				#pragma unroll
				for (int r0 = 0; r0 < 32; r0 += 16) {
					double2 v[8];
					#pragma unroll
					for (int i = 0; i < 8; i++) {
						const int r = r0 + 2 * i + half;
						const int cr = __shfl_sync(0xffffffffu, col, r);
						if (r < cnt) v[i] = __ldg(reinterpret_cast<const double2*>(K + (size_t) cr * n + c0) + l16);
					}
					#pragma unroll
					for (int i = 0; i < 8; i++) {
						const int r = r0 + 2 * i + half;
						if (r < cnt) *reinterpret_cast<double2*>(tile + r * SDDMM_LD + 2 * l16) = v[i];
					}
				}
				__syncwarp();
				if (lane < cnt) {
					const double2* kr = reinterpret_cast<const double2*>(tile + lane * SDDMM_LD);
					const double2* qr = reinterpret_cast<const double2*>(qrow + c0);
					#pragma unroll
					for (int d = 0; d < SDDMM_CH / 2; d++) {
						const double2 a = qr[d], b = kr[d];
						s = __dadd_rn(s, __dmul_rn(a.x, b.x));
						s = __dadd_rn(s, __dmul_rn(a.y, b.y));
					}
				}
			}
			if (p < je) y[p] = s * __ldg(mask_a + p);
		}
	}
}

// ---- softmax over ALL nnz values (softmax(), sddmm_taco_naive.cpp:191-209): max, exp(y - max) and its sum,
// divide.  Partials are combined in a fixed order (deterministic); the reference sums left to right on one
// thread, so the result agrees to rounding, not bit for bit.
template <int THREADS, bool IS_MAX>
__device__ __forceinline__ double block_reduce(double v)
{
	__shared__ double sh[THREADS / 32];
	#pragma unroll
	for (int o = 16; o > 0; o >>= 1) {
		const double w = __shfl_xor_sync(0xffffffffu, v, o);
		v = IS_MAX ? fmax(v, w) : v + w;
	}
	__syncthreads();
	if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
	__syncthreads();
	double r = IS_MAX ? -DBL_MAX : 0.0;
	if (threadIdx.x < 32) {
		r = threadIdx.x < THREADS / 32 ? sh[threadIdx.x] : r;
		#pragma unroll
		for (int o = 16; o > 0; o >>= 1) {
			const double w = __shfl_xor_sync(0xffffffffu, r, o);
			r = IS_MAX ? fmax(r, w) : r + w;
		}
	}
	return r;   // valid in warp 0
}

template <int THREADS>
__global__ void __launch_bounds__(THREADS)
softmax_max_kernel(const double* __restrict__ y, int64_t n, double* __restrict__ part)
{
	double v = -DBL_MAX;
	for (int64_t i = (int64_t) blockIdx.x * THREADS + threadIdx.x; i < n; i += (int64_t) gridDim.x * THREADS) v = fmax(v, y[i]);
	v = block_reduce<THREADS, true>(v);
	if (threadIdx.x == 0) part[blockIdx.x] = v;
}

// part[0..nparts) -> part[nparts] (one block)
template <int THREADS, bool IS_MAX>
__global__ void __launch_bounds__(THREADS)
softmax_combine_kernel(double* __restrict__ part, int nparts)
{
	double v = IS_MAX ? -DBL_MAX : 0.0;
	for (int i = threadIdx.x; i < nparts; i += THREADS) v = IS_MAX ? fmax(v, part[i]) : v + part[i];
	v = block_reduce<THREADS, IS_MAX>(v);
	if (threadIdx.x == 0) part[nparts] = v;
}

template <int THREADS>
__global__ void __launch_bounds__(THREADS)
softmax_exp_kernel(double* __restrict__ y, int64_t n, const double* __restrict__ maxv, double* __restrict__ part)
{
	const double mx = *maxv;
	double s = 0.0;
	for (int64_t i = (int64_t) blockIdx.x * THREADS + threadIdx.x; i < n; i += (int64_t) gridDim.x * THREADS) {
		const double e = exp(y[i] - mx);
		y[i] = e;
		s += e;
	}
	s = block_reduce<THREADS, false>(s);
	if (threadIdx.x == 0) part[blockIdx.x] = s;
}

template <int THREADS>
__global__ void __launch_bounds__(THREADS)
softmax_scale_kernel(double* __restrict__ y, int64_t n, const double* __restrict__ sum)
{
	const double s = *sum;
	for (int64_t i = (int64_t) blockIdx.x * THREADS + threadIdx.x; i < n; i += (int64_t) gridDim.x * THREADS) y[i] = y[i] / s;
}

// Row-wise softmax (one warp per mask row), the normalisation attention uses; not in the reference.
template <int THREADS>
__global__ void __launch_bounds__(THREADS)
softmax_rows_kernel(const int32_t* __restrict__ ia, double* __restrict__ y, int64_t m)
{
	constexpr int NW = THREADS / 32;
	const int lane = threadIdx.x & 31;
	for (int64_t row = (int64_t) blockIdx.x * NW + (threadIdx.x >> 5); row < m; row += (int64_t) gridDim.x * NW) {
		const int js = ia[row], je = ia[row + 1];
		if (js == je) continue;
		double mx = -DBL_MAX;
		for (int p = js + lane; p < je; p += 32) mx = fmax(mx, y[p]);
		#pragma unroll
		for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
		double s = 0.0;
		for (int p = js + lane; p < je; p += 32) {
			const double e = exp(y[p] - mx);
			y[p] = e;
			s += e;
		}
		#pragma unroll
		for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
		for (int p = js + lane; p < je; p += 32) y[p] = y[p] / s;
	}
}

// ---- fp32 softmax (the "_f" pipeline): y in float, running sums in double so that a global sum over 10^7 entries keeps
// its digits; exp in fp32 (expf), like an "_f" build of the reference's commented-out host softmax would compute it.
template <int THREADS>
__global__ void __launch_bounds__(THREADS)
softmax_max_f32_kernel(const float* __restrict__ y, int64_t n, double* __restrict__ part)
{
	double v = -DBL_MAX;
	for (int64_t i = (int64_t) blockIdx.x * THREADS + threadIdx.x; i < n; i += (int64_t) gridDim.x * THREADS) v = fmax(v, (double) y[i]);
	v = block_reduce<THREADS, true>(v);
	if (threadIdx.x == 0) part[blockIdx.x] = v;
}

template <int THREADS>
__global__ void __launch_bounds__(THREADS)
softmax_exp_f32_kernel(float* __restrict__ y, int64_t n, const double* __restrict__ maxv, double* __restrict__ part)
{
	const float mx = (float) *maxv;
	double s = 0.0;
	for (int64_t i = (int64_t) blockIdx.x * THREADS + threadIdx.x; i < n; i += (int64_t) gridDim.x * THREADS) {
		const float e = expf(y[i] - mx);
		y[i] = e;
		s += (double) e;
	}
	s = block_reduce<THREADS, false>(s);
	if (threadIdx.x == 0) part[blockIdx.x] = s;
}

template <int THREADS>
__global__ void __launch_bounds__(THREADS)
softmax_scale_f32_kernel(float* __restrict__ y, int64_t n, const double* __restrict__ sum)
{
	const float s = (float) *sum;
	for (int64_t i = (int64_t) blockIdx.x * THREADS + threadIdx.x; i < n; i += (int64_t) gridDim.x * THREADS) y[i] = y[i] / s;
}

template <int THREADS>
__global__ void __launch_bounds__(THREADS)
softmax_rows_f32_kernel(const int32_t* __restrict__ ia, float* __restrict__ y, int64_t m)
{
	constexpr int NW = THREADS / 32;
	const int lane = threadIdx.x & 31;
	for (int64_t row = (int64_t) blockIdx.x * NW + (threadIdx.x >> 5); row < m; row += (int64_t) gridDim.x * NW) {
		const int js = ia[row], je = ia[row + 1];
		if (js == je) continue;
		float mx = -FLT_MAX;
		for (int p = js + lane; p < je; p += 32) mx = fmaxf(mx, y[p]);
		#pragma unroll
		for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
		float s = 0.f;
		for (int p = js + lane; p < je; p += 32) {
			const float e = expf(y[p] - mx);
			y[p] = e;
			s += e;
		}
		#pragma unroll
		for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
		for (int p = js + lane; p < je; p += 32) y[p] = y[p] / s;
	}
}

} // namespace sb200
