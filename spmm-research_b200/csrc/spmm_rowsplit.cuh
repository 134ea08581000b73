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
//   * a group of G lanes owns one row at a time; a lane owns V of the k = G*V dense columns, so one
//     group-wide load of a B row is k*8 contiguous bytes (k=32: G=8, V=4: every 16-byte load
//     instruction of a group covers exactly one 128-byte line);
//   * who owns which rows is the scheduling, and there are three:
//       spmm_rowsplit_kernel         CTA b owns the rows [cta_rows[b], cta_rows[b+1]) of the reference's
//                                    nnz-balanced split (partition.cuh: cta_rows_nnz_kernel, bit-exact with
//                                    lib/parallel_util.h:141-165);
//       spmm_rowsplit_ticket_kernel  one resident wave of persistent warps draws small row chunks from an
//                                    atomic counter: the GPU sweeps A and C front to back, no tail wave;
//       spmm_rowsplit_staged_kernel  the ticketed kernel with the (col, val) stream of a ticket brought into a
//                                    per-warp, double-buffered shared-memory stage by 1-D bulk copies (TMA,
//                                    SASS UBLKCP + mbarrier) one ticket ahead, so that a warp never waits on
//                                    the HBM latency of the A stream and spends no registers or shuffles on it;
//   * spmm_rowsplit_ring_kernel is the experiment VERDICT r1 asked for: gathered B rows land in a per-lane
//     shared-memory ring through cp.async (LDGSTS) so that the gathers in flight are not paid in registers.
//   All of them run the same per-row arithmetic: same bits.
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>
#include "ptx.cuh"

namespace sb200 {

template <int V> struct Frag { double v[V]; };

// L2 eviction-priority hints (createpolicy + ld/st .L2::cache_hint): B rows "evict last", the single-use A stream and C
// "evict first" -- for matrices whose gather window is about the size of L2 (HINT kernels, option L2_HINT)
__device__ __forceinline__ uint64_t l2_policy_evict_last() { uint64_t p; asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p)); return p; }
__device__ __forceinline__ uint64_t l2_policy_evict_first() { uint64_t p; asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p)); return p; }
__device__ __forceinline__ double2 ldg_hint_v2(const double2* p, uint64_t pol)
{
	double2 v;
	asm volatile("ld.global.nc.L2::cache_hint.v2.f64 {%0, %1}, [%2], %3;" : "=d"(v.x), "=d"(v.y) : "l"(p), "l"(pol));
	return v;
}
__device__ __forceinline__ int ldg_hint_s32(const int* p, uint64_t pol)
{
	int v;
	asm volatile("ld.global.nc.L2::cache_hint.s32 %0, [%1], %2;" : "=r"(v) : "l"(p), "l"(pol));
	return v;
}
__device__ __forceinline__ double ldg_hint_f64(const double* p, uint64_t pol)
{
	double v;
	asm volatile("ld.global.nc.L2::cache_hint.f64 %0, [%1], %2;" : "=d"(v) : "l"(p), "l"(pol));
	return v;
}
__device__ __forceinline__ void stg_hint_v2(double2* p, double2 v, uint64_t pol)
{
	asm volatile("st.global.L2::cache_hint.v2.f64 [%0], {%1, %2}, %3;" ::"l"(p), "d"(v.x), "d"(v.y), "l"(pol) : "memory");
}

// Load this lane's V columns of B row `col`.  Columns are dealt to lanes in pairs:
// lane l holds columns {2l, 2l+1} (+ 2G for the second pair when V == 4), so every
// load instruction of the group covers one contiguous span.
template <int G, int V, int LD = 0, int HINT = 0>
__device__ __forceinline__ Frag<V> load_b(const double* __restrict__ B, int col, int lig, int ldb, uint64_t pol = 0)
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
			const double2 t = HINT ? ldg_hint_v2(reinterpret_cast<const double2*>(p) + i * G + lig, pol)
			                       : __ldg(reinterpret_cast<const double2*>(p) + i * G + lig);
			f.v[2 * i] = t.x; f.v[2 * i + 1] = t.y;
		}
	}
	return f;
}

template <int G, int V, int HINT = 0>
__device__ __forceinline__ void store_c(double* __restrict__ C, size_t row, int lig, const double (&acc)[V], int ldc, uint64_t pol = 0)
{
	double* p = C + row * ldc;
	if constexpr (V == 1) {
		p[lig] = acc[0];
	} else {
		// (streaming stores / loads -- st.global.cs for C, ld.global.cs for the A stream -- were measured in round 2 to keep
		// B in L2 on the large-window workloads: no gain on cfg5-large, 5-8 % slower on cfg2 bw 0.3 / 1; plain accesses stay)
		#pragma unroll
		for (int i = 0; i < V / 2; i++) {
			if constexpr (HINT) stg_hint_v2(reinterpret_cast<double2*>(p) + i * G + lig, make_double2(acc[2 * i], acc[2 * i + 1]), pol);
			else reinterpret_cast<double2*>(p)[i * G + lig] = make_double2(acc[2 * i], acc[2 * i + 1]);
		}
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

// The row loop the static and the ticketed kernel share: rows r0 + gid, r0 + gid + stride, ... below r1.
// (col, val) staging: the G lanes of a group read G consecutive nonzeros of their row with ONE coalesced load each
// (the groups of a warp own consecutive rows, so the warp's loads are contiguous in ja / a) and hand them round
// with shuffles; UNROLL B-row gathers are in flight per group.  All lanes of a warp run the same trip counts (the
// shuffles need the full warp).  The per-row sum runs strictly left to right.
template <int G, int V, int UNROLL, int LD, int HINT = 0>
__device__ __forceinline__ void row_sweep(const int32_t* __restrict__ ia, const int32_t* __restrict__ ja, const double* __restrict__ a,
                                          const double* __restrict__ B, double* __restrict__ C,
                                          int r0, int r1, int stride, int gid, int lig, int ldb, int ldc)
{
	static_assert(G % UNROLL == 0 || UNROLL % G == 0, "UNROLL and G must nest");
	constexpr int U = UNROLL < G ? UNROLL : G;
	uint64_t pol_b = 0, pol_s = 0;
	if constexpr (HINT) { pol_b = l2_policy_evict_last(); pol_s = l2_policy_evict_first(); }
	for (int base = r0; base < r1; base += stride) {
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
			if (lig < rem) {
				if constexpr (HINT) { cj = ldg_hint_s32(ja + j + lig, pol_s); ca = ldg_hint_f64(a + j + lig, pol_s); }
				else { cj = __ldg(ja + j + lig); ca = __ldg(a + j + lig); }
			}
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
					if (u + i < rem) bb[i] = load_b<G, V, LD, HINT>(B, cc[i], lig, ldb, pol_b);
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
		if (live) store_c<G, V, HINT>(C, (size_t) row, lig, acc, ldc, pol_s);
	}
}

// ldb / ldc = row strides of B and C in doubles (G*V for a whole operand; larger when the launch
// works on a column slice of a wider B / C, see launch_spmm's L2-aware k slicing)
template <int G, int V, int UNROLL, int THREADS, int LD = 0, int HINT = 0>
__global__ void __launch_bounds__(THREADS)
spmm_rowsplit_kernel(const int32_t* __restrict__ ia, const int32_t* __restrict__ ja, const double* __restrict__ a,
                     const double* __restrict__ B, double* __restrict__ C,
                     const int32_t* __restrict__ cta_rows, int ldb, int ldc)
{
	row_sweep<G, V, UNROLL, LD, HINT>(ia, ja, a, B, C, cta_rows[blockIdx.x], cta_rows[blockIdx.x + 1], THREADS / G,
	                            threadIdx.x / G, threadIdx.x % G, ldb, ldc);
}

// Row tickets.  `ticket` points at two integers: [0] the ticket counter, [1] the number of warps that have left the
// kernel.  The last warp to leave zeroes both, so the next launch of the handle (stream-ordered) -- or the next replay of
// a CUDA graph that captured this launch -- starts from zero without a memset node in between.
__device__ __forceinline__ void ticket_leave(int* __restrict__ ticket, int lane, int nwarps)
{
	if (lane == 0) {
		const int left = atomicAdd(ticket + 1, 1);
		if (left == nwarps - 1) { ticket[0] = 0; ticket[1] = 0; }      // every warp has drawn its last ticket before it counts itself out
	}
}

template <int G, int V, int UNROLL, int THREADS, int LD = 0>
__global__ void __launch_bounds__(THREADS)
spmm_rowsplit_ticket_kernel(const int32_t* __restrict__ ia, const int32_t* __restrict__ ja, const double* __restrict__ a,
                            const double* __restrict__ B, double* __restrict__ C,
                            int m, int rows_per_ticket, int* __restrict__ ticket, int ldb, int ldc)
{
	constexpr int NGW = 32 / G;              // row groups per warp
	const int lane = threadIdx.x & 31;
	for (;;) {
		int t = 0;
		if (lane == 0) t = atomicAdd(ticket, 1);
		t = __shfl_sync(0xffffffffu, t, 0);
		const int64_t r0l = (int64_t) t * rows_per_ticket;
		if (r0l >= m) break;
		row_sweep<G, V, UNROLL, LD>(ia, ja, a, B, C, (int) r0l, (int) min((int64_t) m, r0l + rows_per_ticket), NGW,
		                            lane / G, lane % G, ldb, ldc);
	}
	ticket_leave(ticket, lane, gridDim.x * (THREADS / 32));
}

// ------------------------------------------------------------------------------------------------
// Row tickets with the A stream staged through shared memory by TMA bulk copies.
//
// A ticket = R = (32/G)*PASSES consecutive rows; its nonzeros [ia[r0], ia[r0+R]) are one contiguous range of ja and
// of a.  Each warp runs a four-stage software pipeline over ITS tickets:
//     A  draw a ticket number (atomicAdd; the value is needed an iteration later)
//     B  load the R+1 row pointers of the ticket drawn one iteration ago (lane l holds ia[r0+l])
//     C  lane 0 issues two cp.async.bulk copies (col_ind range, value range; 16-byte aligned windows) of the ticket
//        whose row pointers arrived, into the warp's free stage, completion on that stage's mbarrier
//     D  wait for the stage of the ticket issued one iteration ago and run its rows out of shared memory:
//        every lane of a group reads (col, val) of the same nonzero (a shared-memory broadcast), gathers its V
//        columns of that B row and multiplies; no shuffles, no warp votes, lane groups advance independently.
// The latency of the counter, of ia and of the (ja, a) stream (HBM) is therefore never on a warp's critical path;
// what remains is the L2 gather of B rows.  A ticket whose nonzeros exceed CAP (or an unaligned operand) runs the
// same loop on global-memory loads.  Same per-row arithmetic and order as row_sweep: same bits.
// ------------------------------------------------------------------------------------------------
template <int G, int PASSES, int CAP>
struct StagedCfg {
	static constexpr int R = (32 / G) * PASSES;                      // rows per ticket
	static constexpr int SJ = CAP + 8, SA = CAP + 4;                 // slack: the copy windows are aligned outwards
	static constexpr int BUF_BYTES = SA * 8 + SJ * 4;
	static_assert(CAP % 4 == 0 && R + 1 <= 32, "CAP must keep the stages 16-byte aligned; row pointers live in one warp");
	static constexpr size_t smem(int threads) { return (size_t) (threads / 32) * (2 * BUF_BYTES + 16); }
};

// sj[jj - offj], sa[jj - offa] hold (col, val) of nonzero jj in the warp's shared-memory stage (offsets = start of the
// copy windows).  Every lane of a group reads the same word: a broadcast.
template <int G, int V, int UNROLL, int LD>
__device__ __forceinline__ void staged_row(const int32_t* __restrict__ sj, const double* __restrict__ sa, const int offj, const int offa,
                                           int j, const int je, const double* __restrict__ B, int lig, int ldb, double (&acc)[V])
{
	constexpr int U = UNROLL;
	for (; j < je; j += U) {
		int cc[U];
		Frag<V> bb[U];
		#pragma unroll
		for (int i = 0; i < U; i++)
			if (j + i < je) cc[i] = sj[j + i - offj];
		// all U gathers are issued before the first multiply.  Predicated like row_sweep: with unconditional loads (an
		// unpredicated main loop + tail was tried) ptxas interleaves the multiplies with the loads under the 64-register
		// cap and a group keeps only ~2 B rows in flight.
		#pragma unroll
		for (int i = 0; i < U; i++)
			if (j + i < je) bb[i] = load_b<G, V, LD>(B, cc[i], lig, ldb);
		// the values are read while the gathers fly: they are not live across the issue of the loads
		#pragma unroll
		for (int i = 0; i < U; i++) {
			if (j + i < je) {
				const double vv = sa[j + i - offa];
				#pragma unroll
				for (int v = 0; v < V; v++) acc[v] = fma(vv, bb[i].v[v], acc[v]);
			}
		}
	}
}

template <int G, int V, int UNROLL, int THREADS, int PASSES, int CAP, int LD = 0>
__global__ void __launch_bounds__(THREADS, 4)      // 64 registers
spmm_rowsplit_staged_kernel(const int32_t* __restrict__ ia, const int32_t* __restrict__ ja, const double* __restrict__ a,
                            const double* __restrict__ B, double* __restrict__ C,
                            int m, int* __restrict__ ticket, int ldb, int ldc)
{
	using Cfg = StagedCfg<G, PASSES, CAP>;
	constexpr int NW = THREADS / 32, NGW = 32 / G, R = Cfg::R;
	extern __shared__ __align__(128) unsigned char staged_smem[];
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const int lig = lane % G, gid = lane / G;
	unsigned char* const wbuf = staged_smem + (size_t) warp * 2 * Cfg::BUF_BYTES;
	const uint32_t wbuf_u = smem_u32(wbuf);
	const uint32_t bar_u = smem_u32(staged_smem + (size_t) NW * 2 * Cfg::BUF_BYTES) + warp * 16;
	int* const counter = ticket;
	if (lane == 0) {
		mbar_init(bar_u, 1);
		mbar_init(bar_u + 8, 1);
		asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
		asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
	}
	__syncwarp();

	int tA = 0;                                  // lane 0: the ticket drawn this iteration
	int tB = -1, tC = -1, tD = -1;               // warp-uniform ticket numbers in stages B, C, D (-1: pipeline filling)
	int iaB = 0, iaC = 0, iaD = 0;               // lane l <= R: ia[min(r0 + l, m)] of that ticket
	bool stagedC = false, stagedD = false;
	uint32_t phase = 0;                          // bit b: parity the next wait on stage b uses
	const int ntickets = (m + R - 1) / R;
	for (int it = 0;; it++) {
		tD = tC; iaD = iaC; stagedD = stagedC;
		tC = tB; iaC = iaB;
		tB = it >= 1 ? __shfl_sync(0xffffffffu, tA, 0) : -1;
		// A
		if (lane == 0) tA = atomicAdd(counter, 1);
		// B
		iaB = 0;
		if (tB >= 0 && tB < ntickets && lane <= R) iaB = __ldg(ia + min(tB * R + lane, m));
		// C
		stagedC = false;
		if (tC >= 0 && tC < ntickets) {
			const int j0 = __shfl_sync(0xffffffffu, iaC, 0), j1 = __shfl_sync(0xffffffffu, iaC, R);
			const int n = j1 - j0;
			if (n > 0 && n <= CAP) {
				stagedC = true;
				if (lane == 0) {
					const int b = it & 1;
					const int jj = j0 & ~3, ja_n = ((j1 + 3) & ~3) - jj;        // col_ind window, multiples of 4 entries
					const int jv = j0 & ~1, a_n = ((j1 + 1) & ~1) - jv;         // value window, multiples of 2 entries
					const uint32_t dst = wbuf_u + b * Cfg::BUF_BYTES;
					mbar_expect_tx(bar_u + b * 8, (uint32_t) (a_n * 8 + ja_n * 4));
					bulk_g2s(dst, a + jv, (uint32_t) a_n * 8, bar_u + b * 8);
					bulk_g2s(dst + Cfg::SA * 8, ja + jj, (uint32_t) ja_n * 4, bar_u + b * 8);
				}
			}
		}
		// D
		if (tD < 0) continue;
		if (tD >= ntickets) break;
		const int b = (it - 1) & 1;
		const int r0 = tD * R;
		const uint32_t bar = bar_u + b * 8;
		const double* sa = reinterpret_cast<const double*>(wbuf + (size_t) b * Cfg::BUF_BYTES);
		const int32_t* sj = reinterpret_cast<const int32_t*>(wbuf + (size_t) b * Cfg::BUF_BYTES + Cfg::SA * 8);
		// Common case: the whole ticket [0, R) is in flight since stage C.  A ticket of more than CAP nonzeros (or of
		// none) was not prefetched: it goes through the same stage in greedy sub-ranges of whole rows that fit
		// (the launcher guarantees max_row <= CAP), copied here and waited for on the spot.
		int rs = 0, re = R;
		bool ready = stagedD;
		do {
			const int j0 = __shfl_sync(0xffffffffu, iaD, rs);
			if (!ready) {
				const unsigned fit = __ballot_sync(0xffffffffu, lane > rs && lane <= R && iaD - j0 <= CAP);
				re = fit ? 31 - __clz(fit) : rs + 1;
				const int j1 = __shfl_sync(0xffffffffu, iaD, re);
				ready = j1 > j0;
				if (ready && lane == 0) {
					const int jj = j0 & ~3, ja_n = ((j1 + 3) & ~3) - jj;
					const int jv = j0 & ~1, a_n = ((j1 + 1) & ~1) - jv;
					const uint32_t dst = wbuf_u + b * Cfg::BUF_BYTES;
					mbar_expect_tx(bar, (uint32_t) (a_n * 8 + ja_n * 4));
					bulk_g2s(dst, a + jv, (uint32_t) a_n * 8, bar);
					bulk_g2s(dst + Cfg::SA * 8, ja + jj, (uint32_t) ja_n * 4, bar);
				}
			}
			if (ready) {
				while (!mbar_try_wait(bar, (phase >> b) & 1)) {}
				phase ^= 1u << b;
			}
			#pragma unroll 1
			for (int p = rs; p < re; p += NGW) {
				const int lr = min(p + gid, R);                      // ticket-local row of my group (R: none)
				const int js = __shfl_sync(0xffffffffu, iaD, lr);
				const int je_row = __shfl_sync(0xffffffffu, iaD, min(lr + 1, R));      // every lane shuffles: lr differs per group
				const int je = lr < re ? je_row : js;
				double acc[V];
				#pragma unroll
				for (int v = 0; v < V; v++) acc[v] = 0.0;
				staged_row<G, V, UNROLL, LD>(sj, sa, j0 & ~3, j0 & ~1, js, je, B, lig, ldb, acc);
				if (lr < re && r0 + lr < m) store_c<G, V>(C, (size_t) (r0 + lr), lig, acc, ldc);
				__syncwarp();
			}
			rs = re;
			ready = false;
		} while (rs < R);
		__syncwarp();                            // every lane is done with stage b before iteration it+1 refills it
	}
	ticket_leave(ticket, lane, gridDim.x * (THREADS / 32));
}

// ------------------------------------------------------------------------------------------------
// Experiment: gathered B rows through a shared-memory ring ("the one experiment not yet run").
// The ticketed row loop, but a group's gathers are cp.async (LDGSTS, 16 bytes per lane and piece) into a per-lane
// ring of 2 stages x H nonzeros instead of loads into registers: a stage is being filled while the previous one is
// consumed, so 4..8 B rows are in flight per group with 8 fewer live doubles per lane.  Each lane reads back exactly
// the bytes it copied, so cp.async.wait_group is all the synchronisation needed.  The ring is drained at the end
// of every row.  Same arithmetic and order: same bits.
// ------------------------------------------------------------------------------------------------
template <int G, int V, int H, int THREADS>
__global__ void __launch_bounds__(THREADS)
spmm_rowsplit_ring_kernel(const int32_t* __restrict__ ia, const int32_t* __restrict__ ja, const double* __restrict__ a,
                          const double* __restrict__ B, double* __restrict__ C,
                          int m, int rows_per_ticket, int* __restrict__ ticket, int ldb, int ldc)
{
	static_assert(G % H == 0 && V % 2 == 0, "a chunk of G nonzeros is a whole number of stages");
	constexpr int NGW = 32 / G, P = V / 2;                  // 16-byte pieces per lane and B row
	extern __shared__ __align__(128) unsigned char ring_smem[];
	const int lane = threadIdx.x & 31, lig = lane % G, gid = lane / G;
	// layout [stage][slot][piece][thread] of 16 bytes: consecutive threads, consecutive 16-byte words (no bank conflicts)
	double2* const ring = reinterpret_cast<double2*>(ring_smem) + threadIdx.x;
	const uint32_t ring_u = smem_u32(ring);
	auto slot_off = [](int stage, int i, int p) { return ((stage * H + i) * P + p) * THREADS; };
	int* const counter = ticket;
	for (;;) {
		int t = 0;
		if (lane == 0) t = atomicAdd(counter, 1);
		t = __shfl_sync(0xffffffffu, t, 0);
		const int64_t r0l = (int64_t) t * rows_per_ticket;
		if (r0l >= m) break;
		const int r0 = (int) r0l, r1 = (int) min((int64_t) m, r0l + rows_per_ticket);
		for (int base = r0; base < r1; base += NGW) {
			const int row = base + gid;
			const bool live = row < r1;
			int j = 0, je = 0;
			if (live) { j = __ldg(ia + row); je = __ldg(ia + row + 1); }
			double acc[V];
			#pragma unroll
			for (int v = 0; v < V; v++) acc[v] = 0.0;
			double pv[H];                                        // values of the stage in flight
			int pn = 0;                                          // its live nonzeros for my row (0: nothing in flight)
			int stage = 0;
			auto consume = [&](int st) {
				#pragma unroll
				for (int i = 0; i < H; i++) {
					if (i < pn) {
						#pragma unroll
						for (int p = 0; p < P; p++) {
							const double2 bb = ring[slot_off(st, i, p)];
							acc[2 * p] = fma(pv[i], bb.x, acc[2 * p]);
							acc[2 * p + 1] = fma(pv[i], bb.y, acc[2 * p + 1]);
						}
					}
				}
			};
			bool inflight = false;                               // warp-uniform: a stage has been committed and not consumed
			while (true) {
				const int rem = je - j;
				const int rem_warp = __reduce_max_sync(0xffffffffu, rem);
				if (rem_warp <= 0) break;
				int cj = 0;
				double ca = 0.0;
				if (lig < rem) { cj = __ldg(ja + j + lig); ca = __ldg(a + j + lig); }
				const int cnt = min(rem_warp, G);
				#pragma unroll
				for (int u = 0; u < G; u += H) {
					if (u >= cnt) break;                             // warp-uniform
					int cc[H];
					double vv[H];
					#pragma unroll
					for (int i = 0; i < H; i++) {
						cc[i] = __shfl_sync(0xffffffffu, cj, u + i, G);
						vv[i] = shfl_f64<G>(ca, u + i);
					}
					#pragma unroll
					for (int i = 0; i < H; i++) {
						if (u + i < rem) {
							const double2* src = reinterpret_cast<const double2*>(B + (size_t) cc[i] * ldb) + lig;
							#pragma unroll
							for (int p = 0; p < P; p++)
								cp_async16(ring_u + (uint32_t) slot_off(stage, i, p) * 16, src + p * G);
						}
					}
					cp_async_commit();
					if (inflight) {                                  // the other stage: everything but the newest group is done
						cp_async_wait<1>();
						consume(stage ^ 1);
					}
					#pragma unroll
					for (int i = 0; i < H; i++) pv[i] = vv[i];
					pn = max(0, min(H, rem - u));
					inflight = true;
					stage ^= 1;
				}
				j += G;
			}
			if (inflight) {
				cp_async_wait<0>();
				consume(stage ^ 1);
			}
			if (live) store_c<G, V>(C, (size_t) row, lig, acc, ldc);
		}
	}
	ticket_leave(ticket, lane, gridDim.x * (THREADS / 32));
}

// Any k: one warp per row, lanes own columns, column chunks of 32 (A re-read per chunk
// out of L1/L2).  Used for k values without a specialised instantiation and for operands that are
// not 16-byte aligned (the specialised kernels move B and C as double2).
template <int THREADS>
__global__ void __launch_bounds__(THREADS)
spmm_rowsplit_generic_kernel(const int32_t* __restrict__ ia, const int32_t* __restrict__ ja, const double* __restrict__ a,
                             const double* __restrict__ B, double* __restrict__ C,
                             const int32_t* __restrict__ cta_rows, int k, int ldb, int ldc)
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
					acc = fma(__ldg(a + j), __ldg(B + (size_t) __ldg(ja + j) * ldb + n), acc);
				C[(size_t) row * ldc + n] = acc;
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

// The same for `rows` consecutive rows of B out of a column-major operand whose columns are `pitch` elements apart:
// x[n*pitch + j] -> B[j*k + n], j < rows (the row-band host pipeline uploads B a few rows at a time).
__global__ void transpose_b_rows_kernel(const double* __restrict__ x, double* __restrict__ B, int64_t pitch, int64_t rows, int k)
{
	__shared__ double tile[32][33];
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
