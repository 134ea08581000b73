// spmm_tf32_blocked.cuh -- fp32 SpMM for BANDED matrices on the 5th-generation tensor cores (tcgen05.mma kind::tf32,
// accumulators in TMEM).  SURVEY 8(f) N3 / north_star kernel (3): "tensor cores only where dense blocks make it a
// genuine contraction" -- this is the one place on the SpMM path where tcgen05 / TMEM applies (there is no FP64 kind).
//
// Idea.  A slab of 128 consecutive rows of a banded matrix (diagonal placement, SURVEY cfg 4) touches one narrow
// window of columns [lo, hi].  Inside that window the slab is treated as a DENSE 128 x W matrix:
//     C[slab, :] = A_dense[128 x W] * B[lo .. hi, :]          (W ~ band width + 128)
// and B[lo .. hi, :] is a contiguous block of the row-major dense operand.  One CTA per slab walks the window in
// chunks of KT = 32 columns:
//   * the 128 threads (one per row) zero a 128 x KT stage and scatter their row's nonzeros of the chunk into it as TF32
//     (cvt.rna), in the UMMA K-major core-matrix layout (8 rows x 16 bytes);
//   * the matching KT rows of B (KT x N floats, contiguous in global memory) are read with coalesced 16-byte loads and
//     transposed 4 x 4 in registers into the same K-major core-matrix layout (the tensor core takes B as N x K, K
//     contiguous; the MN-major form, which would let cp.async drop B rows in unchanged, returned zeros for kind::tf32 with
//     un-swizzled core matrices on this hardware -- scripts/umma_probe.cu, profiles/r2_tf32_probe.md);
//   * one thread issues KT/8 tcgen05.mma (M = 128, N = k, K = 8) that accumulate into N fp32 columns of TMEM, and
//     tcgen05.commit releases the stage through an mbarrier when the tensor core has read it;
//   * producers (8 warps) and the MMA thread (a 9th warp) hand stages over through mbarriers (full: 256 arrivals after a
//     fence.proxy.async; free: tcgen05.commit) -- no CTA-wide barrier in the chunk loop, and the global loads of chunk
//     it+1 are in flight while chunk it is staged;
//   * epilogue: tcgen05.ld (32 lanes x 32 bit, 16 columns per instruction) -> registers -> C.
// Several CTAs are resident per SM (32 KB of shared memory per stage, N of the 512 TMEM columns each), so one slab's
// scatter / epilogue overlaps another's tensor work without warp specialisation.
//
// Numerics: A is rounded to TF32 (10 mantissa bits, round to nearest), B is read by the tensor core as TF32
// (low 13 mantissa bits ignored), products are exact, accumulation is fp32 in TMEM.  Relative error ~1e-3 per term:
// the tests state 4e-3 against the fp32 reference.  Therefore never chosen by AUTO: variant SPMM_B200_BCSR on an fp32 handle.
#pragma once
#include <stdint.h>
#include <limits.h>
#include <cuda_runtime.h>
#include "ptx.cuh"

namespace sb200 {

struct Tf32Plan {
	int2* d_win = nullptr;       // [nslabs] (first, last) column of the slab; first > last: the slab has no nonzero
	int32_t* d_chunk0 = nullptr; // [nslabs + 1] exclusive prefix of the slabs' chunk counts
	int32_t* d_eptr = nullptr;   // [total_chunks + 1] first entry of every (slab, chunk)
	uint2* d_ent = nullptr;      // [nnz] chunk-major copy of A: x = row_in_slab * 32 + column_in_chunk, y = value as TF32 bits
	int nslabs = 0;
	int64_t total_chunks = 0;
	double window_cols = 0;      // sum of the slabs' window widths (dense work = 128 * window_cols * k multiply-adds)
};

constexpr int TF32_M = 128;     // rows per slab = UMMA M
constexpr int TF32_KT = 32;     // columns of A (rows of B) per stage

// per slab: first / last column, number of KT-column chunks of its window
__global__ void __launch_bounds__(TF32_M)
tf32_window_kernel(const int32_t* __restrict__ ia, const int32_t* __restrict__ ja, int m, int2* __restrict__ win,
                   int32_t* __restrict__ nchunks, unsigned long long* __restrict__ total)
{
	__shared__ int s_lo[4], s_hi[4];
	const int r = blockIdx.x * TF32_M + threadIdx.x;
	int lo = INT_MAX, hi = -1;
	if (r < m) {
		const int js = ia[r], je = ia[r + 1];
		if (je > js) { lo = ja[js]; hi = ja[je - 1]; }      // columns are sorted inside a row
	}
	lo = __reduce_min_sync(0xffffffffu, lo);
	hi = __reduce_max_sync(0xffffffffu, hi);
	if ((threadIdx.x & 31) == 0) { s_lo[threadIdx.x >> 5] = lo; s_hi[threadIdx.x >> 5] = hi; }
	__syncthreads();
	if (threadIdx.x == 0) {
		lo = min(min(s_lo[0], s_lo[1]), min(s_lo[2], s_lo[3]));
		hi = max(max(s_hi[0], s_hi[1]), max(s_hi[2], s_hi[3]));
		win[blockIdx.x] = make_int2(lo, hi);
		nchunks[blockIdx.x] = hi >= lo ? (hi - lo) / TF32_KT + 1 : 0;
		if (blockIdx.x == gridDim.x - 1) nchunks[gridDim.x] = 0;
		if (hi >= lo) atomicAdd(total, (unsigned long long) (hi - lo + 1));
	}
}

// Chunk-major copy of A for the tensor-core kernel.  FILL = false: count the entries of every (slab, chunk);
// FILL = true: place them (cursor = running copy of the entry pointers).  The order inside a chunk is whatever the
// atomics give: every entry lands on its own element of the dense tile, so the result does not depend on it.
template <bool FILL>
__global__ void __launch_bounds__(TF32_M)
tf32_bucket_kernel(const int32_t* __restrict__ ia, const int32_t* __restrict__ ja, const float* __restrict__ a, int m,
                   const int2* __restrict__ win, const int32_t* __restrict__ chunk0, int32_t* __restrict__ counter,
                   uint2* __restrict__ ent)
{
	const int r = blockIdx.x * TF32_M + threadIdx.x;
	if (r >= m) return;
	const int lo = win[blockIdx.x].x;
	const int c0 = chunk0[blockIdx.x];
	for (int j = ia[r], je = ia[r + 1]; j < je; j++) {
		const int rel = ja[j] - lo;
		const int chunk = c0 + rel / TF32_KT;
		const int pos = atomicAdd(counter + chunk, 1);
		if constexpr (FILL) {
			uint32_t bits;
			asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(bits) : "f"(a[j]));
			ent[pos] = make_uint2((uint32_t) (threadIdx.x * TF32_KT + rel % TF32_KT), bits);
		}
	}
}

// ---- tcgen05 / TMEM wrappers ------------------------------------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t smem_slot, uint32_t ncols)       // whole warp
{
	asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_slot), "r"(ncols) : "memory");
	asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols)         // whole warp
{
	asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// arrive on an mbarrier once every tcgen05.mma issued so far by this thread has completed (and released its operands)
__device__ __forceinline__ void tc_commit(uint32_t bar)
{
	asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc], TF32 inputs, fp32 accumulate; one thread issues for the CTA
__device__ __forceinline__ void tc_mma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate)
{
	asm volatile("{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n"
	             " tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n}"
	             ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
// 16 consecutive fp32 columns of this thread's TMEM lane (warp w owns lanes 32w .. 32w+31)
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16])
{
	uint32_t r[16];
	asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
	             : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
	               "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
	             : "r"(taddr));
	asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
	#pragma unroll
	for (int i = 0; i < 16; i++) v[i] = __uint_as_float(r[i]);
}
// 16-byte cp.async with zero fill: src_bytes = 0 writes zeros (rows of B beyond the matrix)
__device__ __forceinline__ void cp_async16_zfill(uint32_t dst, const void* src, uint32_t src_bytes)
{
	asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
// bounded spin: a broken pipeline traps instead of hanging the GPU
__device__ __forceinline__ void mbar_wait_bounded(uint32_t bar, uint32_t parity)
{
	for (uint32_t spin = 0; !mbar_try_wait(bar, parity); spin++)
		if (spin > (1u << 24)) __trap();
}

// shared-memory matrix descriptor, no swizzle ("interleave"): 8 x 16-byte core matrices;
// lbo / sbo = byte strides between core matrices (which one is which depends on the operand's major, see below)
__device__ __forceinline__ uint64_t umma_desc(uint32_t smem_addr, uint32_t lbo, uint32_t sbo)
{
	return (uint64_t) ((smem_addr >> 4) & 0x3FFF) | ((uint64_t) ((lbo >> 4) & 0x3FFF) << 16)
	       | ((uint64_t) ((sbo >> 4) & 0x3FFF) << 32) | (1ull << 46);      // bits 46-47: descriptor version 1 (sm_100)
}

template <int N, int STAGES>
struct Tf32Cfg {
	// Both operands K-major, un-swizzled 8 x 16-byte core matrices:
	//   A stage: element (r, kk) at (r/8)*SBO_A + (kk/4)*LBO + (r%8)*16 + (kk%4)*4        (r = row of the slab)
	//   B stage: element (n, kk) at (n/8)*SBO_B + (kk/4)*LBO + (n%8)*16 + (kk%4)*4        (n = dense column)
	// SBO_B is 16 bytes more than a packed group so that the transposing 16-byte stores of a quarter-warp hit 8 different
	// bank groups.
	static constexpr int LBO = 128, SBO_A = (TF32_KT / 4) * 128, SBO_B = (TF32_KT / 4) * 128 + 16;
	static constexpr int A_BYTES = (TF32_M / 8) * SBO_A, B_BYTES = (N / 8) * SBO_B;
	static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
	static constexpr size_t SMEM = (size_t) STAGES * STAGE_BYTES + 128;          // + barriers, TMEM slot, chunk minima
	// instruction descriptor: D = F32 (bits 4-5 = 1), A = B = TF32 (bits 7-9, 10-12 = 2), both K-major (bits 15, 16 = 0),
	// N >> 3 at bits 17-22, M >> 4 at bits 24-28
	static constexpr uint32_t IDESC = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t) (N >> 3) << 17) | ((uint32_t) (TF32_M >> 4) << 24);
};

constexpr int TF32_PRODUCERS = 256;               // warps 0..3: threads 0..127 own a row of the slab each (A scatter); warps 4..7 move B; all 8 read TMEM
constexpr int TF32_BTHREADS = 128;
constexpr int TF32_THREADS = TF32_PRODUCERS + 32;  // + warp 8: one lane issues the MMAs

__device__ __forceinline__ void mbar_arrive(uint32_t bar)
{
	asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}

// 4 rows x 4 columns of B for chunk `kc` into registers: unit = (kk4, n4); a warp reads 512 contiguous bytes of a row
template <int N, int PER_THREAD>
__device__ __forceinline__ void tf32_load_b(const float* __restrict__ B, int kc, int c, int tid, float4 (&br)[PER_THREAD][4])
{
	constexpr int UNITS = (TF32_KT / 4) * (N / 4);
	#pragma unroll
	for (int u = 0; u < PER_THREAD; u++) {
		const int unit = u * TF32_BTHREADS + tid;
		const int n4 = unit % (N / 4), kk4 = unit / (N / 4);
		#pragma unroll
		for (int r = 0; r < 4; r++) {
			const int brow = kc + kk4 * 4 + r;
			br[u][r] = (unit < UNITS && brow < c) ? __ldg(reinterpret_cast<const float4*>(B + (size_t) brow * N) + n4) : make_float4(0.f, 0.f, 0.f, 0.f);
		}
	}
}

template <int N, int STAGES>
__global__ void __launch_bounds__(TF32_THREADS)
spmm_tf32_blocked_kernel(const uint2* __restrict__ ent, const int32_t* __restrict__ eptr, const int32_t* __restrict__ chunk0,
                         const float* __restrict__ B, float* __restrict__ C, const int2* __restrict__ win, int m, int c)
{
	using Cfg = Tf32Cfg<N, STAGES>;
	constexpr int EPT = 3;                          // entries of a chunk a row thread keeps in registers (128 * 3 per chunk)
	constexpr int KT = TF32_KT;
	constexpr int UNITS = (KT / 4) * (N / 4), PER_THREAD = (UNITS + TF32_BTHREADS - 1) / TF32_BTHREADS;
	static_assert(Cfg::STAGE_BYTES % 16 == 0, "stages stay 16-byte aligned");
	extern __shared__ __align__(128) unsigned char tf_smem[];
	const int tid = threadIdx.x, warp = tid >> 5;
	const bool producer = tid < TF32_PRODUCERS, row_thread = tid < TF32_M;
	const int slab = blockIdx.x;
	const int row = slab * TF32_M + (tid & (TF32_M - 1));
	unsigned char* const ctl = tf_smem + (size_t) STAGES * Cfg::STAGE_BYTES;
	const uint32_t smem_u = smem_u32(tf_smem), ctl_u = smem_u32(ctl);
	// no CTA-wide barrier inside the chunk loop: stages are handed over through mbarriers
	const uint32_t bar_full = ctl_u;                  // STAGES barriers, 256 arrivals: every producer has written its part
	const uint32_t bar_free = ctl_u + 8 * STAGES;     // STAGES barriers, tcgen05.commit: the tensor core has read the stage
	const uint32_t bar_done = ctl_u + 16 * STAGES;    // accumulator complete
	volatile uint32_t* const tmem_slot = reinterpret_cast<volatile uint32_t*>(ctl + 16 * STAGES + 8);

	if (warp == 0) tmem_alloc(smem_u32(ctl + 16 * STAGES + 8), N);
	if (tid == 0) {
		#pragma unroll
		for (int s = 0; s < STAGES; s++) { mbar_init(bar_full + 8 * s, TF32_PRODUCERS); mbar_init(bar_free + 8 * s, 1); }
		mbar_init(bar_done, 1);
		asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
	}
	tc_fence_before();
	__syncthreads();
	tc_fence_after();
	const uint32_t tmem = *tmem_slot;

	const int2 w = win[slab];
	const bool any = w.y >= w.x;
	// every chunk of the window is processed (no skipping: in a band nearly every column is occupied); all threads
	// derive the same chunk sequence from the window alone, so nothing has to be agreed on inside the loop
	const int nchunks = any ? (w.y - w.x) / KT + 1 : 0;

	if (producer) {
		int kc = w.x;
		// warps 0..3 stage A, warps 4..7 stage B: the two run side by side and meet at the stage's barrier.  The operands
		// of chunk it+1 are requested before chunk it is handed over, so their latency hides behind the hand-over.
		float4 br[PER_THREAD][4];
		uint2 er[EPT];
		const int btid = tid - TF32_M;
		const int cb = chunk0[slab];
		int e0 = 0, e1 = 0, e2 = 0;                  // entry ranges: [e0, e1) = this chunk, [e1, e2) = the next one
		if (any) {
			if (row_thread) {
				e0 = __ldg(eptr + cb); e1 = __ldg(eptr + cb + 1); e2 = nchunks > 1 ? __ldg(eptr + cb + 2) : e1;
				#pragma unroll
				for (int q = 0; q < EPT; q++)
					if (e0 + q * TF32_M + tid < e1) er[q] = __ldg(ent + e0 + q * TF32_M + tid);
			} else {
				tf32_load_b<N, PER_THREAD>(B, kc, c, btid, br);
			}
		}
		for (int it = 0; it < nchunks; it++, kc += KT) {
			const int stage = it % STAGES;
			if (it >= STAGES) mbar_wait_bounded(bar_free + 8 * stage, (uint32_t) ((it / STAGES) - 1) & 1u);
			unsigned char* const sA = tf_smem + (size_t) stage * Cfg::STAGE_BYTES;
			unsigned char* const sB = sA + Cfg::A_BYTES;
			if (!row_thread) {
				// ---- B: 4 x 4 transpose in registers, four 16-byte stores = half a core matrix (4 columns x 4 k)
				#pragma unroll
				for (int u = 0; u < PER_THREAD; u++) {
					const int unit = u * TF32_BTHREADS + btid;
					if (unit < UNITS) {
						const int n4 = unit % (N / 4), kk4 = unit / (N / 4);
						unsigned char* const dst = sB + (n4 >> 1) * Cfg::SBO_B + kk4 * Cfg::LBO + (n4 & 1) * 64;
						*reinterpret_cast<float4*>(dst) = make_float4(br[u][0].x, br[u][1].x, br[u][2].x, br[u][3].x);
						*reinterpret_cast<float4*>(dst + 16) = make_float4(br[u][0].y, br[u][1].y, br[u][2].y, br[u][3].y);
						*reinterpret_cast<float4*>(dst + 32) = make_float4(br[u][0].z, br[u][1].z, br[u][2].z, br[u][3].z);
						*reinterpret_cast<float4*>(dst + 48) = make_float4(br[u][0].w, br[u][1].w, br[u][2].w, br[u][3].w);
					}
				}
				if (it + 1 < nchunks) tf32_load_b<N, PER_THREAD>(B, kc + KT, c, btid, br);
			} else {
				// ---- A: zero the stage (thread = row), then scatter the chunk's entries from the chunk-major copy of A
				// (coalesced 8-byte loads, already TF32; thread != row here, hence the barrier of the four A warps)
				unsigned char* const myrow = sA + (tid >> 3) * Cfg::SBO_A + (tid & 7) * 16;
				#pragma unroll
				for (int q = 0; q < KT / 4; q++) *reinterpret_cast<uint4*>(myrow + q * Cfg::LBO) = make_uint4(0, 0, 0, 0);
				asm volatile("bar.sync 1, 128;" ::: "memory");
				#pragma unroll
				for (int q = 0; q < EPT; q++) {
					if (e0 + q * TF32_M + tid < e1) {
						const uint32_t r = er[q].x / KT, kk = er[q].x % KT;
						*reinterpret_cast<uint32_t*>(sA + (r >> 3) * Cfg::SBO_A + (kk >> 2) * Cfg::LBO + (r & 7) * 16 + (kk & 3) * 4) = er[q].y;
					}
				}
				for (int e = e0 + EPT * TF32_M + tid; e < e1; e += TF32_M) {        // a chunk with more than 384 entries
					const uint2 v = __ldg(ent + e);
					const uint32_t r = v.x / KT, kk = v.x % KT;
					*reinterpret_cast<uint32_t*>(sA + (r >> 3) * Cfg::SBO_A + (kk >> 2) * Cfg::LBO + (r & 7) * 16 + (kk & 3) * 4) = v.y;
				}
				// next chunk's entries (contiguous with this chunk's) and the pointer after them
				e0 = e1; e1 = e2;
				if (it + 2 < nchunks) e2 = __ldg(eptr + cb + it + 3);
				#pragma unroll
				for (int q = 0; q < EPT; q++)
					if (e0 + q * TF32_M + tid < e1) er[q] = __ldg(ent + e0 + q * TF32_M + tid);
			}
			asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // st.shared data -> visible to the tensor core
			mbar_arrive(bar_full + 8 * stage);
		}
	} else if (tid == TF32_PRODUCERS) {
		// ---- the MMA thread
		for (int it = 0; it < nchunks; it++) {
			const int stage = it % STAGES;
			mbar_wait_bounded(bar_full + 8 * stage, (uint32_t) (it / STAGES) & 1u);
			tc_fence_after();
			const uint32_t sA_u = smem_u + stage * Cfg::STAGE_BYTES, sB_u = sA_u + Cfg::A_BYTES;
			#pragma unroll
			for (int ks = 0; ks < KT / 8; ks++)
				tc_mma_tf32(tmem, umma_desc(sA_u + ks * 2 * Cfg::LBO, Cfg::LBO, Cfg::SBO_A),
				            umma_desc(sB_u + ks * 2 * Cfg::LBO, Cfg::LBO, Cfg::SBO_B), Cfg::IDESC, (it > 0 || ks > 0) ? 1u : 0u);
			tc_commit(bar_free + 8 * stage);
		}
		if (any) tc_commit(bar_done);
	}
	// epilogue: warps 0-3 read the lower half of the columns, warps 4-7 the upper half; thread = row, 16 columns per TMEM load
	if (producer) {
		constexpr int HALF = N / 2;
		const int cbase = (warp >> 2) * HALF;
		if (any) {
			mbar_wait_bounded(bar_done, 0);
			tc_fence_after();
			#pragma unroll 1
			for (int c0 = cbase; c0 < cbase + HALF; c0 += 16) {
				float v[16];
				tmem_ld16(tmem + ((uint32_t) ((warp & 3) * 32) << 16) + c0, v);
				if (row < m) {
					float4* dst = reinterpret_cast<float4*>(C + (size_t) row * N + c0);
					#pragma unroll
					for (int q = 0; q < 4; q++) dst[q] = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
				}
			}
		} else if (row < m) {
			float4* dst = reinterpret_cast<float4*>(C + (size_t) row * N + cbase);
			#pragma unroll 1
			for (int q = 0; q < HALF / 4; q++) dst[q] = make_float4(0.f, 0.f, 0.f, 0.f);
		}
	}
	tc_fence_before();
	__syncthreads();
	if (warp == 0) tmem_dealloc(tmem, N);
}

} // namespace sb200
