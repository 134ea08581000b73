// spmm_bcsr.cuh -- blocked-CSR SpMM (north_star kernel (3)) with 8x1 blocks.
//
// The reference has no BCSR of its own: the only definition is the cuSPARSE BSR wrapper
// SpmvBsrData{blockDim, nnzb, rowPtr, colInd, values} filled by cusparseDcsr2bsr with
// CUSPARSE_DIRECTION_COLUMN (benchmark_code/GPU/spmv_code_cusparse-11.x/inc/spmv_operator.hpp:56-62,
// src/CuSPARSE.cu:41,104-172): block rows, explicit zeros for fill, column-major inside a block.
// Parity for this format is pinned through the SpMM result (compute_csr).
//
// Layout here (same three arrays, rectangular block 8 rows x 1 column):
//   mb = ceil(m/8) block rows;  brow_ptr[mb+1];  bcol[nnzb] = sorted distinct columns of the
//   8 rows;  bval[nnzb][8] = the column's 8 values (zeros where a row has no entry; this IS
//   column-major inside the block);  bmask[nnzb] = which of the 8 rows really hold an entry.
// Why 8x1: the FP64 tensor instruction on sm_100a is mma.sync.m8n8k4 (tcgen05 has no FP64
// kind), D(8x8) += A(8x4) * B(4x8).  M = 8 matrix rows, K = 4 block columns -- and nothing
// requires the 4 columns to be adjacent, each K slot gathers its own B row.  8x1 blocks keep
// the fill ratio at its maximum for an 8-row slab (cfg4: 0.40 against 0.32 for aligned 8x4).
//
// Two kernels over the same arrays:
//   spmm_bcsr_dmma_kernel   tensor pipe: one DMMA per (4 block columns, 8 dense columns);
//                           executes 8 FMAs per stored slot, so it pays only when blocks are
//                           dense ("a genuine contraction");
//   spmm_bcsr_simt_kernel   FP64 FMA pipe: one B-row gather per block column feeds up to 8 row
//                           accumulators held in registers; absent rows are skipped through
//                           the (warp-uniform) mask, so it executes only the real 2*nnz*k flops.
// Both cut the L1/L2 gather traffic of the CSR kernels by the block fill (nnz / nnzb).
#pragma once
#include <stdint.h>
#include <limits.h>
#include <cuda_runtime.h>
#include "spmm_rowsplit.cuh"
#include "ptx.cuh"

namespace sb200 {

struct BcsrPlan {
	int mb = 0;
	int64_t nnzb = 0;
	int32_t* d_brow = nullptr;
	int32_t* d_bcol = nullptr;
	double* d_bval = nullptr;
	uint8_t* d_bmask = nullptr;
};

// ---- CSR -> BCSR(8x1): 8 lanes per block row run an 8-way merge of the sorted column lists ----
// FILL = false: count distinct columns per block row into brow[br]; FILL = true: write the blocks.
template <bool FILL>
__global__ void __launch_bounds__(256)
bcsr_convert_kernel(const int32_t* __restrict__ ia, const int32_t* __restrict__ ja, const double* __restrict__ a,
                    int m, int mb, int32_t* __restrict__ brow,
                    int32_t* __restrict__ bcol, double* __restrict__ bval, uint8_t* __restrict__ bmask)
{
	const int t = blockIdx.x * blockDim.x + threadIdx.x;
	const int br = t >> 3, r = t & 7;
	if (br >= mb) return;                                   // whole 8-lane groups leave together
	const unsigned gmask = 0xFFu << ((threadIdx.x & 31) & ~7);
	const int gshift = (threadIdx.x & 31) & ~7;
	const int row = br * 8 + r;
	int p = 0, pe = 0;
	if (row < m) { p = ia[row]; pe = ia[row + 1]; }
	int out = FILL ? brow[br] : 0;
	int count = 0;
	while (true) {
		const int cur = p < pe ? ja[p] : INT_MAX;
		int mn = cur;
		mn = min(mn, __shfl_xor_sync(gmask, mn, 1));
		mn = min(mn, __shfl_xor_sync(gmask, mn, 2));
		mn = min(mn, __shfl_xor_sync(gmask, mn, 4));
		if (mn == INT_MAX) break;
		const bool hit = cur == mn;
		if (FILL) {
			const unsigned bal = __ballot_sync(gmask, hit);
			bval[(size_t) out * 8 + r] = hit ? a[p] : 0.0;
			if (r == 0) { bcol[out] = mn; bmask[out] = (uint8_t) ((bal >> gshift) & 0xFFu); }
			out++;
		}
		if (hit) p++;
		count++;
	}
	if (!FILL && r == 0) brow[br] = count;
}

// ---- tensor-pipe kernel -------------------------------------------------------------------------
__device__ __forceinline__ void dmma_m8n8k4(double& d0, double& d1, double a, double b)
{
	asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
	             : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

// One warp per block row, all K dense columns (K % 16 == 0).  Fragment roles of lane (g = lane/4,
// t = lane%4):  A[g][t] = bval[e+t][g];  B[t][n] with n = g;  D[g][2t], D[g][2t+1].
// Dense columns are dealt to the two DMMAs of a 16-column chunk in an interleaved way (DMMA #0
// takes the even columns, #1 the odd ones) so that a lane's B operand for both is ONE 16-byte
// load and its four D values are four consecutive columns of C.
template <int K, int WARPS>
__global__ void __launch_bounds__(WARPS * 32)
spmm_bcsr_dmma_kernel(const int32_t* __restrict__ brow, const int32_t* __restrict__ bcol, const double* __restrict__ bval,
                      const double* __restrict__ B, double* __restrict__ C, int m, int mb)
{
	static_assert(K % 16 == 0, "the tensor-pipe kernel needs k to be a multiple of 16");
	constexpr int Q = K / 16;
	const int lane = threadIdx.x & 31;
	const int g = lane >> 2, t = lane & 3;
	const int br = blockIdx.x * WARPS + (threadIdx.x >> 5);
	if (br >= mb) return;
	const int bs = __ldg(brow + br), be = __ldg(brow + br + 1);
	double acc[Q][4];
	#pragma unroll
	for (int q = 0; q < Q; q++) { acc[q][0] = acc[q][1] = acc[q][2] = acc[q][3] = 0.0; }

	for (int e = bs; e < be; e += 4) {
		const int idx = e + t;
		const bool valid = idx < be;
		const int col = __ldg(bcol + (valid ? idx : be - 1));
		const double av = valid ? __ldg(bval + (size_t) idx * 8 + g) : 0.0;
		const double2* bp = reinterpret_cast<const double2*>(B + (size_t) col * K) + g;
		double2 bb[Q];
		#pragma unroll
		for (int q = 0; q < Q; q++) bb[q] = __ldg(bp + q * 8);
		#pragma unroll
		for (int q = 0; q < Q; q++) {
			dmma_m8n8k4(acc[q][0], acc[q][2], av, bb[q].x);     // columns 16q + 4t + {0, 2}
			dmma_m8n8k4(acc[q][1], acc[q][3], av, bb[q].y);     // columns 16q + 4t + {1, 3}
		}
	}
	const int row = br * 8 + g;
	if (row < m) {
		double2* cp = reinterpret_cast<double2*>(C + (size_t) row * K) + 2 * t;
		#pragma unroll
		for (int q = 0; q < Q; q++) {
			cp[q * 8] = make_double2(acc[q][0], acc[q][1]);
			cp[q * 8 + 1] = make_double2(acc[q][2], acc[q][3]);
		}
	}
}

// ---- tensor-pipe kernel, operands gathered by the TMA engine ------------------------------------
// ncu on the kernel above (profiles/r1_bcsr_cfg4.md): the m8n8k4 fragment layout makes the gathered
// row index the fastest-varying lane index, so every quarter-warp of a fragment load touches four
// different B rows -> 16 L1 wavefronts per 512 B, L1 data pipe 73 % busy, 104 registers, 16 warps
// per SM, tensor pipe 34 %.  Here the four B rows (and the 32 A values) of a step are fetched with
// 1-D bulk copies (cp.async.bulk -> UBLKCP) into a per-warp ring of STAGES shared-memory stages,
// completion on an mbarrier, and fragments are read from shared memory without bank conflicts
// (row stride = 8K + 32 bytes).  A warp owns `brw` consecutive block rows and keeps the ring
// flowing across their boundaries; no CTA-wide barrier is used.
template <int K, int WARPS, int STAGES>
struct BcsrTmaCfg {
	static constexpr int ROW_BYTES = K * 8;
	static constexpr int ROW_STRIDE = ROW_BYTES + 32;
	static constexpr int STAGE_BYTES = 4 * ROW_STRIDE + 256;
	static constexpr size_t SMEM = (size_t) WARPS * STAGES * STAGE_BYTES + (size_t) WARPS * STAGES * 8;
};

template <int K, int WARPS, int STAGES>
__global__ void __launch_bounds__(WARPS * 32)
spmm_bcsr_dmma_tma_kernel(const int32_t* __restrict__ brow, const int32_t* __restrict__ bcol, const double* __restrict__ bval,
                          const double* __restrict__ B, double* __restrict__ C, int m, int mb, int brw)
{
	using Cfg = BcsrTmaCfg<K, WARPS, STAGES>;
	constexpr int Q = K / 16;
	extern __shared__ __align__(128) unsigned char bcsr_smem[];
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const int g = lane >> 2, t = lane & 3;
	// Warp gw owns block rows gw, gw + W, gw + 2W, ... (W = warps in the grid): CTAs that run at the
	// same time then work on ADJACENT block rows, whose gathered B rows overlap on banded matrices
	// and stay in L2 (contiguous ownership made the live B window 8x larger and DRAM reads 3.5x).
	const int W = gridDim.x * WARPS;
	const int br0 = blockIdx.x * WARPS + warp;
	const int br1 = min(br0 + (brw - 1) * W + 1, mb);      // exclusive bound of the strided walk
	if (br0 >= mb) return;
	unsigned char* ring = bcsr_smem + (size_t) warp * STAGES * Cfg::STAGE_BYTES;
	const uint32_t ring_u = smem_u32(ring);
	const uint32_t bar_u = smem_u32(bcsr_smem + (size_t) WARPS * STAGES * Cfg::STAGE_BYTES) + warp * STAGES * 8;
	if (lane == 0) {
		#pragma unroll
		for (int s = 0; s < STAGES; s++) mbar_init(bar_u + s * 8, 1);
		asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
		asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
	}
	__syncwarp();

	// producer cursor over the flattened (block row, step) stream; empty block rows produce nothing
	int p_br = br0, p_e = 0, p_be = 0;
	while (p_br < br1) {
		p_e = __ldg(brow + p_br); p_be = __ldg(brow + p_br + 1);
		if (p_e < p_be) break;
		p_br += W;
	}
	int p_col = 0;                                        // lanes 0..3: column of slot `lane` of the next step
	if (p_br < br1 && lane < 4) p_col = __ldg(bcol + min(p_e + lane, p_be - 1));
	int pit = 0;
	auto produce = [&]() {
		const int stage = pit % STAGES;
		const uint32_t dst = ring_u + stage * Cfg::STAGE_BYTES, bar = bar_u + stage * 8;
		// one lane issues all five copies (per-lane issue makes ptxas loop over the active lanes to
		// move each lane's operands into uniform registers: ~100 instructions per step, see profiles/)
		const int c0 = __shfl_sync(0xffffffffu, p_col, 0), c1 = __shfl_sync(0xffffffffu, p_col, 1);
		const int c2 = __shfl_sync(0xffffffffu, p_col, 2), c3 = __shfl_sync(0xffffffffu, p_col, 3);
		if (lane == 0) {
			mbar_expect_tx(bar, 4 * Cfg::ROW_BYTES + 256);
			bulk_g2s(dst, B + (size_t) c0 * K, Cfg::ROW_BYTES, bar);
			bulk_g2s(dst + Cfg::ROW_STRIDE, B + (size_t) c1 * K, Cfg::ROW_BYTES, bar);
			bulk_g2s(dst + 2 * Cfg::ROW_STRIDE, B + (size_t) c2 * K, Cfg::ROW_BYTES, bar);
			bulk_g2s(dst + 3 * Cfg::ROW_STRIDE, B + (size_t) c3 * K, Cfg::ROW_BYTES, bar);
			bulk_g2s(dst + 4 * Cfg::ROW_STRIDE, bval + (size_t) p_e * 8, 256, bar);
		}
		pit++;
		p_e += 4;
		if (p_e >= p_be) {
			p_br += W;
			while (p_br < br1) {
				p_e = __ldg(brow + p_br); p_be = __ldg(brow + p_br + 1);
				if (p_e < p_be) break;
				p_br += W;
			}
		}
		if (p_br < br1 && lane < 4) p_col = __ldg(bcol + min(p_e + lane, p_be - 1));
	};
	#pragma unroll 1
	for (int s = 0; s < STAGES - 1; s++)
		if (p_br < br1) produce();

	int it = 0;
	for (int br = br0; br < br1; br += W) {
		const int bs = __ldg(brow + br), be = __ldg(brow + br + 1);
		double acc[Q][4];
		#pragma unroll
		for (int q = 0; q < Q; q++) { acc[q][0] = acc[q][1] = acc[q][2] = acc[q][3] = 0.0; }
		for (int e = bs; e < be; e += 4) {
			if (p_br < br1) produce();                       // refills the stage consumed one step ago
			const int stage = it % STAGES;
			const uint32_t parity = (it / STAGES) & 1;
			while (!mbar_try_wait(bar_u + stage * 8, parity)) {}
			const unsigned char* st = ring + (size_t) stage * Cfg::STAGE_BYTES;
			const double av = (e + t < be) ? reinterpret_cast<const double*>(st + 4 * Cfg::ROW_STRIDE)[t * 8 + g] : 0.0;
			const double2* bp = reinterpret_cast<const double2*>(st + t * Cfg::ROW_STRIDE) + g;
			#pragma unroll
			for (int q = 0; q < Q; q++) {
				const double2 bb = bp[q * 8];
				dmma_m8n8k4(acc[q][0], acc[q][2], av, bb.x);
				dmma_m8n8k4(acc[q][1], acc[q][3], av, bb.y);
			}
			__syncwarp();                                    // every lane has read the stage
			it++;
		}
		const int row = br * 8 + g;
		if (row < m) {
			double2* cp = reinterpret_cast<double2*>(C + (size_t) row * K) + 2 * t;
			#pragma unroll
			for (int q = 0; q < Q; q++) {
				cp[q * 8] = make_double2(acc[q][0], acc[q][1]);
				cp[q * 8 + 1] = make_double2(acc[q][2], acc[q][3]);
			}
		}
	}
}

// ---- FP64-FMA kernel with register reuse of the gathered B row ----------------------------------
// One warp per block row; lane owns V = K/32 dense columns (K % 32 == 0) of all 8 rows.
// The warp reads 32 block columns' (col, mask) and 4 block columns' 32 values with coalesced
// loads and hands them round with shuffles.
template <int K, int WARPS>
__global__ void __launch_bounds__(WARPS * 32)
spmm_bcsr_simt_kernel(const int32_t* __restrict__ brow, const int32_t* __restrict__ bcol, const double* __restrict__ bval,
                      const uint8_t* __restrict__ bmask, const double* __restrict__ B, double* __restrict__ C, int m, int mb)
{
	static_assert(K % 32 == 0, "the blocked FMA kernel needs k to be a multiple of 32");
	constexpr int V = K / 32;
	constexpr int VV = V >= 2 ? 2 : 1;                       // doubles per load
	const int lane = threadIdx.x & 31;
	const int br = blockIdx.x * WARPS + (threadIdx.x >> 5);
	if (br >= mb) return;
	const int bs = __ldg(brow + br), be = __ldg(brow + br + 1);
	double acc[8][V];
	#pragma unroll
	for (int r = 0; r < 8; r++)
		#pragma unroll
		for (int v = 0; v < V; v++) acc[r][v] = 0.0;

	for (int e0 = bs; e0 < be; e0 += 32) {
		const int n32 = min(32, be - e0);
		int my_col = 0, my_mask = 0;
		if (lane < n32) { my_col = __ldg(bcol + e0 + lane); my_mask = __ldg(bmask + e0 + lane); }
		for (int e4 = 0; e4 < n32; e4 += 4) {
			// 4 block columns x 8 rows = 32 values, one per lane
			const int vidx = (e0 + e4) * 8 + lane;
			const double my_val = (e4 * 8 + lane < n32 * 8) ? __ldg(bval + (size_t) vidx) : 0.0;
			int cc[4], mm[4];
			double bb[4][V];
			#pragma unroll
			for (int i = 0; i < 4; i++) {
				cc[i] = __shfl_sync(0xffffffffu, my_col, e4 + i);
				mm[i] = __shfl_sync(0xffffffffu, my_mask, e4 + i);   // lanes past n32 hold mask 0
			}
			#pragma unroll
			for (int i = 0; i < 4; i++) {
				if (mm[i]) {
					const double* bp = B + (size_t) cc[i] * K;
					if constexpr (VV == 2) {
						#pragma unroll
						for (int v2 = 0; v2 < V / 2; v2++) {
							double2 tmp = __ldg(reinterpret_cast<const double2*>(bp) + v2 * 32 + lane);
							bb[i][2 * v2] = tmp.x; bb[i][2 * v2 + 1] = tmp.y;
						}
					} else {
						bb[i][0] = __ldg(bp + lane);
					}
				}
			}
			#pragma unroll
			for (int i = 0; i < 4; i++) {
				#pragma unroll
				for (int r = 0; r < 8; r++) {
					if (mm[i] & (1 << r)) {                             // warp-uniform
						const double val = shfl_f64<32>(my_val, i * 8 + r);
						#pragma unroll
						for (int v = 0; v < V; v++) acc[r][v] = fma(val, bb[i][v], acc[r][v]);
					}
				}
			}
		}
	}
	#pragma unroll
	for (int r = 0; r < 8; r++) {
		const int row = br * 8 + r;
		if (row < m) {
			double* cp = C + (size_t) row * K;
			if constexpr (VV == 2) {
				#pragma unroll
				for (int v2 = 0; v2 < V / 2; v2++)
					reinterpret_cast<double2*>(cp)[v2 * 32 + lane] = make_double2(acc[r][2 * v2], acc[r][2 * v2 + 1]);
			} else {
				cp[lane] = acc[r][0];
			}
		}
	}
}

} // namespace sb200
