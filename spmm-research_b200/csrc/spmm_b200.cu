// spmm_b200.cu -- the C ABI of libspmm_b200.so (include/spmm_b200.h).
//
// Host side of the drop-in for Matrix_Format::spmm / csr_to_format
// (ref: benchmark_code/CPU/AMD/spmv_code_bench/spmv_kernel.h:9-30).  Owns the
// device copy of A, the CTA work lists, optional device copies of B and C, and
// launches the hand-written sm_100a kernels in spmm_rowsplit.cuh / spmm_merge.cuh.
// There is no CPU fallback: every compute entry point needs a CC 10.x device.
#include "../../include/spmm_b200.h"
#include "partition.cuh"
#include "spmm_rowsplit.cuh"
#include "spmm_merge.cuh"
#include "scan.cuh"
#include "spmm_bcsr.cuh"
#include "coo_to_csr.cuh"
#include "check_accuracy.cuh"

#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <limits.h>
#include <math.h>
#include <vector>
#include <chrono>
#include <algorithm>

using namespace sb200;

static thread_local char g_cuda_err[512] = "";

#define CU_TRY(expr)                                                                              \
	do {                                                                                          \
		cudaError_t _e = (expr);                                                                  \
		if (_e != cudaSuccess) {                                                                  \
			snprintf(g_cuda_err, sizeof(g_cuda_err), "%s:%d: %s -> %s", __FILE__, __LINE__, #expr, \
			         cudaGetErrorString(_e));                                                     \
			return (_e == cudaErrorMemoryAllocation) ? SPMM_B200_ENOMEM : SPMM_B200_ECUDA;        \
		}                                                                                         \
	} while (0)

static constexpr int THREADS = 256;

// Tuning knobs.  Read from the environment ONCE per handle (at creation), never on the hot path; a caller changes
// them afterwards with spmm_b200_set_option(h, "NAME", value).  Names = the SPMM_B200_* variables without prefix.
struct Options {
	int ticket = -1;           // TICKET: rows per ticket of the persistent row-split kernel; -1 = rule, 0 = static ranges
	int ticket_ctas = 4;       // TICKET_CTAS: CTAs per SM of a ticketed grid
	int ticket_min_nnz = 2 << 20; // TICKET_MIN_NNZ: the rule takes tickets from this many nonzeros on
	int stage = 0;             // STAGE: 1 = A stream through shared memory by TMA (staged kernel), 0 = off, -1 = rule, 2 = B-row ring experiment
	int const_ld = 1;          // CONST_LD: compile-time row stride when the operand is whole
	int unroll = 4;            // UNROLL: B-row gathers in flight per lane group
	int kslice = 0;            // KSLICE: 0 = measure, -1 = never, w = force
	int kslice_min_mb = 64;    // KSLICE_MIN_MB: measure when the touched window of B (col_span x 8k bytes) exceeds this
	int pipeline_cols = -1;    // PIPELINE_COLS: host pipeline slice width; 0 = off, -1 = choose
	int pipeline_min_kb = 8 << 10;
	int direct_store = 1;      // DIRECT_STORE
	int rowband = -1;          // ROWBAND: host pipeline over row blocks of a banded matrix; -1 = by the transfer model, 0 = off, R = force R equal blocks, -2 = force the graded blocks
	int rowband_direct = 0;    // ROWBAND_DIRECT: 1 = row-band blocks store C straight into mapped host memory (measured slower than one contiguous copy per block)
	int multi_route = -1;      // MULTI_ROUTE (spmm_b200_run_multi): which devices download C; -1 = measure the box once, 0 = every device its own rows, mask = bit g set: device g downloads
	int upload_window = 1;     // UPLOAD_WINDOW: host-buffer calls upload only the rows of B the matrix's nonzeros reference (first..last touched column)
	int ctas_per_sm = 0;       // CTAS_PER_SM: 0 = by k
	int merge_maxrow = 256;    // MERGE_MAXROW
	int bcsr_fill_pct = 30;    // BCSR_FILL_PCT
	int bcsr_tma = 1, bcsr_stages = 2, bcsr_brw = 0;
	int l2_hint = 0;           // L2_HINT: 1 = static row-split launches carry L2 eviction hints (B evict-last, A / C evict-first)
	int tf32_stages = 2;       // TF32_STAGES: shared-memory stages of the TF32 blocked kernel (2 or 3)
};

static int env_int(const char* name, int dflt)
{
	const char* s = getenv(name);
	return (s && *s) ? atoi(s) : dflt;
}

struct OptionName { const char* name; int Options::*field; };
static const OptionName OPTION_TABLE[] = {
	{"TICKET", &Options::ticket}, {"TICKET_CTAS", &Options::ticket_ctas}, {"TICKET_MIN_NNZ", &Options::ticket_min_nnz}, {"STAGE", &Options::stage},
	{"CONST_LD", &Options::const_ld}, {"UNROLL", &Options::unroll}, {"KSLICE", &Options::kslice}, {"KSLICE_MIN_MB", &Options::kslice_min_mb},
	{"PIPELINE_COLS", &Options::pipeline_cols}, {"PIPELINE_MIN_KB", &Options::pipeline_min_kb},
	{"DIRECT_STORE", &Options::direct_store}, {"ROWBAND", &Options::rowband}, {"ROWBAND_DIRECT", &Options::rowband_direct}, {"UPLOAD_WINDOW", &Options::upload_window}, {"MULTI_ROUTE", &Options::multi_route}, {"CTAS_PER_SM", &Options::ctas_per_sm},
	{"MERGE_MAXROW", &Options::merge_maxrow}, {"BCSR_FILL_PCT", &Options::bcsr_fill_pct},
	{"TF32_STAGES", &Options::tf32_stages}, {"L2_HINT", &Options::l2_hint}, {"BCSR_TMA", &Options::bcsr_tma}, {"BCSR_STAGES", &Options::bcsr_stages}, {"BCSR_BRW", &Options::bcsr_brw},
};

static Options options_from_env()
{
	Options o;
	char name[64];
	for (const OptionName& e : OPTION_TABLE) {
		snprintf(name, sizeof(name), "SPMM_B200_%s", e.name);
		o.*(e.field) = env_int(name, o.*(e.field));
	}
	return o;
}

struct KsliceChoice { int k = 0, best = 0; };     // per-k result of the slice-width measurement

struct spmm_b200 {
	int device = 0;
	int variant = SPMM_B200_ROWSPLIT;     // resolved variant
	int variant_req = SPMM_B200_AUTO;
	int b_layout = SPMM_B200_B_COLMAJOR;
	int k = 0;                            // k the work lists were tuned for
	int sm_count = 0;
	Options opt;
	int64_t m = 0, c = 0, nnz = 0;        // this handle's shard
	int64_t first_row = 0, m_global = 0;
	int rank = 0, nranks = 1;             // shard identity (spmm_b200_create_shard)
	int max_row = 0;                      // longest row of the shard
	int64_t col_min = 0, col_span = 0;    // first touched column, max column - min column + 1 over the shard's nonzeros
	// first / last touched column of each of BAND_FINE equal row blocks (min > max: the block has no nonzero): what the
	// row-band host pipeline (run_rowband) needs to know which rows of B a block of C rows waits for
	int band_min[64], band_max[64];
	int band_blocks = 0;                  // run_rowband: row blocks of the last run (0: the last run was not row-banded)
	int64_t last_h2d_bytes = 0, last_d2h_bytes = 0;   // bytes the last host-buffer call copied over PCIe (spmm_b200_last_transfer_bytes)
	int64_t win_r0 = 0, win_rows = -1;    // row window of the next row-split launches (-1: all rows); set by run_rowband only
	// L2-aware k slicing of the row-split kernel, chosen by timing the candidates on the first calls
	int ks_k = 0;                         // k being measured (or last looked up)
	int ks_best = 0;                      // chosen slice width (== k: one launch); 0 while still measuring
	int ks_try = 0;                       // index of the candidate the next call runs
	int ks_cand[4] = {0, 0, 0, 0};
	float ks_ms[4] = {0.f, 0.f, 0.f, 0.f};
	std::vector<KsliceChoice> ks_done;    // finished measurements: alternating k values do not re-tune

	int32_t* d_ia = nullptr;
	int32_t* d_ja = nullptr;
	double* d_a = nullptr;

	// row-split work list
	int32_t* d_cta_rows = nullptr;
	int n_ctas = 0;
	int* d_ticket = nullptr;              // {ticket counter, warps that left}: reset by the last warp of a launch (spmm_rowsplit_ticket_kernel)
	uint32_t smem_attr = 0;               // bit i: dynamic shared-memory limit of kernel family i raised on THIS device/handle
	// merge-path work list
	MergePlan merge;
	// blocked CSR (8x1)
	BcsrPlan bcsr;

	// buffers for the host-pointer API
	double* d_x = nullptr; size_t cap_x = 0;   // staging for the column-major upload
	double* d_B = nullptr; size_t cap_B = 0;
	double* d_C = nullptr; size_t cap_C = 0;
	int resident_k = 0;
	std::vector<void*> pinned;            // host ranges we registered

	cudaStream_t stream = nullptr;
	cudaEvent_t ev0 = nullptr, ev1 = nullptr;
	bool timed = false;
	// k-sliced host pipeline of spmm_b200_run (upload slice s+1 | kernel slice s | download slice s-1)
	cudaStream_t s_up = nullptr, s_down = nullptr;
	std::vector<cudaEvent_t> ev_slice;    // per slice: uploaded, kernel start, kernel end
	std::vector<cudaEvent_t> ev_tl;       // timeline events of the last pipelined run (spmm_b200_last_timeline)
	int tl_slices = 0;
	int c_slices = 1;                     // >1: d_C holds c_slices blocks of m x (k / c_slices)
	int timed_slices = 0;                 // >0: last_kernel_ms sums the slices' kernel events
	int64_t launches = 0;
};

extern "C" int spmm_b200_set_option(spmm_b200_t* h, const char* name, int value)
{
	if (!h || !name) return SPMM_B200_EINVAL;
	if (!strncmp(name, "SPMM_B200_", 10)) name += 10;
	for (const OptionName& e : OPTION_TABLE)
		if (!strcmp(name, e.name)) {
			h->opt.*(e.field) = value;
			if (e.field == &Options::kslice) { h->ks_k = 0; h->ks_best = 0; h->ks_done.clear(); }
			return SPMM_B200_OK;
		}
	return SPMM_B200_EINVAL;
}

static int check_device(int device, int* sm_count)
{
	int n = 0;
	cudaError_t e = cudaGetDeviceCount(&n);
	if (e != cudaSuccess || n <= 0) {
		snprintf(g_cuda_err, sizeof(g_cuda_err), "cudaGetDeviceCount: %s (%d devices)", cudaGetErrorString(e), n);
		return SPMM_B200_ENODEV;
	}
	if (device < 0 || device >= n) return SPMM_B200_EINVAL;
	// three attributes instead of cudaGetDeviceProperties (which fills ~100 fields and costs milliseconds per call)
	int major = 0, minor = 0, sms = 0;
	CU_TRY(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
	CU_TRY(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, device));
	CU_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
	if (major != 10) {
		snprintf(g_cuda_err, sizeof(g_cuda_err), "device %d is sm_%d%d; this library holds sm_100a code only", device, major, minor);
		return SPMM_B200_ENODEV;
	}
	if (sm_count) *sm_count = sms;
	CU_TRY(cudaSetDevice(device));
	return SPMM_B200_OK;
}

__global__ void rebase_kernel(int32_t* ia, int64_t n, int32_t base)
{
	int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) ia[i] -= base;
}

__global__ void max_row_kernel(const int32_t* __restrict__ ia, int m, int* __restrict__ out)
{
	int best = 0;
	for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < m; i += gridDim.x * blockDim.x)
		best = max(best, ia[i + 1] - ia[i]);
	for (int off = 16; off > 0; off >>= 1) best = max(best, __shfl_xor_sync(0xffffffffu, best, off));
	if ((threadIdx.x & 31) == 0 && best > 0) atomicMax(out, best);
}

__global__ void col_span_kernel(const int32_t* __restrict__ ja, int64_t nnz, int* __restrict__ out_min, int* __restrict__ out_max)
{
	int lo = INT32_MAX, hi = -1;
	for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < nnz; i += (int64_t) gridDim.x * blockDim.x) {
		const int c = ja[i];
		lo = min(lo, c); hi = max(hi, c);
	}
	for (int off = 16; off > 0; off >>= 1) {
		lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, off));
		hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, off));
	}
	if ((threadIdx.x & 31) == 0 && hi >= 0) { atomicMin(out_min, lo); atomicMax(out_max, hi); }
}

static constexpr int BAND_FINE = 64;

// Column range of each of nblk equal row blocks: block b = rows [m*b/nblk, m*(b+1)/nblk); grid (x, nblk).
__global__ void row_block_span_kernel(const int32_t* __restrict__ ia, const int32_t* __restrict__ ja, int64_t m, int nblk,
                                      int* __restrict__ out_min, int* __restrict__ out_max)
{
	const int b = blockIdx.y;
	const int64_t r0 = m * b / nblk, r1 = m * (b + 1) / nblk;
	const int64_t j0 = ia[r0], j1 = ia[r1];
	int lo = INT32_MAX, hi = -1;
	for (int64_t i = j0 + (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < j1; i += (int64_t) gridDim.x * blockDim.x) {
		const int c = ja[i];
		lo = min(lo, c); hi = max(hi, c);
	}
	for (int off = 16; off > 0; off >>= 1) {
		lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, off));
		hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, off));
	}
	if ((threadIdx.x & 31) == 0 && hi >= 0) { atomicMin(out_min + b, lo); atomicMax(out_max + b, hi); }
}

// ------------------------------------------------------------------------------------------
// work lists
// ------------------------------------------------------------------------------------------

static int build_rowsplit_plan(spmm_b200* h)
{
	// One CTA per nnz-balanced row range.  The k = 32 kernel <8,4,4> holds 64 registers -> 4 resident CTAs per SM:
	// 2 waves (8 per SM) is the optimum for a whole matrix, exactly 1 wave (4 per SM) for small shards, where the
	// tail of a second wave is a visible fraction of the run (profiles/r1_shard_probe.md: 1/8 of cfg2 0.128 -> 0.118 ms;
	// cfg2 0.787 -> 0.780 ms).  The small-k kernels (32 registers, 8 resident CTAs) keep 4 waves.
	int dflt = 32;
	if (h->k == 32) dflt = (h->nnz / ((int64_t) h->sm_count * 8) < 8192) ? 4 : 8;      // k = 64, 128 measured flat-to-better with more CTAs (r1_gvu_probe.md)
	const int per_sm = std::max(1, h->opt.ctas_per_sm > 0 ? h->opt.ctas_per_sm : dflt);
	int64_t want = (int64_t) h->sm_count * per_sm;
	int64_t cap = std::max<int64_t>(1, (h->m + 7) / 8);       // at least ~8 rows per CTA
	// few rows (the row cap binds): whole waves of the 4 resident CTAs per SM, so that no partial last wave
	// runs on a fraction of the machine (m = 8192, k = 64, 410 nnz/row: 1024 CTAs 0.141 ms, 592 CTAs 0.123 ms;
	// profiles/r1_pipeline_probe.md)
	const int64_t wave = (int64_t) h->sm_count * 4;
	if (cap < want && h->k >= 64 && cap >= wave && h->opt.ctas_per_sm <= 0) cap = cap / wave * wave;
	h->n_ctas = (int) std::max<int64_t>(1, std::min(want, cap));
	if (h->d_cta_rows) cudaFree(h->d_cta_rows);
	CU_TRY(cudaMalloc(&h->d_cta_rows, sizeof(int32_t) * (h->n_ctas + 1)));
	if (!h->d_ticket) {
		CU_TRY(cudaMalloc(&h->d_ticket, 2 * sizeof(int)));
		CU_TRY(cudaMemsetAsync(h->d_ticket, 0, 2 * sizeof(int), h->stream));
	}
	const int nb = (h->n_ctas + 1 + 255) / 256;
	cta_rows_nnz_kernel<<<nb, 256, 0, h->stream>>>(h->d_ia, (int) h->m, h->nnz, h->n_ctas, h->d_cta_rows);
	h->launches++;
	CU_TRY(cudaGetLastError());
	return SPMM_B200_OK;
}

static int pick_small_group(const spmm_b200* h)
{
	const double avg = h->m > 0 ? (double) h->nnz / (double) h->m : 1.0;
	int g = 2;
	while (g < 32 && g * 2 <= avg) g *= 2;
	return g;
}

// Kernels that need more than 48 KB of dynamic shared memory: the limit is a per-device function attribute, so it is
// raised once per handle (a handle lives on one device) -- never behind a process-wide flag.
enum { ATTR_MERGE = 0, ATTR_BCSR_TMA = 8, ATTR_STAGED = 16, ATTR_RING = 24 };
template <typename Kern>
static cudaError_t raise_smem_limit(spmm_b200* h, int bit, Kern kern, size_t bytes)
{
	if (bytes <= 48 * 1024 || (h->smem_attr >> bit & 1u)) return cudaSuccess;
	cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) bytes);
	if (e == cudaSuccess) h->smem_attr |= 1u << bit;
	return e;
}

// The staged kernel holds a ticket's nonzeros in a CAP-entry stage: it is taken when the mean ticket fills at most
// ~3/4 of it (tickets beyond CAP fall back to global loads inside the kernel, correct but slower).
static constexpr int STAGED_CAP = 272;
template <int G, int V> struct StagedShape { static constexpr int PASSES = G >= 16 ? (G == 32 ? 4 : 2) : 1; };

template <int G, int V>
static void launch_rowsplit_gv(spmm_b200* h, const double* B, double* C, cudaStream_t s, int ldb = G * V, int ldc = G * V)
{
	const Options& o = h->opt;
	constexpr int K = G * V;
	constexpr int NGW = 32 / G;
	const double avg = h->m > 0 ? (double) h->nnz / (double) h->m : 1.0;
	const bool whole = ldb == K && (uint64_t) h->c * K < (1ull << 32) && o.const_ld;   // compile-time row stride, 32-bit offsets
	// Row tickets: ONE resident wave of persistent warps that draw small row chunks from a counter
	// (spmm_rowsplit_ticket_kernel).  The whole GPU then sweeps A and C front to back (DRAM / L2 locality of the
	// A stream) with no tail: cfg2 0.844 -> 0.754 ms on the same box (profiles/r1_ticket_probe.md).  TICKET=0 keeps
	// the static nnz-balanced CTA ranges everywhere, > 0 forces that many rows per ticket.
	int tick = o.ticket;
	if (tick < 0) {
		// measured over bench/medium_subset.txt x k in {4, 8, 32, 128} (profiles/r1_ticket_probe.md): rows of >= 50
		// nonzeros gain 3-15 %, rows of <= 10 lose 3-15 % (a ticket per ~50 nonzeros is too fine, larger chunks lose
		// the locality); small launches: profiles/r2_shard_probe.md
		// k = 32 (4 rows per warp): 4-row tickets = one row per lane group; same time on a whole matrix (cfg2 0.764 vs
		// 0.767 ms), better on the 1/8 shards of the 8-GPU run (0.114 vs 0.120 ms, static ranges 0.124 ms)
		tick = (avg >= 32.0 && h->nnz >= (int64_t) o.ticket_min_nnz) ? (K == 32 ? 4 : 8) : 0;
	}
	// a row window (run_rowband) always runs on tickets: rows [win_r0, win_r0 + win_rows) through a shifted row
	// pointer (the offsets in it stay absolute) and a shifted C
	const bool win = h->win_rows >= 0;
	if (win) {
		if (tick <= 0) tick = K == 32 ? 4 : 8;
		tick = (tick + NGW - 1) / NGW * NGW;
		const int64_t rows = h->win_rows;
		const int32_t* ia = h->d_ia + h->win_r0;
		double* Cw = C + h->win_r0 * ldc;
		const int grid = (int) std::min<int64_t>((rows + tick * (THREADS / 32) - 1) / (tick * (THREADS / 32)) + 1,
		                                         (int64_t) h->sm_count * std::max(1, o.ticket_ctas));
		if (whole)
			spmm_rowsplit_ticket_kernel<G, V, 4, THREADS, K><<<grid, THREADS, 0, s>>>(ia, h->d_ja, h->d_a, B, Cw, (int) rows, tick, h->d_ticket, ldb, ldc);
		else
			spmm_rowsplit_ticket_kernel<G, V, 4, THREADS><<<grid, THREADS, 0, s>>>(ia, h->d_ja, h->d_a, B, Cw, (int) rows, tick, h->d_ticket, ldb, ldc);
		return;
	}
	if constexpr (V >= 2 && K >= 32) {
		// A stream staged through shared memory by TMA bulk copies (spmm_rowsplit_staged_kernel)
		constexpr int PASSES = StagedShape<G, V>::PASSES;
		using Cfg = StagedCfg<G, PASSES, STAGED_CAP>;
		int stage = o.stage;
		if (stage < 0) stage = (tick > 0 && avg * Cfg::R <= 0.75 * STAGED_CAP) ? 1 : 0;
		// a row must fit a stage (the kernel copies whole rows): longer rows keep the plain ticketed kernel
		if (stage == 1 && h->d_ticket && o.unroll == 4 && h->max_row <= STAGED_CAP) {
			const size_t smem = Cfg::smem(THREADS);
			const int per_sm = std::max(1, std::min(o.ticket_ctas, (int) ((227 * 1024) / (smem + 1024))));
			const int grid = (int) std::min<int64_t>((h->m + Cfg::R * (THREADS / 32) - 1) / (Cfg::R * (THREADS / 32)) + 1, (int64_t) h->sm_count * per_sm);
			if (whole) {
				auto kern = spmm_rowsplit_staged_kernel<G, V, 4, THREADS, PASSES, STAGED_CAP, K>;
				raise_smem_limit(h, ATTR_STAGED + (K == 32 ? 0 : K == 64 ? 1 : 2), kern, smem);
				kern<<<grid, THREADS, smem, s>>>(h->d_ia, h->d_ja, h->d_a, B, C, (int) h->m, h->d_ticket, ldb, ldc);
			} else {
				auto kern = spmm_rowsplit_staged_kernel<G, V, 4, THREADS, PASSES, STAGED_CAP>;
				raise_smem_limit(h, ATTR_STAGED + 3 + (K == 32 ? 0 : K == 64 ? 1 : 2), kern, smem);
				kern<<<grid, THREADS, smem, s>>>(h->d_ia, h->d_ja, h->d_a, B, C, (int) h->m, h->d_ticket, ldb, ldc);
			}
			return;
		}
		if constexpr (K == 32) {
			if (stage == 2 && h->d_ticket) {         // experiment: B rows through a cp.async ring (profiles/r2_rowsplit_stage_probe.md)
				constexpr int H = 4;
				const size_t smem = (size_t) 2 * H * (V / 2) * THREADS * 16;
				auto kern = spmm_rowsplit_ring_kernel<G, V, H, THREADS>;
				raise_smem_limit(h, ATTR_RING, kern, smem);
				const int rpt = std::max(NGW, (std::max(tick, 1) + NGW - 1) / NGW * NGW);
				const int grid = (int) std::min<int64_t>((h->m + rpt * (THREADS / 32) - 1) / (rpt * (THREADS / 32)) + 1, (int64_t) h->sm_count * 3);
					kern<<<grid, THREADS, smem, s>>>(h->d_ia, h->d_ja, h->d_a, B, C, (int) h->m, rpt, h->d_ticket, ldb, ldc);
				return;
			}
		}
	}
	if (tick > 0 && h->d_ticket && o.unroll == 4) {
		tick = (tick + NGW - 1) / NGW * NGW;
		const int grid = (int) std::min<int64_t>((h->m + tick * (THREADS / 32) - 1) / (tick * (THREADS / 32)) + 1,
		                                         (int64_t) h->sm_count * std::max(1, o.ticket_ctas));
		if (whole)
			spmm_rowsplit_ticket_kernel<G, V, 4, THREADS, K><<<grid, THREADS, 0, s>>>(h->d_ia, h->d_ja, h->d_a, B, C, (int) h->m, tick, h->d_ticket, ldb, ldc);
		else
			spmm_rowsplit_ticket_kernel<G, V, 4, THREADS><<<grid, THREADS, 0, s>>>(h->d_ia, h->d_ja, h->d_a, B, C, (int) h->m, tick, h->d_ticket, ldb, ldc);
		return;
	}
	if (o.unroll >= 8 && G >= 8)
		spmm_rowsplit_kernel<G, V, 8, THREADS><<<h->n_ctas, THREADS, 0, s>>>(h->d_ia, h->d_ja, h->d_a, B, C, h->d_cta_rows, ldb, ldc);
	else if (o.unroll <= 2)
		spmm_rowsplit_kernel<G, V, 2, THREADS><<<h->n_ctas, THREADS, 0, s>>>(h->d_ia, h->d_ja, h->d_a, B, C, h->d_cta_rows, ldb, ldc);
	else if (whole)
		spmm_rowsplit_kernel<G, V, 4, THREADS, K><<<h->n_ctas, THREADS, 0, s>>>(h->d_ia, h->d_ja, h->d_a, B, C, h->d_cta_rows, ldb, ldc);
	else if (o.l2_hint && V >= 2)
		spmm_rowsplit_kernel<G, V, 4, THREADS, 0, 1><<<h->n_ctas, THREADS, 0, s>>>(h->d_ia, h->d_ja, h->d_a, B, C, h->d_cta_rows, ldb, ldc);
	else
		spmm_rowsplit_kernel<G, V, 4, THREADS><<<h->n_ctas, THREADS, 0, s>>>(h->d_ia, h->d_ja, h->d_a, B, C, h->d_cta_rows, ldb, ldc);
}

template <int K>
static void launch_smallk(const spmm_b200* h, const double* B, double* C, cudaStream_t s)
{
	switch (pick_small_group(h)) {
	case 2:  spmm_smallk_kernel<K, 2, THREADS><<<h->n_ctas, THREADS, 0, s>>>(h->d_ia, h->d_ja, h->d_a, B, C, h->d_cta_rows); break;
	case 4:  spmm_smallk_kernel<K, 4, THREADS><<<h->n_ctas, THREADS, 0, s>>>(h->d_ia, h->d_ja, h->d_a, B, C, h->d_cta_rows); break;
	case 8:  spmm_smallk_kernel<K, 8, THREADS><<<h->n_ctas, THREADS, 0, s>>>(h->d_ia, h->d_ja, h->d_a, B, C, h->d_cta_rows); break;
	case 16: spmm_smallk_kernel<K, 16, THREADS><<<h->n_ctas, THREADS, 0, s>>>(h->d_ia, h->d_ja, h->d_a, B, C, h->d_cta_rows); break;
	default: spmm_smallk_kernel<K, 32, THREADS><<<h->n_ctas, THREADS, 0, s>>>(h->d_ia, h->d_ja, h->d_a, B, C, h->d_cta_rows); break;
	}
}

// ---- merge-path plan: tile coordinates depend on the tile size, which depends on k ----------
template <int G, int V, int IPG>
struct MergeLaunch {
	using Cfg = MergeCfg<G, V, IPG, THREADS>;
	static int items() { return Cfg::ITEMS; }
	static cudaError_t launch(spmm_b200* h, const double* B, double* C, cudaStream_t s, int attr_bit)
	{
		auto kern = spmm_merge_kernel<G, V, IPG, 4, THREADS>;
		cudaError_t e = raise_smem_limit(h, ATTR_MERGE + attr_bit, kern, Cfg::SMEM);
		if (e != cudaSuccess) return e;
		kern<<<h->merge.ntiles, THREADS, Cfg::SMEM, s>>>(h->d_ia, h->d_ja, h->d_a, B, C, h->merge.d_tile,
		                                                (double*) h->merge.d_carry, h->merge.d_carry_row);
		return cudaGetLastError();
	}
};

// k -> (G, V, IPG) of the merge kernel; 0 items = no specialised instantiation
#define MERGE_DISPATCH(k, WHAT)                                    \
	switch (k) {                                                   \
	case 1:   WHAT(1, 1, 16, 0); break;                            \
	case 2:   WHAT(1, 2, 16, 1); break;                            \
	case 4:   WHAT(2, 2, 16, 2); break;                            \
	case 8:   WHAT(4, 2, 32, 3); break;                            \
	case 16:  WHAT(8, 2, 32, 4); break;                            \
	case 32:  WHAT(8, 4, 64, 5); break;                            \
	case 64:  WHAT(16, 4, 128, 6); break;                          \
	case 128: WHAT(32, 4, 128, 7); break;                          \
	default: break;                                                \
	}

static int merge_items_for_k(int k)
{
	int items = 0;
#define WHAT(G, V, IPG, BIT) items = MergeLaunch<G, V, IPG>::items()
	MERGE_DISPATCH(k, WHAT)
#undef WHAT
	return items;
}

static bool merge_supports_k(int k) { return merge_items_for_k(k) > 0; }

static int build_merge_plan(spmm_b200* h, int k, cudaStream_t s)
{
	const int items = merge_items_for_k(k);
	if (items <= 0) return SPMM_B200_ENOSUP;
	MergePlan& p = h->merge;
	if (p.items == items && p.carry_k >= k && p.d_tile) return SPMM_B200_OK;
	const int64_t L = h->m + h->nnz;
	const int ntiles = (int) ((L + items - 1) / items);
	cudaFree(p.d_tile); cudaFree(p.d_carry); cudaFree(p.d_carry_row);
	p.d_tile = nullptr; p.d_carry = nullptr; p.d_carry_row = nullptr;
	CU_TRY(cudaMalloc(&p.d_tile, sizeof(int2) * ((size_t) ntiles + 1)));
	CU_TRY(cudaMalloc(&p.d_carry, sizeof(double) * std::max<size_t>((size_t) ntiles * k, 1)));
	CU_TRY(cudaMalloc(&p.d_carry_row, sizeof(int32_t) * std::max(ntiles, 1)));
	p.ntiles = ntiles; p.items = items; p.carry_k = k;
	tile_merge_kernel<<<(ntiles + 1 + 255) / 256, 256, 0, s>>>(h->d_ia, (int) h->m, h->nnz, items, ntiles, p.d_tile);
	h->launches++;
	CU_TRY(cudaGetLastError());
	return SPMM_B200_OK;
}

static int launch_merge(spmm_b200* h, const double* B, double* C, int k, cudaStream_t s)
{
	int rc = build_merge_plan(h, k, s);     // no-op when the plan matches k (stream-ordered otherwise)
	if (rc) return rc;
	if (h->merge.ntiles == 0) return 0;
	cudaError_t e = cudaErrorInvalidValue;
#define WHAT(G, V, IPG, BIT) e = MergeLaunch<G, V, IPG>::launch(h, B, C, s, BIT)
	MERGE_DISPATCH(k, WHAT)
#undef WHAT
	if (e != cudaSuccess) {
		snprintf(g_cuda_err, sizeof(g_cuda_err), "merge kernel launch (k=%d): %s", k, cudaGetErrorString(e));
		return SPMM_B200_ECUDA;
	}
	const int64_t n = (int64_t) h->merge.ntiles * k;
	spmm_merge_fixup_kernel<THREADS><<<(unsigned) ((n + THREADS - 1) / THREADS), THREADS, 0, s>>>(
		(const double*) h->merge.d_carry, h->merge.d_carry_row, h->merge.ntiles, (int) h->m, k, C);
	e = cudaGetLastError();
	if (e != cudaSuccess) {
		snprintf(g_cuda_err, sizeof(g_cuda_err), "merge fixup launch (k=%d): %s", k, cudaGetErrorString(e));
		return SPMM_B200_ECUDA;
	}
	h->launches += 2;
	return 2;
}

// ---- blocked CSR (8x1): conversion on the device + launch -------------------------------------
static bool bcsr_supports_k(int variant, int k)
{
	if (variant == SPMM_B200_BCSR) return k == 16 || k == 32 || k == 64 || k == 128;
	return k == 32 || k == 64 || k == 128;
}

// Pass 1 of the conversion: distinct columns per 8-row slab -> brow_ptr, nnzb (also AUTO's fill probe).
static int bcsr_count(spmm_b200* h)
{
	BcsrPlan& p = h->bcsr;
	const int mb = (int) ((h->m + 7) / 8);
	p.mb = mb;
	int32_t* d_tmp = nullptr;
	CU_TRY(cudaMalloc(&p.d_brow, sizeof(int32_t) * ((size_t) mb + 1)));
	CU_TRY(cudaMemsetAsync(p.d_brow, 0, sizeof(int32_t) * ((size_t) mb + 1), h->stream));
	CU_TRY(cudaMalloc(&d_tmp, sizeof(int32_t) * scan_tmp_elems(mb + 1)));
	const unsigned nb = (unsigned) (((int64_t) mb * 8 + 255) / 256);
	if (mb > 0) {
		bcsr_convert_kernel<false><<<nb, 256, 0, h->stream>>>(h->d_ia, h->d_ja, h->d_a, (int) h->m, mb, p.d_brow, nullptr, nullptr, nullptr);
		h->launches++;
	}
	h->launches += exclusive_scan_i32(p.d_brow, (int64_t) mb + 1, d_tmp, h->stream);
	int32_t nnzb = 0;
	CU_TRY(cudaMemcpyAsync(&nnzb, p.d_brow + mb, sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
	CU_TRY(cudaStreamSynchronize(h->stream));
	cudaFree(d_tmp);
	p.nnzb = nnzb;
	return SPMM_B200_OK;
}

static void bcsr_release(spmm_b200* h)
{
	BcsrPlan& p = h->bcsr;
	cudaFree(p.d_brow); cudaFree(p.d_bcol); cudaFree(p.d_bval); cudaFree(p.d_bmask);
	p = BcsrPlan();
}

// Pass 2: write bcol / bval / bmask.
static int bcsr_fill(spmm_b200* h)
{
	BcsrPlan& p = h->bcsr;
	const int mb = p.mb;
	const int64_t nnzb = p.nnzb;
	const unsigned nb = (unsigned) (((int64_t) mb * 8 + 255) / 256);
	CU_TRY(cudaMalloc(&p.d_bcol, sizeof(int32_t) * std::max<size_t>((size_t) nnzb, 1)));
	CU_TRY(cudaMalloc(&p.d_bmask, std::max<size_t>((size_t) nnzb, 1)));
	// +4 blocks of slack: the TMA kernel fetches the A values of a step as one 256-byte bulk copy
	CU_TRY(cudaMalloc(&p.d_bval, sizeof(double) * 8 * ((size_t) nnzb + 4)));
	CU_TRY(cudaMemsetAsync(p.d_bval + (size_t) nnzb * 8, 0, sizeof(double) * 32, h->stream));
	if (mb > 0 && nnzb > 0) {
		bcsr_convert_kernel<true><<<nb, 256, 0, h->stream>>>(h->d_ia, h->d_ja, h->d_a, (int) h->m, mb, p.d_brow, p.d_bcol, p.d_bval, p.d_bmask);
		h->launches++;
	}
	CU_TRY(cudaGetLastError());
	CU_TRY(cudaStreamSynchronize(h->stream));
	return SPMM_B200_OK;
}

static int build_bcsr_plan(spmm_b200* h)
{
	int rc = bcsr_count(h);
	if (rc) return rc;
	return bcsr_fill(h);
}

// The block rows a blocked launch covers: all of them, or -- inside run_rowband -- the block rows of the row window
// (window starts are multiples of 8 rows; the block-row pointer and C are shifted, the offsets in brow stay absolute).
struct BcsrRange { const int32_t* brow; double* C; int m, mb; };
static BcsrRange bcsr_range(const spmm_b200* h, double* C, int K)
{
	const BcsrPlan& p = h->bcsr;
	if (h->win_rows < 0) return {p.d_brow, C, (int) h->m, p.mb};
	const int64_t b0 = h->win_r0 / 8, r1 = h->win_r0 + h->win_rows;
	return {p.d_brow + b0, C + h->win_r0 * K, (int) h->win_rows, (int) ((r1 + 7) / 8 - b0)};
}

template <int K, int WARPS, int STAGES>
static void launch_bcsr_tma(spmm_b200* h, const double* B, const BcsrRange& r, int brw, unsigned grid, cudaStream_t s)
{
	using Cfg = BcsrTmaCfg<K, WARPS, STAGES>;
	const BcsrPlan& p = h->bcsr;
	auto kern = spmm_bcsr_dmma_tma_kernel<K, WARPS, STAGES>;
	raise_smem_limit(h, ATTR_BCSR_TMA + (STAGES - 2) + 3 * (K == 128 ? 2 : K == 64 ? 1 : 0), kern, Cfg::SMEM);
	kern<<<grid, WARPS * 32, Cfg::SMEM, s>>>(r.brow, p.d_bcol, p.d_bval, B, r.C, r.m, r.mb, brw);
}

template <int K>
static void launch_bcsr_k(spmm_b200* h, const double* B, double* C, cudaStream_t s)
{
	constexpr int WARPS = 4;
	const BcsrPlan& p = h->bcsr;
	const Options& o = h->opt;
	const BcsrRange r = bcsr_range(h, C, K);
	if (r.mb <= 0) return;
	const unsigned grid = (unsigned) ((r.mb + WARPS - 1) / WARPS);
	// bulk copies of 8K bytes per B row pay from K = 64 on (cfg4: K=32 0.57 ms fragment loads vs 0.63 ms TMA;
	// K=128 2.55 ms vs 1.23 ms); 2 stages = most warps per SM, which is what the tensor pipe needs
	if (h->variant == SPMM_B200_BCSR && (o.bcsr_tma == 2 || (o.bcsr_tma == 1 && K >= 64))) {
		if constexpr (K >= 32) {
			// block rows per warp: enough to keep the ring flowing, few enough for >= 8 CTAs per SM
			const int brw = o.bcsr_brw > 0 ? o.bcsr_brw : (int) std::max<int64_t>(1, std::min<int64_t>(8, r.mb / ((int64_t) h->sm_count * WARPS * 8)));
			const unsigned g2 = (unsigned) ((r.mb + (int64_t) WARPS * brw - 1) / ((int64_t) WARPS * brw));
			if (o.bcsr_stages <= 2) launch_bcsr_tma<K, WARPS, 2>(h, B, r, brw, g2, s);
			else if (o.bcsr_stages == 3) launch_bcsr_tma<K, WARPS, 3>(h, B, r, brw, g2, s);
			else launch_bcsr_tma<K, WARPS, 4>(h, B, r, brw, g2, s);
			return;
		}
	}
	if (h->variant == SPMM_B200_BCSR)
		spmm_bcsr_dmma_kernel<K, WARPS><<<grid, WARPS * 32, 0, s>>>(r.brow, p.d_bcol, p.d_bval, B, r.C, r.m, r.mb);
	else if constexpr (K % 32 == 0)
		spmm_bcsr_simt_kernel<K, WARPS><<<grid, WARPS * 32, 0, s>>>(r.brow, p.d_bcol, p.d_bval, p.d_bmask, B, r.C, r.m, r.mb);
}

static int launch_bcsr(spmm_b200* h, const double* B, double* C, int k, cudaStream_t s)
{
	switch (k) {
	case 16:  launch_bcsr_k<16>(h, B, C, s); break;
	case 32:  launch_bcsr_k<32>(h, B, C, s); break;
	case 64:  launch_bcsr_k<64>(h, B, C, s); break;
	case 128: launch_bcsr_k<128>(h, B, C, s); break;
	default: return SPMM_B200_ENOSUP;
	}
	cudaError_t e = cudaGetLastError();
	if (e != cudaSuccess) {
		snprintf(g_cuda_err, sizeof(g_cuda_err), "bcsr kernel launch (k=%d): %s", k, cudaGetErrorString(e));
		return SPMM_B200_ECUDA;
	}
	h->launches += 1;
	return 1;
}

// k for which the row-split family has a specialised (double2) kernel
static bool rowsplit_specialised(int k) { return k == 1 || k == 2 || k == 4 || k == 8 || k == 16 || k == 32 || k == 64 || k == 128; }

static void launch_rowsplit_width(spmm_b200* h, const double* B, double* C, int w, cudaStream_t s, int ldb, int ldc)
{
	switch (w) {
	case 8:  launch_rowsplit_gv<4, 2>(h, B, C, s, ldb, ldc); break;
	case 16: launch_rowsplit_gv<8, 2>(h, B, C, s, ldb, ldc); break;
	case 32: launch_rowsplit_gv<8, 4>(h, B, C, s, ldb, ldc); break;
	default: launch_rowsplit_gv<16, 4>(h, B, C, s, ldb, ldc); break;      // 64
	}
}

// Launches the SpMM kernel(s) of the handle's variant on stream s: C (row stride ldc) = A * B (row stride ldb), both
// row-major with k columns.  Returns kernels launched or <0.
static int launch_spmm(spmm_b200* h, const double* B, double* C, int k, cudaStream_t s, int ldb = 0, int ldc = 0)
{
	if (h->m == 0) return 0;
	if (ldb <= 0) ldb = k;
	if (ldc <= 0) ldc = k;
	// Every specialised kernel moves B and C as double2: operands that are not 16-byte aligned (a tensor view at an
	// odd offset, an odd row stride) take the generic kernel instead of faulting.
	const bool aligned = k == 1 || ((((uintptr_t) B | (uintptr_t) C) & 15) == 0 && ldb % 2 == 0 && ldc % 2 == 0);
	const bool whole = ldb == k && ldc == k;
	if (aligned && whole) {
		// merge-path has no instantiation for odd k: those run on the row-split generic kernel
		if (h->variant == SPMM_B200_MERGE && merge_supports_k(k)) return launch_merge(h, B, C, k, s);
		// blocked variants need k % 16 (tensor pipe) / k % 32 (FMA pipe); other k run on the CSR kernels
		if ((h->variant == SPMM_B200_BCSR || h->variant == SPMM_B200_BCSR_FMA) && bcsr_supports_k(h->variant, k))
			return launch_bcsr(h, B, C, k, s);
	}
	if (!aligned || !rowsplit_specialised(k) || (!whole && k < 8)) {
		spmm_rowsplit_generic_kernel<THREADS><<<h->n_ctas, THREADS, 0, s>>>(h->d_ia, h->d_ja, h->d_a, B, C, h->d_cta_rows, k, ldb, ldc);
		cudaError_t e = cudaGetLastError();
		if (e != cudaSuccess) {
			snprintf(g_cuda_err, sizeof(g_cuda_err), "generic spmm launch (k=%d): %s", k, cudaGetErrorString(e));
			return SPMM_B200_ECUDA;
		}
		h->launches += 1;
		return 1;
	}
	// L2-aware k slicing: run the k columns as k/w launches on w-column slices of the same row-major B and C,
	// so that the B rows a launch gathers (touched window x 8w bytes) stay in L2.  It pays when the window is
	// far larger than L2 AND rows are long (every slice re-reads A and pays the per-row overhead again):
	// profiles/r1_kslice_probe.md, r1_sweep_medium.md.  KSLICE=w forces a width, =0 (default) lets the
	// handle time the candidates on its first calls (spmm_b200_run_device), =-1 turns slicing off.
	int w = h->opt.kslice;
	if (w == 0 && h->ks_k == k) w = h->ks_best > 0 ? h->ks_best : h->ks_cand[h->ks_try / 2];
	if (whole && w > 0 && w < k && k % w == 0 && (w == 8 || w == 16 || w == 32 || w == 64)) {
		for (int off = 0; off < k; off += w) launch_rowsplit_width(h, B + off, C + off, w, s, k, k);
		cudaError_t e = cudaGetLastError();
		if (e != cudaSuccess) {
			snprintf(g_cuda_err, sizeof(g_cuda_err), "sliced spmm launch (k=%d, w=%d): %s", k, w, cudaGetErrorString(e));
			return SPMM_B200_ECUDA;
		}
		h->launches += k / w;
		return k / w;
	}
	switch (k) {
	case 1:   launch_smallk<1>(h, B, C, s); break;
	case 2:   launch_smallk<2>(h, B, C, s); break;
	case 4:   launch_rowsplit_gv<2, 2>(h, B, C, s, ldb, ldc); break;      // 16 rows per warp: cfg2 0.496 -> 0.380 ms, cfg5-large 0.440 -> 0.284 ms vs the small-k kernel
	// lane-group shapes from the sweep in profiles/r1_gvu_probe.md: V = 4 columns per lane (two 16-byte loads)
	// makes a group of 8 lanes cover one 128-byte line per load and halves the shuffle traffic per nonzero
	// (cfg2: k=32 0.935 -> 0.786 ms, k=64 1.563 -> 1.336 ms); V = 8 costs too many registers
	case 128: launch_rowsplit_gv<32, 4>(h, B, C, s, ldb, ldc); break;
	default:  launch_rowsplit_width(h, B, C, k, s, ldb, ldc); break;      // 8, 16, 32, 64
	}
	cudaError_t e = cudaGetLastError();
	if (e != cudaSuccess) {
		snprintf(g_cuda_err, sizeof(g_cuda_err), "spmm kernel launch (k=%d): %s", k, cudaGetErrorString(e));
		return SPMM_B200_ECUDA;
	}
	h->launches += 1;
	return 1;
}

// ------------------------------------------------------------------------------------------
// construction
// ------------------------------------------------------------------------------------------

static int create_from_slice(spmm_b200_t** out, const int32_t* row_ptr_slice, const int32_t* col_ind, const double* values,
                             int64_t m_local, int64_t c, int64_t nnz_local, int32_t base, int k,
                             int device, int variant, int b_layout, int sm_count, int64_t first_row, int64_t m_global)
{
	spmm_b200* h = new (std::nothrow) spmm_b200();
	if (!h) return SPMM_B200_ENOMEM;
	h->device = device; h->variant_req = variant; h->b_layout = b_layout; h->k = k;
	h->sm_count = sm_count;
	h->opt = options_from_env();
	h->m = m_local; h->c = c; h->nnz = nnz_local; h->first_row = first_row; h->m_global = m_global;
	*out = h;   // from here on the caller destroys on error

	CU_TRY(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
	CU_TRY(cudaEventCreate(&h->ev0));
	CU_TRY(cudaEventCreate(&h->ev1));
	CU_TRY(cudaMalloc(&h->d_ia, sizeof(int32_t) * (m_local + 1)));
	// 16 entries of slack: the staged kernel's bulk copies read 16-byte aligned windows that may end past nnz
	CU_TRY(cudaMalloc(&h->d_ja, sizeof(int32_t) * (nnz_local + 16)));
	CU_TRY(cudaMalloc(&h->d_a, sizeof(double) * (nnz_local + 16)));
	CU_TRY(cudaMemsetAsync(h->d_ja + nnz_local, 0, sizeof(int32_t) * 16, h->stream));
	CU_TRY(cudaMemsetAsync(h->d_a + nnz_local, 0, sizeof(double) * 16, h->stream));
	CU_TRY(cudaMemcpyAsync(h->d_ia, row_ptr_slice, sizeof(int32_t) * (m_local + 1), cudaMemcpyHostToDevice, h->stream));
	if (nnz_local > 0) {
		CU_TRY(cudaMemcpyAsync(h->d_ja, col_ind + base, sizeof(int32_t) * nnz_local, cudaMemcpyHostToDevice, h->stream));
		CU_TRY(cudaMemcpyAsync(h->d_a, values + base, sizeof(double) * nnz_local, cudaMemcpyHostToDevice, h->stream));
	}
	if (base != 0) {
		rebase_kernel<<<(unsigned) ((m_local + 1 + 255) / 256), 256, 0, h->stream>>>(h->d_ia, m_local + 1, base);
		h->launches++;
	}
	// row-length statistics for AUTO
	if (m_local > 0) {
		int* d_max = nullptr;
		CU_TRY(cudaMalloc(&d_max, sizeof(int)));
		CU_TRY(cudaMemsetAsync(d_max, 0, sizeof(int), h->stream));
		max_row_kernel<<<std::min<int64_t>(1024, (m_local + 255) / 256), 256, 0, h->stream>>>(h->d_ia, (int) m_local, d_max);
		h->launches++;
		CU_TRY(cudaMemcpyAsync(&h->max_row, d_max, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
		CU_TRY(cudaStreamSynchronize(h->stream));
		cudaFree(d_max);
	}
	if (nnz_local > 0) {
		// touched window of B (for the L2-aware k slicing)
		int* d_mm = nullptr;
		int mm[2] = {INT32_MAX, -1};
		CU_TRY(cudaMalloc(&d_mm, 2 * sizeof(int)));
		CU_TRY(cudaMemcpyAsync(d_mm, mm, sizeof(mm), cudaMemcpyHostToDevice, h->stream));
		col_span_kernel<<<(unsigned) std::min<int64_t>((int64_t) sm_count * 8, (nnz_local + 255) / 256), 256, 0, h->stream>>>(h->d_ja, nnz_local, d_mm, d_mm + 1);
		h->launches++;
		CU_TRY(cudaMemcpyAsync(mm, d_mm, sizeof(mm), cudaMemcpyDeviceToHost, h->stream));
		CU_TRY(cudaStreamSynchronize(h->stream));
		cudaFree(d_mm);
		h->col_span = mm[1] >= mm[0] ? (int64_t) mm[1] - mm[0] + 1 : 0;
		h->col_min = mm[1] >= mm[0] ? mm[0] : 0;
	}
	for (int b = 0; b < BAND_FINE; b++) { h->band_min[b] = INT32_MAX; h->band_max[b] = -1; }
	if (nnz_local > 0 && m_local >= BAND_FINE) {
		int* d_bb = nullptr;
		CU_TRY(cudaMalloc(&d_bb, 2 * BAND_FINE * sizeof(int)));
		CU_TRY(cudaMemcpyAsync(d_bb, h->band_min, BAND_FINE * sizeof(int), cudaMemcpyHostToDevice, h->stream));
		CU_TRY(cudaMemcpyAsync(d_bb + BAND_FINE, h->band_max, BAND_FINE * sizeof(int), cudaMemcpyHostToDevice, h->stream));
		const unsigned gx = (unsigned) std::max<int64_t>(1, std::min<int64_t>((int64_t) sm_count * 8 / BAND_FINE + 1, (nnz_local / BAND_FINE + 255) / 256));
		row_block_span_kernel<<<dim3(gx, BAND_FINE), 256, 0, h->stream>>>(h->d_ia, h->d_ja, m_local, BAND_FINE, d_bb, d_bb + BAND_FINE);
		h->launches++;
		CU_TRY(cudaMemcpyAsync(h->band_min, d_bb, BAND_FINE * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
		CU_TRY(cudaMemcpyAsync(h->band_max, d_bb + BAND_FINE, BAND_FINE * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
		CU_TRY(cudaStreamSynchronize(h->stream));
		cudaFree(d_bb);
	}
	if (variant == SPMM_B200_AUTO) {
		// row-split serialises a row on one lane group: switch to merge-path when one row is far
		// longer than the mean (SURVEY 8(d) cfg 3: 200 040 nnz against a mean of 20)
		const double avg = m_local > 0 ? (double) nnz_local / (double) m_local : 0.0;
		// measured (gpurun_out/r2_*): cfg3b (max 866, mean 20) merge 1.24 ms vs row-split 2.03 ms;
		// cfg2 (max 131, mean 50) row-split 0.91 ms vs merge 0.96 ms
		const int thr = h->opt.merge_maxrow;
		h->variant = (h->max_row >= thr && h->max_row > 8.0 * avg) ? SPMM_B200_MERGE : SPMM_B200_ROWSPLIT;
	} else {
		h->variant = variant;
	}
	int rc = build_rowsplit_plan(h);        // always: the fallback for k without a specialised instantiation
	if (rc) return rc;
	if (variant == SPMM_B200_AUTO && h->variant == SPMM_B200_ROWSPLIT && k >= 32 && bcsr_supports_k(SPMM_B200_BCSR, k)
	    && m_local >= 8 && nnz_local > 0) {
		// "tensor cores only where dense blocks make it a genuine contraction": probe the 8x1 block
		// fill with pass 1 of the conversion.  Break-even measured on B200: fill 0.15
		// at k=128, 0.25 at k=32; below 0.30 the CSR kernels stay.
		rc = bcsr_count(h);
		if (rc) return rc;
		const double fill = h->bcsr.nnzb > 0 ? (double) nnz_local / (8.0 * (double) h->bcsr.nnzb) : 0.0;
		const double thr = h->opt.bcsr_fill_pct / 100.0;
		if (fill >= thr) {
			rc = bcsr_fill(h);
			if (rc) return rc;
			h->variant = SPMM_B200_BCSR;
		} else {
			bcsr_release(h);
		}
	} else if (h->variant == SPMM_B200_BCSR || h->variant == SPMM_B200_BCSR_FMA) {
		rc = build_bcsr_plan(h);
		if (rc) return rc;
	}
	if (h->variant == SPMM_B200_MERGE && merge_supports_k(k)) {
		rc = build_merge_plan(h, k, h->stream);
		if (rc) return rc;
	}
	CU_TRY(cudaStreamSynchronize(h->stream));
	return SPMM_B200_OK;
}

static int check_sizes(const void* rp, const void* ci, const void* va, int64_t m, int64_t c, int64_t nnz, int k)
{
	if (!rp || m < 0 || c < 0 || nnz < 0 || k < 1) return SPMM_B200_EINVAL;
	if (nnz > 0 && (!ci || !va)) return SPMM_B200_EINVAL;
	if (m >= INT32_MAX || c >= INT32_MAX || nnz >= INT32_MAX) return SPMM_B200_ERANGE;
	return SPMM_B200_OK;
}

extern "C" int spmm_b200_create(spmm_b200_t** h, const int32_t* row_ptr, const int32_t* col_ind, const double* values,
                                int64_t m, int64_t c, int64_t nnz, int k, int device, int variant, int b_layout)
{
	if (!h) return SPMM_B200_EINVAL;
	*h = nullptr;
	int rc = check_sizes(row_ptr, col_ind, values, m, c, nnz, k);
	if (rc) return rc;
	if (variant < SPMM_B200_ROWSPLIT || variant > SPMM_B200_BCSR_FMA) return SPMM_B200_EINVAL;
	if (b_layout != SPMM_B200_B_COLMAJOR && b_layout != SPMM_B200_B_ROWMAJOR) return SPMM_B200_EINVAL;
	int sms = 0;
	rc = check_device(device, &sms);
	if (rc) return rc;
	rc = create_from_slice(h, row_ptr, col_ind, values, m, c, nnz, 0, k, device, variant, b_layout, sms, 0, m);
	if (rc) { spmm_b200_destroy(*h); *h = nullptr; }
	return rc;
}

extern "C" int spmm_b200_create_shard(spmm_b200_t** h, const int32_t* row_ptr, const int32_t* col_ind, const double* values,
                                      int64_t m, int64_t c, int64_t nnz, int k, int device, int variant, int b_layout,
                                      int rank, int nranks, int64_t* first_row, int64_t* last_row)
{
	if (!h) return SPMM_B200_EINVAL;
	*h = nullptr;
	int rc = check_sizes(row_ptr, col_ind, values, m, c, nnz, k);
	if (rc) return rc;
	if (nranks < 1 || rank < 0 || rank >= nranks) return SPMM_B200_EINVAL;
	int sms = 0;
	rc = check_device(device, &sms);
	if (rc) return rc;
	// shard boundaries: the reference's nnz-balanced split with P = nranks, searched on the device
	std::vector<int64_t> bounds(nranks + 1);
	rc = spmm_b200_partition(row_ptr, m, nnz, nranks, SPMM_B200_PART_NNZ, bounds.data(), device);
	if (rc) return rc;
	const int64_t r0 = bounds[rank], r1 = std::max(bounds[rank + 1], bounds[rank]);
	if (first_row) *first_row = r0;
	if (last_row) *last_row = r1;
	const int32_t base = row_ptr[r0];
	const int64_t nnz_local = (int64_t) row_ptr[r1] - base;
	rc = create_from_slice(h, row_ptr + r0, col_ind, values, r1 - r0, c, nnz_local, base, k, device, variant, b_layout, sms, r0, m);
	if (rc) { spmm_b200_destroy(*h); *h = nullptr; }
	else { (*h)->rank = rank; (*h)->nranks = nranks; }
	return rc;
}

extern "C" void spmm_b200_destroy(spmm_b200_t* h)
{
	if (!h) return;
	cudaSetDevice(h->device);
	if (h->stream) cudaStreamSynchronize(h->stream);
	for (void* p : h->pinned) cudaHostUnregister(p);
	cudaFree(h->d_ia); cudaFree(h->d_ja); cudaFree(h->d_a);
	cudaFree(h->d_cta_rows);
	cudaFree(h->merge.d_tile); cudaFree(h->merge.d_carry); cudaFree(h->merge.d_carry_row);
	cudaFree(h->bcsr.d_brow); cudaFree(h->bcsr.d_bcol); cudaFree(h->bcsr.d_bval); cudaFree(h->bcsr.d_bmask);
	cudaFree(h->d_x); cudaFree(h->d_B); cudaFree(h->d_C); cudaFree(h->d_ticket);
	if (h->ev0) cudaEventDestroy(h->ev0);
	if (h->ev1) cudaEventDestroy(h->ev1);
	for (cudaEvent_t e : h->ev_slice) cudaEventDestroy(e);
	for (cudaEvent_t e : h->ev_tl) cudaEventDestroy(e);
	if (h->s_up) cudaStreamDestroy(h->s_up);
	if (h->s_down) cudaStreamDestroy(h->s_down);
	if (h->stream) cudaStreamDestroy(h->stream);
	delete h;
}

// ------------------------------------------------------------------------------------------
// the hot call
// ------------------------------------------------------------------------------------------

static int ensure(double** p, size_t* cap, size_t want)
{
	if (*cap >= want) return SPMM_B200_OK;
	if (*p) cudaFree(*p);
	*p = nullptr; *cap = 0;
	CU_TRY(cudaMalloc(p, sizeof(double) * std::max<size_t>(want, 1)));
	*cap = want;
	return SPMM_B200_OK;
}

// Page-lock a caller buffer so that the copies run at full PCIe rate.  Only on explicit request
// (spmm_b200_register_host): the caller promises the range stays mapped until it is unregistered
// or the handle is destroyed (true for the harness's x and y, spmv_bench.cpp:896-907; NOT true
// for short-lived buffers, where a stale registration would send DMA to recycled pages).
extern "C" int spmm_b200_register_host(spmm_b200_t* h, const void* p, size_t bytes)
{
	if (!h || !p || bytes == 0) return SPMM_B200_EINVAL;
	if (std::find(h->pinned.begin(), h->pinned.end(), (void*) p) != h->pinned.end()) return SPMM_B200_OK;
	CU_TRY(cudaSetDevice(h->device));
	cudaPointerAttributes at;
	if (cudaPointerGetAttributes(&at, p) == cudaSuccess && at.type != cudaMemoryTypeUnregistered) return SPMM_B200_OK;
	cudaGetLastError();
	CU_TRY(cudaHostRegister((void*) p, bytes, cudaHostRegisterDefault));
	h->pinned.push_back((void*) p);
	return SPMM_B200_OK;
}

extern "C" int spmm_b200_unregister_host(spmm_b200_t* h, const void* p)
{
	if (!h || !p) return SPMM_B200_EINVAL;
	auto it = std::find(h->pinned.begin(), h->pinned.end(), (void*) p);
	if (it == h->pinned.end()) return SPMM_B200_EINVAL;
	h->pinned.erase(it);
	CU_TRY(cudaSetDevice(h->device));
	CU_TRY(cudaHostUnregister((void*) p));
	return SPMM_B200_OK;
}

// Rows of B the handle's nonzeros can read: [lo, lo + n) = first..last touched column (col_span_kernel at create).
// A row of B no nonzero references cannot change C, so host-buffer calls do not move it over PCIe: the reference's
// generator with bw = 0.05 puts every nonzero into the middle 5.2 % of the columns (cfg2: 13 of 256 MB).
static void b_window(const spmm_b200* h, int64_t& lo, int64_t& n)
{
	lo = 0; n = h->c;
	if (!h->opt.upload_window) return;
	if (h->nnz == 0 || h->col_span <= 0) { n = 0; return; }
	lo = h->col_min; n = std::min<int64_t>(h->col_span, h->c - lo);
}

// Upload rows [lo, lo + n) of the column block [col0, col0 + w) of the host operand (k columns in all) into the
// row-major device block Bs (c x w, row stride w) on stream s; column-major operands pass through d_x and a transpose.
static int upload_b_rows(spmm_b200* h, const double* x_host, int k, int col0, int w, int64_t lo, int64_t n, double* Bs, cudaStream_t s)
{
	if (n <= 0 || w <= 0) return SPMM_B200_OK;
	const int64_t c = h->c;
	if (h->b_layout == SPMM_B200_B_ROWMAJOR || k == 1) {
		if (w == k)
			CU_TRY(cudaMemcpyAsync(Bs + lo * k, x_host + lo * k, sizeof(double) * n * k, cudaMemcpyHostToDevice, s));
		else
			CU_TRY(cudaMemcpy2DAsync(Bs + lo * w, sizeof(double) * w, x_host + lo * k + col0, sizeof(double) * k, sizeof(double) * w, (size_t) n,
			                         cudaMemcpyHostToDevice, s));
	} else {
		double* xs = h->d_x + (size_t) col0 * c + lo;
		if (n == c)
			CU_TRY(cudaMemcpyAsync(xs, x_host + (size_t) col0 * c, sizeof(double) * c * w, cudaMemcpyHostToDevice, s));
		else
			CU_TRY(cudaMemcpy2DAsync(xs, sizeof(double) * c, x_host + (size_t) col0 * c + lo, sizeof(double) * c, sizeof(double) * n, (size_t) w,
			                         cudaMemcpyHostToDevice, s));
		dim3 grid((unsigned) ((n + 31) / 32), (unsigned) ((w + 31) / 32)), block(32, 8);
		transpose_b_rows_kernel<<<grid, block, 0, s>>>(xs, Bs + lo * w, c, n, w);
		CU_TRY(cudaGetLastError());
		h->launches++;
	}
	h->last_h2d_bytes += (int64_t) sizeof(double) * n * w;
	return SPMM_B200_OK;
}

static int upload_b(spmm_b200* h, const double* x_host, int k)
{
	const size_t nb = (size_t) h->c * k;
	int rc = ensure(&h->d_B, &h->cap_B, nb);
	if (rc) return rc;
	if (h->b_layout != SPMM_B200_B_ROWMAJOR && k != 1) {
		rc = ensure(&h->d_x, &h->cap_x, nb);
		if (rc) return rc;
	}
	int64_t lo, n;
	b_window(h, lo, n);
	rc = upload_b_rows(h, x_host, k, 0, k, lo, n, h->d_B, h->stream);
	if (rc) return rc;
	h->resident_k = k;
	return SPMM_B200_OK;
}

extern "C" int spmm_b200_transpose_b_device(spmm_b200_t* h, const double* x_dev, double* b_dev, int k, void* stream)
{
	if (!h || !x_dev || !b_dev || k < 1) return SPMM_B200_EINVAL;
	if (h->c == 0) return SPMM_B200_OK;
	CU_TRY(cudaSetDevice(h->device));
	dim3 grid((unsigned) ((h->c + 31) / 32), (unsigned) ((k + 31) / 32)), block(32, 8);
	transpose_b_kernel<<<grid, block, 0, (cudaStream_t) stream>>>(x_dev, b_dev, h->c, k);
	h->launches++;
	CU_TRY(cudaGetLastError());
	return SPMM_B200_OK;
}

static int run_between(spmm_b200* h, const double* b_dev, double* c_dev, int k, cudaStream_t s, cudaEvent_t e0, cudaEvent_t e1)
{
	CU_TRY(cudaEventRecord(e0, s));
	int n = launch_spmm(h, b_dev, c_dev, k, s);
	if (n < 0) return n;
	CU_TRY(cudaEventRecord(e1, s));
	return SPMM_B200_OK;
}

extern "C" int spmm_b200_run_device(spmm_b200_t* h, const double* b_dev, double* c_dev, int k, void* stream)
{
	if (!h || k < 1) return SPMM_B200_EINVAL;
	if ((h->c > 0 && h->nnz > 0 && !b_dev) || (h->m > 0 && !c_dev)) return SPMM_B200_EINVAL;
	CU_TRY(cudaSetDevice(h->device));
	// L2-aware k slicing: when the touched window of B (col_span x 8k bytes) is far beyond L2, the first calls with
	// this k run one candidate slice width each and are timed (they block until their kernels finish; every
	// candidate computes the same bits), then the fastest is kept -- per k, so alternating k values do not re-tune.
	// Never while the stream is being captured into a graph (the measurement synchronises).
	// (k >= 32 and 64 MB since round 2: gathers read by all SMs see roughly half of the 126 MB L2 -- cfg5-large's
	// 107 MB slice window still misses, profiles/r2_cfg5_large_sliced.md -- so a 77 MB window at k = 32 is worth a try too)
	bool tunable = h->variant == SPMM_B200_ROWSPLIT && k >= 32 && k % 16 == 0 && h->opt.kslice == 0
	               && (double) h->col_span * 8.0 * k > h->opt.kslice_min_mb * 1e6 && h->nnz > 0;
	if (tunable && h->ks_k != k) {
		h->ks_k = k; h->ks_best = 0; h->ks_try = 0;
		for (const KsliceChoice& d : h->ks_done)
			if (d.k == k) h->ks_best = d.best;
		int n = 0;
		h->ks_cand[n++] = k;
		for (int w = 64; w >= 16 && n < 4; w /= 2)
			if (w < k && k % w == 0) h->ks_cand[n++] = w;
		for (; n < 4; n++) h->ks_cand[n] = 0;
	} else if (!tunable && h->ks_k == k) {
		h->ks_k = 0;
	}
	bool measuring = tunable && h->ks_best == 0;
	if (measuring) {
		cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
		if (cudaStreamIsCapturing((cudaStream_t) stream, &cap) != cudaSuccess) cudaGetLastError();
		if (cap != cudaStreamCaptureStatusNone) { measuring = false; h->ks_k = 0; }     // unsliced inside a capture
	}
	int rc = run_between(h, b_dev, c_dev, k, (cudaStream_t) stream, h->ev0, h->ev1);
	if (rc) return rc;
	h->timed = true;
	h->timed_slices = 0;
	if (measuring) {
		float ms = 0.f;
		CU_TRY(cudaEventSynchronize(h->ev1));
		CU_TRY(cudaEventElapsedTime(&ms, h->ev0, h->ev1));
		// every candidate runs twice; the second run (warm caches, plan already resident) is the one that counts
		if (h->ks_try & 1) h->ks_ms[h->ks_try / 2] = ms;
		h->ks_try++;
		const int ncand = h->ks_try / 2;
		if ((h->ks_try & 1) == 0 && (ncand >= 4 || h->ks_cand[ncand] == 0)) {
			int best = 0;
			for (int i = 1; i < ncand; i++)
				if (h->ks_ms[i] < h->ks_ms[best]) best = i;
			h->ks_best = h->ks_cand[best];
			KsliceChoice d; d.k = k; d.best = h->ks_best;
			h->ks_done.push_back(d);
		}
	}
	return SPMM_B200_OK;
}

extern "C" int spmm_b200_set_b(spmm_b200_t* h, const double* x_host, int k)
{
	if (!h || !x_host || k < 1) return SPMM_B200_EINVAL;
	CU_TRY(cudaSetDevice(h->device));
	int rc = upload_b(h, x_host, k);
	if (rc) return rc;
	CU_TRY(cudaStreamSynchronize(h->stream));
	return SPMM_B200_OK;
}

extern "C" int spmm_b200_run_resident(spmm_b200_t* h, int k)
{
	if (!h || k < 1 || h->resident_k != k) return SPMM_B200_EINVAL;
	int rc = ensure(&h->d_C, &h->cap_C, (size_t) h->m * k);
	if (rc) return rc;
	h->c_slices = 1;
	rc = spmm_b200_run_device(h, h->d_B, h->d_C, k, h->stream);
	if (rc) return rc;
	CU_TRY(cudaStreamSynchronize(h->stream));
	return SPMM_B200_OK;
}

// d_C -> y_host (row-major m x k).  After a k-sliced run d_C holds c_slices blocks of m x k/c_slices.
static int download_c(spmm_b200* h, double* y_host, int k, cudaStream_t s)
{
	if (h->m == 0) return SPMM_B200_OK;
	if (h->c_slices <= 1) {
		CU_TRY(cudaMemcpyAsync(y_host, h->d_C, sizeof(double) * h->m * k, cudaMemcpyDeviceToHost, s));
		return SPMM_B200_OK;
	}
	const int ks = k / h->c_slices;
	for (int i = 0; i < h->c_slices; i++)
		CU_TRY(cudaMemcpy2DAsync(y_host + (size_t) i * ks, sizeof(double) * k, h->d_C + (size_t) i * h->m * ks, sizeof(double) * ks,
		                         sizeof(double) * ks, (size_t) h->m, cudaMemcpyDeviceToHost, s));
	return SPMM_B200_OK;
}

extern "C" int spmm_b200_fetch_c(spmm_b200_t* h, double* y_host, int k)
{
	if (!h || !y_host || k < 1 || h->cap_C < (size_t) h->m * k) return SPMM_B200_EINVAL;
	if (h->c_slices > 1 && k % h->c_slices) return SPMM_B200_EINVAL;
	if (h->c_slices == 0) return SPMM_B200_EINVAL;      // the last run stored C straight into its (mapped) y_host
	CU_TRY(cudaSetDevice(h->device));
	int rc = download_c(h, y_host, k, h->stream);
	if (rc) return rc;
	CU_TRY(cudaStreamSynchronize(h->stream));
	return SPMM_B200_OK;
}

// ---- k-sliced host pipeline --------------------------------------------------------------------
// spmm_b200_run is PCIe-bound (cfg2, k=32: 4.6 ms up + 0.9 ms kernel + 4.5 ms down).  The link is full
// duplex (measured 50 GB/s each way at once, profiles/r1_pcie_probe.md) and the oracle layout of B is k
// contiguous vectors, so the columns are cut into slices of 16: slice s+1 uploads while the kernel
// runs on slice s and C[:, slice s-1] downloads.  A 16-column slice of the row-major C is a 128-byte
// wide 2-D copy, which the copy engine still moves at 52.6 GB/s (64 bytes: 31 GB/s -- hence 16).
static int slice_width(const spmm_b200* h, int k)
{
	int w = h->opt.pipeline_cols;                                    // 0 = off, -1 = choose
	if (w < 0) {
		// at most 4 slices: every slice re-reads A, and wider slices run the better kernels
		// (cfg4, k=128: 27.8 ms with 32-column slices, 29.3 ms with 16, 38.8 ms unsliced)
		w = 16;
		while (k / w > 4 && k % (2 * w) == 0) w *= 2;
	}
	if (w <= 0 || k < 2 * w || k % w) return 0;
	// small operands are latency-bound, not bandwidth-bound: one upload, one kernel, one download
	const size_t min_bytes = (size_t) h->opt.pipeline_min_kb << 10;
	if ((size_t) h->c * k * sizeof(double) < min_bytes) return 0;
	return w;
}

// Timeline of the last pipelined run: events t0 (start) and, per slice, {H2D done, replication done, transpose done,
// kernel start, kernel end, D2H done}; read with spmm_b200_last_timeline.
enum { TL_H2D = 0, TL_BCAST, TL_TRANSPOSE, TL_K0, TL_K1, TL_D2H, TL_PER_SLICE };

static int run_pipelined(spmm_b200* h, const double* x_host, double* y_host, int k, int ks,
                         spmm_b200_bcast_fn bcast = nullptr, void* bcast_ctx = nullptr, int parts = 1, bool sync = true)
{
	const int S = k / ks;
	const size_t nb = (size_t) h->c * k;
	int rc = ensure(&h->d_B, &h->cap_B, nb);
	if (rc) return rc;
	if (h->b_layout == SPMM_B200_B_COLMAJOR) {
		rc = ensure(&h->d_x, &h->cap_x, nb);
		if (rc) return rc;
	}
	rc = ensure(&h->d_C, &h->cap_C, (size_t) h->m * k);
	if (rc) return rc;
	if (!h->s_up) CU_TRY(cudaStreamCreateWithFlags(&h->s_up, cudaStreamNonBlocking));
	if (!h->s_down) CU_TRY(cudaStreamCreateWithFlags(&h->s_down, cudaStreamNonBlocking));
	while (h->ev_slice.size() < (size_t) 3 * S) {
		cudaEvent_t e;
		CU_TRY(cudaEventCreate(&e));
		h->ev_slice.push_back(e);
	}
	while (h->ev_tl.size() < (size_t) TL_PER_SLICE * S + 1) {
		cudaEvent_t e;
		CU_TRY(cudaEventCreate(&e));
		h->ev_tl.push_back(e);
	}
	cudaEvent_t* tl = h->ev_tl.data() + 1;
	h->tl_slices = 0;
	h->resident_k = 0;                       // d_B is about to hold slices, not the c x k operand
	const size_t bslice = (size_t) h->c * ks, cslice = (size_t) h->m * ks;
	CU_TRY(cudaEventRecord(h->ev_tl[0], h->s_up));
	int64_t wlo, wn;
	b_window(h, wlo, wn);
	for (int s = 0; s < S; s++) {
		double* Bs = h->d_B + s * bslice;
		if (!bcast) {
			// no replication step: this handle uploads exactly the rows of B its own nonzeros read
			if (x_host) {
				rc = upload_b_rows(h, x_host, k, s * ks, ks, wlo, wn, Bs, h->s_up);
				if (rc) return rc;
			}
			CU_TRY(cudaEventRecord(tl[TL_PER_SLICE * s + TL_H2D], h->s_up));
			CU_TRY(cudaEventRecord(tl[TL_PER_SLICE * s + TL_BCAST], h->s_up));
		} else if (h->b_layout == SPMM_B200_B_COLMAJOR) {
			// columns [s*ks, (s+1)*ks) are contiguous in x: one H2D, then transpose to row-major c x ks
			if (x_host && parts > 1) {
				// every rank holds x: upload only part `rank` of the slice over this GPU's own PCIe link;
				// the replication callback then fills in the other ranks' parts over NVLink
				const size_t per = (bslice + parts - 1) / parts;
				const size_t lo = std::min(bslice, per * (size_t) h->rank), hi = std::min(bslice, lo + per);
				if (hi > lo) {
					CU_TRY(cudaMemcpyAsync(h->d_x + s * bslice + lo, x_host + s * bslice + lo, (hi - lo) * sizeof(double), cudaMemcpyHostToDevice, h->s_up));
					h->last_h2d_bytes += (int64_t) ((hi - lo) * sizeof(double));
				}
			} else if (x_host) {
				CU_TRY(cudaMemcpyAsync(h->d_x + s * bslice, x_host + s * bslice, bslice * sizeof(double), cudaMemcpyHostToDevice, h->s_up));
				h->last_h2d_bytes += (int64_t) (bslice * sizeof(double));
			}
			CU_TRY(cudaEventRecord(tl[TL_PER_SLICE * s + TL_H2D], h->s_up));
			if (bcast && bslice > 0 && bcast(bcast_ctx, h->d_x + s * bslice, bslice, parts, h->s_up)) return SPMM_B200_ECOMM;
			CU_TRY(cudaEventRecord(tl[TL_PER_SLICE * s + TL_BCAST], h->s_up));
			rc = spmm_b200_transpose_b_device(h, h->d_x + s * bslice, Bs, ks, h->s_up);
			if (rc) return rc;
		} else {
			if (x_host) {
				CU_TRY(cudaMemcpy2DAsync(Bs, sizeof(double) * ks, x_host + (size_t) s * ks, sizeof(double) * k, sizeof(double) * ks,
				                         (size_t) h->c, cudaMemcpyHostToDevice, h->s_up));
				h->last_h2d_bytes += (int64_t) (bslice * sizeof(double));
			}
			CU_TRY(cudaEventRecord(tl[TL_PER_SLICE * s + TL_H2D], h->s_up));
			if (bcast && bslice > 0 && bcast(bcast_ctx, Bs, bslice, 1, h->s_up)) return SPMM_B200_ECOMM;
			CU_TRY(cudaEventRecord(tl[TL_PER_SLICE * s + TL_BCAST], h->s_up));
		}
		CU_TRY(cudaEventRecord(tl[TL_PER_SLICE * s + TL_TRANSPOSE], h->s_up));
		CU_TRY(cudaEventRecord(h->ev_slice[3 * s], h->s_up));
	}
	// Direct store: when y_host is page-locked and mapped into the device's address space, the row-split kernel
	// writes its w-column slice of every C row (w*8 contiguous bytes, ldc = k) straight into host memory, so the
	// download of slice s is the kernel itself and overlaps the upload of slice s+1 with no copy-engine pass and no
	// staging in d_C (profiles/r1_pcie_probe.md: SM stores into mapped host memory run at the 2-D copy's rate).
	double* y_map = nullptr;
	if (y_host && h->m > 0 && S > 1 && h->variant == SPMM_B200_ROWSPLIT && (ks == 8 || ks == 16 || ks == 32 || ks == 64)
	    && h->opt.direct_store) {
		cudaPointerAttributes at;
		if (cudaPointerGetAttributes(&at, y_host) == cudaSuccess && at.type == cudaMemoryTypeHost && at.devicePointer)
			y_map = (double*) at.devicePointer;
		else
			cudaGetLastError();
		if (((uintptr_t) y_map & 15) != 0 || k % 2) y_map = nullptr;     // the kernels store C as double2: staged download instead
	}
	for (int s = 0; s < S; s++) {
		CU_TRY(cudaStreamWaitEvent(h->stream, h->ev_slice[3 * s], 0));
		if (y_map) {
			CU_TRY(cudaEventRecord(h->ev_slice[3 * s + 1], h->stream));
			CU_TRY(cudaEventRecord(tl[TL_PER_SLICE * s + TL_K0], h->stream));
			launch_rowsplit_width(h, h->d_B + s * bslice, y_map + (size_t) s * ks, ks, h->stream, ks, k);
			CU_TRY(cudaGetLastError());
			h->launches += 1;
			CU_TRY(cudaEventRecord(h->ev_slice[3 * s + 2], h->stream));
			CU_TRY(cudaEventRecord(tl[TL_PER_SLICE * s + TL_K1], h->stream));
			CU_TRY(cudaEventRecord(tl[TL_PER_SLICE * s + TL_D2H], h->stream));
			continue;
		}
		CU_TRY(cudaEventRecord(tl[TL_PER_SLICE * s + TL_K0], h->stream));
		rc = run_between(h, h->d_B + s * bslice, h->d_C + s * cslice, ks, h->stream, h->ev_slice[3 * s + 1], h->ev_slice[3 * s + 2]);
		if (rc) return rc;
		CU_TRY(cudaEventRecord(tl[TL_PER_SLICE * s + TL_K1], h->stream));
		if (y_host && h->m > 0) {
			CU_TRY(cudaStreamWaitEvent(h->s_down, h->ev_slice[3 * s + 2], 0));
			CU_TRY(cudaMemcpy2DAsync(y_host + (size_t) s * ks, sizeof(double) * k, h->d_C + s * cslice, sizeof(double) * ks,
			                         sizeof(double) * ks, (size_t) h->m, cudaMemcpyDeviceToHost, h->s_down));
		}
		CU_TRY(cudaEventRecord(tl[TL_PER_SLICE * s + TL_D2H], h->s_down));
	}
	if (y_host) h->last_d2h_bytes += (int64_t) sizeof(double) * h->m * k;
	h->c_slices = y_map ? 0 : S;            // 0: C went straight to the host, d_C holds nothing to fetch
	h->timed = true;
	h->timed_slices = S;
	h->tl_slices = S;
	if (!sync) return SPMM_B200_OK;          // the multi-GPU caller synchronises all devices at the end
	CU_TRY(cudaStreamSynchronize(h->stream));
	CU_TRY(cudaStreamSynchronize(h->s_down));
	return SPMM_B200_OK;
}

// ------------------------------------------------------------------------------------------
// Row-band host pipeline.  The k-slice pipeline above leaves a whole slice of B un-overlapped at the head and a whole
// slice of C at the tail (cfg2: 2.3 + 2.5 of 7.9 ms) and moves C in w*8-byte pieces.  When the matrix is banded -- the
// rows of C block b only read B rows up to band_max(b) -- the call can instead walk down the rows: upload the B rows
// block b adds to the window (all k columns), run the row-split kernel on the block's rows alone, download its C rows
// as ONE contiguous range.  Head and tail shrink to 1/R of the transfers and both directions of the link stay busy in
// between.  Whether a matrix is banded enough is decided by a transfer model over the per-block column ranges taken
// at create (row_block_span_kernel); ROWBAND=0 turns it off, ROWBAND=R forces R blocks.
// ------------------------------------------------------------------------------------------

struct BandStep { int64_t r0, r1; int64_t up[2][2]; int nup; };      // rows of C; up to two row ranges of B to upload first

// The steps of a row-band run whose block b covers the fine row blocks [cut[b], cut[b+1]) (cut[0] = 0, last = BAND_FINE).
static bool rowband_steps(const spmm_b200* h, const std::vector<int>& cut, std::vector<BandStep>& steps)
{
	if (cut.size() < 2 || cut.front() != 0 || cut.back() != BAND_FINE || h->m < BAND_FINE) return false;
	steps.clear();
	int64_t u_lo = 0, u_hi = 0;
	bool have = false;
	for (size_t b = 0; b + 1 < cut.size(); b++) {
		const int f0 = cut[b], f1 = cut[b + 1];
		if (f1 <= f0) return false;
		BandStep st;
		// block starts on multiples of 8 rows (the blocked variants work on 8-row block rows); the rows this moves across
		// a fine-block border have their columns in the neighbouring fine block, which an earlier step uploaded
		st.r0 = (h->m * f0 / BAND_FINE) & ~(int64_t) 7;
		st.r1 = f1 == BAND_FINE ? h->m : (h->m * f1 / BAND_FINE) & ~(int64_t) 7;
		st.nup = 0;
		int64_t lo = INT64_MAX, hi = -1;
		for (int f = f0; f < f1; f++)
			if (h->band_max[f] >= h->band_min[f]) { lo = std::min<int64_t>(lo, h->band_min[f]); hi = std::max<int64_t>(hi, (int64_t) h->band_max[f] + 1); }
		if (hi > lo) {
			if (!have) { st.up[st.nup][0] = lo; st.up[st.nup][1] = hi; st.nup++; u_lo = lo; u_hi = hi; have = true; }
			else {
				if (hi > u_hi) { st.up[st.nup][0] = u_hi; st.up[st.nup][1] = hi; st.nup++; u_hi = hi; }
				if (lo < u_lo) { st.up[st.nup][0] = lo; st.up[st.nup][1] = u_lo; st.nup++; u_lo = lo; }
			}
		}
		steps.push_back(st);
	}
	return true;
}

static std::vector<int> equal_cuts(int R)
{
	std::vector<int> cut;
	for (int b = 0; b <= R; b++) cut.push_back(BAND_FINE * b / R);
	return cut;
}

// Transfer model (one PCIe Gen5 x16 link; profiles/r1_pcie_probe.md, r2_e2e_probe.md, r2_rowband_probe.md): 55 GB/s one
// way for contiguous copies, ~42 GB/s per direction while both are busy, 48 GB/s for the w*8-byte pieces the k-slice
// kernels store into host memory, ~10 us per pipeline step for the hand-overs.
static double xfer_ms(double up, double down, double down_rate)
{
	const double one = 55e6, both = 42e6;        // bytes per ms
	const double shared = std::min(up, down);
	return shared / both + (up - shared) / one + (down - shared) / down_rate;
}

// Three queues (upload, kernel, download), each in block order; a block's kernel waits for its upload, its download
// for its kernel.  Both directions share the link when the uploads are comparable to the downloads.
static double rowband_model_ms(const spmm_b200* h, int k, const std::vector<BandStep>& steps)
{
	double up_all = 0.0;
	for (const BandStep& st : steps)
		for (int u = 0; u < st.nup; u++) up_all += (double) (st.up[u][1] - st.up[u][0]) * k * 8.0;
	const double down_all = (double) h->m * k * 8.0;
	const double rate = up_all > 0.25 * down_all ? 44e6 : 55e6;
	const double kernel_all = (double) h->nnz * k * 8.0 / 12e9;          // ms: B rows gathered at ~12 TB/s
	double t_up = 0.0, t_k = 0.0, t_d = 0.0;
	for (const BandStep& st : steps) {
		double up = 0.0;
		for (int u = 0; u < st.nup; u++) up += (double) (st.up[u][1] - st.up[u][0]) * k * 8.0;
		const double rows = (double) (st.r1 - st.r0);
		t_up += up / rate + (up > 0.0 ? 0.004 : 0.0);
		t_k = std::max(t_up, t_k) + kernel_all * rows / (double) h->m + 0.005;
		t_d = std::max(t_k, t_d) + rows * k * 8.0 / rate + 0.008;
	}
	return t_d;
}

static double kslice_model_ms(const spmm_b200* h, int k, int ks)
{
	int64_t lo, n;
	b_window(h, lo, n);
	const double bb = (double) n * k * 8.0, cb = (double) h->m * k * 8.0;
	if (ks <= 0) return bb / 55e6 + cb / 55e6;
	const int S = k / ks;
	return bb / S / 55e6 + (S - 1) * (xfer_ms(bb / S, cb / S, 48e6) + 0.012) + cb / S / 48e6;
}

// Row blocks for this call, 0 = use the k-slice pipeline.
static int rowband_blocks(const spmm_b200* h, int k, int ks, const double* x_host, const double* y_host, std::vector<BandStep>& steps)
{
	const int want = h->opt.rowband;
	if (want == 0 || !x_host || !y_host || h->m < BAND_FINE || h->nnz == 0) return 0;
	// kernels that can run on a row window: the row-split ticket kernel and the blocked (8-row block rows) kernels
	if (h->variant == SPMM_B200_ROWSPLIT) {
		if (!(k == 8 || k == 16 || k == 32 || k == 64 || k == 128) || !h->d_ticket || h->opt.unroll != 4) return 0;
	} else if (h->variant == SPMM_B200_BCSR || h->variant == SPMM_B200_BCSR_FMA) {
		if (!bcsr_supports_k(h->variant, k)) return 0;
	} else {
		return 0;
	}
	if (want > 0) return rowband_steps(h, equal_cuts(std::min(want, BAND_FINE)), steps) ? (int) steps.size() : 0;
	const std::vector<int> graded = {0, 1, 3, 8, 16, 32, 48, BAND_FINE};
	if (want == -2) return rowband_steps(h, graded, steps) ? (int) steps.size() : 0;      // probe: force the graded blocks
	int64_t wlo, wn;
	b_window(h, wlo, wn);
	if ((size_t) (wn + h->m) * k * sizeof(double) < 2 * ((size_t) h->opt.pipeline_min_kb << 10)) return 0;
	const double t_ks = kslice_model_ms(h, k, ks);
	int best = 0;
	double t_best = 0.97 * t_ks;            // near a tie the contiguous copies of the row blocks have been the faster ones (cfg2 on 1 and 2 GPUs)
	std::vector<BandStep> cand;
	// equal blocks for a diagonal band (every block brings its own rows of B); a small first block and few large ones
	// after it when the uploads are short (a column stripe): the download chain then starts early and has few seams
	for (const std::vector<int>& cut : {equal_cuts(8), equal_cuts(16), graded}) {
		if (!rowband_steps(h, cut, cand)) continue;
		const double t = rowband_model_ms(h, k, cand);
		if (t < t_best) { t_best = t; best = (int) cand.size(); steps = cand; }
	}
	return best;
}

struct RowWindowGuard {            // the row window never outlives run_rowband, whatever path leaves it
	spmm_b200* h;
	~RowWindowGuard() { h->win_r0 = 0; h->win_rows = -1; }
};

static int run_rowband(spmm_b200* h, const double* x_host, double* y_host, int k, const std::vector<BandStep>& steps, bool sync = true)
{
	const int R = (int) steps.size();
	const bool colmajor = h->b_layout == SPMM_B200_B_COLMAJOR;
	int rc = ensure(&h->d_B, &h->cap_B, (size_t) h->c * k);
	if (rc) return rc;
	if (colmajor) {
		rc = ensure(&h->d_x, &h->cap_x, (size_t) h->c * k);
		if (rc) return rc;
	}
	if (!h->s_up) CU_TRY(cudaStreamCreateWithFlags(&h->s_up, cudaStreamNonBlocking));
	if (!h->s_down) CU_TRY(cudaStreamCreateWithFlags(&h->s_down, cudaStreamNonBlocking));
	while (h->ev_slice.size() < (size_t) 3 * R) {
		cudaEvent_t e;
		CU_TRY(cudaEventCreate(&e));
		h->ev_slice.push_back(e);
	}
	while (h->ev_tl.size() < (size_t) TL_PER_SLICE * R + 1) {
		cudaEvent_t e;
		CU_TRY(cudaEventCreate(&e));
		h->ev_tl.push_back(e);
	}
	cudaEvent_t* tl = h->ev_tl.data() + 1;
	h->tl_slices = 0;
	h->band_blocks = 0;
	h->resident_k = 0;                       // d_B holds only the rows of B this matrix reads
	// C rows go down as ONE contiguous copy per block (cfg2, 8 blocks: 5.02 ms); ROWBAND_DIRECT=1 lets the kernel store
	// them straight into a page-locked, mapped y_host instead (5.46 ms: SM stores reach 50, the copy engine 55 GB/s)
	double* y_map = nullptr;
	if (h->opt.rowband_direct) {
		cudaPointerAttributes at;
		if (cudaPointerGetAttributes(&at, y_host) == cudaSuccess && at.type == cudaMemoryTypeHost && at.devicePointer)
			y_map = (double*) at.devicePointer;
		else
			cudaGetLastError();
		if (((uintptr_t) y_map & 15) != 0) y_map = nullptr;
	}
	if (!y_map) {
		rc = ensure(&h->d_C, &h->cap_C, (size_t) h->m * k);
		if (rc) return rc;
	}
	CU_TRY(cudaEventRecord(h->ev_tl[0], h->s_up));
	RowWindowGuard guard{h};
	double* const c_base = y_map ? y_map : h->d_C;
	for (int b = 0; b < R; b++) {
		const BandStep& st = steps[b];
		for (int u = 0; u < st.nup; u++) {
			rc = upload_b_rows(h, x_host, k, 0, k, st.up[u][0], st.up[u][1] - st.up[u][0], h->d_B, h->s_up);
			if (rc) return rc;
		}
		CU_TRY(cudaEventRecord(tl[TL_PER_SLICE * b + TL_H2D], h->s_up));
		CU_TRY(cudaEventRecord(tl[TL_PER_SLICE * b + TL_BCAST], h->s_up));
		CU_TRY(cudaEventRecord(tl[TL_PER_SLICE * b + TL_TRANSPOSE], h->s_up));
		CU_TRY(cudaEventRecord(h->ev_slice[3 * b], h->s_up));
		// block by block on all three streams: the first kernel is enqueued right behind the first upload
		CU_TRY(cudaStreamWaitEvent(h->stream, h->ev_slice[3 * b], 0));
		CU_TRY(cudaEventRecord(h->ev_slice[3 * b + 1], h->stream));
		CU_TRY(cudaEventRecord(tl[TL_PER_SLICE * b + TL_K0], h->stream));
		h->win_r0 = st.r0; h->win_rows = st.r1 - st.r0;
		if (h->win_rows > 0) {
			const int n = launch_spmm(h, h->d_B, c_base, k, h->stream);
			if (n < 0) return n;
		}
		CU_TRY(cudaEventRecord(h->ev_slice[3 * b + 2], h->stream));
		CU_TRY(cudaEventRecord(tl[TL_PER_SLICE * b + TL_K1], h->stream));
		if (y_map) {
			CU_TRY(cudaEventRecord(tl[TL_PER_SLICE * b + TL_D2H], h->stream));
		} else {
			CU_TRY(cudaStreamWaitEvent(h->s_down, h->ev_slice[3 * b + 2], 0));
			if (st.r1 > st.r0)
				CU_TRY(cudaMemcpyAsync(y_host + st.r0 * k, h->d_C + st.r0 * k, sizeof(double) * (st.r1 - st.r0) * k, cudaMemcpyDeviceToHost, h->s_down));
			CU_TRY(cudaEventRecord(tl[TL_PER_SLICE * b + TL_D2H], h->s_down));
		}
	}
	h->last_d2h_bytes += (int64_t) sizeof(double) * h->m * k;
	h->c_slices = y_map ? 0 : 1;
	h->timed = true;
	h->timed_slices = R;
	h->tl_slices = R;
	h->band_blocks = R;
	if (!sync) return SPMM_B200_OK;
	CU_TRY(cudaStreamSynchronize(h->stream));
	CU_TRY(cudaStreamSynchronize(h->s_down));
	return SPMM_B200_OK;
}

// out[0] = number of slices S of the last pipelined host-buffer run; then per slice six times in ms since the
// run's first enqueue: H2D done, replication done, transpose done, kernel start, kernel end, D2H done.
extern "C" int spmm_b200_last_timeline(const spmm_b200_t* h, double* out, int max_values)
{
	if (!h || !out || max_values < 1) return SPMM_B200_EINVAL;
	const int S = h->tl_slices;
	out[0] = S;
	int n = 1;
	for (int i = 0; i < S * TL_PER_SLICE && n < max_values; i++, n++) {
		float ms = 0.f;
		if (cudaEventElapsedTime(&ms, h->ev_tl[0], h->ev_tl[1 + i]) != cudaSuccess) { cudaGetLastError(); ms = -1.f; }
		out[n] = ms;
	}
	return n;
}

extern "C" int spmm_b200_run(spmm_b200_t* h, const double* x_host, double* y_host, int k, int copy_back)
{
	if (!h || !x_host || k < 1 || (copy_back && !y_host)) return SPMM_B200_EINVAL;
	CU_TRY(cudaSetDevice(h->device));
	const int ks = slice_width(h, k);
	h->last_h2d_bytes = h->last_d2h_bytes = 0;
	{
		std::vector<BandStep> steps;
		if (copy_back && rowband_blocks(h, k, ks, x_host, y_host, steps) > 0) return run_rowband(h, x_host, y_host, k, steps);
	}
	h->band_blocks = 0;
	if (ks > 0) return run_pipelined(h, x_host, copy_back ? y_host : nullptr, k, ks);
	int rc = upload_b(h, x_host, k);
	if (rc) return rc;
	rc = ensure(&h->d_C, &h->cap_C, (size_t) h->m * k);
	if (rc) return rc;
	h->c_slices = 1;
	rc = spmm_b200_run_device(h, h->d_B, h->d_C, k, h->stream);
	if (rc) return rc;
	if (copy_back) {
		rc = download_c(h, y_host, k, h->stream);
		if (rc) return rc;
		h->last_d2h_bytes += (int64_t) sizeof(double) * h->m * k;
	}
	CU_TRY(cudaStreamSynchronize(h->stream));
	return SPMM_B200_OK;
}

// Multi-GPU form of the hot call: the root rank holds x on the host, every rank gets the
// replica through the caller's broadcast (ncclBroadcast / torch.distributed.broadcast on the stream
// handed to it), slice by slice, so that upload, broadcast, kernel and download of C overlap.
extern "C" int spmm_b200_run_replicated(spmm_b200_t* h, const double* x_host, double* y_host, int k, int copy_back,
                                        spmm_b200_bcast_fn bcast, void* bcast_ctx, int x_on_every_rank)
{
	if (!h || k < 1 || (copy_back && !y_host) || (!x_host && !bcast)) return SPMM_B200_EINVAL;
	if (x_on_every_rank && !x_host) return SPMM_B200_EINVAL;
	CU_TRY(cudaSetDevice(h->device));
	h->last_h2d_bytes = h->last_d2h_bytes = 0;
	h->band_blocks = 0;
	int ks = slice_width(h, k);
	if (ks <= 0) ks = k;
	// split uploads need the oracle layout (a slice of x is one contiguous range) and a communicator
	const int parts = (x_on_every_rank && bcast && h->nranks > 1 && h->b_layout == SPMM_B200_B_COLMAJOR) ? h->nranks : 1;
	if (x_on_every_rank && parts == 1 && h->rank != 0) x_host = nullptr;      // fall back: only the root uploads
	return run_pipelined(h, x_host, copy_back ? y_host : nullptr, k, ks, bcast, bcast_ctx, parts);
}

// One process per GPU with routed downloads (include/spmm_b200.h): own window up, kernel on the shard, then either
// send C to the downloader rank or carry own + clients' rows into the shared y.
extern "C" int spmm_b200_run_routed(spmm_b200_t* h, const double* x_host, double* y_all, int k, int my_rank, int downloader_rank,
                                    const int* client_ranks, const int64_t* client_first_row, const int64_t* client_rows, int nclients,
                                    spmm_b200_xfer_fn xfer, void* xfer_ctx)
{
	if (!h || !x_host || !y_all || k < 1 || nclients < 0) return SPMM_B200_EINVAL;
	const bool download = downloader_rank == my_rank;
	if ((!download || nclients > 0) && !xfer) return SPMM_B200_EINVAL;
	if (nclients > 0 && (!download || !client_ranks || !client_first_row || !client_rows)) return SPMM_B200_EINVAL;
	CU_TRY(cudaSetDevice(h->device));
	h->last_h2d_bytes = h->last_d2h_bytes = 0;
	h->band_blocks = 0; h->tl_slices = 0;
	int rc = ensure(&h->d_B, &h->cap_B, (size_t) h->c * k);
	if (rc) return rc;
	if (h->b_layout == SPMM_B200_B_COLMAJOR && k != 1) {
		rc = ensure(&h->d_x, &h->cap_x, (size_t) h->c * k);
		if (rc) return rc;
	}
	size_t relay = 0;
	for (int i = 0; i < nclients; i++) {
		if (client_rows[i] < 0 || client_first_row[i] < 0) return SPMM_B200_EINVAL;
		relay += (size_t) client_rows[i] * k;
	}
	rc = ensure(&h->d_C, &h->cap_C, (size_t) h->m * k + relay);        // own rows, then the clients' rows behind them
	if (rc) return rc;
	if (!h->s_up) CU_TRY(cudaStreamCreateWithFlags(&h->s_up, cudaStreamNonBlocking));
	if (!h->s_down) CU_TRY(cudaStreamCreateWithFlags(&h->s_down, cudaStreamNonBlocking));
	while (h->ev_slice.size() < (size_t) 3 + nclients) {
		cudaEvent_t e;
		CU_TRY(cudaEventCreate(&e));
		h->ev_slice.push_back(e);
	}
	h->resident_k = 0;
	int64_t lo, n;
	b_window(h, lo, n);
	rc = upload_b_rows(h, x_host, k, 0, k, lo, n, h->d_B, h->s_up);
	if (rc) return rc;
	CU_TRY(cudaEventRecord(h->ev_slice[0], h->s_up));
	CU_TRY(cudaStreamWaitEvent(h->stream, h->ev_slice[0], 0));
	CU_TRY(cudaEventRecord(h->ev_slice[1], h->stream));
	if (h->m > 0) {
		const int nl = launch_spmm(h, h->d_B, h->d_C, k, h->stream);
		if (nl < 0) return nl;
	}
	CU_TRY(cudaEventRecord(h->ev_slice[2], h->stream));
	h->c_slices = 1; h->timed = true; h->timed_slices = 1;
	if (!download) {
		if (h->m > 0 && xfer(xfer_ctx, h->d_C, (size_t) h->m * k, downloader_rank, 1, h->stream)) return SPMM_B200_ECOMM;
		CU_TRY(cudaStreamSynchronize(h->stream));
		return SPMM_B200_OK;
	}
	CU_TRY(cudaStreamWaitEvent(h->s_down, h->ev_slice[2], 0));
	if (h->m > 0) {
		CU_TRY(cudaMemcpyAsync(y_all + (size_t) h->first_row * k, h->d_C, sizeof(double) * h->m * k, cudaMemcpyDeviceToHost, h->s_down));
		h->last_d2h_bytes += (int64_t) sizeof(double) * h->m * k;
	}
	double* dst = h->d_C + (size_t) h->m * k;
	for (int i = 0; i < nclients; i++) {
		const size_t cnt = (size_t) client_rows[i] * k;
		if (cnt == 0) continue;
		if (xfer(xfer_ctx, dst, cnt, client_ranks[i], 0, h->stream)) return SPMM_B200_ECOMM;
		CU_TRY(cudaEventRecord(h->ev_slice[3 + i], h->stream));
		CU_TRY(cudaStreamWaitEvent(h->s_down, h->ev_slice[3 + i], 0));
		CU_TRY(cudaMemcpyAsync(y_all + (size_t) client_first_row[i] * k, dst, sizeof(double) * cnt, cudaMemcpyDeviceToHost, h->s_down));
		h->last_d2h_bytes += (int64_t) (sizeof(double) * cnt);
		dst += cnt;
	}
	CU_TRY(cudaStreamSynchronize(h->stream));
	CU_TRY(cudaStreamSynchronize(h->s_down));
	return SPMM_B200_OK;
}

// ------------------------------------------------------------------------------------------
// partitioning (device search, host pointers in and out)
// ------------------------------------------------------------------------------------------

static size_t partition_out_len(int mode, int P)
{
	switch (mode) {
	case SPMM_B200_PART_NNZ:       return (size_t) P + 1;
	case SPMM_B200_PART_ROWS:      return 2 * (size_t) P;
	case SPMM_B200_PART_MERGE:     return 2 * ((size_t) P + 1);
	default:                       return 4 * (size_t) P;        // SPMM_B200_PART_NNZ_CARRY
	}
}

extern "C" int spmm_b200_partition(const int32_t* row_ptr, int64_t m, int64_t nnz, int P, int mode, int64_t* out, int device)
{
	if (!out || P < 1 || m < 0 || nnz < 0) return SPMM_B200_EINVAL;
	if (mode != SPMM_B200_PART_ROWS && !row_ptr) return SPMM_B200_EINVAL;
	if (mode < SPMM_B200_PART_ROWS || mode > SPMM_B200_PART_NNZ_CARRY) return SPMM_B200_EINVAL;
	int rc = check_device(device, nullptr);
	if (rc) return rc;
	const size_t n_out = partition_out_len(mode, P);
	int32_t* d_ia = nullptr;
	int64_t* d_out = nullptr;
	CU_TRY(cudaMalloc(&d_out, sizeof(int64_t) * n_out));
	if (mode != SPMM_B200_PART_ROWS) {
		cudaError_t e = cudaMalloc(&d_ia, sizeof(int32_t) * (m + 1));
		if (e != cudaSuccess) { cudaFree(d_out); CU_TRY(e); }
		e = cudaMemcpy(d_ia, row_ptr, sizeof(int32_t) * (m + 1), cudaMemcpyHostToDevice);
		if (e != cudaSuccess) { cudaFree(d_out); cudaFree(d_ia); CU_TRY(e); }
	}
	const unsigned nb = (unsigned) ((P + 1 + 255) / 256);
	if (mode == SPMM_B200_PART_NNZ) partition_nnz_kernel<<<nb, 256>>>(d_ia, m, nnz, P, d_out);
	else if (mode == SPMM_B200_PART_ROWS) partition_rows_kernel<<<nb, 256>>>(0, m, P, d_out);
	else if (mode == SPMM_B200_PART_MERGE) partition_merge_kernel<<<nb, 256>>>(d_ia, m, nnz, P, d_out);
	else partition_nnz_carry_kernel<<<nb, 256>>>(d_ia, m, nnz, P, d_out);
	cudaError_t e = cudaGetLastError();
	if (e == cudaSuccess) e = cudaMemcpy(out, d_out, sizeof(int64_t) * n_out, cudaMemcpyDeviceToHost);
	cudaFree(d_out); cudaFree(d_ia);
	CU_TRY(e);
	return SPMM_B200_OK;
}

// Host twin of spmm_b200_partition: the same integer functions (partition.cuh, __host__ __device__)
// evaluated on the CPU.  Shard planning only -- O(P log m) integer work, no SpMM arithmetic --
// so that a launcher can size per-rank buffers before any device exists (and so that the
// multi-rank plumbing can be tested with gloo on a CPU box).
extern "C" int spmm_b200_partition_host(const int32_t* row_ptr, int64_t m, int64_t nnz, int P, int mode, int64_t* out)
{
	if (!out || P < 1 || m < 0 || nnz < 0) return SPMM_B200_EINVAL;
	if (mode < SPMM_B200_PART_ROWS || mode > SPMM_B200_PART_NNZ_CARRY) return SPMM_B200_EINVAL;
	if (mode != SPMM_B200_PART_ROWS && !row_ptr) return SPMM_B200_EINVAL;
	if (mode == SPMM_B200_PART_NNZ) {
		for (int64_t t = 0; t <= P; t++) out[t] = nnz_split_start(row_ptr, m, nnz, P, t);
	} else if (mode == SPMM_B200_PART_ROWS) {
		for (int64_t t = 0; t < P; t++) equal_split(0, m, P, t, &out[2 * t], &out[2 * t + 1]);
	} else if (mode == SPMM_B200_PART_MERGE) {
		const int64_t L = m + nnz, items = (L + P - 1) / P;
		for (int64_t t = 0; t <= P; t++) {
			int64_t d = items * t; if (d > L) d = L;
			merge_path_search(d, row_ptr + 1, m, nnz, &out[2 * t], &out[2 * t + 1]);
		}
	} else {
		for (int64_t t = 0; t < P; t++) nnz_carry_split(row_ptr, m, nnz, P, t, out + 4 * t);
	}
	return SPMM_B200_OK;
}

// ------------------------------------------------------------------------------------------
// COO -> CSR (host pointers in and out, all work on the device)
// ------------------------------------------------------------------------------------------

namespace {
struct DevBuf {
	void* p = nullptr;
	~DevBuf() { if (p) cudaFree(p); }
	template <typename T> T* as() { return static_cast<T*>(p); }
};
}

extern "C" int spmm_b200_coo_to_csr(const int32_t* R, const int32_t* C, const double* V, int64_t m, int64_t n, int64_t nnz,
                                    int32_t* row_ptr, int32_t* col_idx, double* values, int sort_columns, int transpose, int device)
{
	if (m < 0 || n < 0 || nnz < 0 || !row_ptr) return SPMM_B200_EINVAL;
	if (nnz > 0 && (!R || !C || !col_idx)) return SPMM_B200_EINVAL;
	if ((V == nullptr) != (values == nullptr) && nnz > 0) return SPMM_B200_EINVAL;
	if (m >= INT32_MAX || n >= INT32_MAX || nnz >= INT32_MAX) return SPMM_B200_ERANGE;
	int sms = 0;
	int rc = check_device(device, &sms);
	if (rc) return rc;
	const int64_t nrows = transpose ? n : m;                 // csr_gen.c:170-173
	const int32_t* h_rows = transpose ? C : R;
	const int32_t* h_cols = transpose ? R : C;
	for (int64_t i = 0; i < nnz; i++)                         // the device histogram must stay in bounds
		if (h_rows[i] < 0 || h_rows[i] >= nrows) return SPMM_B200_EINVAL;

	DevBuf d_rows, d_cols, d_vals, d_ptr, d_cursor, d_keys, d_tmp, d_long, d_nlong, d_colout, d_valout;
	const size_t nz = (size_t) std::max<int64_t>(nnz, 1);
	CU_TRY(cudaMalloc(&d_rows.p, nz * 4));
	CU_TRY(cudaMalloc(&d_cols.p, nz * 4));
	if (V) CU_TRY(cudaMalloc(&d_vals.p, nz * 8));
	CU_TRY(cudaMalloc(&d_ptr.p, ((size_t) nrows + 1) * 4));
	CU_TRY(cudaMalloc(&d_cursor.p, ((size_t) nrows + 1) * 4));
	CU_TRY(cudaMalloc(&d_keys.p, nz * 8));
	CU_TRY(cudaMalloc(&d_tmp.p, scan_tmp_elems(nrows + 1) * 4));
	CU_TRY(cudaMalloc(&d_long.p, (size_t) std::max<int64_t>(nrows, 1) * 4));
	CU_TRY(cudaMalloc(&d_nlong.p, 4));
	CU_TRY(cudaMalloc(&d_colout.p, nz * 4));
	if (V) CU_TRY(cudaMalloc(&d_valout.p, nz * 8));
	cudaStream_t s = nullptr;
	if (nnz > 0) {
		CU_TRY(cudaMemcpyAsync(d_rows.p, h_rows, nnz * 4, cudaMemcpyHostToDevice, s));
		CU_TRY(cudaMemcpyAsync(d_cols.p, h_cols, nnz * 4, cudaMemcpyHostToDevice, s));
		if (V) CU_TRY(cudaMemcpyAsync(d_vals.p, V, nnz * 8, cudaMemcpyHostToDevice, s));
	}
	CU_TRY(cudaMemsetAsync(d_ptr.p, 0, ((size_t) nrows + 1) * 4, s));
	CU_TRY(cudaMemsetAsync(d_nlong.p, 0, 4, s));
	const unsigned nb = (unsigned) std::min<int64_t>((int64_t) sms * 16, (nnz + 255) / 256);
	if (nnz > 0) coo_histogram_kernel<<<nb, 256, 0, s>>>(d_rows.as<int32_t>(), nnz, d_ptr.as<int32_t>());
	exclusive_scan_i32(d_ptr.as<int32_t>(), nrows + 1, d_tmp.as<int32_t>(), s);
	CU_TRY(cudaMemcpyAsync(d_cursor.p, d_ptr.p, ((size_t) nrows + 1) * 4, cudaMemcpyDeviceToDevice, s));
	int32_t n_long = 0;
	if (nnz > 0) {
		coo_scatter_kernel<<<nb, 256, 0, s>>>(d_rows.as<int32_t>(), d_cols.as<int32_t>(), nnz, d_cursor.as<int32_t>(),
		                                     d_keys.as<unsigned long long>(), sort_columns ? 1 : 0);
		coo_sort_rows_warp_kernel<<<(unsigned) ((nrows + 7) / 8), 256, 0, s>>>(d_ptr.as<int32_t>(), (int) nrows, d_keys.as<unsigned long long>());
		coo_list_long_rows_kernel<<<(unsigned) ((nrows + 255) / 256), 256, 0, s>>>(d_ptr.as<int32_t>(), (int) nrows, d_long.as<int32_t>(), d_nlong.as<int32_t>());
		CU_TRY(cudaMemcpyAsync(&n_long, d_nlong.p, 4, cudaMemcpyDeviceToHost, s));
		CU_TRY(cudaStreamSynchronize(s));
		if (n_long > 0)
			coo_sort_rows_cta_kernel<<<(unsigned) n_long, 256, 0, s>>>(d_ptr.as<int32_t>(), d_long.as<int32_t>(), d_keys.as<unsigned long long>());
		coo_gather_kernel<<<nb, 256, 0, s>>>(d_keys.as<unsigned long long>(), nnz, d_cols.as<int32_t>(), d_vals.as<double>(),
		                                    d_colout.as<int32_t>(), d_valout.as<double>());
		CU_TRY(cudaGetLastError());
		CU_TRY(cudaMemcpyAsync(col_idx, d_colout.p, nnz * 4, cudaMemcpyDeviceToHost, s));
		if (V) CU_TRY(cudaMemcpyAsync(values, d_valout.p, nnz * 8, cudaMemcpyDeviceToHost, s));
	}
	CU_TRY(cudaMemcpyAsync(row_ptr, d_ptr.p, ((size_t) nrows + 1) * 4, cudaMemcpyDeviceToHost, s));
	CU_TRY(cudaStreamSynchronize(s));
	CU_TRY(cudaGetLastError());
	return SPMM_B200_OK;
}

// ------------------------------------------------------------------------------------------
// accuracy check (CheckAccuracy twin): gold SpMM in double-double + the reference's error metrics
// ------------------------------------------------------------------------------------------

extern "C" int spmm_b200_check_accuracy(spmm_b200_t* h, const double* x_host, const double* y_host, int k,
                                        double* stats, double* y_gold_host)
{
	if (!h || !x_host || !y_host || !stats || k < 1) return SPMM_B200_EINVAL;
	CU_TRY(cudaSetDevice(h->device));
	int rc = upload_b(h, x_host, k);             // row-major c x k on the device
	if (rc) return rc;
	const int64_t n = h->m * (int64_t) k;
	DevBuf d_gold, d_lo, d_test, d_part;
	const int nblocks = (int) std::max<int64_t>(1, std::min<int64_t>((int64_t) h->sm_count * 4, (n + 255) / 256));
	CU_TRY(cudaMalloc(&d_gold.p, sizeof(double) * std::max<int64_t>(n, 1)));
	CU_TRY(cudaMalloc(&d_lo.p, sizeof(double) * std::max<int64_t>(n, 1)));
	CU_TRY(cudaMalloc(&d_test.p, sizeof(double) * std::max<int64_t>(n, 1)));
	CU_TRY(cudaMalloc(&d_part.p, sizeof(double) * ACC_NPART * nblocks));
	std::vector<double> part((size_t) ACC_NPART * nblocks, 0.0);
	if (n > 0) {
		CU_TRY(cudaMemcpyAsync(d_test.p, y_host, sizeof(double) * n, cudaMemcpyHostToDevice, h->stream));
		gold_spmm_kernel<<<(unsigned) ((n + 255) / 256), 256, 0, h->stream>>>(h->d_ia, h->d_ja, h->d_a, h->d_B, d_gold.as<double>(), d_lo.as<double>(), h->m, k);
		accuracy_stats_kernel<256><<<nblocks, 256, 0, h->stream>>>(d_gold.as<double>(), d_lo.as<double>(), d_test.as<double>(), n, 1e-10, d_part.as<double>());
		h->launches += 2;
		CU_TRY(cudaGetLastError());
		CU_TRY(cudaMemcpyAsync(part.data(), d_part.p, sizeof(double) * part.size(), cudaMemcpyDeviceToHost, h->stream));
		if (y_gold_host) CU_TRY(cudaMemcpyAsync(y_gold_host, d_gold.p, sizeof(double) * n, cudaMemcpyDeviceToHost, h->stream));
	}
	CU_TRY(cudaStreamSynchronize(h->stream));
	double acc[ACC_NPART] = {0, 0, 0, 0, 0, 0, 0};
	for (int b = 0; b < nblocks; b++)
		for (int i = 0; i < ACC_NPART; i++) {
			const double v = part[(size_t) b * ACC_NPART + i];
			acc[i] = (i == 0 || i == 2) ? std::max(acc[i], v) : acc[i] + v;
		}
	const double N = (double) n;
	stats[0] = acc[0];
	stats[1] = acc[1] / N;                        // mae        (array_metrics.c ARRAY_METRICS_mae_output)
	stats[2] = acc[2];                            // max_ae
	stats[3] = acc[3] / N;                        // mse
	stats[4] = 100.0 * (acc[4] / N);              // mape = 100 * mare
	stats[5] = 100.0 * (acc[5] / N);              // smape = 100 * smare
	stats[6] = acc[6] / N;                        // lnQ_error
	const long double e = stats[6];               // ARRAY_METRICS_mlare_output: extended precision for the power
	stats[7] = (double) log10l(fabsl(powl(10, e) - 1));
	stats[8] = pow(10, stats[7]);                 // gmare
	stats[9] = (acc[0] <= 1e-10) ? 1.0 : 0.0;
	return SPMM_B200_OK;
}

// ------------------------------------------------------------------------------------------
// introspection
// ------------------------------------------------------------------------------------------

extern "C" int64_t spmm_b200_rows(const spmm_b200_t* h) { return h ? h->m : -1; }
extern "C" int64_t spmm_b200_nnz(const spmm_b200_t* h) { return h ? h->nnz : -1; }
extern "C" int spmm_b200_variant(const spmm_b200_t* h) { return h ? h->variant : SPMM_B200_EINVAL; }
extern "C" int64_t spmm_b200_bcsr_blocks(const spmm_b200_t* h) { return h ? h->bcsr.nnzb : -1; }
extern "C" int spmm_b200_kslice(const spmm_b200_t* h) { return h ? h->ks_best : SPMM_B200_EINVAL; }
extern "C" int spmm_b200_last_rowband(const spmm_b200_t* h) { return h ? h->band_blocks : SPMM_B200_EINVAL; }
extern "C" int spmm_b200_col_window(const spmm_b200_t* h, int64_t* first_col, int64_t* ncols)
{
	if (!h) return SPMM_B200_EINVAL;
	if (first_col) *first_col = h->col_min;
	if (ncols) *ncols = h->col_span;
	return SPMM_B200_OK;
}
extern "C" int spmm_b200_last_transfer_bytes(const spmm_b200_t* h, int64_t* h2d_bytes, int64_t* d2h_bytes)
{
	if (!h) return SPMM_B200_EINVAL;
	if (h2d_bytes) *h2d_bytes = h->last_h2d_bytes;
	if (d2h_bytes) *d2h_bytes = h->last_d2h_bytes;
	return SPMM_B200_OK;
}
extern "C" int64_t spmm_b200_launch_count(const spmm_b200_t* h) { return h ? h->launches : -1; }

extern "C" double spmm_b200_last_kernel_ms(const spmm_b200_t* h)
{
	if (!h || !h->timed) return -1.0;
	float ms = 0.f;
	if (h->timed_slices > 0) {               // k-sliced run: sum of the slices' kernel times
		double sum = 0.0;
		for (int s = 0; s < h->timed_slices; s++) {
			if (cudaEventSynchronize(h->ev_slice[3 * s + 2]) != cudaSuccess) return -1.0;
			if (cudaEventElapsedTime(&ms, h->ev_slice[3 * s + 1], h->ev_slice[3 * s + 2]) != cudaSuccess) return -1.0;
			sum += ms;
		}
		return sum;
	}
	if (cudaEventSynchronize(h->ev1) != cudaSuccess) return -1.0;
	if (cudaEventElapsedTime(&ms, h->ev0, h->ev1) != cudaSuccess) return -1.0;
	return (double) ms;
}

extern "C" int spmm_b200_stats_labels(char* buf, long buf_n)
{
	if (!buf || buf_n <= 0) return 0;
	return snprintf(buf, buf_n, ",b200_variant,b200_ctas,b200_max_row,b200_kernel_ms,b200_kernel_gflops,b200_hbm_roofline_frac_8TBs");
}

extern "C" int spmm_b200_stats(const spmm_b200_t* h, char* buf, long buf_n)
{
	if (!h || !buf || buf_n <= 0) return 0;
	const double ms = spmm_b200_last_kernel_ms(h);
	const int k = h->resident_k > 0 ? h->resident_k : h->k;
	const double gflops = ms > 0 ? 2.0 * h->nnz * k / (ms * 1e6) : 0.0;
	const double bytes = 12.0 * h->nnz + 4.0 * (h->m + 1) + 8.0 * k * (double) h->c + 8.0 * k * (double) h->m;
	const double frac = ms > 0 ? (bytes / 8.0e12) / (ms * 1e-3) : 0.0;
	static const char* names[] = {"rowsplit", "merge", "bcsr_dmma", "auto", "bcsr_fma"};
	return snprintf(buf, buf_n, ",%s,%d,%d,%g,%g,%g", names[h->variant >= 0 && h->variant <= 4 ? h->variant : 3],
	                h->variant == SPMM_B200_MERGE ? h->merge.ntiles : h->n_ctas, h->max_row, ms, gflops, frac);
}

extern "C" const char* spmm_b200_strerror(int code)
{
	switch (code) {
	case SPMM_B200_OK:     return "ok";
	case SPMM_B200_EINVAL: return "invalid argument";
	case SPMM_B200_ENODEV: return "no usable sm_100 CUDA device (this library has no CPU fallback)";
	case SPMM_B200_ECUDA:  return "CUDA runtime error";
	case SPMM_B200_ENOMEM: return "out of memory";
	case SPMM_B200_ERANGE: return "size exceeds the int32 index contract";
	case SPMM_B200_ENOSUP: return "variant not supported for this input";
	case SPMM_B200_ECOMM:  return "the caller's broadcast callback failed";
	case SPMM_B200_EIO:    return "cannot open or read the matrix file";
	case SPMM_B200_EFORMAT: return "malformed matrix file";
	case SPMM_B200_ESCRATCH: return "input needs more scratch than the fixed no-allocation array holds";
	default:               return "unknown error";
	}
}

extern "C" const char* spmm_b200_last_cuda_error(void) { return g_cuda_err; }

extern "C" int spmm_b200_device_count(void)
{
	int n = 0;
	if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
	return n;
}

// ------------------------------------------------------------------------------------------
// single-process multi-GPU entry points
// ------------------------------------------------------------------------------------------
#include "spmm_b200_multi.inl"

// ------------------------------------------------------------------------------------------
// fp32 entry points (the reference's "_f" builds)
// ------------------------------------------------------------------------------------------
#include "spmm_b200_f32.inl"

// ------------------------------------------------------------------------------------------
// sparse-attention pipeline entry points (include/spmm_b200_pipeline.h)
// ------------------------------------------------------------------------------------------
#include "spmm_b200_pipeline.inl"

// ------------------------------------------------------------------------------------------
// Allocation-free entry point: spmm_b200_spmm_merge_noalloc (include/spmm_b200.h).
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It runs the merge-path SpMM -- the one
// kernel family here whose scratch (tile start coordinates, per-tile carries, carry rows) depends on the input -- with
// that scratch in ONE file-scope device array of SPMM_B200_NOALLOC_SCRATCH_BYTES (8 MiB) that every GPU which loads the
// library holds until exit.  The caller passes CSR, B and C as device pointers.  Nothing here calls cudaMalloc / cudaFree
// or any allocating routine: three kernels and two memset nodes on the caller's stream.
//
// g_noalloc_scratch is, in HOST code, only a stand-in for the symbol: its device address is asked for once per device,
// with that device current (cudaGetSymbolAddress), and that address is what the kernels receive.
// Every call on a device shares the array and finds the previous call's contents in it: the parts a call reads before
// writing are cleared first, and calls that overlap on the GPU (two streams, two threads) RACE on it.
// ------------------------------------------------------------------------------------------
#include <mutex>

static __device__ __align__(256) unsigned char g_noalloc_scratch[SPMM_B200_NOALLOC_SCRATCH_BYTES];

static constexpr int NOALLOC_MAX_DEVICES = 64;
static std::mutex g_noalloc_mu;
static unsigned char* g_noalloc_ptr[NOALLOC_MAX_DEVICES];      // device address of g_noalloc_scratch on device d (null: not asked yet)
static uint32_t g_noalloc_attr[NOALLOC_MAX_DEVICES];           // bit i: shared-memory limit of merge instantiation i raised on device d

template <int G, int V, int IPG>
struct MergeLaunchNoAlloc {
	using Cfg = MergeCfg<G, V, IPG, THREADS>;
	static cudaError_t launch(int dev, int bit, const int32_t* ia, const int32_t* ja, const double* a, const double* B, double* C,
	                          int ntiles, const int2* tile, double* carry, int32_t* carry_row, cudaStream_t s)
	{
		auto kern = spmm_merge_kernel<G, V, IPG, 4, THREADS>;
		if (Cfg::SMEM > 48 * 1024) {
			std::lock_guard<std::mutex> lock(g_noalloc_mu);
			if (!(g_noalloc_attr[dev] >> bit & 1u)) {
				cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) Cfg::SMEM);
				if (e != cudaSuccess) return e;
				g_noalloc_attr[dev] |= 1u << bit;
			}
		}
		kern<<<ntiles, THREADS, Cfg::SMEM, s>>>(ia, ja, a, B, C, tile, carry, carry_row);
		return cudaGetLastError();
	}
};

static size_t round256(size_t v) { return (v + 255) & ~(size_t) 255; }

extern "C" int spmm_b200_spmm_merge_noalloc(const int32_t* row_ptr_dev, const int32_t* col_ind_dev, const double* values_dev,
                                            int64_t m, int64_t c, int64_t nnz, const double* b_dev, double* c_dev, int k, void* stream)
{
	int rc = check_sizes(row_ptr_dev, col_ind_dev, values_dev, m, c, nnz, k);
	if (rc) return rc;
	if ((nnz > 0 && !b_dev) || (m > 0 && !c_dev)) return SPMM_B200_EINVAL;
	const int items = merge_items_for_k(k);
	if (items <= 0) {
		snprintf(g_cuda_err, sizeof(g_cuda_err), "spmm_b200_spmm_merge_noalloc: k = %d has no merge-path instantiation (1, 2, 4, 8, 16, 32, 64, 128)", k);
		return SPMM_B200_ENOSUP;
	}
	// the kernels move B and C as 16-byte pieces from k = 2 on
	if (k >= 2 && ((((uintptr_t) b_dev | (uintptr_t) c_dev) & 15) != 0)) return SPMM_B200_EINVAL;
	// scratch layout: tile start coordinates | per-tile carries (k values each) | per-tile carry rows
	const int64_t ntiles64 = (m + nnz + items - 1) / items;
	const size_t off_carry = round256(sizeof(int2) * ((size_t) ntiles64 + 1));
	const size_t off_row = off_carry + round256(sizeof(double) * (size_t) ntiles64 * k);
	const size_t need = off_row + round256(sizeof(int32_t) * (size_t) ntiles64);
	if (need > (size_t) SPMM_B200_NOALLOC_SCRATCH_BYTES) {
		snprintf(g_cuda_err, sizeof(g_cuda_err),
		         "spmm_b200_spmm_merge_noalloc: m + nnz = %lld at k = %d needs %zu bytes of scratch (%lld tiles of %d items), the fixed array holds %zu; "
		         "use spmm_b200_create + spmm_b200_run_device for this size",
		         (long long) (m + nnz), k, need, (long long) ntiles64, items, (size_t) SPMM_B200_NOALLOC_SCRATCH_BYTES);
		return SPMM_B200_ESCRATCH;
	}
	int dev = -1;
	if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); return SPMM_B200_ENODEV; }
	if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) return SPMM_B200_ENOSUP;
	unsigned char* scratch = nullptr;
	{
		// once per device, with that device current: the DEVICE address of the array (the name itself is a host-side stand-in)
		std::lock_guard<std::mutex> lock(g_noalloc_mu);
		if (!g_noalloc_ptr[dev]) {
			void* p = nullptr;
			cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
			if (e != cudaSuccess || !p) {
				snprintf(g_cuda_err, sizeof(g_cuda_err), "cudaGetSymbolAddress(g_noalloc_scratch) on device %d: %s", dev, cudaGetErrorString(e));
				cudaGetLastError();
				return SPMM_B200_ECUDA;
			}
			g_noalloc_ptr[dev] = (unsigned char*) p;
		}
		scratch = g_noalloc_ptr[dev];
	}
	const int ntiles = (int) ntiles64;
	if (m == 0 || ntiles == 0) return SPMM_B200_OK;
	cudaStream_t s = (cudaStream_t) stream;
	int2* tile = reinterpret_cast<int2*>(scratch);
	double* carry = reinterpret_cast<double*>(scratch + off_carry);
	int32_t* carry_row = reinterpret_cast<int32_t*>(scratch + off_row);
	// the array still holds the previous call's carries: every tile rewrites its own, but start from zero all the same
	// so that no call can depend on what another left behind
	CU_TRY(cudaMemsetAsync(scratch + off_carry, 0, sizeof(double) * (size_t) ntiles * k, s));
	CU_TRY(cudaMemsetAsync(scratch + off_row, 0, sizeof(int32_t) * (size_t) ntiles, s));
	tile_merge_kernel<<<(ntiles + 1 + 255) / 256, 256, 0, s>>>(row_ptr_dev, (int) m, nnz, items, ntiles, tile);
	CU_TRY(cudaGetLastError());
	cudaError_t e = cudaErrorInvalidValue;
#define WHAT(G, V, IPG, BIT) e = MergeLaunchNoAlloc<G, V, IPG>::launch(dev, BIT, row_ptr_dev, col_ind_dev, values_dev, b_dev, c_dev, ntiles, tile, carry, carry_row, s)
	MERGE_DISPATCH(k, WHAT)
#undef WHAT
	if (e != cudaSuccess) {
		snprintf(g_cuda_err, sizeof(g_cuda_err), "no-alloc merge kernel launch (k=%d): %s", k, cudaGetErrorString(e));
		return SPMM_B200_ECUDA;
	}
	const int64_t n = (int64_t) ntiles * k;
	spmm_merge_fixup_kernel<THREADS><<<(unsigned) ((n + THREADS - 1) / THREADS), THREADS, 0, s>>>(carry, carry_row, ntiles, (int) m, k, c_dev);
	CU_TRY(cudaGetLastError());
	return SPMM_B200_OK;
}
