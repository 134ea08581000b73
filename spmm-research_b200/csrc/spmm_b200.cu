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

struct spmm_b200 {
	int device = 0;
	int variant = SPMM_B200_ROWSPLIT;     // resolved variant
	int variant_req = SPMM_B200_AUTO;
	int b_layout = SPMM_B200_B_COLMAJOR;
	int k = 0;                            // k the work lists were tuned for
	int sm_count = 148;
	int64_t m = 0, c = 0, nnz = 0;        // this handle's shard
	int64_t first_row = 0, m_global = 0;
	int rank = 0, nranks = 1;             // shard identity (spmm_b200_create_shard)
	int max_row = 0;                      // longest row of the shard
	int64_t col_span = 0;                 // max column - min column + 1 over the shard's nonzeros
	// L2-aware k slicing of the row-split kernel, chosen by timing the candidates on the first calls
	int ks_k = 0;                         // k the choice below is for
	int ks_best = 0;                      // chosen slice width (== k: one launch); 0 while still measuring
	int ks_try = 0;                       // index of the candidate the next call runs
	int ks_cand[4] = {0, 0, 0, 0};
	float ks_ms[4] = {0.f, 0.f, 0.f, 0.f};

	int32_t* d_ia = nullptr;
	int32_t* d_ja = nullptr;
	double* d_a = nullptr;

	// row-split work list
	int32_t* d_cta_rows = nullptr;
	int n_ctas = 0;
	int* d_ticket = nullptr;              // row-ticket counter of the persistent row-split experiment
	// row-padded copy of A for the vectorised row-split kernel (rows padded to multiples of 4 entries)
	int32_t* d_iap = nullptr;
	int32_t* d_jap = nullptr;
	double* d_ap = nullptr;
	int64_t nnz_pad = 0;
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
	int c_slices = 1;                     // >1: d_C holds c_slices blocks of m x (k / c_slices)
	int timed_slices = 0;                 // >0: last_kernel_ms sums the slices' kernel events
	int64_t launches = 0;
};

static int env_int(const char* name, int dflt)
{
	const char* s = getenv(name);
	return (s && *s) ? atoi(s) : dflt;
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
	const int per_sm = std::max(1, env_int("SPMM_B200_CTAS_PER_SM", dflt));
	int64_t want = (int64_t) h->sm_count * per_sm;
	int64_t cap = std::max<int64_t>(1, (h->m + 7) / 8);       // at least ~8 rows per CTA
	// few rows (the row cap binds): whole waves of the 4 resident CTAs per SM, so that no partial last wave
	// runs on a fraction of the machine (m = 8192, k = 64, 410 nnz/row: 1024 CTAs 0.141 ms, 592 CTAs 0.123 ms;
	// profiles/r1_pipeline_probe.md)
	const int64_t wave = (int64_t) h->sm_count * 4;
	if (cap < want && h->k >= 64 && cap >= wave && !env_int("SPMM_B200_CTAS_PER_SM", 0)) cap = cap / wave * wave;
	h->n_ctas = (int) std::max<int64_t>(1, std::min(want, cap));
	if (h->d_cta_rows) cudaFree(h->d_cta_rows);
	CU_TRY(cudaMalloc(&h->d_cta_rows, sizeof(int32_t) * (h->n_ctas + 1)));
	if (!h->d_ticket) CU_TRY(cudaMalloc(&h->d_ticket, sizeof(int)));
	const int nb = (h->n_ctas + 1 + 255) / 256;
	cta_rows_nnz_kernel<<<nb, 256, 0, h->stream>>>(h->d_ia, (int) h->m, h->nnz, h->n_ctas, h->d_cta_rows);
	h->launches++;
	CU_TRY(cudaGetLastError());
	return SPMM_B200_OK;
}

// Row-padded copy of A for spmm_rowsplit_vec_kernel: iap = exclusive scan of the padded row lengths.
static int build_padded_plan(spmm_b200* h)
{
	if (h->d_iap || h->m == 0) return SPMM_B200_OK;
	if (h->nnz + 3 * h->m >= INT32_MAX) return SPMM_B200_ENOSUP;
	const int m = (int) h->m;
	int32_t* d_tmp = nullptr;
	CU_TRY(cudaMalloc(&h->d_iap, sizeof(int32_t) * ((size_t) m + 1)));
	CU_TRY(cudaMalloc(&d_tmp, sizeof(int32_t) * scan_tmp_elems(m + 1)));
	pad_len_kernel<<<(m + 1 + 255) / 256, 256, 0, h->stream>>>(h->d_ia, m, h->d_iap);
	h->launches += 1 + exclusive_scan_i32(h->d_iap, (int64_t) m + 1, d_tmp, h->stream);
	int32_t total = 0;
	CU_TRY(cudaMemcpyAsync(&total, h->d_iap + m, sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
	CU_TRY(cudaStreamSynchronize(h->stream));
	cudaFree(d_tmp);
	h->nnz_pad = total;
	CU_TRY(cudaMalloc(&h->d_jap, sizeof(int32_t) * std::max<size_t>((size_t) total, 4)));
	CU_TRY(cudaMalloc(&h->d_ap, sizeof(double) * std::max<size_t>((size_t) total, 4)));
	pad_fill_kernel<<<(unsigned) (((int64_t) m * 32 + 255) / 256), 256, 0, h->stream>>>(h->d_ia, h->d_iap, h->d_ja, h->d_a, m, h->d_jap, h->d_ap);
	h->launches++;
	CU_TRY(cudaGetLastError());
	CU_TRY(cudaStreamSynchronize(h->stream));
	return SPMM_B200_OK;
}

template <int G, int V>
static void launch_rowsplit_vec(const spmm_b200* h, const double* B, double* C, cudaStream_t s, int ldb = G * V, int ldc = G * V)
{
	spmm_rowsplit_vec_kernel<G, V, THREADS><<<h->n_ctas, THREADS, 0, s>>>(h->d_ia, h->d_iap, h->d_jap, h->d_ap, B, C, h->d_cta_rows, ldb, ldc);
}

static int pick_small_group(const spmm_b200* h)
{
	const double avg = h->m > 0 ? (double) h->nnz / (double) h->m : 1.0;
	int g = 2;
	while (g < 32 && g * 2 <= avg) g *= 2;
	return g;
}

template <int G, int V>
static void launch_rowsplit_gv(const spmm_b200* h, const double* B, double* C, cudaStream_t s, int ldb = G * V, int ldc = G * V)
{
	// B-row gathers in flight per lane group; tuning knob, default from the ncu study in profiles/
	static const int unroll = env_int("SPMM_B200_UNROLL", 4);
	constexpr int K = G * V;
	// Default: ONE resident wave of persistent warps that draw small row chunks from a ticket counter
	// (spmm_rowsplit_ticket_kernel).  The whole GPU then sweeps A and C front to back (DRAM / L2 locality of the
	// A stream) with no tail: cfg2 0.844 -> 0.754 ms on the same box (profiles/r1_ticket_probe.md).  Chunk = 8 rows
	// (rounded to a multiple of the rows a warp works on at once).  SPMM_B200_TICKET=0 keeps the static
	// nnz-balanced CTA ranges everywhere, > 0 forces that many rows per ticket.
	int tick = env_int("SPMM_B200_TICKET", -1);
	if (tick < 0) {
		// measured over bench/medium_subset.txt x k in {4, 8, 32, 128} (profiles/r1_ticket_probe.md): rows of >= 50
		// nonzeros gain 3-15 %, rows of <= 10 lose 3-15 % (a ticket per ~50 nonzeros is too fine, larger chunks lose
		// the locality), launches of a few tens of microseconds lose to the counter reset and the atomics
		const double avg = h->m > 0 ? (double) h->nnz / (double) h->m : 1.0;
		tick = (avg >= 32.0 && h->nnz >= (8 << 20)) ? 8 : 0;
	}
	if (tick > 0 && h->d_ticket && unroll == 4) {
		constexpr int NGW = 32 / G;
		tick = (tick + NGW - 1) / NGW * NGW;
		cudaMemsetAsync(h->d_ticket, 0, sizeof(int), s);
		const int grid = (int) std::min<int64_t>((h->m + tick * (THREADS / 32) - 1) / (tick * (THREADS / 32)) + 1,
		                                         (int64_t) h->sm_count * std::max(1, env_int("SPMM_B200_TICKET_CTAS", 4)));
		if (ldb == K && (uint64_t) h->c * K < (1ull << 32) && env_int("SPMM_B200_CONST_LD", 1))
			spmm_rowsplit_ticket_kernel<G, V, 4, THREADS, K><<<grid, THREADS, 0, s>>>(h->d_ia, h->d_ja, h->d_a, B, C, (int) h->m, tick, h->d_ticket, ldb, ldc);
		else
			spmm_rowsplit_ticket_kernel<G, V, 4, THREADS><<<grid, THREADS, 0, s>>>(h->d_ia, h->d_ja, h->d_a, B, C, (int) h->m, tick, h->d_ticket, ldb, ldc);
		return;
	}
	if (unroll >= 8 && G >= 8)
		spmm_rowsplit_kernel<G, V, 8, THREADS><<<h->n_ctas, THREADS, 0, s>>>(h->d_ia, h->d_ja, h->d_a, B, C, h->d_cta_rows, ldb, ldc);
	else if (unroll <= 2)
		spmm_rowsplit_kernel<G, V, 2, THREADS><<<h->n_ctas, THREADS, 0, s>>>(h->d_ia, h->d_ja, h->d_a, B, C, h->d_cta_rows, ldb, ldc);
	else if (ldb == K && (uint64_t) h->c * K < (1ull << 32) && env_int("SPMM_B200_CONST_LD", 1))
		// whole operand: compile-time row stride, 32-bit element offsets
		spmm_rowsplit_kernel<G, V, 4, THREADS, K><<<h->n_ctas, THREADS, 0, s>>>(h->d_ia, h->d_ja, h->d_a, B, C, h->d_cta_rows, ldb, ldc);
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
	static cudaError_t launch(const spmm_b200* h, const double* B, double* C, cudaStream_t s)
	{
		auto kern = spmm_merge_kernel<G, V, IPG, 4, THREADS>;
		static bool attr_set = false;
		if (!attr_set && Cfg::SMEM > 48 * 1024) {
			cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) Cfg::SMEM);
			if (e != cudaSuccess) return e;
			attr_set = true;
		}
		kern<<<h->merge.ntiles, THREADS, Cfg::SMEM, s>>>(h->d_ia, h->d_ja, h->d_a, B, C, h->merge.d_tile,
		                                                h->merge.d_carry, h->merge.d_carry_row);
		return cudaGetLastError();
	}
};

// k -> (G, V, IPG) of the merge kernel; 0 items = no specialised instantiation
#define MERGE_DISPATCH(k, WHAT)                                    \
	switch (k) {                                                   \
	case 1:   WHAT(1, 1, 16); break;                               \
	case 2:   WHAT(1, 2, 16); break;                               \
	case 4:   WHAT(2, 2, 16); break;                               \
	case 8:   WHAT(4, 2, 32); break;                               \
	case 16:  WHAT(8, 2, 32); break;                               \
	case 32:  WHAT(8, 4, 64); break;                               \
	case 64:  WHAT(16, 4, 128); break;                             \
	case 128: WHAT(32, 4, 128); break;                             \
	default: break;                                                \
	}

static int merge_items_for_k(int k)
{
	int items = 0;
#define WHAT(G, V, IPG) items = MergeLaunch<G, V, IPG>::items()
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
#define WHAT(G, V, IPG) e = MergeLaunch<G, V, IPG>::launch(h, B, C, s)
	MERGE_DISPATCH(k, WHAT)
#undef WHAT
	if (e != cudaSuccess) {
		snprintf(g_cuda_err, sizeof(g_cuda_err), "merge kernel launch (k=%d): %s", k, cudaGetErrorString(e));
		return SPMM_B200_ECUDA;
	}
	const int64_t n = (int64_t) h->merge.ntiles * k;
	spmm_merge_fixup_kernel<THREADS><<<(unsigned) ((n + THREADS - 1) / THREADS), THREADS, 0, s>>>(
		h->merge.d_carry, h->merge.d_carry_row, h->merge.ntiles, (int) h->m, k, C);
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

template <int K, int WARPS, int STAGES>
static void launch_bcsr_tma(const BcsrPlan& p, const double* B, double* C, int m, int brw, unsigned grid, cudaStream_t s)
{
	using Cfg = BcsrTmaCfg<K, WARPS, STAGES>;
	auto kern = spmm_bcsr_dmma_tma_kernel<K, WARPS, STAGES>;
	static bool attr_set = false;
	if (!attr_set) {
		cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) Cfg::SMEM);
		attr_set = true;
	}
	kern<<<grid, WARPS * 32, Cfg::SMEM, s>>>(p.d_brow, p.d_bcol, p.d_bval, B, C, m, p.mb, brw);
}

template <int K>
static void launch_bcsr_k(const spmm_b200* h, const double* B, double* C, cudaStream_t s)
{
	constexpr int WARPS = 4;
	const BcsrPlan& p = h->bcsr;
	const unsigned grid = (unsigned) ((p.mb + WARPS - 1) / WARPS);
	static const int use_tma = env_int("SPMM_B200_BCSR_TMA", 1);
	// bulk copies of 8K bytes per B row pay from K = 64 on (cfg4: K=32 0.57 ms fragment loads vs 0.63 ms TMA;
	// K=128 2.55 ms vs 1.23 ms); 2 stages = most warps per SM, which is what the tensor pipe needs
	if (h->variant == SPMM_B200_BCSR && (use_tma == 2 || (use_tma == 1 && K >= 64)) && ((uintptr_t) B % 16) == 0) {
		static const int stages = env_int("SPMM_B200_BCSR_STAGES", 2);
		static const int brw_env = env_int("SPMM_B200_BCSR_BRW", 0);
		// block rows per warp: enough to keep the ring flowing, few enough for >= 8 CTAs per SM
		const int brw = brw_env > 0 ? brw_env : (int) std::max<int64_t>(1, std::min<int64_t>(8, p.mb / ((int64_t) h->sm_count * WARPS * 8)));
		const unsigned g2 = (unsigned) ((p.mb + (int64_t) WARPS * brw - 1) / ((int64_t) WARPS * brw));
		if (stages <= 2) launch_bcsr_tma<K, WARPS, 2>(p, B, C, (int) h->m, brw, g2, s);
		else if (stages == 3) launch_bcsr_tma<K, WARPS, 3>(p, B, C, (int) h->m, brw, g2, s);
		else launch_bcsr_tma<K, WARPS, 4>(p, B, C, (int) h->m, brw, g2, s);
	} else if (h->variant == SPMM_B200_BCSR)
		spmm_bcsr_dmma_kernel<K, WARPS><<<grid, WARPS * 32, 0, s>>>(p.d_brow, p.d_bcol, p.d_bval, B, C, (int) h->m, p.mb);
	else if constexpr (K % 32 == 0)
		spmm_bcsr_simt_kernel<K, WARPS><<<grid, WARPS * 32, 0, s>>>(p.d_brow, p.d_bcol, p.d_bval, p.d_bmask, B, C, (int) h->m, p.mb);
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

// Launches the SpMM kernel(s) of the handle's variant on stream s.  Returns kernels launched or <0.
static int launch_spmm(spmm_b200* h, const double* B, double* C, int k, cudaStream_t s)
{
	if (h->m == 0) return 0;
	// merge-path has no instantiation for odd k: those run on the row-split generic kernel
	if (h->variant == SPMM_B200_MERGE && merge_supports_k(k)) return launch_merge(h, B, C, k, s);
	// blocked variants need k % 16 (tensor pipe) / k % 32 (FMA pipe); other k run on the CSR kernels
	if ((h->variant == SPMM_B200_BCSR || h->variant == SPMM_B200_BCSR_FMA) && bcsr_supports_k(h->variant, k))
		return launch_bcsr(h, B, C, k, s);
	int n = 1;
	// L2-aware k slicing: run the k columns as k/w launches on w-column slices of the same row-major B and C,
	// so that the B rows a launch gathers (touched window x 8w bytes) stay in L2.  It pays when the window is
	// far larger than L2 AND rows are long (every slice re-reads A and pays the per-row overhead again):
	// profiles/r1_kslice_probe.md, r1_sweep_medium.md.  SPMM_B200_KSLICE=w forces a width, =0 (default) lets the
	// handle time the candidates on its first calls (spmm_b200_run_device), =-1 turns slicing off.
	int w = env_int("SPMM_B200_KSLICE", 0);
	if (w == 0 && h->ks_k == k) w = h->ks_best > 0 ? h->ks_best : h->ks_cand[h->ks_try / 2];
	if (w > 0 && w < k && k % w == 0 && (w == 8 || w == 16 || w == 32 || w == 64)) {
		for (int off = 0; off < k; off += w) {
			switch (w) {
			case 8:  launch_rowsplit_gv<4, 2>(h, B + off, C + off, s, k, k); break;
			case 16: launch_rowsplit_gv<8, 2>(h, B + off, C + off, s, k, k); break;
			case 32: launch_rowsplit_gv<8, 4>(h, B + off, C + off, s, k, k); break;
			default: launch_rowsplit_gv<16, 4>(h, B + off, C + off, s, k, k); break;
			}
		}
		cudaError_t e = cudaGetLastError();
		if (e != cudaSuccess) {
			snprintf(g_cuda_err, sizeof(g_cuda_err), "sliced spmm launch (k=%d, w=%d): %s", k, w, cudaGetErrorString(e));
			return SPMM_B200_ECUDA;
		}
		h->launches += k / w;
		return k / w;
	}
	const int vec = env_int("SPMM_B200_VEC", 0);
	if (vec && h->d_iap && (k == 16 || k == 32 || k == 64 || k == 128) && ((uintptr_t) B % 16) == 0) {
		switch (k) {
		case 16:  if (vec == 2) launch_rowsplit_vec<8, 2>(h, B, C, s); else launch_rowsplit_vec<4, 4>(h, B, C, s); break;
		case 32:  if (vec == 2) launch_rowsplit_vec<16, 2>(h, B, C, s); else launch_rowsplit_vec<8, 4>(h, B, C, s); break;
		case 64:  if (vec == 2) launch_rowsplit_vec<32, 2>(h, B, C, s); else launch_rowsplit_vec<16, 4>(h, B, C, s); break;
		default:  launch_rowsplit_vec<32, 4>(h, B, C, s); break;
		}
		cudaError_t e = cudaGetLastError();
		if (e != cudaSuccess) {
			snprintf(g_cuda_err, sizeof(g_cuda_err), "vec spmm launch (k=%d): %s", k, cudaGetErrorString(e));
			return SPMM_B200_ECUDA;
		}
		h->launches += 1;
		return 1;
	}
	// experiment knob: lane-group shape "GxVxU" (G lanes per row, V columns per lane, U gathers in flight)
	const int gvu = env_int("SPMM_B200_GVU", 0);       // e.g. 842 = G 8, V 4, U 2;  1682 = G 16, V 8, U 2
	if (gvu > 0) {
		const int G = gvu / 100, V = (gvu / 10) % 10, U = gvu % 10;
		bool ok = true;
#define TRY_GVU(g, v, u) else if (G == g && V == v && U == u && k == g * v) \
		spmm_rowsplit_kernel<g, v, u, THREADS><<<h->n_ctas, THREADS, 0, s>>>(h->d_ia, h->d_ja, h->d_a, B, C, h->d_cta_rows, k, k)
		if (false) {}
		TRY_GVU(8, 4, 2); TRY_GVU(8, 4, 4); TRY_GVU(8, 4, 8);
		TRY_GVU(8, 8, 2); TRY_GVU(8, 8, 4); TRY_GVU(16, 4, 2); TRY_GVU(16, 4, 4);
		TRY_GVU(16, 8, 2); TRY_GVU(16, 8, 4); TRY_GVU(32, 4, 2); TRY_GVU(32, 4, 4);
		TRY_GVU(8, 2, 4); TRY_GVU(8, 2, 8); TRY_GVU(4, 4, 4); TRY_GVU(2, 2, 2); TRY_GVU(4, 1, 4); TRY_GVU(2, 1, 2); TRY_GVU(4, 2, 4); TRY_GVU(2, 4, 2);
		else ok = false;
#undef TRY_GVU
		if (ok) {
			h->launches += 1;
			return cudaGetLastError() == cudaSuccess ? 1 : SPMM_B200_ECUDA;
		}
	}
	switch (k) {
	case 1:   launch_smallk<1>(h, B, C, s); break;
	case 2:   launch_smallk<2>(h, B, C, s); break;
	case 4:   launch_rowsplit_gv<2, 2>(h, B, C, s); break;      // 16 rows per warp: cfg2 0.496 -> 0.380 ms, cfg5-large 0.440 -> 0.284 ms vs the small-k kernel
	// lane-group shapes from the sweep in profiles/r1_gvu_probe.md: V = 4 columns per lane (two 16-byte loads)
	// makes a group of 8 lanes cover one 128-byte line per load and halves the shuffle traffic per nonzero
	// (cfg2: k=32 0.935 -> 0.786 ms, k=64 1.563 -> 1.336 ms); V = 8 costs too many registers
	case 8:   launch_rowsplit_gv<4, 2>(h, B, C, s); break;
	case 16:  launch_rowsplit_gv<8, 2>(h, B, C, s); break;
	case 32:  launch_rowsplit_gv<8, 4>(h, B, C, s); break;
	case 64:  launch_rowsplit_gv<16, 4>(h, B, C, s); break;
	case 128: launch_rowsplit_gv<32, 4>(h, B, C, s); break;
	default:
		spmm_rowsplit_generic_kernel<THREADS><<<h->n_ctas, THREADS, 0, s>>>(h->d_ia, h->d_ja, h->d_a, B, C, h->d_cta_rows, k);
		break;
	}
	cudaError_t e = cudaGetLastError();
	if (e != cudaSuccess) {
		snprintf(g_cuda_err, sizeof(g_cuda_err), "spmm kernel launch (k=%d): %s", k, cudaGetErrorString(e));
		return SPMM_B200_ECUDA;
	}
	h->launches += n;
	return n;
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
	h->m = m_local; h->c = c; h->nnz = nnz_local; h->first_row = first_row; h->m_global = m_global;
	*out = h;   // from here on the caller destroys on error

	CU_TRY(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
	CU_TRY(cudaEventCreate(&h->ev0));
	CU_TRY(cudaEventCreate(&h->ev1));
	CU_TRY(cudaMalloc(&h->d_ia, sizeof(int32_t) * (m_local + 1)));
	CU_TRY(cudaMalloc(&h->d_ja, sizeof(int32_t) * std::max<int64_t>(nnz_local, 1)));
	CU_TRY(cudaMalloc(&h->d_a, sizeof(double) * std::max<int64_t>(nnz_local, 1)));
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
		col_span_kernel<<<(unsigned) std::min<int64_t>(148 * 8, (nnz_local + 255) / 256), 256, 0, h->stream>>>(h->d_ja, nnz_local, d_mm, d_mm + 1);
		h->launches++;
		CU_TRY(cudaMemcpyAsync(mm, d_mm, sizeof(mm), cudaMemcpyDeviceToHost, h->stream));
		CU_TRY(cudaStreamSynchronize(h->stream));
		cudaFree(d_mm);
		h->col_span = mm[1] >= mm[0] ? (int64_t) mm[1] - mm[0] + 1 : 0;
	}
	if (variant == SPMM_B200_AUTO) {
		// row-split serialises a row on one lane group: switch to merge-path when one row is far
		// longer than the mean (SURVEY 8(d) cfg 3: 200 040 nnz against a mean of 20)
		const double avg = m_local > 0 ? (double) nnz_local / (double) m_local : 0.0;
		// measured (gpurun_out/r2_*): cfg3b (max 866, mean 20) merge 1.24 ms vs row-split 2.03 ms;
		// cfg2 (max 131, mean 50) row-split 0.91 ms vs merge 0.96 ms
		const int thr = env_int("SPMM_B200_MERGE_MAXROW", 256);
		h->variant = (h->max_row >= thr && h->max_row > 8.0 * avg) ? SPMM_B200_MERGE : SPMM_B200_ROWSPLIT;
	} else {
		h->variant = variant;
	}
	int rc = build_rowsplit_plan(h);        // always: the fallback for k without a specialised instantiation
	if (rc) return rc;
	if (env_int("SPMM_B200_VEC", 0) && k >= 16) {
		rc = build_padded_plan(h);
		if (rc && rc != SPMM_B200_ENOSUP) return rc;
	}
	if (variant == SPMM_B200_AUTO && h->variant == SPMM_B200_ROWSPLIT && k >= 32 && bcsr_supports_k(SPMM_B200_BCSR, k)
	    && m_local >= 8 && nnz_local > 0) {
		// "tensor cores only where dense blocks make it a genuine contraction": probe the 8x1 block
		// fill with pass 1 of the conversion.  Break-even measured on B200: fill 0.15
		// at k=128, 0.25 at k=32; below 0.30 the CSR kernels stay.
		rc = bcsr_count(h);
		if (rc) return rc;
		const double fill = h->bcsr.nnzb > 0 ? (double) nnz_local / (8.0 * (double) h->bcsr.nnzb) : 0.0;
		const double thr = env_int("SPMM_B200_BCSR_FILL_PCT", 30) / 100.0;
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
	cudaFree(h->d_iap); cudaFree(h->d_jap); cudaFree(h->d_ap);
	cudaFree(h->merge.d_tile); cudaFree(h->merge.d_carry); cudaFree(h->merge.d_carry_row);
	cudaFree(h->bcsr.d_brow); cudaFree(h->bcsr.d_bcol); cudaFree(h->bcsr.d_bval); cudaFree(h->bcsr.d_bmask);
	cudaFree(h->d_x); cudaFree(h->d_B); cudaFree(h->d_C); cudaFree(h->d_ticket);
	if (h->ev0) cudaEventDestroy(h->ev0);
	if (h->ev1) cudaEventDestroy(h->ev1);
	for (cudaEvent_t e : h->ev_slice) cudaEventDestroy(e);
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

static int upload_b(spmm_b200* h, const double* x_host, int k)
{
	const size_t nb = (size_t) h->c * k;
	int rc = ensure(&h->d_B, &h->cap_B, nb);
	if (rc) return rc;
	if (h->b_layout == SPMM_B200_B_ROWMAJOR || k == 1) {
		CU_TRY(cudaMemcpyAsync(h->d_B, x_host, nb * sizeof(double), cudaMemcpyHostToDevice, h->stream));
	} else {
		rc = ensure(&h->d_x, &h->cap_x, nb);
		if (rc) return rc;
		CU_TRY(cudaMemcpyAsync(h->d_x, x_host, nb * sizeof(double), cudaMemcpyHostToDevice, h->stream));
		rc = spmm_b200_transpose_b_device(h, h->d_x, h->d_B, k, h->stream);
		if (rc) return rc;
	}
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
	// candidate computes the same bits), then the fastest is kept.
	const bool tunable = h->variant == SPMM_B200_ROWSPLIT && k >= 64 && k % 16 == 0 && env_int("SPMM_B200_KSLICE", 0) == 0
	                     && (double) h->col_span * 8.0 * k > 192e6 && h->nnz > 0;
	if (tunable && h->ks_k != k) {
		h->ks_k = k; h->ks_best = 0; h->ks_try = 0;
		int n = 0;
		h->ks_cand[n++] = k;
		for (int w = 64; w >= 16 && n < 4; w /= 2)
			if (w < k && k % w == 0) h->ks_cand[n++] = w;
		for (; n < 4; n++) h->ks_cand[n] = 0;
	} else if (!tunable && h->ks_k == k) {
		h->ks_k = 0;
	}
	int rc = run_between(h, b_dev, c_dev, k, (cudaStream_t) stream, h->ev0, h->ev1);
	if (rc) return rc;
	h->timed = true;
	h->timed_slices = 0;
	if (tunable && h->ks_best == 0) {
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
	int w = env_int("SPMM_B200_PIPELINE_COLS", -1);                  // 0 = off, -1 = choose
	if (w < 0) {
		// at most 4 slices: every slice re-reads A, and wider slices run the better kernels
		// (cfg4, k=128: 27.8 ms with 32-column slices, 29.3 ms with 16, 38.8 ms unsliced)
		w = 16;
		while (k / w > 4 && k % (2 * w) == 0) w *= 2;
	}
	if (w <= 0 || k < 2 * w || k % w) return 0;
	// small operands are latency-bound, not bandwidth-bound: one upload, one kernel, one download
	const size_t min_bytes = (size_t) env_int("SPMM_B200_PIPELINE_MIN_KB", 8 << 10) << 10;
	if ((size_t) h->c * k * sizeof(double) < min_bytes) return 0;
	return w;
}

static int run_pipelined(spmm_b200* h, const double* x_host, double* y_host, int k, int ks,
                         spmm_b200_bcast_fn bcast = nullptr, void* bcast_ctx = nullptr, int parts = 1)
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
	h->resident_k = 0;                       // d_B is about to hold slices, not the c x k operand
	const size_t bslice = (size_t) h->c * ks, cslice = (size_t) h->m * ks;
	for (int s = 0; s < S; s++) {
		double* Bs = h->d_B + s * bslice;
		if (h->b_layout == SPMM_B200_B_COLMAJOR) {
			// columns [s*ks, (s+1)*ks) are contiguous in x: one H2D, then transpose to row-major c x ks
			if (x_host && parts > 1) {
				// every rank holds x: upload only part `rank` of the slice over this GPU's own PCIe link;
				// the replication callback then fills in the other ranks' parts over NVLink
				const size_t per = (bslice + parts - 1) / parts;
				const size_t lo = std::min(bslice, per * (size_t) h->rank), hi = std::min(bslice, lo + per);
				if (hi > lo)
					CU_TRY(cudaMemcpyAsync(h->d_x + s * bslice + lo, x_host + s * bslice + lo, (hi - lo) * sizeof(double), cudaMemcpyHostToDevice, h->s_up));
			} else if (x_host) {
				CU_TRY(cudaMemcpyAsync(h->d_x + s * bslice, x_host + s * bslice, bslice * sizeof(double), cudaMemcpyHostToDevice, h->s_up));
			}
			if (bcast && bslice > 0 && bcast(bcast_ctx, h->d_x + s * bslice, bslice, parts, h->s_up)) return SPMM_B200_ECOMM;
			rc = spmm_b200_transpose_b_device(h, h->d_x + s * bslice, Bs, ks, h->s_up);
			if (rc) return rc;
		} else {
			if (x_host)
				CU_TRY(cudaMemcpy2DAsync(Bs, sizeof(double) * ks, x_host + (size_t) s * ks, sizeof(double) * k, sizeof(double) * ks,
				                         (size_t) h->c, cudaMemcpyHostToDevice, h->s_up));
			if (bcast && bslice > 0 && bcast(bcast_ctx, Bs, bslice, 1, h->s_up)) return SPMM_B200_ECOMM;
		}
		CU_TRY(cudaEventRecord(h->ev_slice[3 * s], h->s_up));
	}
	// Direct store: when y_host is page-locked and mapped into the device's address space, the row-split kernel
	// writes its w-column slice of every C row (w*8 contiguous bytes, ldc = k) straight into host memory, so the
	// download of slice s is the kernel itself and overlaps the upload of slice s+1 with no copy-engine pass and no
	// staging in d_C (profiles/r1_pcie_probe.md: SM stores into mapped host memory run at the 2-D copy's rate).
	double* y_map = nullptr;
	if (y_host && h->m > 0 && h->variant == SPMM_B200_ROWSPLIT && (ks == 8 || ks == 16 || ks == 32 || ks == 64)
	    && env_int("SPMM_B200_DIRECT_STORE", 1)) {
		cudaPointerAttributes at;
		if (cudaPointerGetAttributes(&at, y_host) == cudaSuccess && at.type == cudaMemoryTypeHost && at.devicePointer)
			y_map = (double*) at.devicePointer;
		else
			cudaGetLastError();
	}
	for (int s = 0; s < S; s++) {
		CU_TRY(cudaStreamWaitEvent(h->stream, h->ev_slice[3 * s], 0));
		if (y_map) {
			const double* Bs = h->d_B + s * bslice;
			double* Cs = y_map + (size_t) s * ks;
			CU_TRY(cudaEventRecord(h->ev_slice[3 * s + 1], h->stream));
			switch (ks) {
			case 8:  launch_rowsplit_gv<4, 2>(h, Bs, Cs, h->stream, ks, k); break;
			case 16: launch_rowsplit_gv<8, 2>(h, Bs, Cs, h->stream, ks, k); break;
			case 32: launch_rowsplit_gv<8, 4>(h, Bs, Cs, h->stream, ks, k); break;
			default: launch_rowsplit_gv<16, 4>(h, Bs, Cs, h->stream, ks, k); break;
			}
			CU_TRY(cudaGetLastError());
			h->launches += 1;
			CU_TRY(cudaEventRecord(h->ev_slice[3 * s + 2], h->stream));
			continue;
		}
		rc = run_between(h, h->d_B + s * bslice, h->d_C + s * cslice, ks, h->stream, h->ev_slice[3 * s + 1], h->ev_slice[3 * s + 2]);
		if (rc) return rc;
		if (y_host && h->m > 0) {
			CU_TRY(cudaStreamWaitEvent(h->s_down, h->ev_slice[3 * s + 2], 0));
			CU_TRY(cudaMemcpy2DAsync(y_host + (size_t) s * ks, sizeof(double) * k, h->d_C + s * cslice, sizeof(double) * ks,
			                         sizeof(double) * ks, (size_t) h->m, cudaMemcpyDeviceToHost, h->s_down));
		}
	}
	h->c_slices = y_map ? 0 : S;            // 0: C went straight to the host, d_C holds nothing to fetch
	h->timed = true;
	h->timed_slices = S;
	CU_TRY(cudaStreamSynchronize(h->stream));
	CU_TRY(cudaStreamSynchronize(h->s_down));
	return SPMM_B200_OK;
}

extern "C" int spmm_b200_run(spmm_b200_t* h, const double* x_host, double* y_host, int k, int copy_back)
{
	if (!h || !x_host || k < 1 || (copy_back && !y_host)) return SPMM_B200_EINVAL;
	CU_TRY(cudaSetDevice(h->device));
	const int ks = slice_width(h, k);
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
	int ks = slice_width(h, k);
	if (ks <= 0) ks = k;
	// split uploads need the oracle layout (a slice of x is one contiguous range) and a communicator
	const int parts = (x_on_every_rank && bcast && h->nranks > 1 && h->b_layout == SPMM_B200_B_COLMAJOR) ? h->nranks : 1;
	if (x_on_every_rank && parts == 1 && h->rank != 0) x_host = nullptr;      // fall back: only the root uploads
	return run_pipelined(h, x_host, copy_back ? y_host : nullptr, k, ks, bcast, bcast_ctx, parts);
}

// ------------------------------------------------------------------------------------------
// partitioning (device search, host pointers in and out)
// ------------------------------------------------------------------------------------------

extern "C" int spmm_b200_partition(const int32_t* row_ptr, int64_t m, int64_t nnz, int P, int mode, int64_t* out, int device)
{
	if (!out || P < 1 || m < 0 || nnz < 0) return SPMM_B200_EINVAL;
	if (mode != SPMM_B200_PART_ROWS && !row_ptr) return SPMM_B200_EINVAL;
	if (mode < SPMM_B200_PART_ROWS || mode > SPMM_B200_PART_MERGE) return SPMM_B200_EINVAL;
	int rc = check_device(device, nullptr);
	if (rc) return rc;
	const size_t n_out = (mode == SPMM_B200_PART_NNZ) ? (size_t) P + 1 : (mode == SPMM_B200_PART_ROWS ? 2 * (size_t) P : 2 * ((size_t) P + 1));
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
	else partition_merge_kernel<<<nb, 256>>>(d_ia, m, nnz, P, d_out);
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
	if (mode < SPMM_B200_PART_ROWS || mode > SPMM_B200_PART_MERGE) return SPMM_B200_EINVAL;
	if (mode != SPMM_B200_PART_ROWS && !row_ptr) return SPMM_B200_EINVAL;
	if (mode == SPMM_B200_PART_NNZ) {
		for (int64_t t = 0; t <= P; t++) out[t] = nnz_split_start(row_ptr, m, nnz, P, t);
	} else if (mode == SPMM_B200_PART_ROWS) {
		for (int64_t t = 0; t < P; t++) equal_split(0, m, P, t, &out[2 * t], &out[2 * t + 1]);
	} else {
		const int64_t L = m + nnz, items = (L + P - 1) / P;
		for (int64_t t = 0; t <= P; t++) {
			int64_t d = items * t; if (d > L) d = L;
			merge_path_search(d, row_ptr + 1, m, nnz, &out[2 * t], &out[2 * t + 1]);
		}
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
	int rc = check_device(device, nullptr);
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
	const unsigned nb = (unsigned) std::min<int64_t>(148 * 16, (nnz + 255) / 256);
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
// fp32 entry points (the reference's "_f" builds)
// ------------------------------------------------------------------------------------------
#include "spmm_b200_f32.inl"

// ------------------------------------------------------------------------------------------
// sparse-attention pipeline entry points (include/spmm_b200_pipeline.h)
// ------------------------------------------------------------------------------------------
#include "spmm_b200_pipeline.inl"
