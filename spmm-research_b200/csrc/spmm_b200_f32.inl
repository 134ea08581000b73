// spmm_b200_f32.cu -- fp32 entry points of the C ABI (include/spmm_b200.h, "_f32" functions).
//
// The reference builds every kernel twice, "_d" (ValueType = double) and "_f" (ValueType = float,
// make.sh:98-108); this file is the "_f" side of the drop-in: same plug-in interface
// (spmv_kernel.h:9-30), int32 CSR indices, fp32 values / B / C.  Variants: row-split (spmm_f32.cuh), merge-path
// (the kernel of spmm_merge.cuh instantiated for float: skewed rows no longer serialise on one lane group), and the
// TF32 blocked kernel on tcgen05 / TMEM for banded matrices (spmm_tf32_blocked.cuh, explicit request only: it
// rounds the operands to TF32).  CTA work lists, merge tiles and shard boundaries come from the same bit-exact
// partitioners as the fp64 path (partition.cuh).  No CPU fallback.
// Included at the end of spmm_b200.cu (one translation unit: the kernels of partition.cuh must not be
// defined twice, and CU_TRY / g_cuda_err are shared).
#include "spmm_f32.cuh"
#include "spmm_tf32_blocked.cuh"

struct spmm_b200_f32 {
	int device = 0, b_layout = SPMM_B200_B_COLMAJOR, k = 0, sm_count = 0;
	int variant = SPMM_B200_ROWSPLIT, variant_req = SPMM_B200_AUTO;
	int max_row = 0;
	Options opt;
	MergePlan merge;
	Tf32Plan tf32;
	uint32_t smem_attr = 0;
	int64_t m = 0, c = 0, nnz = 0, first_row = 0;
	int64_t col_min = 0, col_span = 0;       // first touched column, last - first + 1 (window upload of spmm_b200_run_f32)
	int64_t last_h2d_bytes = 0;
	int32_t* d_ia = nullptr;
	int32_t* d_ja = nullptr;
	float* d_a = nullptr;
	int32_t* d_cta_rows = nullptr;
	int n_ctas = 0;
	int* d_ticket = nullptr;
	float* d_x = nullptr; size_t cap_x = 0;
	float* d_B = nullptr; size_t cap_B = 0;
	float* d_C = nullptr; size_t cap_C = 0;
	cudaStream_t stream = nullptr;
	cudaEvent_t ev0 = nullptr, ev1 = nullptr;
	bool timed = false;
	int64_t launches = 0;
};

__global__ void rebase_f32_kernel(int32_t* ia, int64_t n, int32_t base)
{
	int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) ia[i] -= base;
}

static int device_ok(int device, int* sms)
{
	int n = 0;
	if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0) { cudaGetLastError(); return SPMM_B200_ENODEV; }
	if (device < 0 || device >= n) return SPMM_B200_EINVAL;
	int major = 0;
	CU_TRY(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
	if (major != 10) return SPMM_B200_ENODEV;
	CU_TRY(cudaDeviceGetAttribute(sms, cudaDevAttrMultiProcessorCount, device));
	CU_TRY(cudaSetDevice(device));
	return SPMM_B200_OK;
}

// ---- TF32 blocked kernel (tcgen05 / TMEM): per-slab column windows + launch --------------------------------
static void tf32_release_plan(spmm_b200_f32* h)
{
	cudaFree(h->tf32.d_win); cudaFree(h->tf32.d_chunk0); cudaFree(h->tf32.d_eptr); cudaFree(h->tf32.d_ent);
	h->tf32 = Tf32Plan();
}

// Per-slab column windows, then the chunk-major copy of A (count -> scan -> place), all on the device.
static int tf32_build_plan(spmm_b200_f32* h)
{
	Tf32Plan& p = h->tf32;
	p.nslabs = (int) ((h->m + TF32_M - 1) / TF32_M);
	if (p.nslabs == 0) return SPMM_B200_OK;
	unsigned long long* d_total = nullptr;
	unsigned long long total = 0;
	int32_t* d_tmp = nullptr;
	int32_t* d_cursor = nullptr;
	int32_t nchunks = 0;
	CU_TRY(cudaMalloc(&p.d_win, sizeof(int2) * p.nslabs));
	CU_TRY(cudaMalloc(&p.d_chunk0, sizeof(int32_t) * ((size_t) p.nslabs + 1)));
	CU_TRY(cudaMalloc(&d_total, sizeof(unsigned long long)));
	CU_TRY(cudaMemsetAsync(d_total, 0, sizeof(unsigned long long), h->stream));
	tf32_window_kernel<<<p.nslabs, TF32_M, 0, h->stream>>>(h->d_ia, h->d_ja, (int) h->m, p.d_win, p.d_chunk0, d_total);
	h->launches++;
	CU_TRY(cudaGetLastError());
	CU_TRY(cudaMemcpyAsync(&total, d_total, sizeof(total), cudaMemcpyDeviceToHost, h->stream));
	CU_TRY(cudaStreamSynchronize(h->stream));
	cudaFree(d_total);
	p.window_cols = (double) total;
	// a matrix whose windows add up to more chunks than fit an int32 (or a few hundred MB of pointers) is not banded:
	// the handle then stays on the row-split kernel (tf32_supports() is false without a plan)
	if (p.window_cols / TF32_KT + p.nslabs > 64e6) { tf32_release_plan(h); return SPMM_B200_OK; }
	CU_TRY(cudaMalloc(&d_tmp, sizeof(int32_t) * scan_tmp_elems((int64_t) p.nslabs + 1)));
	h->launches += exclusive_scan_i32(p.d_chunk0, (int64_t) p.nslabs + 1, d_tmp, h->stream);
	CU_TRY(cudaMemcpyAsync(&nchunks, p.d_chunk0 + p.nslabs, sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
	CU_TRY(cudaStreamSynchronize(h->stream));
	cudaFree(d_tmp); d_tmp = nullptr;
	p.total_chunks = nchunks;
	CU_TRY(cudaMalloc(&p.d_eptr, sizeof(int32_t) * ((size_t) nchunks + 1)));
	CU_TRY(cudaMalloc(&d_cursor, sizeof(int32_t) * ((size_t) nchunks + 1)));
	CU_TRY(cudaMalloc(&p.d_ent, sizeof(uint2) * std::max<int64_t>(h->nnz, 1)));
	CU_TRY(cudaMalloc(&d_tmp, sizeof(int32_t) * scan_tmp_elems((int64_t) nchunks + 1)));
	CU_TRY(cudaMemsetAsync(p.d_eptr, 0, sizeof(int32_t) * ((size_t) nchunks + 1), h->stream));
	tf32_bucket_kernel<false><<<p.nslabs, TF32_M, 0, h->stream>>>(h->d_ia, h->d_ja, h->d_a, (int) h->m, p.d_win, p.d_chunk0, p.d_eptr, nullptr);
	h->launches += 1 + exclusive_scan_i32(p.d_eptr, (int64_t) nchunks + 1, d_tmp, h->stream);
	CU_TRY(cudaMemcpyAsync(d_cursor, p.d_eptr, sizeof(int32_t) * ((size_t) nchunks + 1), cudaMemcpyDeviceToDevice, h->stream));
	tf32_bucket_kernel<true><<<p.nslabs, TF32_M, 0, h->stream>>>(h->d_ia, h->d_ja, h->d_a, (int) h->m, p.d_win, p.d_chunk0, d_cursor, p.d_ent);
	h->launches++;
	CU_TRY(cudaGetLastError());
	CU_TRY(cudaStreamSynchronize(h->stream));
	cudaFree(d_tmp); cudaFree(d_cursor);
	return SPMM_B200_OK;
}

static bool tf32_supports(const spmm_b200_f32* h, int k) { return h->tf32.d_ent && (k == 32 || k == 64 || k == 128); }

template <int N, int STAGES>
static cudaError_t tf32_launch_ns(spmm_b200_f32* h, const float* B, float* C, cudaStream_t s, int bit)
{
	using Cfg = Tf32Cfg<N, STAGES>;
	auto kern = spmm_tf32_blocked_kernel<N, STAGES>;
	if (Cfg::SMEM > 48 * 1024 && !(h->smem_attr >> bit & 1u)) {
		cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) Cfg::SMEM);
		if (e != cudaSuccess) return e;
		h->smem_attr |= 1u << bit;
	}
	kern<<<h->tf32.nslabs, TF32_THREADS, Cfg::SMEM, s>>>(h->tf32.d_ent, h->tf32.d_eptr, h->tf32.d_chunk0, B, C, h->tf32.d_win, (int) h->m, (int) h->c);
	return cudaGetLastError();
}

template <int N>
static cudaError_t tf32_launch_n(spmm_b200_f32* h, const float* B, float* C, cudaStream_t s, int bit)
{
	// 2 stages = 3 resident slabs per SM at k = 128 (the default); 3 stages = deeper hand-over, 2 slabs per SM
	return h->opt.tf32_stages >= 3 ? tf32_launch_ns<N, 3>(h, B, C, s, bit + 3) : tf32_launch_ns<N, 2>(h, B, C, s, bit);
}

static int tf32_launch(spmm_b200_f32* h, const float* B, float* C, int k, cudaStream_t s)
{
	cudaError_t e = k == 32 ? tf32_launch_n<32>(h, B, C, s, 16) : k == 64 ? tf32_launch_n<64>(h, B, C, s, 17) : tf32_launch_n<128>(h, B, C, s, 18);
	if (e != cudaSuccess) {
		snprintf(g_cuda_err, sizeof(g_cuda_err), "tf32 blocked launch (k=%d): %s", k, cudaGetErrorString(e));
		return SPMM_B200_ECUDA;
	}
	h->launches += 1;
	return 1;
}

// k -> (G, V, IPG) of the float merge kernel (V = 4 floats = one 16-byte piece per lane and load)
#define MERGE_F32_DISPATCH(k, WHAT)                                \
	switch (k) {                                                   \
	case 1:   WHAT(1, 1, 16, 0); break;                            \
	case 2:   WHAT(1, 2, 16, 1); break;                            \
	case 4:   WHAT(1, 4, 16, 2); break;                            \
	case 8:   WHAT(2, 4, 16, 3); break;                            \
	case 16:  WHAT(4, 4, 32, 4); break;                            \
	case 32:  WHAT(8, 4, 64, 5); break;                            \
	case 64:  WHAT(16, 4, 128, 6); break;                          \
	case 128: WHAT(32, 4, 128, 7); break;                          \
	default: break;                                                \
	}

static int merge_items_f32(int k)
{
	int items = 0;
#define WHAT(G, V, IPG, BIT) items = MergeCfg<G, V, IPG, THREADS, float>::ITEMS
	MERGE_F32_DISPATCH(k, WHAT)
#undef WHAT
	return items;
}

static int build_merge_plan_f32(spmm_b200_f32* h, int k, cudaStream_t s)
{
	const int items = merge_items_f32(k);
	if (items <= 0) return SPMM_B200_ENOSUP;
	MergePlan& p = h->merge;
	if (p.items == items && p.carry_k >= k && p.d_tile) return SPMM_B200_OK;
	const int64_t L = h->m + h->nnz;
	const int ntiles = (int) ((L + items - 1) / items);
	cudaFree(p.d_tile); cudaFree(p.d_carry); cudaFree(p.d_carry_row);
	p.d_tile = nullptr; p.d_carry = nullptr; p.d_carry_row = nullptr;
	CU_TRY(cudaMalloc(&p.d_tile, sizeof(int2) * ((size_t) ntiles + 1)));
	CU_TRY(cudaMalloc(&p.d_carry, sizeof(float) * std::max<size_t>((size_t) ntiles * k, 1)));
	CU_TRY(cudaMalloc(&p.d_carry_row, sizeof(int32_t) * std::max(ntiles, 1)));
	p.ntiles = ntiles; p.items = items; p.carry_k = k;
	tile_merge_kernel<<<(ntiles + 1 + 255) / 256, 256, 0, s>>>(h->d_ia, (int) h->m, h->nnz, items, ntiles, p.d_tile);
	h->launches++;
	CU_TRY(cudaGetLastError());
	return SPMM_B200_OK;
}

template <int G, int V, int IPG>
static cudaError_t launch_merge_f32_gv(spmm_b200_f32* h, const float* B, float* C, cudaStream_t s, int bit)
{
	using Cfg = MergeCfg<G, V, IPG, THREADS, float>;
	auto kern = spmm_merge_kernel<G, V, IPG, 4, THREADS, float>;
	if (Cfg::SMEM > 48 * 1024 && !(h->smem_attr >> bit & 1u)) {
		cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) Cfg::SMEM);
		if (e != cudaSuccess) return e;
		h->smem_attr |= 1u << bit;
	}
	kern<<<h->merge.ntiles, THREADS, Cfg::SMEM, s>>>(h->d_ia, h->d_ja, h->d_a, B, C, h->merge.d_tile,
	                                                (float*) h->merge.d_carry, h->merge.d_carry_row);
	return cudaGetLastError();
}

static int launch_merge_f32(spmm_b200_f32* h, const float* B, float* C, int k, cudaStream_t s)
{
	int rc = build_merge_plan_f32(h, k, s);
	if (rc) return rc;
	if (h->merge.ntiles == 0) return 0;
	cudaError_t e = cudaErrorInvalidValue;
#define WHAT(G, V, IPG, BIT) e = launch_merge_f32_gv<G, V, IPG>(h, B, C, s, BIT)
	MERGE_F32_DISPATCH(k, WHAT)
#undef WHAT
	if (e == cudaSuccess) {
		const int64_t n = (int64_t) h->merge.ntiles * k;
		spmm_merge_fixup_kernel<THREADS, float><<<(unsigned) ((n + THREADS - 1) / THREADS), THREADS, 0, s>>>(
			(const float*) h->merge.d_carry, h->merge.d_carry_row, h->merge.ntiles, (int) h->m, k, C);
		e = cudaGetLastError();
	}
	if (e != cudaSuccess) {
		snprintf(g_cuda_err, sizeof(g_cuda_err), "fp32 merge launch (k=%d): %s", k, cudaGetErrorString(e));
		return SPMM_B200_ECUDA;
	}
	h->launches += 2;
	return 2;
}

static int create_slice(spmm_b200_f32** out, const int32_t* ia_slice, const int32_t* ja, const float* a, int64_t m, int64_t c,
                        int64_t nnz, int32_t base, int k, int device, int variant, int b_layout, int sms, int64_t first_row)
{
	spmm_b200_f32* h = new (std::nothrow) spmm_b200_f32();
	if (!h) return SPMM_B200_ENOMEM;
	*out = h;
	h->device = device; h->b_layout = b_layout; h->k = k; h->sm_count = sms;
	h->variant_req = variant;
	h->opt = options_from_env();
	h->m = m; h->c = c; h->nnz = nnz; h->first_row = first_row;
	CU_TRY(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
	CU_TRY(cudaEventCreate(&h->ev0));
	CU_TRY(cudaEventCreate(&h->ev1));
	CU_TRY(cudaMalloc(&h->d_ia, sizeof(int32_t) * (m + 1)));
	CU_TRY(cudaMalloc(&h->d_ja, sizeof(int32_t) * std::max<int64_t>(nnz, 1)));
	CU_TRY(cudaMalloc(&h->d_a, sizeof(float) * std::max<int64_t>(nnz, 1)));
	CU_TRY(cudaMemcpyAsync(h->d_ia, ia_slice, sizeof(int32_t) * (m + 1), cudaMemcpyHostToDevice, h->stream));
	if (nnz > 0) {
		CU_TRY(cudaMemcpyAsync(h->d_ja, ja + base, sizeof(int32_t) * nnz, cudaMemcpyHostToDevice, h->stream));
		CU_TRY(cudaMemcpyAsync(h->d_a, a + base, sizeof(float) * nnz, cudaMemcpyHostToDevice, h->stream));
	}
	if (base != 0) {
		rebase_f32_kernel<<<(unsigned) ((m + 1 + 255) / 256), 256, 0, h->stream>>>(h->d_ia, m + 1, base);
		h->launches++;
	}
	// CTA work list: the reference's nnz-balanced row split with P = 32 CTAs per SM
	const int64_t want = (int64_t) sms * 32, cap = std::max<int64_t>(1, (m + 7) / 8);
	h->n_ctas = (int) std::max<int64_t>(1, std::min(want, cap));
	CU_TRY(cudaMalloc(&h->d_cta_rows, sizeof(int32_t) * (h->n_ctas + 1)));
	CU_TRY(cudaMalloc(&h->d_ticket, sizeof(int)));
	cta_rows_nnz_kernel<<<(h->n_ctas + 1 + 255) / 256, 256, 0, h->stream>>>(h->d_ia, (int) m, nnz, h->n_ctas, h->d_cta_rows);
	h->launches++;
	CU_TRY(cudaGetLastError());
	if (m > 0) {
		int* d_max = nullptr;
		CU_TRY(cudaMalloc(&d_max, sizeof(int)));
		CU_TRY(cudaMemsetAsync(d_max, 0, sizeof(int), h->stream));
		max_row_kernel<<<(unsigned) std::min<int64_t>(1024, (m + 255) / 256), 256, 0, h->stream>>>(h->d_ia, (int) m, d_max);
		h->launches++;
		CU_TRY(cudaMemcpyAsync(&h->max_row, d_max, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
		CU_TRY(cudaStreamSynchronize(h->stream));
		cudaFree(d_max);
	}
	if (nnz > 0) {
		// touched window of B, as in the fp64 handle (create_from_slice)
		int* d_mm = nullptr;
		int mm[2] = {INT32_MAX, -1};
		CU_TRY(cudaMalloc(&d_mm, 2 * sizeof(int)));
		CU_TRY(cudaMemcpyAsync(d_mm, mm, sizeof(mm), cudaMemcpyHostToDevice, h->stream));
		col_span_kernel<<<(unsigned) std::min<int64_t>((int64_t) h->sm_count * 8, (nnz + 255) / 256), 256, 0, h->stream>>>(h->d_ja, nnz, d_mm, d_mm + 1);
		h->launches++;
		CU_TRY(cudaMemcpyAsync(mm, d_mm, sizeof(mm), cudaMemcpyDeviceToHost, h->stream));
		CU_TRY(cudaStreamSynchronize(h->stream));
		cudaFree(d_mm);
		h->col_span = mm[1] >= mm[0] ? (int64_t) mm[1] - mm[0] + 1 : 0;
		h->col_min = mm[1] >= mm[0] ? mm[0] : 0;
	}
	if (variant == SPMM_B200_AUTO) {
		// the fp64 rule (create_from_slice): merge-path when one row is far longer than the mean.  AUTO never picks the
		// TF32 blocked kernel: it rounds A and B to 10 mantissa bits, which the caller has to ask for.
		const double avg = m > 0 ? (double) nnz / (double) m : 0.0;
		h->variant = (h->max_row >= h->opt.merge_maxrow && h->max_row > 8.0 * avg) ? SPMM_B200_MERGE : SPMM_B200_ROWSPLIT;
	} else {
		h->variant = variant == SPMM_B200_BCSR_FMA ? SPMM_B200_ROWSPLIT : variant;
	}
	if (h->variant == SPMM_B200_MERGE && merge_items_f32(k) > 0) {
		int rc = build_merge_plan_f32(h, k, h->stream);
		if (rc) return rc;
	}
	if (h->variant == SPMM_B200_BCSR) {
		int rc = tf32_build_plan(h);
		if (rc) return rc;
		if (!h->tf32.d_ent && h->m > 0) h->variant = SPMM_B200_ROWSPLIT;      // not a banded matrix (see tf32_build_plan)
	}
	CU_TRY(cudaStreamSynchronize(h->stream));
	return SPMM_B200_OK;
}

static int check_args(const void* rp, const void* ci, const void* va, int64_t m, int64_t c, int64_t nnz, int k, int b_layout)
{
	if (!rp || m < 0 || c < 0 || nnz < 0 || k < 1) return SPMM_B200_EINVAL;
	if (nnz > 0 && (!ci || !va)) return SPMM_B200_EINVAL;
	if (b_layout != SPMM_B200_B_COLMAJOR && b_layout != SPMM_B200_B_ROWMAJOR) return SPMM_B200_EINVAL;
	if (m >= INT32_MAX || c >= INT32_MAX || nnz >= INT32_MAX) return SPMM_B200_ERANGE;
	return SPMM_B200_OK;
}

extern "C" int spmm_b200_create_f32(spmm_b200_f32_t** h, const int32_t* row_ptr, const int32_t* col_ind, const float* values,
                                    int64_t m, int64_t c, int64_t nnz, int k, int device, int variant, int b_layout)
{
	if (!h) return SPMM_B200_EINVAL;
	*h = nullptr;
	int rc = check_args(row_ptr, col_ind, values, m, c, nnz, k, b_layout);
	if (rc) return rc;
	if (variant < SPMM_B200_ROWSPLIT || variant > SPMM_B200_BCSR_FMA) return SPMM_B200_EINVAL;
	int sms = 0;
	rc = device_ok(device, &sms);
	if (rc) return rc;
	rc = create_slice(h, row_ptr, col_ind, values, m, c, nnz, 0, k, device, variant, b_layout, sms, 0);
	if (rc) { spmm_b200_destroy_f32(*h); *h = nullptr; }
	return rc;
}

extern "C" int spmm_b200_create_shard_f32(spmm_b200_f32_t** h, const int32_t* row_ptr, const int32_t* col_ind, const float* values,
                                          int64_t m, int64_t c, int64_t nnz, int k, int device, int variant, int b_layout,
                                          int rank, int nranks, int64_t* first_row, int64_t* last_row)
{
	if (!h) return SPMM_B200_EINVAL;
	*h = nullptr;
	int rc = check_args(row_ptr, col_ind, values, m, c, nnz, k, b_layout);
	if (rc) return rc;
	if (variant < SPMM_B200_ROWSPLIT || variant > SPMM_B200_BCSR_FMA) return SPMM_B200_EINVAL;
	if (nranks < 1 || rank < 0 || rank >= nranks) return SPMM_B200_EINVAL;
	int sms = 0;
	rc = device_ok(device, &sms);
	if (rc) return rc;
	std::vector<int64_t> bounds(nranks + 1);
	rc = spmm_b200_partition_host(row_ptr, m, nnz, nranks, SPMM_B200_PART_NNZ, bounds.data());   // same integers as the device search
	if (rc) return rc;
	const int64_t r0 = bounds[rank], r1 = std::max(bounds[rank + 1], bounds[rank]);
	if (first_row) *first_row = r0;
	if (last_row) *last_row = r1;
	const int32_t base = row_ptr[r0];
	rc = create_slice(h, row_ptr + r0, col_ind, values, r1 - r0, c, (int64_t) row_ptr[r1] - base, base, k, device, variant, b_layout, sms, r0);
	if (rc) { spmm_b200_destroy_f32(*h); *h = nullptr; }
	return rc;
}

extern "C" void spmm_b200_destroy_f32(spmm_b200_f32_t* h)
{
	if (!h) return;
	cudaSetDevice(h->device);
	if (h->stream) cudaStreamSynchronize(h->stream);
	cudaFree(h->d_ia); cudaFree(h->d_ja); cudaFree(h->d_a); cudaFree(h->d_cta_rows); cudaFree(h->d_ticket);
	cudaFree(h->d_x); cudaFree(h->d_B); cudaFree(h->d_C);
	cudaFree(h->merge.d_tile); cudaFree(h->merge.d_carry); cudaFree(h->merge.d_carry_row);
	tf32_release_plan(h);
	if (h->ev0) cudaEventDestroy(h->ev0);
	if (h->ev1) cudaEventDestroy(h->ev1);
	if (h->stream) cudaStreamDestroy(h->stream);
	delete h;
}

template <int G, int V, int U>
static void launch_gv(const spmm_b200_f32* h, const float* B, float* C, cudaStream_t s)
{
	// long rows and enough of them: persistent warps with 8-row tickets, as in the fp64 kernel (launch_rowsplit_gv)
	int tick = h->opt.ticket;
	if (tick < 0) {
		const double avg = h->m > 0 ? (double) h->nnz / (double) h->m : 1.0;
		// measured on cfg2 (profiles/r1_f32_probe.md): k = 64 0.784 -> 0.698 ms, k = 128 1.339 -> 1.267 ms, k = 32 0.466 -> 0.493 ms
		tick = (avg >= 32.0 && h->nnz >= (8 << 20) && G * V >= 64) ? 8 : 0;
	}
	if (tick > 0 && h->d_ticket) {
		constexpr int NGW = 32 / G;
		tick = (tick + NGW - 1) / NGW * NGW;
		cudaMemsetAsync(h->d_ticket, 0, sizeof(int), s);
		const int grid = (int) std::min<int64_t>((h->m + tick * (THREADS / 32) - 1) / (tick * (THREADS / 32)) + 1, (int64_t) h->sm_count * 4);
		spmm_rowsplit_f32_kernel<G, V, U, THREADS, true><<<grid, THREADS, 0, s>>>(h->d_ia, h->d_ja, h->d_a, B, C, h->d_cta_rows, G * V, G * V,
		                                                                         (int) h->m, tick, h->d_ticket);
		return;
	}
	spmm_rowsplit_f32_kernel<G, V, U, THREADS><<<h->n_ctas, THREADS, 0, s>>>(h->d_ia, h->d_ja, h->d_a, B, C, h->d_cta_rows, G * V, G * V);
}

extern "C" int spmm_b200_run_device_f32(spmm_b200_f32_t* h, const float* b_dev, float* c_dev, int k, void* stream)
{
	if (!h || k < 1) return SPMM_B200_EINVAL;
	if ((h->c > 0 && h->nnz > 0 && !b_dev) || (h->m > 0 && !c_dev)) return SPMM_B200_EINVAL;
	CU_TRY(cudaSetDevice(h->device));
	cudaStream_t s = (cudaStream_t) stream;
	CU_TRY(cudaEventRecord(h->ev0, s));
	if (h->m > 0 && h->variant == SPMM_B200_MERGE && merge_items_f32(k) > 0
	    && (k < 4 || (((uintptr_t) b_dev | (uintptr_t) c_dev) % 16) == 0)) {
		int n = launch_merge_f32(h, b_dev, c_dev, k, s);
		if (n < 0) return n;
	} else if (h->m > 0 && h->variant == SPMM_B200_BCSR && tf32_supports(h, k)
	           && (((uintptr_t) b_dev | (uintptr_t) c_dev) % 16) == 0) {
		int n = tf32_launch(h, b_dev, c_dev, k, s);
		if (n < 0) return n;
	} else if (h->m > 0) {
		const bool aligned = ((uintptr_t) b_dev % 16) == 0 && ((uintptr_t) c_dev % 16) == 0;
		// a group of 8 lanes covers one 128-byte line per 16-byte load (see profiles/r1_gvu_probe.md)
		if (aligned && k == 16) launch_gv<4, 4, 4>(h, b_dev, c_dev, s);
		else if (aligned && k == 32) launch_gv<8, 4, 4>(h, b_dev, c_dev, s);
		else if (aligned && k == 64) launch_gv<8, 8, 4>(h, b_dev, c_dev, s);
		else if (aligned && k == 128) launch_gv<16, 8, 4>(h, b_dev, c_dev, s);
		else if (aligned && k == 256) launch_gv<32, 8, 4>(h, b_dev, c_dev, s);
		else spmm_rowsplit_generic_f32_kernel<THREADS><<<h->n_ctas, THREADS, 0, s>>>(h->d_ia, h->d_ja, h->d_a, b_dev, c_dev, h->d_cta_rows, k);
		h->launches++;
		CU_TRY(cudaGetLastError());
	}
	CU_TRY(cudaEventRecord(h->ev1, s));
	h->timed = true;
	return SPMM_B200_OK;
}

static int ensure32(float** p, size_t* cap, size_t want)
{
	if (*cap >= want) return SPMM_B200_OK;
	if (*p) cudaFree(*p);
	*p = nullptr; *cap = 0;
	CU_TRY(cudaMalloc(p, sizeof(float) * std::max<size_t>(want, 1)));
	*cap = want;
	return SPMM_B200_OK;
}

extern "C" int spmm_b200_run_f32(spmm_b200_f32_t* h, const float* x_host, float* y_host, int k, int copy_back)
{
	if (!h || !x_host || k < 1 || (copy_back && !y_host)) return SPMM_B200_EINVAL;
	CU_TRY(cudaSetDevice(h->device));
	const size_t nb = (size_t) h->c * k;
	int rc = ensure32(&h->d_B, &h->cap_B, nb);
	if (rc) return rc;
	rc = ensure32(&h->d_C, &h->cap_C, (size_t) h->m * k);
	if (rc) return rc;
	// Only the rows of B between the first and the last touched column travel (a row no nonzero references cannot change
	// C; the fp64 twin is upload_b_rows).  The TF32 blocked kernel reads whole 32-row chunks of B, zeros
	// of A included, so it keeps the full upload.
	int64_t lo = 0, n = h->c;
	if (h->opt.upload_window && h->variant != SPMM_B200_BCSR) {
		if (h->nnz == 0 || h->col_span <= 0) n = 0;
		else { lo = h->col_min; n = std::min<int64_t>(h->col_span, h->c - lo); }
	}
	h->last_h2d_bytes = 0;
	if (n > 0) {
		if (h->b_layout == SPMM_B200_B_ROWMAJOR || k == 1) {
			CU_TRY(cudaMemcpyAsync(h->d_B + lo * k, x_host + lo * k, sizeof(float) * n * k, cudaMemcpyHostToDevice, h->stream));
		} else {
			rc = ensure32(&h->d_x, &h->cap_x, nb);
			if (rc) return rc;
			if (n == h->c)
				CU_TRY(cudaMemcpyAsync(h->d_x, x_host, nb * sizeof(float), cudaMemcpyHostToDevice, h->stream));
			else
				CU_TRY(cudaMemcpy2DAsync(h->d_x + lo, sizeof(float) * h->c, x_host + lo, sizeof(float) * h->c, sizeof(float) * n, (size_t) k,
				                         cudaMemcpyHostToDevice, h->stream));
			dim3 grid((unsigned) ((n + 31) / 32), (unsigned) ((k + 31) / 32)), block(32, 8);
			transpose_b_rows_f32_kernel<<<grid, block, 0, h->stream>>>(h->d_x + lo, h->d_B + lo * k, h->c, n, k);
			CU_TRY(cudaGetLastError());
			h->launches++;
		}
		h->last_h2d_bytes = (int64_t) sizeof(float) * n * k;
	}
	rc = spmm_b200_run_device_f32(h, h->d_B, h->d_C, k, h->stream);
	if (rc) return rc;
	if (copy_back && h->m > 0)
		CU_TRY(cudaMemcpyAsync(y_host, h->d_C, sizeof(float) * h->m * k, cudaMemcpyDeviceToHost, h->stream));
	CU_TRY(cudaStreamSynchronize(h->stream));
	return SPMM_B200_OK;
}

extern "C" int spmm_b200_fetch_c_f32(spmm_b200_f32_t* h, float* y_host, int k)
{
	if (!h || !y_host || k < 1 || h->cap_C < (size_t) h->m * k) return SPMM_B200_EINVAL;
	CU_TRY(cudaSetDevice(h->device));
	if (h->m > 0) CU_TRY(cudaMemcpyAsync(y_host, h->d_C, sizeof(float) * h->m * k, cudaMemcpyDeviceToHost, h->stream));
	CU_TRY(cudaStreamSynchronize(h->stream));
	return SPMM_B200_OK;
}

extern "C" int64_t spmm_b200_rows_f32(const spmm_b200_f32_t* h) { return h ? h->m : -1; }
extern "C" int spmm_b200_variant_f32(const spmm_b200_f32_t* h) { return h ? h->variant : SPMM_B200_EINVAL; }
extern "C" int64_t spmm_b200_last_h2d_bytes_f32(const spmm_b200_f32_t* h) { return h ? h->last_h2d_bytes : -1; }

extern "C" int spmm_b200_set_option_f32(spmm_b200_f32_t* h, const char* name, int value)
{
	if (!h || !name) return SPMM_B200_EINVAL;
	if (!strncmp(name, "SPMM_B200_", 10)) name += 10;
	for (const OptionName& e : OPTION_TABLE)
		if (!strcmp(name, e.name)) { h->opt.*(e.field) = value; return SPMM_B200_OK; }
	return SPMM_B200_EINVAL;
}
extern "C" int64_t spmm_b200_nnz_f32(const spmm_b200_f32_t* h) { return h ? h->nnz : -1; }
extern "C" int64_t spmm_b200_launch_count_f32(const spmm_b200_f32_t* h) { return h ? h->launches : -1; }

extern "C" double spmm_b200_last_kernel_ms_f32(const spmm_b200_f32_t* h)
{
	if (!h || !h->timed) return -1.0;
	float ms = 0.f;
	if (cudaEventSynchronize(h->ev1) != cudaSuccess) return -1.0;
	if (cudaEventElapsedTime(&ms, h->ev0, h->ev1) != cudaSuccess) return -1.0;
	return (double) ms;
}
