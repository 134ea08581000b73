// spmm_b200_pipeline.inl -- host side of include/spmm_b200_pipeline.h, included by spmm_b200.cu.
// The three projections and the final product run on the SpMM handles above (row-major dense operands, the
// layout mkl_dcsrmm's 'C' indexing gives the reference, sddmm_taco_naive.cpp:219-260); SDDMM and softmax are
// the kernels of sddmm.cuh.  y is written straight into the value array of the mask's SpMM handle, so the
// final product needs no copy.  SDDMM and the final SpMM stay two kernels: fusing them would save one write
// and one read of y (16 bytes per nonzero) next to the 8n bytes per nonzero the V gather moves.
#include "sddmm.cuh"
#include "../../include/spmm_b200_pipeline.h"

struct spmm_b200_pipeline {
	int device = 0;
	int n = 0;
	int64_t m = 0, nnz = 0;
	const int32_t* mask_ia_host = nullptr;   // identity of the arrays the format was built from
	const int32_t* mask_ja_host = nullptr;
	spmm_b200* mask = nullptr;               // SpMM handle on the mask; its d_a holds y after sddmm()
	double* d_mask_a = nullptr;              // the mask's own values
	struct Weight {
		spmm_b200* h = nullptr;
		const int32_t* ia = nullptr; const int32_t* ja = nullptr; const double* a = nullptr;
		int64_t m = 0, k = 0, nnz = 0;
		uint64_t sig = 0;                    // pipe_weight_signature of the arrays that were uploaded
		double* d_out = nullptr;             // K, Q or V: m x n row-major
		bool have = false;
	} w[3];
	double* d_x = nullptr; size_t cap_x = 0;
	double* d_tmp = nullptr; size_t cap_tmp = 0;
	double* d_final = nullptr;
	double* d_part = nullptr;
	int* d_ticket = nullptr;                 // row ticket counter of the persistent SDDMM / fused kernels
	bool have_y = false, have_final = false;
	cudaStream_t stream = nullptr;
	cudaEvent_t ev0 = nullptr, ev1 = nullptr;
	// attention(): the three projections run side by side, V's overlaps the SDDMM (fork / join on events)
	cudaStream_t s_side[3] = {nullptr, nullptr, nullptr};
	cudaEvent_t ev_fork = nullptr, ev_done[3] = {nullptr, nullptr, nullptr};
	bool timed = false;
	int64_t own_launches = 0;
	// knobs, read from the environment once at creation (spmm_b200_pipeline_set_option changes them afterwards)
	struct Opt { int sddmm_group = 1, ticket_rev = 1, sddmm_tune = 0, sddmm_staged = 1, fork = 1, fused = 1; } opt;
};

struct PipeOptionName { const char* name; int spmm_b200_pipeline::Opt::*field; };
static const PipeOptionName PIPE_OPTION_TABLE[] = {
	{"SDDMM_GROUP", &spmm_b200_pipeline::Opt::sddmm_group}, {"TICKET_REV", &spmm_b200_pipeline::Opt::ticket_rev},
	{"SDDMM_TUNE", &spmm_b200_pipeline::Opt::sddmm_tune}, {"SDDMM_STAGED", &spmm_b200_pipeline::Opt::sddmm_staged},
	{"PIPELINE_FORK", &spmm_b200_pipeline::Opt::fork}, {"ATTENTION_FUSED", &spmm_b200_pipeline::Opt::fused},
};

extern "C" int spmm_b200_pipeline_set_option(spmm_b200_pipeline_t* p, const char* name, int value)
{
	if (!p || !name) return SPMM_B200_EINVAL;
	if (!strncmp(name, "SPMM_B200_", 10)) name += 10;
	for (const PipeOptionName& e : PIPE_OPTION_TABLE)
		if (!strcmp(name, e.name)) { p->opt.*(e.field) = value; return SPMM_B200_OK; }
	return SPMM_B200_EINVAL;
}

// Cheap fingerprint of a host CSR weight (sizes + up to 1024 strided samples of the arrays): a weight that was refilled
// in place behind the same pointers is then uploaded again instead of being taken for the resident copy.
static uint64_t pipe_weight_signature(const int32_t* ia, const int32_t* ja, const double* a, int64_t m, int64_t nnz)
{
	uint64_t h = 1469598103934665603ull;
	auto mix = [&h](uint64_t v) { h = (h ^ v) * 1099511628211ull; };
	mix((uint64_t) m); mix((uint64_t) nnz);
	const int64_t sm = std::max<int64_t>(1, (m + 1) / 1024), sn = std::max<int64_t>(1, nnz / 1024);
	for (int64_t i = 0; i <= m; i += sm) mix((uint64_t) (uint32_t) ia[i]);
	for (int64_t j = 0; j < nnz; j += sn) {
		uint64_t bits;
		memcpy(&bits, a + j, sizeof(bits));
		mix(bits); mix((uint64_t) (uint32_t) ja[j]);
	}
	return h;
}

static constexpr int PIPE_PARTS = 1184;      // 148 SMs x 8 blocks: partials of the global softmax

static int pipe_index(char type)
{
	return type == 'K' ? 0 : type == 'Q' ? 1 : type == 'V' ? 2 : -1;
}

// SpMM handle whose value array can be swapped: never a blocked variant.
static int pipe_make_handle(spmm_b200** h, const int32_t* ia, const int32_t* ja, const double* a,
                            int64_t m, int64_t c, int64_t nnz, int n, int device, bool plain_values)
{
	int rc = spmm_b200_create(h, ia, ja, a, m, c, nnz, n, device, SPMM_B200_AUTO, SPMM_B200_B_ROWMAJOR);
	if (rc) return rc;
	if (plain_values && ((*h)->variant == SPMM_B200_BCSR || (*h)->variant == SPMM_B200_BCSR_FMA)) {
		spmm_b200_destroy(*h);
		*h = nullptr;
		rc = spmm_b200_create(h, ia, ja, a, m, c, nnz, n, device, SPMM_B200_ROWSPLIT, SPMM_B200_B_ROWMAJOR);
	}
	return rc;
}

extern "C" void spmm_b200_pipeline_destroy(spmm_b200_pipeline_t* p)
{
	if (!p) return;
	cudaSetDevice(p->device);
	if (p->stream) cudaStreamSynchronize(p->stream);
	if (p->mask) spmm_b200_destroy(p->mask);
	for (auto& w : p->w) {
		if (w.h) spmm_b200_destroy(w.h);
		if (w.d_out) cudaFree(w.d_out);
	}
	if (p->d_mask_a) cudaFree(p->d_mask_a);
	if (p->d_x) cudaFree(p->d_x);
	if (p->d_tmp) cudaFree(p->d_tmp);
	if (p->d_final) cudaFree(p->d_final);
	if (p->d_part) cudaFree(p->d_part);
	if (p->d_ticket) cudaFree(p->d_ticket);
	if (p->ev0) cudaEventDestroy(p->ev0);
	if (p->ev1) cudaEventDestroy(p->ev1);
	if (p->ev_fork) cudaEventDestroy(p->ev_fork);
	for (int i = 0; i < 3; i++) {
		if (p->ev_done[i]) cudaEventDestroy(p->ev_done[i]);
		if (p->s_side[i]) { cudaStreamSynchronize(p->s_side[i]); cudaStreamDestroy(p->s_side[i]); }
	}
	if (p->stream) cudaStreamDestroy(p->stream);
	delete p;
}

extern "C" int spmm_b200_pipeline_create(spmm_b200_pipeline_t** out,
                                         const int32_t* mask_ia, const int32_t* mask_ja, const double* mask_a,
                                         int64_t m, int64_t nnz, int n, int device)
{
	if (!out) return SPMM_B200_EINVAL;
	*out = nullptr;
	if (m < 0 || nnz < 0 || n < 1 || !mask_ia || (nnz > 0 && (!mask_ja || !mask_a))) return SPMM_B200_EINVAL;
	if (m >= INT32_MAX || nnz >= INT32_MAX) return SPMM_B200_ERANGE;
	int rc = check_device(device, nullptr);
	if (rc) return rc;
	spmm_b200_pipeline* p = new (std::nothrow) spmm_b200_pipeline;
	if (!p) return SPMM_B200_ENOMEM;
	p->device = device; p->n = n; p->m = m; p->nnz = nnz;
	p->mask_ia_host = mask_ia; p->mask_ja_host = mask_ja;
#define PIPE_TRY(expr) do { cudaError_t _e = (expr); if (_e != cudaSuccess) { \
		snprintf(g_cuda_err, sizeof(g_cuda_err), "%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
		spmm_b200_pipeline_destroy(p); return _e == cudaErrorMemoryAllocation ? SPMM_B200_ENOMEM : SPMM_B200_ECUDA; } } while (0)
	PIPE_TRY(cudaSetDevice(device));
	PIPE_TRY(cudaStreamCreateWithFlags(&p->stream, cudaStreamNonBlocking));
	PIPE_TRY(cudaEventCreate(&p->ev0));
	PIPE_TRY(cudaEventCreate(&p->ev1));
	PIPE_TRY(cudaEventCreateWithFlags(&p->ev_fork, cudaEventDisableTiming));
	for (int i = 0; i < 3; i++) {
		PIPE_TRY(cudaStreamCreateWithFlags(&p->s_side[i], cudaStreamNonBlocking));
		PIPE_TRY(cudaEventCreateWithFlags(&p->ev_done[i], cudaEventDisableTiming));
	}
	rc = pipe_make_handle(&p->mask, mask_ia, mask_ja, mask_a, m, m, nnz, n, device, true);
	if (rc) { spmm_b200_pipeline_destroy(p); return rc; }
	PIPE_TRY(cudaMalloc(&p->d_mask_a, sizeof(double) * std::max<int64_t>(nnz, 1)));
	if (nnz > 0) PIPE_TRY(cudaMemcpy(p->d_mask_a, mask_a, sizeof(double) * nnz, cudaMemcpyHostToDevice));
	PIPE_TRY(cudaMalloc(&p->d_final, sizeof(double) * std::max<int64_t>(m * n, 1)));
	PIPE_TRY(cudaMalloc(&p->d_part, sizeof(double) * (PIPE_PARTS + 2)));
	PIPE_TRY(cudaMalloc(&p->d_ticket, sizeof(int)));
	{
		char name[64];
		for (const PipeOptionName& e : PIPE_OPTION_TABLE) {
			snprintf(name, sizeof(name), "SPMM_B200_%s", e.name);
			p->opt.*(e.field) = env_int(name, p->opt.*(e.field));
		}
	}
#undef PIPE_TRY
	*out = p;
	return SPMM_B200_OK;
}

extern "C" int spmm_b200_pipeline_set_weight(spmm_b200_pipeline_t* p, char type, int64_t m, int64_t k, int64_t nnz,
                                             const int32_t* ia, const int32_t* ja, const double* a)
{
	const int idx = pipe_index(type);
	if (!p || idx < 0 || !ia || m != p->m || k < 0 || nnz < 0 || (nnz > 0 && (!ja || !a))) return SPMM_B200_EINVAL;
	auto& w = p->w[idx];
	const uint64_t sig = pipe_weight_signature(ia, ja, a, m, nnz);
	if (w.h && w.ia == ia && w.ja == ja && w.a == a && w.m == m && w.k == k && w.nnz == nnz && w.sig == sig) return SPMM_B200_OK;
	CU_TRY(cudaSetDevice(p->device));
	if (w.h) { p->own_launches += w.h->launches; spmm_b200_destroy(w.h); w.h = nullptr; }
	w.have = false;
	int rc = pipe_make_handle(&w.h, ia, ja, a, m, k, nnz, p->n, p->device, false);
	if (rc) return rc;
	if (!w.d_out) CU_TRY(cudaMalloc(&w.d_out, sizeof(double) * std::max<int64_t>(m * p->n, 1)));
	w.ia = ia; w.ja = ja; w.a = a; w.m = m; w.k = k; w.nnz = nnz; w.sig = sig;
	return SPMM_B200_OK;
}

// Forget what is resident for `type` ('K', 'Q', 'V'; '*': all three): the next spmm / set_weight with the same
// host pointers uploads the arrays again.  For callers that refill a weight in place.
extern "C" int spmm_b200_pipeline_invalidate(spmm_b200_pipeline_t* p, char type)
{
	if (!p) return SPMM_B200_EINVAL;
	if (type != '*' && pipe_index(type) < 0) return SPMM_B200_EINVAL;
	for (int i = 0; i < 3; i++)
		if (type == '*' || pipe_index(type) == i) { p->w[i].ia = nullptr; p->w[i].ja = nullptr; p->w[i].a = nullptr; p->w[i].sig = 0; }
	return SPMM_B200_OK;
}

static int pipe_sddmm_launch(spmm_b200_pipeline* p, int mode)
{
	if (mode != SPMM_B200_SDDMM_REFERENCE && mode != SPMM_B200_SDDMM_ATTENTION && mode != SPMM_B200_SDDMM_ATTENTION_ORDERED) return SPMM_B200_EINVAL;
	if (!p->w[0].have || !p->w[1].have) return SPMM_B200_EINVAL;
	if (p->m > 0 && p->nnz > 0) {
		const int grid = (int) std::min<int64_t>((p->m + THREADS / 32 - 1) / (THREADS / 32), (int64_t) p->mask->sm_count * 64);
		if (mode == SPMM_B200_SDDMM_REFERENCE)
			sb200::sddmm_rowdot_kernel<THREADS><<<grid, THREADS, 0, p->stream>>>(p->mask->d_ia, p->d_mask_a, p->w[1].d_out, p->w[0].d_out,
			                                                                    p->mask->d_a, p->m, p->n);
		else {
			// attention mode: the group kernel for the widths it is instantiated for (tree-ordered sum), else -- and
			// always in ORDERED mode -- a kernel that keeps the CPU loop's left-to-right order
			const bool fast = mode == SPMM_B200_SDDMM_ATTENTION && p->opt.sddmm_group;
			const int gw = (int) std::min<int64_t>((p->m + THREADS / 32 - 1) / (THREADS / 32), (int64_t) p->mask->sm_count * 64);
			// one resident wave of persistent CTAs (MINB per SM); rows are handed out by the ticket counter
#define SDDMM_GROUP(G, V, U, MINB) do { \
				CU_TRY(cudaMemsetAsync(p->d_ticket, 0, sizeof(int), p->stream)); \
				sb200::sddmm_attention_group_kernel<G, V, U, THREADS, MINB><<<(int) std::min<int64_t>(gw, (int64_t) p->mask->sm_count * MINB), THREADS, 0, p->stream>>>( \
				p->mask->d_ia, p->mask->d_ja, p->d_mask_a, p->w[1].d_out, p->w[0].d_out, p->mask->d_a, p->m, p->d_ticket, p->opt.ticket_rev); } while (0)
			// shapes measured on m = 8192 / 16384 (profiles/r1_pipeline_probe.md): U = 2 rows in flight; with the
			// persistent ticketed grid n = 64 runs best at 126 registers / 2 CTAs per SM, n = 32 at 64 registers / 4
			const int tune = p->opt.sddmm_tune;     // experiment knob for n = 64: U*10 + CTAs per SM
			if (fast && p->n == 16) SDDMM_GROUP(4, 4, 2, 4);
			else if (fast && p->n == 32) SDDMM_GROUP(8, 4, 2, 4);
			else if (fast && p->n == 64 && tune == 42) SDDMM_GROUP(16, 4, 4, 2);
			else if (fast && p->n == 64 && tune == 24) SDDMM_GROUP(16, 4, 2, 4);
			else if (fast && p->n == 64) SDDMM_GROUP(16, 4, 2, 2);
			else if (fast && p->n == 128) SDDMM_GROUP(16, 8, 2, 2);
#undef SDDMM_GROUP
			else if (p->n % sb200::SDDMM_CH == 0 && p->n <= 1024 && p->opt.sddmm_staged) {
				// K rows staged through shared memory (sddmm.cuh): 4 warps per CTA, 32 x 34 + n doubles per warp
				constexpr int T = 128;
				const size_t smem = sizeof(double) * (T / 32) * (32 * sb200::SDDMM_LD + p->n);
				if (smem > 48 * 1024)       // per device and cheap: no process-wide flag
					CU_TRY(cudaFuncSetAttribute(sb200::sddmm_attention_staged_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
				const int g2 = (int) std::min<int64_t>((p->m + T / 32 - 1) / (T / 32), (int64_t) p->mask->sm_count * 64);
				sb200::sddmm_attention_staged_kernel<T><<<g2, T, smem, p->stream>>>(p->mask->d_ia, p->mask->d_ja, p->d_mask_a,
				                                                                  p->w[1].d_out, p->w[0].d_out, p->mask->d_a, p->m, p->n);
			} else
				sb200::sddmm_attention_kernel<THREADS><<<grid, THREADS, 0, p->stream>>>(p->mask->d_ia, p->mask->d_ja, p->d_mask_a,
				                                                                       p->w[1].d_out, p->w[0].d_out, p->mask->d_a, p->m, p->n);
		}
		CU_TRY(cudaGetLastError());
		p->own_launches += 1;
	}
	p->have_y = true;
	return SPMM_B200_OK;
}

static int pipe_softmax_launch(spmm_b200_pipeline* p, int rowwise)
{
	if (!p->have_y) return SPMM_B200_EINVAL;
	if (p->nnz == 0) return SPMM_B200_OK;
	double* y = p->mask->d_a;
	if (rowwise) {
		const int grid = (int) std::min<int64_t>((p->m + THREADS / 32 - 1) / (THREADS / 32), (int64_t) p->mask->sm_count * 64);
		sb200::softmax_rows_kernel<THREADS><<<grid, THREADS, 0, p->stream>>>(p->mask->d_ia, y, p->m);
		p->own_launches += 1;
	} else {
		const int grid = (int) std::min<int64_t>((p->nnz + THREADS - 1) / THREADS, PIPE_PARTS);
		sb200::softmax_max_kernel<THREADS><<<grid, THREADS, 0, p->stream>>>(y, p->nnz, p->d_part);
		sb200::softmax_combine_kernel<THREADS, true><<<1, THREADS, 0, p->stream>>>(p->d_part, grid);
		// the maximum moves to part[PIPE_PARTS + 1] so that the sums can reuse part[0..grid]
		CU_TRY(cudaMemcpyAsync(p->d_part + PIPE_PARTS + 1, p->d_part + grid, sizeof(double), cudaMemcpyDeviceToDevice, p->stream));
		sb200::softmax_exp_kernel<THREADS><<<grid, THREADS, 0, p->stream>>>(y, p->nnz, p->d_part + PIPE_PARTS + 1, p->d_part);
		sb200::softmax_combine_kernel<THREADS, false><<<1, THREADS, 0, p->stream>>>(p->d_part, grid);
		sb200::softmax_scale_kernel<THREADS><<<grid, THREADS, 0, p->stream>>>(y, p->nnz, p->d_part + grid);
		p->own_launches += 5;
	}
	CU_TRY(cudaGetLastError());
	return SPMM_B200_OK;
}

extern "C" int spmm_b200_pipeline_spmm(spmm_b200_pipeline_t* p, char type, int64_t m, int64_t k, int n,
                                       const int32_t* ia, const int32_t* ja, const double* a,
                                       const double* x_host, double* y_host)
{
	if (!p || !ia || m < 0 || k < 0 || n != p->n) return SPMM_B200_EINVAL;
	if (k > 0 && !x_host) return SPMM_B200_EINVAL;
	CU_TRY(cudaSetDevice(p->device));
	const int idx = pipe_index(type);
	int rc;
	if (idx >= 0) {
		// read nnz from the caller's row pointer exactly as csr_to_format's callers do
		const int64_t nnz = ia[m];
		rc = spmm_b200_pipeline_set_weight(p, type, m, k, nnz, ia, ja, a);
		if (rc) return rc;
		auto& w = p->w[idx];
		rc = ensure(&p->d_x, &p->cap_x, (size_t) k * n);
		if (rc) return rc;
		if (k > 0) CU_TRY(cudaMemcpyAsync(p->d_x, x_host, sizeof(double) * k * n, cudaMemcpyHostToDevice, p->stream));
		CU_TRY(cudaEventRecord(p->ev0, p->stream));
		if (m > 0) {
			rc = spmm_b200_run_device(w.h, p->d_x, w.d_out, n, p->stream);
			if (rc) return rc;
		}
		w.have = true;
		CU_TRY(cudaEventRecord(p->ev1, p->stream));
		if (y_host && m > 0) CU_TRY(cudaMemcpyAsync(y_host, w.d_out, sizeof(double) * m * n, cudaMemcpyDeviceToHost, p->stream));
		CU_TRY(cudaStreamSynchronize(p->stream));
		p->timed = true;
		return SPMM_B200_OK;
	}
	if (!y_host && m > 0) return SPMM_B200_EINVAL;
	const int64_t nnz = ia[m];
	if (nnz > 0 && (!ja || !a)) return SPMM_B200_EINVAL;
	rc = ensure(&p->d_x, &p->cap_x, (size_t) k * n);
	if (rc) return rc;
	if (k > 0) CU_TRY(cudaMemcpyAsync(p->d_x, x_host, sizeof(double) * k * n, cudaMemcpyHostToDevice, p->stream));
	if (ia == p->mask_ia_host && ja == p->mask_ja_host && m == p->m && k == p->m && nnz == p->nnz) {
		// the harness's spmm('final', Mask->m, Mask->m, n, Mask->csr_ia, Mask->csr_ja, y, V, y_final)
		// (sddmm_bench.cpp:934): the mask's structure is resident, only the values and V travel
		if (nnz > 0) CU_TRY(cudaMemcpyAsync(p->mask->d_a, a, sizeof(double) * nnz, cudaMemcpyHostToDevice, p->stream));
		p->have_y = true;
		CU_TRY(cudaEventRecord(p->ev0, p->stream));
		if (m > 0) {
			rc = spmm_b200_run_device(p->mask, p->d_x, p->d_final, n, p->stream);
			if (rc) return rc;
		}
		CU_TRY(cudaEventRecord(p->ev1, p->stream));
		p->have_final = true;
		if (m > 0) CU_TRY(cudaMemcpyAsync(y_host, p->d_final, sizeof(double) * m * n, cudaMemcpyDeviceToHost, p->stream));
		CU_TRY(cudaStreamSynchronize(p->stream));
		p->timed = true;
		return SPMM_B200_OK;
	}
	// any other matrix: a one-off product
	spmm_b200* h = nullptr;
	rc = pipe_make_handle(&h, ia, ja, a, m, k, nnz, n, p->device, false);
	if (rc) return rc;
	rc = ensure(&p->d_tmp, &p->cap_tmp, (size_t) m * n);
	if (!rc) {
		cudaEventRecord(p->ev0, p->stream);
		if (m > 0) rc = spmm_b200_run_device(h, p->d_x, p->d_tmp, n, p->stream);
		cudaEventRecord(p->ev1, p->stream);
	}
	if (!rc && m > 0 && cudaMemcpyAsync(y_host, p->d_tmp, sizeof(double) * m * n, cudaMemcpyDeviceToHost, p->stream) != cudaSuccess) rc = SPMM_B200_ECUDA;
	if (cudaStreamSynchronize(p->stream) != cudaSuccess && !rc) rc = SPMM_B200_ECUDA;
	p->own_launches += h->launches;
	spmm_b200_destroy(h);
	p->timed = rc == SPMM_B200_OK;
	return rc;
}

extern "C" int spmm_b200_pipeline_sddmm(spmm_b200_pipeline_t* p, double* y_host, int mode)
{
	if (!p) return SPMM_B200_EINVAL;
	CU_TRY(cudaSetDevice(p->device));
	CU_TRY(cudaEventRecord(p->ev0, p->stream));
	int rc = pipe_sddmm_launch(p, mode);
	if (rc) return rc;
	CU_TRY(cudaEventRecord(p->ev1, p->stream));
	if (y_host && p->nnz > 0) CU_TRY(cudaMemcpyAsync(y_host, p->mask->d_a, sizeof(double) * p->nnz, cudaMemcpyDeviceToHost, p->stream));
	CU_TRY(cudaStreamSynchronize(p->stream));
	p->timed = true;
	return SPMM_B200_OK;
}

extern "C" int spmm_b200_pipeline_softmax(spmm_b200_pipeline_t* p, double* y_host, int rowwise)
{
	if (!p) return SPMM_B200_EINVAL;
	CU_TRY(cudaSetDevice(p->device));
	CU_TRY(cudaEventRecord(p->ev0, p->stream));
	int rc = pipe_softmax_launch(p, rowwise);
	if (rc) return rc;
	CU_TRY(cudaEventRecord(p->ev1, p->stream));
	if (y_host && p->nnz > 0) CU_TRY(cudaMemcpyAsync(y_host, p->mask->d_a, sizeof(double) * p->nnz, cudaMemcpyDeviceToHost, p->stream));
	CU_TRY(cudaStreamSynchronize(p->stream));
	p->timed = true;
	return SPMM_B200_OK;
}

extern "C" int spmm_b200_pipeline_final(spmm_b200_pipeline_t* p, double* y_final_host)
{
	if (!p || !p->have_y || !p->w[2].have) return SPMM_B200_EINVAL;
	CU_TRY(cudaSetDevice(p->device));
	CU_TRY(cudaEventRecord(p->ev0, p->stream));
	if (p->m > 0) {
		int rc = spmm_b200_run_device(p->mask, p->w[2].d_out, p->d_final, p->n, p->stream);
		if (rc) return rc;
	}
	p->have_final = true;
	CU_TRY(cudaEventRecord(p->ev1, p->stream));
	if (y_final_host && p->m > 0)
		CU_TRY(cudaMemcpyAsync(y_final_host, p->d_final, sizeof(double) * p->m * p->n, cudaMemcpyDeviceToHost, p->stream));
	CU_TRY(cudaStreamSynchronize(p->stream));
	p->timed = true;
	return SPMM_B200_OK;
}

extern "C" int spmm_b200_pipeline_attention(spmm_b200_pipeline_t* p, const double* x_host, double* y_final_host,
                                            int mode, int softmax)
{
	if (!p || softmax < 0 || softmax > 2) return SPMM_B200_EINVAL;
	for (auto& w : p->w)
		if (!w.h || w.k != p->w[0].k) return SPMM_B200_EINVAL;
	const int64_t k = p->w[0].k;
	if (k > 0 && !x_host) return SPMM_B200_EINVAL;
	CU_TRY(cudaSetDevice(p->device));
	int rc = ensure(&p->d_x, &p->cap_x, (size_t) k * p->n);
	if (rc) return rc;
	if (k > 0) CU_TRY(cudaMemcpyAsync(p->d_x, x_host, sizeof(double) * k * p->n, cudaMemcpyHostToDevice, p->stream));
	CU_TRY(cudaEventRecord(p->ev0, p->stream));
	// K | Q | V on three side streams; the SDDMM waits for K and Q only, the final product for V
	const bool fork = p->m > 0 && p->opt.fork;
	if (fork) CU_TRY(cudaEventRecord(p->ev_fork, p->stream));
	for (int i = 0; i < 3; i++) {
		auto& w = p->w[i];
		if (p->m > 0) {
			cudaStream_t s = fork ? p->s_side[i] : p->stream;
			if (fork) CU_TRY(cudaStreamWaitEvent(s, p->ev_fork, 0));
			rc = spmm_b200_run_device(w.h, p->d_x, w.d_out, p->n, s);
			if (rc) return rc;
			if (fork) CU_TRY(cudaEventRecord(p->ev_done[i], s));
		}
		w.have = true;
	}
	if (fork) {
		CU_TRY(cudaStreamWaitEvent(p->stream, p->ev_done[0], 0));
		CU_TRY(cudaStreamWaitEvent(p->stream, p->ev_done[1], 0));
	}
	// attention-mode SDDMM + row-wise softmax + final product as ONE kernel (sddmm.cuh: attention_fused_kernel);
	// y is not materialised on this path (spmm_b200_pipeline_device_ptr(p, 'y') returns NULL afterwards)
	const bool fused = mode == SPMM_B200_SDDMM_ATTENTION && softmax == 2 && p->m > 0 && p->nnz > 0
	                   && (p->n == 32 || p->n == 64 || p->n == 128) && p->opt.fused;
	if (fused) {
		if (fork) CU_TRY(cudaStreamWaitEvent(p->stream, p->ev_done[2], 0));
		const int gw = (int) std::min<int64_t>((p->m + THREADS / 32 - 1) / (THREADS / 32), (int64_t) p->mask->sm_count * 64);
#define ATT_FUSED(G, V, U, MINB) do { \
			CU_TRY(cudaMemsetAsync(p->d_ticket, 0, sizeof(int), p->stream)); \
			sb200::attention_fused_kernel<G, V, U, THREADS, MINB><<<(int) std::min<int64_t>(gw, (int64_t) p->mask->sm_count * MINB), THREADS, 0, p->stream>>>( \
			p->mask->d_ia, p->mask->d_ja, p->d_mask_a, p->w[1].d_out, p->w[0].d_out, p->w[2].d_out, p->d_final, p->m, p->d_ticket, p->opt.ticket_rev); } while (0)
		// 128 registers (2 CTAs per SM) measured best: tighter caps spill part[], looser ones lose the second CTA;
		// n = 128 keeps one K / V row in flight per group to stay inside them (profiles/r1_pipeline_probe.md)
		if (p->n == 32) ATT_FUSED(8, 4, 2, 2);
		else if (p->n == 64) ATT_FUSED(16, 4, 2, 2);
		else ATT_FUSED(16, 8, 1, 2);
#undef ATT_FUSED
		CU_TRY(cudaGetLastError());
		p->own_launches += 1;
		p->have_y = false;
	} else {
		rc = pipe_sddmm_launch(p, mode);
		if (rc) return rc;
		if (softmax) {
			rc = pipe_softmax_launch(p, softmax == 2);
			if (rc) return rc;
		}
		if (fork) CU_TRY(cudaStreamWaitEvent(p->stream, p->ev_done[2], 0));
		if (p->m > 0) {
			rc = spmm_b200_run_device(p->mask, p->w[2].d_out, p->d_final, p->n, p->stream);
			if (rc) return rc;
		}
	}
	p->have_final = true;
	CU_TRY(cudaEventRecord(p->ev1, p->stream));
	if (y_final_host && p->m > 0)
		CU_TRY(cudaMemcpyAsync(y_final_host, p->d_final, sizeof(double) * p->m * p->n, cudaMemcpyDeviceToHost, p->stream));
	CU_TRY(cudaStreamSynchronize(p->stream));
	p->timed = true;
	return SPMM_B200_OK;
}

extern "C" const double* spmm_b200_pipeline_device_ptr(const spmm_b200_pipeline_t* p, char which)
{
	if (!p) return nullptr;
	const int idx = pipe_index(which);
	if (idx >= 0) return p->w[idx].have ? p->w[idx].d_out : nullptr;
	if (which == 'y') return p->have_y ? p->mask->d_a : nullptr;
	if (which == 'f') return p->have_final ? p->d_final : nullptr;
	return nullptr;
}

extern "C" double spmm_b200_pipeline_last_kernel_ms(const spmm_b200_pipeline_t* p)
{
	if (!p || !p->timed) return -1.0;
	float ms = 0.f;
	if (cudaEventElapsedTime(&ms, p->ev0, p->ev1) != cudaSuccess) return -1.0;
	return (double) ms;
}

extern "C" int64_t spmm_b200_pipeline_launch_count(const spmm_b200_pipeline_t* p)
{
	if (!p) return 0;
	int64_t n = p->own_launches + (p->mask ? p->mask->launches : 0);
	for (auto& w : p->w) n += w.h ? w.h->launches : 0;
	return n;
}

// ------------------------------------------------------------------------------------------
// fp32 pipeline (the pipeline bench's "_f" build): the step-by-step interface on the fp32 SpMM handles
// (spmm_b200_f32.inl) and the ordered fp32 SDDMM kernels of sddmm.cuh.  No softmax / fused path yet.
// ------------------------------------------------------------------------------------------
struct spmm_b200_pipeline_f32 {
	int device = 0, n = 0;
	int64_t m = 0, nnz = 0;
	const int32_t* mask_ia_host = nullptr;
	const int32_t* mask_ja_host = nullptr;
	spmm_b200_f32* mask = nullptr;           // its d_a holds y after sddmm
	float* d_mask_a = nullptr;
	struct Weight {
		spmm_b200_f32* h = nullptr;
		const int32_t* ia = nullptr; const int32_t* ja = nullptr; const float* a = nullptr;
		int64_t m = 0, k = 0, nnz = 0;
		float* d_out = nullptr;
		bool have = false;
	} w[3];
	float* d_x = nullptr; size_t cap_x = 0;
	float* d_tmp = nullptr; size_t cap_tmp = 0;
	float* d_final = nullptr;
	double* d_part = nullptr;                // partials of the global softmax (allocated on first use)
	bool have_y = false;
	cudaStream_t stream = nullptr;
	int64_t own_launches = 0;
};

extern "C" void spmm_b200_pipeline_destroy_f32(spmm_b200_pipeline_f32_t* p)
{
	if (!p) return;
	cudaSetDevice(p->device);
	if (p->stream) cudaStreamSynchronize(p->stream);
	if (p->d_part) cudaFree(p->d_part);
	if (p->mask) spmm_b200_destroy_f32(p->mask);
	for (auto& w : p->w) {
		if (w.h) spmm_b200_destroy_f32(w.h);
		if (w.d_out) cudaFree(w.d_out);
	}
	if (p->d_mask_a) cudaFree(p->d_mask_a);
	if (p->d_x) cudaFree(p->d_x);
	if (p->d_tmp) cudaFree(p->d_tmp);
	if (p->d_final) cudaFree(p->d_final);
	if (p->stream) cudaStreamDestroy(p->stream);
	delete p;
}

extern "C" int spmm_b200_pipeline_create_f32(spmm_b200_pipeline_f32_t** out,
                                             const int32_t* mask_ia, const int32_t* mask_ja, const float* mask_a,
                                             int64_t m, int64_t nnz, int n, int device)
{
	if (!out) return SPMM_B200_EINVAL;
	*out = nullptr;
	if (m < 0 || nnz < 0 || n < 1 || !mask_ia || (nnz > 0 && (!mask_ja || !mask_a))) return SPMM_B200_EINVAL;
	if (m >= INT32_MAX || nnz >= INT32_MAX) return SPMM_B200_ERANGE;
	int rc = check_device(device, nullptr);
	if (rc) return rc;
	spmm_b200_pipeline_f32* p = new (std::nothrow) spmm_b200_pipeline_f32;
	if (!p) return SPMM_B200_ENOMEM;
	p->device = device; p->n = n; p->m = m; p->nnz = nnz;
	p->mask_ia_host = mask_ia; p->mask_ja_host = mask_ja;
	cudaError_t e = cudaStreamCreateWithFlags(&p->stream, cudaStreamNonBlocking);
	if (e == cudaSuccess) e = cudaMalloc(&p->d_mask_a, sizeof(float) * std::max<int64_t>(nnz, 1));
	if (e == cudaSuccess && nnz > 0) e = cudaMemcpy(p->d_mask_a, mask_a, sizeof(float) * nnz, cudaMemcpyHostToDevice);
	if (e == cudaSuccess) e = cudaMalloc(&p->d_final, sizeof(float) * std::max<int64_t>(m * n, 1));
	if (e != cudaSuccess) {
		snprintf(g_cuda_err, sizeof(g_cuda_err), "spmm_b200_pipeline_create_f32: %s", cudaGetErrorString(e));
		spmm_b200_pipeline_destroy_f32(p);
		return e == cudaErrorMemoryAllocation ? SPMM_B200_ENOMEM : SPMM_B200_ECUDA;
	}
	rc = spmm_b200_create_f32(&p->mask, mask_ia, mask_ja, mask_a, m, m, nnz, n, device, SPMM_B200_ROWSPLIT, SPMM_B200_B_ROWMAJOR);
	if (rc) { spmm_b200_pipeline_destroy_f32(p); return rc; }
	*out = p;
	return SPMM_B200_OK;
}

extern "C" int spmm_b200_pipeline_spmm_f32(spmm_b200_pipeline_f32_t* p, char type, int64_t m, int64_t k, int n,
                                           const int32_t* ia, const int32_t* ja, const float* a,
                                           const float* x_host, float* y_host)
{
	if (!p || !ia || m < 0 || k < 0 || n != p->n) return SPMM_B200_EINVAL;
	if (k > 0 && !x_host) return SPMM_B200_EINVAL;
	CU_TRY(cudaSetDevice(p->device));
	const int idx = pipe_index(type);
	const int64_t nnz = ia[m];
	if (nnz > 0 && (!ja || !a)) return SPMM_B200_EINVAL;
	int rc = ensure32(&p->d_x, &p->cap_x, (size_t) k * n);
	if (rc) return rc;
	if (k > 0) CU_TRY(cudaMemcpyAsync(p->d_x, x_host, sizeof(float) * k * n, cudaMemcpyHostToDevice, p->stream));
	if (idx >= 0) {
		if (m != p->m) return SPMM_B200_EINVAL;
		auto& w = p->w[idx];
		if (!(w.h && w.ia == ia && w.ja == ja && w.a == a && w.m == m && w.k == k && w.nnz == nnz)) {
			if (w.h) { p->own_launches += w.h->launches; spmm_b200_destroy_f32(w.h); w.h = nullptr; }
			w.have = false;
			rc = spmm_b200_create_f32(&w.h, ia, ja, a, m, k, nnz, n, p->device, SPMM_B200_ROWSPLIT, SPMM_B200_B_ROWMAJOR);
			if (rc) return rc;
			if (!w.d_out) CU_TRY(cudaMalloc(&w.d_out, sizeof(float) * std::max<int64_t>(m * n, 1)));
			w.ia = ia; w.ja = ja; w.a = a; w.m = m; w.k = k; w.nnz = nnz;
		}
		if (m > 0) {
			rc = spmm_b200_run_device_f32(w.h, p->d_x, w.d_out, n, p->stream);
			if (rc) return rc;
		}
		w.have = true;
		if (y_host && m > 0) CU_TRY(cudaMemcpyAsync(y_host, w.d_out, sizeof(float) * m * n, cudaMemcpyDeviceToHost, p->stream));
		CU_TRY(cudaStreamSynchronize(p->stream));
		return SPMM_B200_OK;
	}
	if (!y_host && m > 0) return SPMM_B200_EINVAL;
	if (ia == p->mask_ia_host && ja == p->mask_ja_host && m == p->m && k == p->m && nnz == p->nnz) {
		// spmm('final', mask, y, V, y_final) with host y and V: only the values and V travel
		if (nnz > 0) CU_TRY(cudaMemcpyAsync(p->mask->d_a, a, sizeof(float) * nnz, cudaMemcpyHostToDevice, p->stream));
		p->have_y = true;
		if (m > 0) {
			rc = spmm_b200_run_device_f32(p->mask, p->d_x, p->d_final, n, p->stream);
			if (rc) return rc;
			CU_TRY(cudaMemcpyAsync(y_host, p->d_final, sizeof(float) * m * n, cudaMemcpyDeviceToHost, p->stream));
		}
		CU_TRY(cudaStreamSynchronize(p->stream));
		return SPMM_B200_OK;
	}
	spmm_b200_f32* h = nullptr;
	rc = spmm_b200_create_f32(&h, ia, ja, a, m, k, nnz, n, p->device, SPMM_B200_ROWSPLIT, SPMM_B200_B_ROWMAJOR);
	if (rc) return rc;
	rc = ensure32(&p->d_tmp, &p->cap_tmp, (size_t) m * n);
	if (!rc && m > 0) rc = spmm_b200_run_device_f32(h, p->d_x, p->d_tmp, n, p->stream);
	if (!rc && m > 0 && cudaMemcpyAsync(y_host, p->d_tmp, sizeof(float) * m * n, cudaMemcpyDeviceToHost, p->stream) != cudaSuccess) rc = SPMM_B200_ECUDA;
	if (cudaStreamSynchronize(p->stream) != cudaSuccess && !rc) rc = SPMM_B200_ECUDA;
	p->own_launches += h->launches;
	spmm_b200_destroy_f32(h);
	return rc;
}

extern "C" int spmm_b200_pipeline_sddmm_f32(spmm_b200_pipeline_f32_t* p, float* y_host, int mode)
{
	if (!p || mode < SPMM_B200_SDDMM_REFERENCE || mode > SPMM_B200_SDDMM_ATTENTION_ORDERED) return SPMM_B200_EINVAL;
	if (!p->w[0].have || !p->w[1].have) return SPMM_B200_EINVAL;
	CU_TRY(cudaSetDevice(p->device));
	if (p->m > 0 && p->nnz > 0) {
		const int grid = (int) std::min<int64_t>((p->m + THREADS / 32 - 1) / (THREADS / 32), (int64_t) p->mask->sm_count * 64);
		if (mode == SPMM_B200_SDDMM_REFERENCE)
			sb200::sddmm_rowdot_f32_kernel<THREADS><<<grid, THREADS, 0, p->stream>>>(p->mask->d_ia, p->d_mask_a, p->w[1].d_out, p->w[0].d_out,
			                                                                        p->mask->d_a, p->m, p->n);
		else      // both attention modes: the ordered kernel (no tree-ordered fp32 kernel yet)
			sb200::sddmm_attention_f32_kernel<THREADS><<<grid, THREADS, 0, p->stream>>>(p->mask->d_ia, p->mask->d_ja, p->d_mask_a,
			                                                                           p->w[1].d_out, p->w[0].d_out, p->mask->d_a, p->m, p->n);
		CU_TRY(cudaGetLastError());
		p->own_launches += 1;
	}
	p->have_y = true;
	if (y_host && p->nnz > 0) CU_TRY(cudaMemcpyAsync(y_host, p->mask->d_a, sizeof(float) * p->nnz, cudaMemcpyDeviceToHost, p->stream));
	CU_TRY(cudaStreamSynchronize(p->stream));
	return SPMM_B200_OK;
}

// softmax over the mask's nonzeros in place (y lives in the mask handle's value array): rowwise != 0 = per mask row (what
// attention uses), 0 = over all nonzeros at once (the reference's commented-out host softmax(y), sddmm_bench.cpp)
extern "C" int spmm_b200_pipeline_softmax_f32(spmm_b200_pipeline_f32_t* p, float* y_host, int rowwise)
{
	if (!p || !p->have_y) return SPMM_B200_EINVAL;
	CU_TRY(cudaSetDevice(p->device));
	if (p->nnz > 0) {
		float* y = p->mask->d_a;
		if (rowwise) {
			const int grid = (int) std::min<int64_t>((p->m + THREADS / 32 - 1) / (THREADS / 32), (int64_t) p->mask->sm_count * 64);
			sb200::softmax_rows_f32_kernel<THREADS><<<grid, THREADS, 0, p->stream>>>(p->mask->d_ia, y, p->m);
			p->own_launches += 1;
		} else {
			if (!p->d_part) CU_TRY(cudaMalloc(&p->d_part, sizeof(double) * (PIPE_PARTS + 2)));
			const int grid = (int) std::min<int64_t>((p->nnz + THREADS - 1) / THREADS, PIPE_PARTS);
			sb200::softmax_max_f32_kernel<THREADS><<<grid, THREADS, 0, p->stream>>>(y, p->nnz, p->d_part);
			sb200::softmax_combine_kernel<THREADS, true><<<1, THREADS, 0, p->stream>>>(p->d_part, grid);
			CU_TRY(cudaMemcpyAsync(p->d_part + PIPE_PARTS + 1, p->d_part + grid, sizeof(double), cudaMemcpyDeviceToDevice, p->stream));
			sb200::softmax_exp_f32_kernel<THREADS><<<grid, THREADS, 0, p->stream>>>(y, p->nnz, p->d_part + PIPE_PARTS + 1, p->d_part);
			sb200::softmax_combine_kernel<THREADS, false><<<1, THREADS, 0, p->stream>>>(p->d_part, grid);
			sb200::softmax_scale_f32_kernel<THREADS><<<grid, THREADS, 0, p->stream>>>(y, p->nnz, p->d_part + grid);
			p->own_launches += 5;
		}
		CU_TRY(cudaGetLastError());
		if (y_host) CU_TRY(cudaMemcpyAsync(y_host, y, sizeof(float) * p->nnz, cudaMemcpyDeviceToHost, p->stream));
	}
	CU_TRY(cudaStreamSynchronize(p->stream));
	return SPMM_B200_OK;
}

extern "C" int spmm_b200_pipeline_final_f32(spmm_b200_pipeline_f32_t* p, float* y_final_host)
{
	if (!p || !p->have_y || !p->w[2].have) return SPMM_B200_EINVAL;
	CU_TRY(cudaSetDevice(p->device));
	if (p->m > 0) {
		int rc = spmm_b200_run_device_f32(p->mask, p->w[2].d_out, p->d_final, p->n, p->stream);
		if (rc) return rc;
		if (y_final_host) CU_TRY(cudaMemcpyAsync(y_final_host, p->d_final, sizeof(float) * p->m * p->n, cudaMemcpyDeviceToHost, p->stream));
	}
	CU_TRY(cudaStreamSynchronize(p->stream));
	return SPMM_B200_OK;
}

extern "C" int64_t spmm_b200_pipeline_launch_count_f32(const spmm_b200_pipeline_f32_t* p)
{
	if (!p) return 0;
	int64_t n = p->own_launches + (p->mask ? p->mask->launches : 0);
	for (auto& w : p->w) n += w.h ? w.h->launches : 0;
	return n;
}
