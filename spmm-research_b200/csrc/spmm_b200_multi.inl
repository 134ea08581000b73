// spmm_b200_multi.inl -- single-process multi-GPU entry points of the C ABI (include/spmm_b200.h, "_multi").
//
// the reference's harness is ONE C process (spmv_bench.cpp), so the drop-in reaches 2, 4 or 8 GPUs from
// one host thread: one shard handle per device (rows cut by the bit-exact loop_partitioner_balance_prefix_sums,
// partition.cuh), all work enqueued asynchronously on per-device streams, one wait at the end.  The replication of B
// is the only exchange on the path.  It is not a library broadcast: every device uploads 1/N of a column slice of x
// over ITS PCIe link into its own staging buffer, and then one kernel per device (gather_transpose_b_kernel) reads all
// N parts straight out of the peers' memory over NVLink / NVSwitch (peer loads) while writing the row-major operand
// the SpMM kernels gather from -- all-gather and layout transpose in one pass, no intermediate copy.
// Included at the end of spmm_b200.cu (one translation unit).

static constexpr int MULTI_MAX = 16;

struct PeerParts {
	const double* base[MULTI_MAX];   // base[r]: device r's staging copy of the slice (only part r of it is valid there)
	int64_t per;                     // part length in doubles: part r = [r*per, min((r+1)*per, count))
};

// B[j*k + n] = x[n*c + j] for one column slice (k columns), where element i = n*c + j of the slice is read from the
// device that uploaded it: parts.base[i / per][i].  Reads are coalesced along j, writes along n (32x33 tile).
__global__ void gather_transpose_b_kernel(PeerParts parts, double* __restrict__ B, int64_t c, int k)
{
	__shared__ double tile[32][33];
	const int64_t j0 = (int64_t) blockIdx.x * 32;
	const int n0 = blockIdx.y * 32;
	for (int dy = threadIdx.y; dy < 32; dy += blockDim.y) {
		const int n = n0 + dy;
		const int64_t j = j0 + threadIdx.x;
		if (n < k && j < c) {
			const int64_t i = (int64_t) n * c + j;
			tile[dy][threadIdx.x] = parts.base[i / parts.per][i];
		}
	}
	__syncthreads();
	for (int dy = threadIdx.y; dy < 32; dy += blockDim.y) {
		const int64_t j = j0 + dy;
		const int n = n0 + threadIdx.x;
		if (n < k && j < c) B[j * k + n] = tile[threadIdx.x][dy];
	}
}

struct spmm_b200_multi {
	int n = 0;
	int k = 0, b_layout = SPMM_B200_B_COLMAJOR;
	int64_t m = 0, c = 0, nnz = 0;
	bool peer = false;                           // every pair of devices can read each other's memory
	std::vector<spmm_b200*> sh;                  // shard handle per device
	std::vector<int> dev;
	std::vector<int64_t> bounds;                 // n + 1 row starts
	std::vector<std::vector<cudaEvent_t>> ev_part;   // [device][slice]: that device's part of the slice is uploaded
	std::vector<void*> pinned;
	int resident_k = 0;
	// download routing (multi_choose_routes): route[g] = the device whose PCIe link carries device g's rows of C
	std::vector<int> route;
	int route_mask = -2;                         // the MULTI_ROUTE value the routes were made for (-2: not yet)
	std::vector<double*> relay;                  // relay[p]: buffer on device p that its clients' kernels store C into over NVLink
	std::vector<size_t> relay_cap;
	std::vector<double> route_probe;             // GB/s of the candidate downloader sets, as measured (spmm_b200_multi_routes)
};

extern "C" void spmm_b200_destroy_multi(spmm_b200_multi_t* mh)
{
	if (!mh) return;
	for (int g = 0; g < (int) mh->sh.size(); g++) {
		if (g < (int) mh->ev_part.size()) {
			cudaSetDevice(mh->dev[g]);
			for (cudaEvent_t e : mh->ev_part[g]) cudaEventDestroy(e);
		}
		if (g < (int) mh->relay.size() && mh->relay[g]) { cudaSetDevice(mh->dev[g]); cudaFree(mh->relay[g]); }
		spmm_b200_destroy(mh->sh[g]);
	}
	for (void* p : mh->pinned) cudaHostUnregister(p);
	delete mh;
}

extern "C" int spmm_b200_create_multi(spmm_b200_multi_t** out, const int32_t* row_ptr, const int32_t* col_ind, const double* values,
                                      int64_t m, int64_t c, int64_t nnz, int k, int ngpus, const int* devices, int variant, int b_layout)
{
	if (!out) return SPMM_B200_EINVAL;
	*out = nullptr;
	int rc = check_sizes(row_ptr, col_ind, values, m, c, nnz, k);
	if (rc) return rc;
	if (ngpus < 1 || ngpus > MULTI_MAX) return SPMM_B200_EINVAL;
	if (variant < SPMM_B200_ROWSPLIT || variant > SPMM_B200_BCSR_FMA) return SPMM_B200_EINVAL;
	if (b_layout != SPMM_B200_B_COLMAJOR && b_layout != SPMM_B200_B_ROWMAJOR) return SPMM_B200_EINVAL;
	int have = 0;
	if (cudaGetDeviceCount(&have) != cudaSuccess || have <= 0) { cudaGetLastError(); return SPMM_B200_ENODEV; }
	spmm_b200_multi* mh = new (std::nothrow) spmm_b200_multi();
	if (!mh) return SPMM_B200_ENOMEM;
	mh->n = ngpus; mh->k = k; mh->b_layout = b_layout; mh->m = m; mh->c = c; mh->nnz = nnz;
	for (int g = 0; g < ngpus; g++) {
		const int d = devices ? devices[g] : g;
		if (d < 0 || d >= have || std::find(mh->dev.begin(), mh->dev.end(), d) != mh->dev.end()) { delete mh; return SPMM_B200_EINVAL; }
		mh->dev.push_back(d);
	}
	// shard boundaries: the reference's nnz-balanced split with P = ngpus (same integers as spmm_b200_partition)
	mh->bounds.resize(ngpus + 1);
	rc = spmm_b200_partition_host(row_ptr, m, nnz, ngpus, SPMM_B200_PART_NNZ, mh->bounds.data());
	if (rc) { delete mh; return rc; }
	for (int g = 0; g < ngpus; g++) mh->bounds[g + 1] = std::max(mh->bounds[g + 1], mh->bounds[g]);
	// peer access: every device reads every other device's staging buffer in the gather kernel
	mh->peer = ngpus > 1;
	for (int g = 0; g < ngpus && mh->peer; g++)
		for (int p = 0; p < ngpus && mh->peer; p++) {
			if (p == g) continue;
			int can = 0;
			if (cudaDeviceCanAccessPeer(&can, mh->dev[g], mh->dev[p]) != cudaSuccess || !can) { cudaGetLastError(); mh->peer = false; }
		}
	for (int g = 0; g < ngpus && mh->peer; g++) {
		if (cudaSetDevice(mh->dev[g]) != cudaSuccess) { cudaGetLastError(); mh->peer = false; break; }
		for (int p = 0; p < ngpus; p++) {
			if (p == g) continue;
			cudaError_t e = cudaDeviceEnablePeerAccess(mh->dev[p], 0);
			if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
			else if (e != cudaSuccess) { cudaGetLastError(); mh->peer = false; break; }
		}
	}
	mh->sh.assign(ngpus, nullptr);
	mh->ev_part.resize(ngpus);
	*out = mh;
	for (int g = 0; g < ngpus; g++) {
		int sms = 0;
		rc = check_device(mh->dev[g], &sms);
		if (rc) break;
		const int64_t r0 = mh->bounds[g], r1 = mh->bounds[g + 1];
		const int32_t base = row_ptr[r0];
		rc = create_from_slice(&mh->sh[g], row_ptr + r0, col_ind, values, r1 - r0, c, (int64_t) row_ptr[r1] - base, base, k,
		                       mh->dev[g], variant, b_layout, sms, r0, m);
		if (rc) break;
		mh->sh[g]->rank = g; mh->sh[g]->nranks = ngpus;
	}
	if (rc) { spmm_b200_destroy_multi(mh); *out = nullptr; }
	return rc;
}

extern "C" int spmm_b200_register_host_multi(spmm_b200_multi_t* mh, const void* p, size_t bytes)
{
	if (!mh || !p || bytes == 0) return SPMM_B200_EINVAL;
	if (std::find(mh->pinned.begin(), mh->pinned.end(), (void*) p) != mh->pinned.end()) return SPMM_B200_OK;
	cudaPointerAttributes at;
	if (cudaPointerGetAttributes(&at, p) == cudaSuccess && at.type != cudaMemoryTypeUnregistered) return SPMM_B200_OK;
	cudaGetLastError();
	CU_TRY(cudaSetDevice(mh->dev[0]));
	CU_TRY(cudaHostRegister((void*) p, bytes, cudaHostRegisterPortable | cudaHostRegisterMapped));
	mh->pinned.push_back((void*) p);
	return SPMM_B200_OK;
}

// Enqueue upload + exchange + transpose of column slice s (ks columns, of S) on every device's upload stream.
static int multi_replicate_slice(spmm_b200_multi* mh, const double* x_host, int k, int ks, int s)
{
	const int N = mh->n;
	const size_t bslice = (size_t) mh->c * ks;
	const bool split = mh->peer && mh->b_layout == SPMM_B200_B_COLMAJOR && N > 1 && bslice >= (size_t) N;
	const size_t per = split ? (bslice + N - 1) / N : bslice;
	for (int g = 0; g < N; g++) {
		spmm_b200* h = mh->sh[g];
		CU_TRY(cudaSetDevice(h->device));
		if (mh->b_layout == SPMM_B200_B_COLMAJOR) {
			const size_t lo = split ? std::min(bslice, per * (size_t) g) : 0, hi = split ? std::min(bslice, lo + per) : bslice;
			if (hi > lo)
				CU_TRY(cudaMemcpyAsync(h->d_x + s * bslice + lo, x_host + s * bslice + lo, (hi - lo) * sizeof(double), cudaMemcpyHostToDevice, h->s_up));
		} else {
			CU_TRY(cudaMemcpy2DAsync(h->d_B + s * bslice, sizeof(double) * ks, x_host + (size_t) s * ks, sizeof(double) * k, sizeof(double) * ks,
			                         (size_t) mh->c, cudaMemcpyHostToDevice, h->s_up));
		}
		CU_TRY(cudaEventRecord(mh->ev_part[g][s], h->s_up));
	}
	for (int g = 0; g < N; g++) {
		spmm_b200* h = mh->sh[g];
		CU_TRY(cudaSetDevice(h->device));
		if (mh->b_layout == SPMM_B200_B_COLMAJOR && mh->c > 0) {
			PeerParts parts;
			parts.per = (int64_t) std::max<size_t>(per, 1);
			for (int p = 0; p < MULTI_MAX; p++) parts.base[p] = mh->sh[split ? std::min(p, N - 1) : g]->d_x + s * bslice;
			if (split)
				for (int p = 0; p < N; p++)
					if (p != g) CU_TRY(cudaStreamWaitEvent(h->s_up, mh->ev_part[p][s], 0));
			dim3 grid((unsigned) ((mh->c + 31) / 32), (unsigned) ((ks + 31) / 32)), block(32, 8);
			gather_transpose_b_kernel<<<grid, block, 0, h->s_up>>>(parts, h->d_B + s * bslice, mh->c, ks);
			h->launches++;
			CU_TRY(cudaGetLastError());
		}
		CU_TRY(cudaEventRecord(h->ev_slice[3 * s], h->s_up));
	}
	return SPMM_B200_OK;
}

static int multi_prepare(spmm_b200_multi* mh, int k, int S)
{
	for (int g = 0; g < mh->n; g++) {
		spmm_b200* h = mh->sh[g];
		CU_TRY(cudaSetDevice(h->device));
		const size_t nb = (size_t) h->c * k;
		int rc = ensure(&h->d_B, &h->cap_B, nb);
		if (rc) return rc;
		if (mh->b_layout == SPMM_B200_B_COLMAJOR) {
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
		while (mh->ev_part[g].size() < (size_t) S) {
			cudaEvent_t e;
			CU_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
			mh->ev_part[g].push_back(e);
		}
		h->resident_k = 0;
		h->tl_slices = 0;
	}
	mh->resident_k = 0;
	return SPMM_B200_OK;
}

// A staging buffer (d_x) may only be refilled once every peer has finished reading the previous contents: the
// run ends with all streams of all devices synchronised, so consecutive calls are ordered by the host.
static int multi_sync(spmm_b200_multi* mh)
{
	for (int g = 0; g < mh->n; g++) {
		spmm_b200* h = mh->sh[g];
		CU_TRY(cudaSetDevice(h->device));
		if (h->s_up) CU_TRY(cudaStreamSynchronize(h->s_up));
		CU_TRY(cudaStreamSynchronize(h->stream));
		if (h->s_down) CU_TRY(cudaStreamSynchronize(h->s_down));
	}
	return SPMM_B200_OK;
}

// ------------------------------------------------------------------------------------------
// Download routing.  The PCIe links of an 8-GPU box do not end in equal shares of the host: on the pool's boxes four
// of the eight devices sit behind ONE ~75 GB/s path while the other four sustain 44 GB/s each, and with all eight
// writing at once the box moves 125 GB/s -- less than the good four alone (177 GB/s; profiles/r2_multi_gpu.md, section
// 6).  C is row-local on each device but any device can carry it to the host: a device that does not download stores
// its rows of C straight into a relay buffer on a partner (the SpMM kernel's C pointer is peer memory: NVLink stores),
// and the partner's copy engine moves them into the caller's y.  Which devices download is MEASURED once per handle
// (a few 16 MB copies per candidate set: all, either half, evens, odds); a subset is used only when it beats "all" by
// 20 %.  MULTI_ROUTE=0 turns it off, a bit mask forces the set.
// ------------------------------------------------------------------------------------------

static int multi_probe_set(spmm_b200_multi* mh, unsigned mask, double* const* dbuf, char* hbuf, size_t bytes, double* gbs)
{
	for (int g = 0; g < mh->n; g++) { CU_TRY(cudaSetDevice(mh->dev[g])); CU_TRY(cudaStreamSynchronize(mh->sh[g]->stream)); }
	const auto t0 = std::chrono::steady_clock::now();
	int active = 0;
	for (int rep = 0; rep < 2; rep++)
		for (int g = 0; g < mh->n; g++) {
			if (!(mask >> g & 1u)) continue;
			CU_TRY(cudaSetDevice(mh->dev[g]));
			CU_TRY(cudaMemcpyAsync(hbuf + (size_t) g * bytes, dbuf[g], bytes, cudaMemcpyDeviceToHost, mh->sh[g]->stream));
			active += rep == 0;
		}
	for (int g = 0; g < mh->n; g++) {
		if (!(mask >> g & 1u)) continue;
		CU_TRY(cudaSetDevice(mh->dev[g]));
		CU_TRY(cudaStreamSynchronize(mh->sh[g]->stream));
	}
	const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
	*gbs = 2.0 * active * (double) bytes / sec / 1e9;
	return SPMM_B200_OK;
}

static void multi_routes_from_mask(spmm_b200_multi* mh, unsigned mask)
{
	std::vector<int> down;
	for (int g = 0; g < mh->n; g++) if (mask >> g & 1u) down.push_back(g);
	mh->route.assign(mh->n, 0);
	int next = 0;
	for (int g = 0; g < mh->n; g++) {
		if (down.empty() || (mask >> g & 1u)) { mh->route[g] = g; continue; }
		mh->route[g] = down[next++ % down.size()];
	}
}

static int multi_choose_routes(spmm_b200_multi* mh)
{
	const int want = mh->sh[0]->opt.multi_route;
	if (mh->route_mask == want && (int) mh->route.size() == mh->n) return SPMM_B200_OK;
	const unsigned all = mh->n >= 32 ? ~0u : (1u << mh->n) - 1u;
	mh->route_mask = want;
	mh->route_probe.clear();
	multi_routes_from_mask(mh, all);
	if (want == 0 || !mh->peer) return SPMM_B200_OK;
	if (want > 0) { if (((unsigned) want & all) != 0) multi_routes_from_mask(mh, (unsigned) want & all); return SPMM_B200_OK; }
	if (mh->n < 4) return SPMM_B200_OK;          // two links have never been seen to hurt each other
	// measure
	const size_t bytes = (size_t) 16 << 20;
	std::vector<double*> dbuf(mh->n, nullptr);
	char* hbuf = nullptr;
	int rc = SPMM_B200_OK;
	if (cudaHostAlloc((void**) &hbuf, bytes * mh->n, cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); return SPMM_B200_OK; }
	for (int g = 0; g < mh->n && !rc; g++) {
		if (cudaSetDevice(mh->dev[g]) != cudaSuccess || cudaMalloc((void**) &dbuf[g], bytes) != cudaSuccess) { cudaGetLastError(); rc = SPMM_B200_ENOMEM; }
	}
	unsigned lower = 0, upper = 0, even = 0, odd = 0;
	for (int g = 0; g < mh->n; g++) {
		(g < mh->n / 2 ? lower : upper) |= 1u << g;
		(g % 2 == 0 ? even : odd) |= 1u << g;
	}
	const unsigned cand[5] = {all, lower, upper, even, odd};
	double rate[5] = {0, 0, 0, 0, 0}, warm = 0;
	if (!rc) rc = multi_probe_set(mh, all, dbuf.data(), hbuf, bytes, &warm);
	for (int i = 0; i < 5 && !rc; i++) rc = multi_probe_set(mh, cand[i], dbuf.data(), hbuf, bytes, &rate[i]);
	for (int g = 0; g < mh->n; g++) if (dbuf[g]) { cudaSetDevice(mh->dev[g]); cudaFree(dbuf[g]); }
	cudaFreeHost(hbuf);
	if (rc) return rc == SPMM_B200_ENOMEM ? SPMM_B200_OK : rc;      // no room for the probe: every device downloads its own rows
	mh->route_probe.assign(rate, rate + 5);
	int best = 0;
	for (int i = 1; i < 5; i++) if (rate[i] > rate[best]) best = i;
	if (best != 0 && rate[best] > 1.2 * rate[0]) multi_routes_from_mask(mh, cand[best]);
	return SPMM_B200_OK;
}

// out[0..n): route[g]; then the measured GB/s of the candidate sets {all, lower half, upper half, evens, odds} (0 when
// nothing was measured).  Returns the number of doubles written.
extern "C" int spmm_b200_multi_routes(const spmm_b200_multi_t* mh, double* out, int max_values)
{
	if (!mh || !out || max_values < 1) return SPMM_B200_EINVAL;
	int n = 0;
	for (int g = 0; g < mh->n && n < max_values; g++) out[n++] = g < (int) mh->route.size() ? mh->route[g] : g;
	for (size_t i = 0; i < 5 && n < max_values; i++) out[n++] = i < mh->route_probe.size() ? mh->route_probe[i] : 0.0;
	return n;
}

// Own windows + routed downloads: per device one upload of its window of B, one kernel on its shard whose C pointer is
// its own d_C (downloaders) or the partner's relay buffer (peer stores), then the downloaders' copy engines carry first
// their own rows and then their clients' rows into y.
static int run_multi_routed(spmm_b200_multi* mh, const double* x_host, double* y_host, int k)
{
	int rc = multi_prepare(mh, k, 1);
	if (rc) return rc;
	mh->relay.resize(mh->n, nullptr);
	mh->relay_cap.resize(mh->n, 0);
	std::vector<size_t> off(mh->n, 0);
	for (int p = 0; p < mh->n; p++) {
		size_t need = 0;
		for (int g = 0; g < mh->n; g++)
			if (mh->route[g] == p && g != p) { off[g] = need; need += (size_t) mh->sh[g]->m * k; }
		if (need > mh->relay_cap[p]) {
			CU_TRY(cudaSetDevice(mh->dev[p]));
			if (mh->relay[p]) cudaFree(mh->relay[p]);
			mh->relay[p] = nullptr; mh->relay_cap[p] = 0;
			CU_TRY(cudaMalloc((void**) &mh->relay[p], need * sizeof(double)));
			mh->relay_cap[p] = need;
		}
	}
	for (int g = 0; g < mh->n; g++) {
		spmm_b200* h = mh->sh[g];
		CU_TRY(cudaSetDevice(h->device));
		h->last_h2d_bytes = h->last_d2h_bytes = 0;
		h->band_blocks = 0;
		int64_t lo, n;
		b_window(h, lo, n);
		rc = upload_b_rows(h, x_host, k, 0, k, lo, n, h->d_B, h->s_up);
		if (rc) return rc;
		CU_TRY(cudaEventRecord(h->ev_slice[0], h->s_up));
		CU_TRY(cudaStreamWaitEvent(h->stream, h->ev_slice[0], 0));
		CU_TRY(cudaEventRecord(h->ev_slice[1], h->stream));
		const int p = mh->route[g];
		if (h->m > 0) {
			const int nl = launch_spmm(h, h->d_B, p == g ? h->d_C : mh->relay[p] + off[g], k, h->stream);
			if (nl < 0) return nl;
		}
		CU_TRY(cudaEventRecord(h->ev_slice[2], h->stream));
		h->c_slices = p == g ? 1 : 0;
		h->timed = true; h->timed_slices = 1;
	}
	for (int p = 0; p < mh->n; p++) {
		if (mh->route[p] != p) continue;
		spmm_b200* hp = mh->sh[p];
		CU_TRY(cudaSetDevice(hp->device));
		CU_TRY(cudaStreamWaitEvent(hp->s_down, hp->ev_slice[2], 0));
		if (hp->m > 0) {
			CU_TRY(cudaMemcpyAsync(y_host + (size_t) hp->first_row * k, hp->d_C, sizeof(double) * hp->m * k, cudaMemcpyDeviceToHost, hp->s_down));
			hp->last_d2h_bytes += (int64_t) sizeof(double) * hp->m * k;
		}
		for (int g = 0; g < mh->n; g++) {
			if (mh->route[g] != p || g == p || mh->sh[g]->m == 0) continue;
			CU_TRY(cudaStreamWaitEvent(hp->s_down, mh->sh[g]->ev_slice[2], 0));
			CU_TRY(cudaMemcpyAsync(y_host + (size_t) mh->sh[g]->first_row * k, mh->relay[p] + off[g], sizeof(double) * mh->sh[g]->m * k,
			                       cudaMemcpyDeviceToHost, hp->s_down));
			hp->last_d2h_bytes += (int64_t) sizeof(double) * mh->sh[g]->m * k;
		}
	}
	return multi_sync(mh);
}

extern "C" int spmm_b200_run_multi(spmm_b200_multi_t* mh, const double* x_host, double* y_host, int k, int copy_back)
{
	if (!mh || !x_host || k < 1 || (copy_back && !y_host)) return SPMM_B200_EINVAL;
	int ks = slice_width(mh->sh[0], k);
	// Own windows: a shard reads only the rows of B between its first and last touched column.  When no shard's window
	// is larger than ~the 1/N share of B it would upload for the peer gather, every device uploads its own window over
	// its own link and runs its own host pipeline (row blocks or column slices) -- no replication step at all.
	{
		int64_t worst = 0;
		for (int g = 0; g < mh->n; g++) {
			int64_t lo, n;
			b_window(mh->sh[g], lo, n);
			worst = std::max(worst, n);
		}
		// (a diagonal band gives each shard c/N rows of B plus the band: the same bytes as its share of a gather, minus the gather)
		if (copy_back && mh->sh[0]->opt.upload_window && worst * mh->n * 4 <= mh->c * 5) {
			int rc = multi_choose_routes(mh);
			if (rc) return rc;
			bool routed = false;
			for (int g = 0; g < mh->n; g++) routed = routed || mh->route[g] != g;
			if (routed) return run_multi_routed(mh, x_host, y_host, k);
			rc = multi_prepare(mh, k, 1);
			if (rc) return rc;
			std::vector<BandStep> steps;
			for (int g = 0; g < mh->n; g++) {
				spmm_b200* h = mh->sh[g];
				CU_TRY(cudaSetDevice(h->device));
				h->last_h2d_bytes = h->last_d2h_bytes = 0;
				h->band_blocks = 0;
				if (h->m == 0) continue;
				double* y_rows = y_host + (size_t) h->first_row * k;
				if (rowband_blocks(h, k, ks, x_host, y_rows, steps) > 0) rc = run_rowband(h, x_host, y_rows, k, steps, false);
				else rc = run_pipelined(h, x_host, y_rows, k, ks > 0 ? ks : k, nullptr, nullptr, 1, false);
				if (rc) return rc;
			}
			return multi_sync(mh);
		}
	}
	if (ks <= 0) ks = k;
	const int S = k / ks;
	int rc = multi_prepare(mh, k, S);
	if (rc) return rc;
	for (int s = 0; s < S; s++) {
		rc = multi_replicate_slice(mh, x_host, k, ks, s);
		if (rc) return rc;
	}
	const size_t bslice = (size_t) mh->c * ks;
	// y_host mapped for the devices (page-locked + portable): the sliced row-split kernels store C straight into it
	double* y_map = nullptr;
	if (copy_back && S > 1 && mh->sh[0]->opt.direct_store) {
		cudaPointerAttributes at;
		if (cudaPointerGetAttributes(&at, y_host) == cudaSuccess && at.type == cudaMemoryTypeHost && at.devicePointer)
			y_map = (double*) at.devicePointer;
		else
			cudaGetLastError();
		if (((uintptr_t) y_map & 15) != 0 || k % 2) y_map = nullptr;
	}
	for (int s = 0; s < S; s++) {
		for (int g = 0; g < mh->n; g++) {
			spmm_b200* h = mh->sh[g];
			CU_TRY(cudaSetDevice(h->device));
			CU_TRY(cudaStreamWaitEvent(h->stream, h->ev_slice[3 * s], 0));
			if (h->m == 0) continue;
			const size_t cslice = (size_t) h->m * ks;
			double* y_rows = copy_back ? y_host + (size_t) h->first_row * k : nullptr;
			const bool direct = y_map && h->variant == SPMM_B200_ROWSPLIT && (ks == 8 || ks == 16 || ks == 32 || ks == 64);
			if (direct) {
				CU_TRY(cudaEventRecord(h->ev_slice[3 * s + 1], h->stream));
				launch_rowsplit_width(h, h->d_B + s * bslice, y_map + (size_t) h->first_row * k + (size_t) s * ks, ks, h->stream, ks, k);
				CU_TRY(cudaGetLastError());
				h->launches += 1;
				CU_TRY(cudaEventRecord(h->ev_slice[3 * s + 2], h->stream));
				h->c_slices = 0;
				continue;
			}
			rc = run_between(h, h->d_B + s * bslice, h->d_C + s * cslice, ks, h->stream, h->ev_slice[3 * s + 1], h->ev_slice[3 * s + 2]);
			if (rc) return rc;
			h->c_slices = S;
			if (y_rows) {
				CU_TRY(cudaStreamWaitEvent(h->s_down, h->ev_slice[3 * s + 2], 0));
				if (S == 1)
					CU_TRY(cudaMemcpyAsync(y_rows, h->d_C, sizeof(double) * h->m * k, cudaMemcpyDeviceToHost, h->s_down));
				else
					CU_TRY(cudaMemcpy2DAsync(y_rows + (size_t) s * ks, sizeof(double) * k, h->d_C + s * cslice, sizeof(double) * ks,
					                         sizeof(double) * ks, (size_t) h->m, cudaMemcpyDeviceToHost, h->s_down));
			}
		}
	}
	for (int g = 0; g < mh->n; g++) { mh->sh[g]->timed = true; mh->sh[g]->timed_slices = mh->sh[g]->m > 0 ? S : 0; }
	rc = multi_sync(mh);
	if (rc) return rc;
	if (S == 1) {
		for (int g = 0; g < mh->n; g++) mh->sh[g]->resident_k = k;
		mh->resident_k = k;
	}
	return SPMM_B200_OK;
}

extern "C" int spmm_b200_set_b_multi(spmm_b200_multi_t* mh, const double* x_host, int k)
{
	if (!mh || !x_host || k < 1) return SPMM_B200_EINVAL;
	int rc = multi_prepare(mh, k, 1);
	if (rc) return rc;
	rc = multi_replicate_slice(mh, x_host, k, k, 0);
	if (rc) return rc;
	rc = multi_sync(mh);
	if (rc) return rc;
	for (int g = 0; g < mh->n; g++) mh->sh[g]->resident_k = k;
	mh->resident_k = k;
	return SPMM_B200_OK;
}

extern "C" int spmm_b200_run_resident_multi(spmm_b200_multi_t* mh, int k)
{
	if (!mh || k < 1 || mh->resident_k != k) return SPMM_B200_EINVAL;
	for (int g = 0; g < mh->n; g++) {
		spmm_b200* h = mh->sh[g];
		h->c_slices = 1;
		int rc = spmm_b200_run_device(h, h->d_B, h->d_C, k, h->stream);     // asynchronous: all devices run side by side
		if (rc) return rc;
	}
	for (int g = 0; g < mh->n; g++) {
		CU_TRY(cudaSetDevice(mh->sh[g]->device));
		CU_TRY(cudaStreamSynchronize(mh->sh[g]->stream));
	}
	return SPMM_B200_OK;
}

extern "C" int spmm_b200_fetch_c_multi(spmm_b200_multi_t* mh, double* y_host, int k)
{
	if (!mh || !y_host || k < 1) return SPMM_B200_EINVAL;
	for (int g = 0; g < mh->n; g++) {
		spmm_b200* h = mh->sh[g];
		if (h->cap_C < (size_t) h->m * k || h->c_slices == 0 || (h->c_slices > 1 && k % h->c_slices)) return SPMM_B200_EINVAL;
		CU_TRY(cudaSetDevice(h->device));
		int rc = download_c(h, y_host + (size_t) h->first_row * k, k, h->stream);
		if (rc) return rc;
	}
	for (int g = 0; g < mh->n; g++) {
		CU_TRY(cudaSetDevice(mh->sh[g]->device));
		CU_TRY(cudaStreamSynchronize(mh->sh[g]->stream));
	}
	return SPMM_B200_OK;
}

extern "C" int spmm_b200_multi_ngpus(const spmm_b200_multi_t* mh) { return mh ? mh->n : SPMM_B200_EINVAL; }
extern "C" int spmm_b200_multi_peer_access(const spmm_b200_multi_t* mh) { return mh ? (mh->peer ? 1 : 0) : SPMM_B200_EINVAL; }
extern "C" spmm_b200_t* spmm_b200_multi_shard(spmm_b200_multi_t* mh, int g) { return (mh && g >= 0 && g < mh->n) ? mh->sh[g] : nullptr; }
extern "C" int64_t spmm_b200_multi_first_row(const spmm_b200_multi_t* mh, int g) { return (mh && g >= 0 && g <= mh->n) ? mh->bounds[g] : -1; }

extern "C" double spmm_b200_multi_last_kernel_ms(const spmm_b200_multi_t* mh)
{
	if (!mh) return -1.0;
	double worst = -1.0;
	for (int g = 0; g < mh->n; g++) {
		cudaSetDevice(mh->sh[g]->device);
		worst = std::max(worst, spmm_b200_last_kernel_ms(mh->sh[g]));
	}
	return worst;
}

extern "C" int64_t spmm_b200_multi_launch_count(const spmm_b200_multi_t* mh)
{
	if (!mh) return -1;
	int64_t n = 0;
	for (int g = 0; g < mh->n; g++) n += mh->sh[g]->launches;
	return n;
}
