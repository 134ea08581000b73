// matrix_io.cpp -- on-disk matrix formats feeding the SpMM path, host side of the C ABI.
//
//   spmm_b200_read_mtx   Matrix Market coordinate files -> COO (R, C, V as double), replacing
//                        mtx_read(file, expand_symmetry=1, pattern_dummy_vals=1) plus the value
//                        conversion the harness applies after it
//                        (ref: lib/storage_formats/matrix_market/matrix_market.c:153-322,
//                         matrix_market_gen.c:66-160, spmv_bench.cpp:723-763).
//   spmm_b200_read_smtx  DLMC ".smtx" files (CSR as text: "m, k, nnz" / row offsets / column indices)
//                        -> CSR, replacing smtx_read (ref: lib/storage_formats/dlcm_matrices/
//                        dlcm_matrix.c:150-265, dlcm_matrix_gen.c:54-118).  The reference fills the values
//                        with rand() seeded from time(NULL); here they come from a caller-supplied seed.
//
// Parsing is plain byte work on the host (strtol / strtod like the reference's gen_strtonum,
// lib/genlib.h:432-455), so that indices AND values are bit-identical with the reference's arrays.
// The arrays go to the device through spmm_b200_coo_to_csr / spmm_b200_create.
#include "../../include/spmm_b200.h"

#include <errno.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <vector>

namespace {

struct FileText {
	std::vector<char> buf;              // whole file + '\0'
	std::vector<char*> lines;           // start of every line ('\n' replaced by '\0')
};

// Same line model as the reference's file_to_lines(): split on '\n'; a trailing fragment without
// '\n' is a line; empty lines are kept out (the reference's atoms never hold them for these files).
int read_lines(const char* path, FileText& t)
{
	FILE* f = fopen(path, "rb");
	if (!f) return SPMM_B200_EIO;
	if (fseek(f, 0, SEEK_END) != 0) { fclose(f); return SPMM_B200_EIO; }
	long n = ftell(f);
	if (n < 0) { fclose(f); return SPMM_B200_EIO; }
	rewind(f);
	t.buf.resize((size_t) n + 1);
	if (n > 0 && fread(t.buf.data(), 1, (size_t) n, f) != (size_t) n) { fclose(f); return SPMM_B200_EIO; }
	fclose(f);
	t.buf[(size_t) n] = '\0';
	char* p = t.buf.data();
	char* end = p + n;
	while (p < end) {
		char* nl = (char*) memchr(p, '\n', (size_t) (end - p));
		char* stop = nl ? nl : end;
		if (nl) *nl = '\0';
		if (stop > p && stop[-1] == '\r') stop[-1] = '\0';
		if (*p != '\0') t.lines.push_back(p);
		p = stop + 1;
	}
	return SPMM_B200_OK;
}

template <typename T> T* alloc_out(size_t n)
{
	return (T*) aligned_alloc(64, ((std::max<size_t>(n, 1) * sizeof(T) + 63) / 64) * 64);
}

struct Owned {                          // frees on early return
	void* p[3] = {nullptr, nullptr, nullptr};
	~Owned() { for (void* q : p) free(q); }
	void release() { p[0] = p[1] = p[2] = nullptr; }
};

} // namespace

extern "C" void spmm_b200_free(void* p) { free(p); }

extern "C" int spmm_b200_read_mtx(const char* path, int expand_symmetry, int64_t* m_out, int64_t* n_out, int64_t* nnz_out,
                                  int32_t** R_out, int32_t** C_out, double** V_out, char* field_out16)
{
	if (!path || !m_out || !n_out || !nnz_out || !R_out || !C_out || !V_out) return SPMM_B200_EINVAL;
	*R_out = nullptr; *C_out = nullptr; *V_out = nullptr;
	FileText t;
	int rc = read_lines(path, t);
	if (rc) return rc;
	if (t.lines.empty()) return SPMM_B200_EFORMAT;

	// ---- header (matrix_market.c:153-215): a file that does not start with %%MatrixMarket is taken as
	// "coordinate real general" with the size line first
	std::string format = "coordinate", field = "real";
	bool symmetric = false, skew = false;
	size_t li = 0;
	{
		char tag[64] = "", object[64] = "", fmt[64] = "", fld[64] = "", sym[64] = "";
		int got = sscanf(t.lines[0], "%63s %63s %63s %63s %63s", tag, object, fmt, fld, sym);
		if (got >= 1 && !strcmp(tag, "%%MatrixMarket")) {
			if (got < 5) return SPMM_B200_EFORMAT;
			if (strcmp(object, "matrix") || (strcmp(fmt, "coordinate") && strcmp(fmt, "array"))) return SPMM_B200_EFORMAT;
			format = fmt; field = fld;
			if (!strcmp(sym, "symmetric") || !strcmp(sym, "Hermitian")) symmetric = true;
			else if (!strcmp(sym, "skew-symmetric")) skew = true;
			else if (strcmp(sym, "general")) return SPMM_B200_EFORMAT;
			li = 1;
		}
	}
	while (li < t.lines.size() && t.lines[li][0] == '%') li++;
	if (li >= t.lines.size()) return SPMM_B200_EFORMAT;
	// "array" files: the reference's parser reads N values from each of the first M lines while its header
	// check demands M*N lines (matrix_market.c:204-212, matrix_market_gen.c:33-60) -- it only works for N = 1
	// and returns no indices (R = C = NULL), which the harness cannot convert.  Not supported.
	if (format != "coordinate") return SPMM_B200_ENOSUP;
	const bool is_real = field == "real", is_int = field == "integer", is_cplx = field == "complex", is_pat = field == "pattern";
	if (!is_real && !is_int && !is_cplx && !is_pat) return SPMM_B200_EFORMAT;

	long M = 0, N = 0, entries = 0;
	if (sscanf(t.lines[li++], "%ld%ld%ld", &M, &N, &entries) != 3) return SPMM_B200_EFORMAT;
	if (M < 0 || N < 0 || entries < 0) return SPMM_B200_EFORMAT;
	if ((size_t) entries != t.lines.size() - li) return SPMM_B200_EFORMAT;      // matrix_market.c:213-214
	if (M >= INT32_MAX || N >= INT32_MAX || 2 * (int64_t) entries >= INT32_MAX) return SPMM_B200_ERANGE;

	const bool sym = symmetric || skew;
	const size_t cap = sym ? 2 * (size_t) entries : (size_t) entries;
	Owned own;
	int32_t* R = alloc_out<int32_t>(cap); own.p[0] = R;
	int32_t* C = alloc_out<int32_t>(cap); own.p[1] = C;
	double* V = alloc_out<double>(cap);   own.p[2] = V;
	if (!R || !C || !V) return SPMM_B200_ENOMEM;

	// ---- entries (matrix_market_gen.c:88-113); values converted like spmv_bench.cpp:730-763:
	// integer -> (double) int, complex -> |z|, pattern -> 1.0
	std::vector<double> im;                       // imaginary parts, only to mirror Hermitian/skew entries
	if (is_cplx) im.resize(cap);
	size_t non_diag = 0;
	for (long i = 0; i < entries; i++) {
		char* p = t.lines[li + i];
		char* e = nullptr;
		errno = 0;
		long r = strtol(p, &e, 10);
		if (e == p) return SPMM_B200_EFORMAT;
		p = e;
		long c = strtol(p, &e, 10);
		if (e == p) return SPMM_B200_EFORMAT;
		p = e;
		if (r < 1 || r > M || c < 1 || c > N) return SPMM_B200_EFORMAT;
		R[i] = (int32_t) (r - 1);                  // base-1 -> base-0
		C[i] = (int32_t) (c - 1);
		if (is_real) {
			V[i] = strtod(p, &e);
			if (e == p) return SPMM_B200_EFORMAT;
		} else if (is_int) {
			long v = strtol(p, &e, 10);
			if (e == p) return SPMM_B200_EFORMAT;
			V[i] = (double) (int) v;
		} else if (is_cplx) {
			double re = strtod(p, &e);
			if (e == p) return SPMM_B200_EFORMAT;
			p = e;
			double imv = strtod(p, &e);
			if (e == p) return SPMM_B200_EFORMAT;
			V[i] = re; im[i] = imv;
		} else {
			V[i] = 1.0;
		}
		if (R[i] != C[i]) non_diag++;
	}
	size_t nnz = (size_t) entries;
	if (expand_symmetry && sym) {
		// mirrored entries follow the file's entries, in file order (matrix_market_gen.c:116-156: per-thread
		// offsets are a prefix sum over contiguous chunks, i.e. file order whatever the thread count)
		size_t j = (size_t) entries;
		for (long i = 0; i < entries; i++) {
			if (R[i] == C[i]) continue;
			R[j] = C[i];
			C[j] = R[i];
			if (is_cplx) { V[j] = skew ? -V[i] : V[i]; im[j] = skew ? im[i] : -im[i]; }   // -conj(z) : conj(z)
			else V[j] = skew ? -V[i] : V[i];
			j++;
		}
		nnz = (size_t) entries + non_diag;
	}
	if (is_cplx)
		for (size_t i = 0; i < nnz; i++) V[i] = hypot(V[i], im[i]);      // cabs()
	if (field_out16) snprintf(field_out16, 16, "%s", field.c_str());
	*m_out = M; *n_out = N; *nnz_out = (int64_t) nnz;
	*R_out = R; *C_out = C; *V_out = V;
	own.release();
	return SPMM_B200_OK;
}

// splitmix64 -> uniform double in [-1, 1): the seeded stand-in for the reference's time-seeded rand()
static inline double smtx_value(uint64_t seed, uint64_t j)
{
	uint64_t z = seed + 0x9e3779b97f4a7c15ull * (j + 1);
	z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
	z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
	z ^= z >> 31;
	return (double) (z >> 11) * (2.0 / 9007199254740992.0) - 1.0;
}

extern "C" int spmm_b200_read_smtx(const char* path, uint64_t seed, int64_t* m_out, int64_t* n_out, int64_t* nnz_out,
                                   int32_t** row_ptr_out, int32_t** col_ind_out, double** values_out)
{
	if (!path || !m_out || !n_out || !nnz_out || !row_ptr_out || !col_ind_out || !values_out) return SPMM_B200_EINVAL;
	*row_ptr_out = nullptr; *col_ind_out = nullptr; *values_out = nullptr;
	FileText t;
	int rc = read_lines(path, t);
	if (rc) return rc;
	long M = 0, K = 0, nnz = 0;
	if (t.lines.empty() || sscanf(t.lines[0], "%ld,%ld,%ld", &M, &K, &nnz) != 3) return SPMM_B200_EFORMAT;   // dlcm_matrix.c:219-221
	if (M < 0 || K < 0 || nnz < 0) return SPMM_B200_EFORMAT;
	if (M >= INT32_MAX || K >= INT32_MAX || nnz >= INT32_MAX) return SPMM_B200_ERANGE;
	if (t.lines.size() < (nnz > 0 ? 3u : 2u)) return SPMM_B200_EFORMAT;
	Owned own;
	int32_t* ia = alloc_out<int32_t>((size_t) M + 1); own.p[0] = ia;
	int32_t* ja = alloc_out<int32_t>((size_t) nnz);   own.p[1] = ja;
	double* a = alloc_out<double>((size_t) nnz);      own.p[2] = a;
	if (!ia || !ja || !a) return SPMM_B200_ENOMEM;
	char* p = t.lines[1];                          // line 1: m+1 row offsets (dlcm_matrix_gen.c:77-86)
	for (long j = 0; j <= M; j++) {
		char* e = nullptr;
		long v = strtol(p, &e, 10);
		if (e == p) return SPMM_B200_EFORMAT;
		ia[j] = (int32_t) v;
		p = e;
	}
	if (ia[0] != 0 || ia[M] != nnz) return SPMM_B200_EFORMAT;
	for (long j = 0; j < M; j++)
		if (ia[j + 1] < ia[j]) return SPMM_B200_EFORMAT;
	if (nnz > 0) {
		p = t.lines[2];                            // line 2: nnz column indices (dlcm_matrix_gen.c:92-101)
		for (long j = 0; j < nnz; j++) {
			char* e = nullptr;
			long v = strtol(p, &e, 10);
			if (e == p || v < 0 || v >= K) return SPMM_B200_EFORMAT;
			ja[j] = (int32_t) v;
			p = e;
		}
	}
	for (long j = 0; j < nnz; j++) a[j] = smtx_value(seed, (uint64_t) j);
	*m_out = M; *n_out = K; *nnz_out = nnz;
	*row_ptr_out = ia; *col_ind_out = ja; *values_out = a;
	own.release();
	return SPMM_B200_OK;
}
