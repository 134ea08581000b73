// sddmm_kernel_b200.cpp -- the reference-side plug-in translation unit of the sparse-attention pipeline.
//
// Implements the pipeline bench's link-time kernel interface
//   struct Matrix_Format { virtual void spmm(type, m, k, n, ia, ja, a, x, y, threads); virtual void sddmm(y, threads);
//                          statistics_start(); statistics_print_data(); }
//   struct Matrix_Format * csr_to_format(row_ptr, col_ind, values, m, nnz, n);
//   int statistics_print_labels(buf, buf_n);
// (ref: benchmark_code/CPU/AMD/pipeline_code_bench/sddmm_kernel.h:9-33) on top of the C ABI in
// include/spmm_b200_pipeline.h.  Compiled against the reference's own headers and linked under the UNCHANGED
// pipeline harness objects exactly like sddmm_taco_naive.cpp (Makefile_in:166); see oracle/Makefile target
// `pipeline_harness` and INTEGRATION.md.  "_d" builds use the fp64 entry points, "_f" builds (ValueType = float,
// DOUBLE == 0) the fp32 ones (spmm_b200_pipeline_*_f32).
//
// Environment (all optional):
//   SPMM_B200_DEVICE      = CUDA device ordinal (default 0)
//   SPMM_B200_SDDMM_MODE  = 0 (default): the reference's SDDMM, both factors at the output row
//                           (sddmm_taco_naive.cpp:128); 1: K at row ja[p] (tree-ordered sum); 2: same, ordered
//   SPMM_B200_RESIDENT    = 0 (default): spmm('final', mask, y, V, y_final) reads its host arguments on every call,
//                           like the reference's Matrix_Format::spmm.  1 (opt-in): when `a` and `x` are the very
//                           buffers the preceding sddmm(y) and spmm('V', ..., V) calls wrote, the copies already on
//                           the device are used -- ONLY valid if the harness does not touch y or V in between (it
//                           would, e.g., with its commented-out host softmax(y) re-enabled): pointer identity is
//                           all that is checked.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "macros/cpp_defines.h"
#include "sddmm_bench_common.h"
#include "sddmm_kernel.h"
#include "spmm_b200_pipeline.h"

#if DOUBLE == 0
	typedef spmm_b200_pipeline_f32_t pipe_t;
	#define PIPE(fn) spmm_b200_pipeline_ ## fn ## _f32
#else
	typedef spmm_b200_pipeline_t pipe_t;
	#define PIPE(fn) spmm_b200_pipeline_ ## fn
#endif

static void die(const char * where, int code)
{
	// reference convention: message on stdout (stderr is the CSV channel), then exit(1)
	printf("spmm_b200 pipeline: %s failed: %s [%s]\n", where, spmm_b200_strerror(code), spmm_b200_last_cuda_error());
	exit(1);
}

struct B200Pipeline : Matrix_Format
{
	pipe_t * handle;
	INT_T * mask_ia;
	INT_T * mask_ja;
	ValueType * last_y;      // host buffer the last sddmm() filled
	ValueType * last_v;      // host buffer the last spmm('V') filled
	int mode;
	int resident;

	B200Pipeline(long m, long n, long nnz) : Matrix_Format(m, n, nnz),
		handle(NULL), mask_ia(NULL), mask_ja(NULL), last_y(NULL), last_v(NULL), mode(0), resident(0) {}

	~B200Pipeline()
	{
		PIPE(destroy)(handle);
	}

	void spmm(char type, INT_T m, INT_T k, INT_T n, INT_T * ia, INT_T * ja, ValueType * a, ValueType * x, ValueType * y, int num_threads);
	void sddmm(ValueType * y, int num_threads);
	void statistics_start();
	int statistics_print_data(char * buf, long buf_n);
};

void
B200Pipeline::spmm(char type, INT_T m, INT_T k, INT_T n, INT_T * ia, INT_T * ja, ValueType * a, ValueType * x, ValueType * y,
	__attribute__((unused)) int num_threads)
{
	int rc;
	if (resident && type != 'K' && type != 'Q' && type != 'V' && ia == mask_ia && ja == mask_ja && a == last_y && x == last_v
			&& last_y != NULL && last_v != NULL)
		rc = PIPE(final)(handle, y);
	else
		rc = PIPE(spmm)(handle, type, m, k, n, ia, ja, a, x, y);
	if (rc)
		die("spmm", rc);
	if (type == 'V')
		last_v = y;
}

void
B200Pipeline::sddmm(ValueType * y, __attribute__((unused)) int num_threads)
{
	int rc = PIPE(sddmm)(handle, y, mode);
	if (rc)
		die("sddmm", rc);
	last_y = y;
}

struct Matrix_Format *
csr_to_format(INT_T * row_ptr, INT_T * col_ind, ValueType * values, long m, long nnz, long n)
{
	struct B200Pipeline * p = new B200Pipeline(m, n, nnz);
	const char * s;
	int device = 0;
	if ((s = getenv("SPMM_B200_DEVICE")) != NULL && *s)
		device = atoi(s);
	if ((s = getenv("SPMM_B200_SDDMM_MODE")) != NULL && *s)
		p->mode = atoi(s);
	if ((s = getenv("SPMM_B200_RESIDENT")) != NULL && *s)
		p->resident = atoi(s);
	p->format_name = (char *) (DOUBLE == 0 ? "B200_CSR_PIPELINE_F32" : "B200_CSR_PIPELINE");
	p->mask_ia = row_ptr;
	p->mask_ja = col_ind;
	int rc = PIPE(create)(&p->handle, row_ptr, col_ind, values, m, nnz, (int) n, device);
	if (rc)
		die("csr_to_format", rc);
	p->mem_footprint = p->csr_mem_footprint;
	return p;
}

void
B200Pipeline::statistics_start()
{
}

int
statistics_print_labels(__attribute__((unused)) char * buf, __attribute__((unused)) long buf_n)
{
	return 0;
}

int
B200Pipeline::statistics_print_data(__attribute__((unused)) char * buf, __attribute__((unused)) long buf_n)
{
	return 0;
}
