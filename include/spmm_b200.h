/*
 * spmm_b200.h -- C ABI of libspmm_b200.so: fp64 CSR sparse x dense multiply
 * C = A*B (SpMV = k 1) on NVIDIA B200 (sm_100a), hand-written CUDA.
 *
 * This is the drop-in boundary for ONE path of ioannatas/SpMM-Research: the
 * plug-in interface  `struct Matrix_Format { virtual void spmm(x, y, k) ... }`
 * + `csr_to_format(row_ptr, col_ind, values, m, n, nnz, k)`
 * (ref: benchmark_code/CPU/AMD/spmv_code_bench/spmv_kernel.h:9-30).  A ~100-line
 * C++ translation unit (spmm-research_b200/csrc/spmm_kernel_b200.cpp) implements
 * that interface on top of the functions below and links under the reference's
 * UNCHANGED harness (spmv_bench.cpp) like any other kernel TU (Makefile_in:52-53).
 *
 * Conventions
 *   - plain pointers and sizes, no C++/torch types; every call returns 0 or a
 *     negative SPMM_B200_E* code and never calls exit() (the reference's
 *     convention "print + exit(1)" is applied by the C++ shim, not here);
 *   - no CPU fallback: without a usable sm_100 device every compute entry point
 *     fails with SPMM_B200_ENODEV / SPMM_B200_ECUDA;
 *   - int32 CSR indices (INT_T=int32_t), fp64 values (ValueType=double), 0-based,
 *     columns sorted inside a row as the generator / coo_to_csr produce them;
 *   - dense operand layouts follow the oracle compute_csr()
 *     (ref: spmm_kernel_csr.cpp:88,93):  B = x = k contiguous vectors of length c
 *     (SPMM_B200_B_COLMAJOR, x[n*c + j]);  C = y = row-major m x k (y[i*k + n]).
 *     On the device B is kept row-major (c x k) so one gathered B row is one
 *     coalesced read; the library transposes on upload.
 */
#ifndef SPMM_B200_H
#define SPMM_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* A handle owns device buffers, streams and row-ticket counters: calls on ONE handle must not overlap (one host
 * thread / one stream at a time).  Different handles are independent.
 * Alignment: the specialised kernels move B and C as 16-byte words.  Device operands handed to
 * spmm_b200_run_device that are not 16-byte aligned (or whose k is odd) are still accepted -- they run on the
 * generic kernel (slower, same result); a page-locked y that is not 16-byte aligned takes the staged download. */
typedef struct spmm_b200 spmm_b200_t;

/* kernel variants (north_star (2),(3)) */
enum {
	SPMM_B200_ROWSPLIT = 0,   /* sub-warp group per row, CTAs own nnz-balanced row ranges   */
	SPMM_B200_MERGE    = 1,   /* merge-path tiles with k-wide carries (skewed rows)          */
	SPMM_B200_BCSR     = 2,   /* blocked CSR on the FP64 tensor pipe (dense blocks only)     */
	SPMM_B200_AUTO     = 3,   /* pick from the row-length statistics of A                    */
	SPMM_B200_BCSR_FMA = 4    /* same 8x1 blocked CSR on the FP64 FMA pipe, absent rows skipped */
};

/* layout of the dense operand handed to spmm_b200_run / spmm_b200_set_b */
enum {
	SPMM_B200_B_COLMAJOR = 0, /* oracle layout x[n*c + j]  (spmm_kernel_csr.cpp:88)          */
	SPMM_B200_B_ROWMAJOR = 1  /* x[j*k + n]  (mkl_sparse_d_mm ROW_MAJOR, spmm_kernel_mkl_ie.cpp:143-146) */
};

/* partition modes of spmm_b200_partition (all bit-exact with the reference) */
enum {
	SPMM_B200_PART_ROWS   = 0, /* loop_partitioner_balance_iterations   lib/parallel_util.h:47-93,116-126 : out[2t],out[2t+1] = [start,end) */
	SPMM_B200_PART_NNZ    = 1, /* loop_partitioner_balance_prefix_sums  lib/parallel_util.h:141-165       : out[0..P] row starts, out[P]=m  */
	SPMM_B200_PART_MERGE  = 2, /* MergePathSearch at diagonals t*ceil((m+nnz)/P), ARM/spmv_kernel_merge.cpp:214-234,262-272 : out[2t]=row, out[2t+1]=nnz idx, t=0..P */
	SPMM_B200_PART_NNZ_CARRY = 3 /* perfect nnz balance with row carry, spmv_kernel_csr.cpp:68-80 (CUSTOM_VECTOR_PERFECT_NNZ_BALANCE):
	                                out[4t..4t+3] = thread_i_s, thread_i_e, thread_j_s, thread_j_e of worker t = 0..P-1 */
};

enum {
	SPMM_B200_OK      =  0,
	SPMM_B200_EINVAL  = -1,   /* bad argument (null pointer, negative size, k mismatch ...)   */
	SPMM_B200_ENODEV  = -2,   /* no CUDA device / not compute capability 10.x                 */
	SPMM_B200_ECUDA   = -3,   /* a CUDA runtime call failed (see spmm_b200_last_cuda_error)   */
	SPMM_B200_ENOMEM  = -4,   /* device or host allocation failed                             */
	SPMM_B200_ERANGE  = -5,   /* sizes overflow the int32 index contract (nnz, m, c < 2^31)  */
	SPMM_B200_ENOSUP  = -6,   /* variant not applicable to this matrix / k                    */
	SPMM_B200_ECOMM   = -7,   /* the caller-supplied broadcast callback returned non-zero     */
	SPMM_B200_EIO     = -8,   /* a matrix file could not be opened / read                     */
	SPMM_B200_EFORMAT = -9,   /* a matrix file is malformed (the reference would exit(1))     */
	SPMM_B200_ESCRATCH = -10  /* spmm_b200_spmm_merge_noalloc: the input needs more scratch than the fixed array holds */
};

/* ---- format construction: replaces csr_to_format() (spmv_kernel.h:29; call site spmv_bench.cpp:996).
 * Copies the host CSR arrays to `device`; the caller keeps ownership of the host
 * arrays (the C++ shim frees them in its destructor as every reference kernel does,
 * spmm_kernel_csr.cpp:34-39).  k is the number of dense columns the handle is tuned
 * for (NUM_COLS); spmm_b200_run accepts any k >= 1. */
int spmm_b200_create(spmm_b200_t **h,
                     const int32_t *row_ptr, const int32_t *col_ind, const double *values,
                     int64_t m, int64_t c, int64_t nnz, int k,
                     int device, int variant, int b_layout);

/* Same, for shard `rank` of `nranks` nnz-balanced row shards (rows
 * [s_rank, s_rank+1) of loop_partitioner_balance_prefix_sums(P = nranks), computed
 * on the device, bit-exact).  Only that shard's slice of A is uploaded, row_ptr is
 * rebased; C is row-local (m_shard x k).  first_row/last_row report the shard. */
int spmm_b200_create_shard(spmm_b200_t **h,
                           const int32_t *row_ptr, const int32_t *col_ind, const double *values,
                           int64_t m, int64_t c, int64_t nnz, int k,
                           int device, int variant, int b_layout,
                           int rank, int nranks, int64_t *first_row, int64_t *last_row);

/* ---- the hot call: replaces Matrix_Format::spmm(x, y, k) (spmv_kernel.h:18; call sites
 * spmv_bench.cpp:318,372).  HOST buffers: x (c*k doubles, layout = b_layout of the
 * handle), y (m_local*k doubles, row-major).  Uploads x (H2D), runs the kernel(s),
 * and when copy_back != 0 downloads y (D2H) before returning.  With
 * copy_back == 0 the result stays on the device (spmm_b200_fetch_c reads it).
 * The call is PCIe-bound; for k >= 32 (k % 16 == 0) and operands of >= 8 MB the columns are cut into
 * 16-column slices so that the upload of slice s+1, the kernel on slice s and the download of
 * C[:, slice s-1] overlap (SPMM_B200_PIPELINE_COLS=0 turns it off).  Results are identical.
 * When y_host is page-locked and mapped for the device (cudaHostAlloc / torch pin_memory / spmm_b200_register_host)
 * and the row-split kernel runs the slices, each slice's kernel stores its part of every C row straight into
 * y_host (no staging copy; spmm_b200_fetch_c has nothing to fetch afterwards and returns SPMM_B200_EINVAL).
 * SPMM_B200_DIRECT_STORE=0 keeps the staged download. */
int spmm_b200_run(spmm_b200_t *h, const double *x_host, double *y_host, int k, int copy_back);

/* Multi-GPU form of the hot call (one process per GPU, handle from spmm_b200_create_shard).  The only
 * collective on this path is the replication of B; the library does not link a communication
 * library, the caller supplies it:  `bcast(ctx, dev_ptr, count, parts, stream)` must enqueue on `stream`
 * (a cudaStream_t) the replication of `count` doubles at `dev_ptr` and return 0:
 *   parts == 1: broadcast from rank 0, e.g. ncclBroadcast(p, p, count, ncclDouble, 0, comm, stream);
 *   parts == P: the range is cut into P parts of ceil(count/P) doubles (the last ones shorter or empty), part r
 *               is valid on rank r; broadcast every part from its owner (ncclBroadcast with root r inside one
 *               ncclGroupStart/End, or ncclAllGather when count % P == 0).
 * x_on_every_rank == 0: x_host is the host B on rank 0 and NULL elsewhere; rank 0 uploads all of it.
 * x_on_every_rank != 0: every rank passes the same host B (SPMD launchers generate their inputs on every rank);
 *   each rank uploads only ITS part over its own PCIe link and the parts are exchanged over NVLink
 *   (oracle layout only; otherwise it falls back to the rank-0 upload).
 * y_host receives this rank's row-local C.  For k >= 32 the columns go through in slices so that upload,
 * replication, kernel and the download of C overlap; with bcast == NULL it is spmm_b200_run. */
typedef int (*spmm_b200_bcast_fn)(void *ctx, double *dev_ptr, size_t count, int parts, void *stream);
int spmm_b200_run_replicated(spmm_b200_t *h, const double *x_host, double *y_host, int k, int copy_back,
                             spmm_b200_bcast_fn bcast, void *bcast_ctx, int x_on_every_rank);

/* One process per GPU on a box whose PCIe links are unequal ("download routing"): the ranks share ONE y
 * (m_global x k, e.g. POSIX shared memory mapped and page-locked by every rank) and the caller has decided -- by
 * measurement, see bench.py -- which ranks carry C to the host.  Every rank uploads its own window of B and runs its
 * shard; a rank with downloader_rank != its own sends its rows of C to that rank and moves nothing over PCIe on the way
 * down; a downloader stores its own rows into y_all at its first row and then, for each of its `nclients` clients,
 * receives the client's rows into a relay buffer and stores them at the client's first row.
 * `xfer(ctx, dev_ptr, count, peer_rank, is_send, stream)` must enqueue on `stream` the send (is_send != 0) or receive of
 * `count` doubles at `dev_ptr` to / from `peer_rank` and return 0 (ncclSend / ncclRecv, torch.distributed.send / recv).
 * Requires x on every rank (x_host != NULL).  Same bits as spmm_b200_run on the shard. */
typedef int (*spmm_b200_xfer_fn)(void *ctx, double *dev_ptr, size_t count, int peer_rank, int is_send, void *stream);
int spmm_b200_run_routed(spmm_b200_t *h, const double *x_host, double *y_all, int k, int my_rank, int downloader_rank,
                         const int *client_ranks, const int64_t *client_first_row, const int64_t *client_rows, int nclients,
                         spmm_b200_xfer_fn xfer, void *xfer_ctx);

/* Optional: page-lock a caller-owned host range (x or y) so that spmm_b200_run copies at full
 * PCIe rate.  The caller promises the range stays mapped until spmm_b200_unregister_host /
 * spmm_b200_destroy -- true for the harness's x and y (spmv_bench.cpp:896-907).  Ranges that
 * are already page-locked (cudaHostAlloc, torch pin_memory) are accepted and left alone. */
int spmm_b200_register_host(spmm_b200_t *h, const void *p, size_t bytes);
int spmm_b200_unregister_host(spmm_b200_t *h, const void *p);

/* Tuning knobs: every SPMM_B200_<NAME> environment variable is read ONCE, when a handle is
 * created; afterwards spmm_b200_set_option(h, "<NAME>", value) changes it for that handle ("SPMM_B200_" prefix
 * optional).  Unknown names: SPMM_B200_EINVAL. */
int spmm_b200_set_option(spmm_b200_t *h, const char *name, int value);

/* Event-stamped timeline of the last pipelined host-buffer run (spmm_b200_run / _run_replicated): out[0] = number of
 * steps S (column slices, or row blocks when the run was row-banded, see spmm_b200_last_rowband), then per step six
 * times in ms since the first enqueue: H2D done, replication done, transpose done, kernel start, kernel end, D2H
 * done.  Returns the number of doubles written. */
int spmm_b200_last_timeline(const spmm_b200_t *h, double *out, int max_values);

/* Upload B once (x is constant across the reference's timed loop,
 * spmv_bench.cpp:895-903) and run on the resident copy. */
int spmm_b200_set_b(spmm_b200_t *h, const double *x_host, int k);
int spmm_b200_run_resident(spmm_b200_t *h, int k);
int spmm_b200_fetch_c(spmm_b200_t *h, double *y_host, int k);

/* Device-pointer variant for callers that own device memory and streams
 * (bench.py / torch.distributed ranks): b_dev = row-major c x k, c_dev =
 * row-major m_local x k, both on the handle's device; `stream` is a
 * cudaStream_t (NULL = default stream).  Asynchronous -- except that, when the B rows the matrix touches
 * (column span x 8k bytes) exceed 64 MB and k >= 32, the first calls with a given k (two per candidate, at most 8) each run and
 * time one candidate column-slice width of the row-split kernel (same bits for every candidate) and BLOCK
 * until their kernels finish; the fastest is kept per k (spmm_b200_kslice).  The measurement is skipped while
 * `stream` is being captured into a CUDA graph; SPMM_B200_KSLICE=-1 turns it off. */
int spmm_b200_run_device(spmm_b200_t *h, const double *b_dev, double *c_dev, int k, void *stream);

/* Column-major (oracle layout) -> row-major transpose of B on the device. */
int spmm_b200_transpose_b_device(spmm_b200_t *h, const double *x_colmajor_dev, double *b_rowmajor_dev,
                                 int k, void *stream);

/* ---- single-process multi-GPU (csr_to_format(..., ngpus), 8(e)): ONE host thread drives `ngpus`
 * devices of one box, which is how the reference's C harness reaches more than one GPU.  Rows are cut by
 * loop_partitioner_balance_prefix_sums with P = ngpus (bit-exact); device g holds shard g of A.  The only exchange is
 * the replication of B: every device uploads 1/ngpus of each column slice of x over its own PCIe link, then ONE
 * kernel per device gathers all parts out of its peers' memory over NVLink (peer loads) and writes the row-major
 * operand -- the all-gather and the layout transpose are the same pass.  C is row-local and lands in rows
 * [first_row_g, last_row_g) of the caller's single y.  devices == NULL: ordinals 0..ngpus-1.  ngpus <= 16.
 * Without peer access between a pair of devices every device uploads all of x itself. */
typedef struct spmm_b200_multi spmm_b200_multi_t;
int spmm_b200_create_multi(spmm_b200_multi_t **mh,
                           const int32_t *row_ptr, const int32_t *col_ind, const double *values,
                           int64_t m, int64_t c, int64_t nnz, int k,
                           int ngpus, const int *devices, int variant, int b_layout);
/* Matrix_Format::spmm with HOST buffers over all shards; x: c*k, y: m*k (whole matrix). */
int spmm_b200_run_multi(spmm_b200_multi_t *mh, const double *x_host, double *y_host, int k, int copy_back);
/* x constant across calls (spmv_bench.cpp:895-903): replicate once, run on the resident copies, fetch when needed */
int spmm_b200_set_b_multi(spmm_b200_multi_t *mh, const double *x_host, int k);
int spmm_b200_run_resident_multi(spmm_b200_multi_t *mh, int k);
int spmm_b200_fetch_c_multi(spmm_b200_multi_t *mh, double *y_host, int k);
/* page-lock a caller buffer for every device (cudaHostRegisterPortable | Mapped) */
int spmm_b200_register_host_multi(spmm_b200_multi_t *mh, const void *p, size_t bytes);
int spmm_b200_multi_ngpus(const spmm_b200_multi_t *mh);
spmm_b200_t *spmm_b200_multi_shard(spmm_b200_multi_t *mh, int g);   /* shard handle (introspection, set_option) */
int64_t spmm_b200_multi_first_row(const spmm_b200_multi_t *mh, int g);  /* g = ngpus: m */
double  spmm_b200_multi_last_kernel_ms(const spmm_b200_multi_t *mh);    /* max over the shards */
int64_t spmm_b200_multi_launch_count(const spmm_b200_multi_t *mh);
int     spmm_b200_multi_peer_access(const spmm_b200_multi_t *mh);      /* 1: B is exchanged over NVLink */
/* Download routes of spmm_b200_run_multi: out[g] = the device whose PCIe link carries device g's rows of C (g itself
 * unless the box's links were measured to be unequal: then devices behind the slow path store C into a partner over
 * NVLink), followed by the measured GB/s of the five candidate downloader sets {all, lower half, upper half, evens,
 * odds} (zeros when nothing was measured).  Option MULTI_ROUTE (on shard 0): -1 measure once, 0 off, mask = forced set.
 * Returns the number of doubles written. */
int     spmm_b200_multi_routes(const spmm_b200_multi_t *mh, double *out, int max_values);
void    spmm_b200_destroy_multi(spmm_b200_multi_t *mh);

/* ---- partitioning, testable without an SpMM kernel (host pointers in and out,
 * the search itself runs on the device).  `out` sizes: see the enum above. */
int spmm_b200_partition(const int32_t *row_ptr, int64_t m, int64_t nnz, int P, int mode,
                        int64_t *out, int device);

/* Same integers computed on the host (shard planning before a device exists; no SpMM arithmetic). */
int spmm_b200_partition_host(const int32_t *row_ptr, int64_t m, int64_t nnz, int P, int mode, int64_t *out);

/* ---- COO -> CSR: replaces coo_to_csr(R, C, V, m, n, nnz, row_ptr, col_idx, values, sort_columns,
 * transpose) (lib/storage_formats/csr/csr_gen.c:162-217; call site spmv_bench.cpp:823) including
 * csr_sort_columns (csr_gen.c:82-156).  Host pointers; histogram, scan, scatter and the per-row
 * column sort run on the device.  row_ptr has m+1 entries (n+1 with transpose).  V/values may both
 * be NULL.  Entries of a row are ordered by (column, position in the input) -- identical to the
 * reference whenever no (row, col) pair repeats; with sort_columns == 0 by input position. */
int spmm_b200_coo_to_csr(const int32_t *R, const int32_t *C, const double *V, int64_t m, int64_t n, int64_t nnz,
                         int32_t *row_ptr, int32_t *col_idx, double *values, int sort_columns, int transpose,
                         int device);

/* ---- on-disk formats (host parsing; the arrays then go through spmm_b200_coo_to_csr / spmm_b200_create).
 * Output arrays are allocated by the library (64-byte aligned); release them with spmm_b200_free.
 *
 * spmm_b200_read_mtx replaces mtx_read(file, expand_symmetry, pattern_dummy_vals = 1)
 * (lib/storage_formats/matrix_market/matrix_market.c:249-322) together with the value conversion the
 * harness applies to its result (spmv_bench.cpp:723-763): Matrix Market "coordinate" files with field
 * real | integer | complex | pattern and symmetry general | symmetric | Hermitian | skew-symmetric ->
 * 0-based COO with fp64 values (integer -> (double), complex -> |z|, pattern -> 1.0); with
 * expand_symmetry the mirrored off-diagonal entries follow the file's entries in file order, negated for
 * skew-symmetric.  Indices and values are bit-identical with the reference's arrays (same strtol/strtod).
 * field16 (may be NULL) receives the field name.  "array" files: SPMM_B200_ENOSUP. */
int spmm_b200_read_mtx(const char *path, int expand_symmetry, int64_t *m, int64_t *n, int64_t *nnz,
                       int32_t **R, int32_t **C, double **V, char *field16);

/* spmm_b200_read_smtx replaces smtx_read (lib/storage_formats/dlcm_matrices/dlcm_matrix.c:257-324,
 * dlcm_matrix_gen.c:54-118; call site spmv_bench.cpp:682-695): DLMC ".smtx" text = "m, k, nnz" / m+1 row
 * offsets / nnz column indices -> CSR.  The file holds no values; the reference draws them from rand()
 * seeded with time(NULL) (not reproducible), here values[j] = uniform [-1, 1) from splitmix64(seed, j). */
int spmm_b200_read_smtx(const char *path, uint64_t seed, int64_t *m, int64_t *n, int64_t *nnz,
                        int32_t **row_ptr, int32_t **col_ind, double **values);
void spmm_b200_free(void *p);

/* ---- accuracy check: replaces CheckAccuracy() (spmv_bench.cpp:121-206; call site :480).  Computes the
 * gold result of y = A*x on the device in double-double arithmetic (fp64 products as in the reference,
 * summed with ~106 significant bits -- the reference uses __float128 with Kahan compensation) and the reference's
 * error statistics of y_host against it.  x_host: c*k (layout of the handle), y_host: m_local*k row-major.
 * stats[0] = max relative difference over entries with gold > 1e-10 (the reference's "Test failed" figure),
 * [1] mae, [2] max_ae, [3] mse, [4] mape, [5] smape, [6] lnQ_error, [7] mlare, [8] gmare
 * (lib/array_metrics.c), [9] = 1.0 if stats[0] <= 1e-10 else 0.0.  y_gold_host (may be NULL) receives the
 * gold result rounded to fp64. */
enum { SPMM_B200_NSTATS = 10 };
int spmm_b200_check_accuracy(spmm_b200_t *h, const double *x_host, const double *y_host, int k,
                             double *stats, double *y_gold_host);

/* ---- fp32 ("_f" builds of the reference, make.sh:98-108: ValueType = float).  Same plug-in interface
 * (spmv_kernel.h:9-30) and layouts as the fp64 entry points above, int32 indices, fp32 values / x / y, the same
 * bit-exact partitioners for CTA work lists, merge tiles and shards.  `variant`:
 *   SPMM_B200_ROWSPLIT  row-split CSR, the row sum runs left to right in fp32 with fused multiply-add
 *                       (bit-identical to the reference's _f build of compute_csr);
 *   SPMM_B200_MERGE     merge-path tiles with k-wide carries (skewed rows; sums are re-associated at the carries);
 *   SPMM_B200_AUTO      merge-path when one row is far longer than the mean, else row-split;
 *   SPMM_B200_BCSR      blocked tensor-core kernel for BANDED matrices: 128-row slabs x their column window as dense
 *                       TF32 products on tcgen05.mma with fp32 accumulators in TMEM (k = 32, 64, 128; other k fall back
 *                       to row-split).  A and B are rounded to TF32 (10 mantissa bits): results agree with the fp32
 *                       reference to ~1e-3 relative, NOT bit for bit -- which is why AUTO never picks it
 *                       (the stand-in for mkl_sparse_s_mm-class kernels, spmm_kernel_mkl_ie.cpp:143-146).
 *   SPMM_B200_BCSR_FMA  = SPMM_B200_ROWSPLIT (no fp32 FMA-pipe blocked kernel). */
typedef struct spmm_b200_f32 spmm_b200_f32_t;
int spmm_b200_create_f32(spmm_b200_f32_t **h, const int32_t *row_ptr, const int32_t *col_ind, const float *values,
                         int64_t m, int64_t c, int64_t nnz, int k, int device, int variant, int b_layout);
int spmm_b200_create_shard_f32(spmm_b200_f32_t **h, const int32_t *row_ptr, const int32_t *col_ind, const float *values,
                               int64_t m, int64_t c, int64_t nnz, int k, int device, int variant, int b_layout,
                               int rank, int nranks, int64_t *first_row, int64_t *last_row);
int spmm_b200_run_f32(spmm_b200_f32_t *h, const float *x_host, float *y_host, int k, int copy_back);
int spmm_b200_run_device_f32(spmm_b200_f32_t *h, const float *b_dev, float *c_dev, int k, void *stream);
int spmm_b200_fetch_c_f32(spmm_b200_f32_t *h, float *y_host, int k);
int     spmm_b200_variant_f32(const spmm_b200_f32_t *h);      /* the variant the handle resolved to */
int     spmm_b200_set_option_f32(spmm_b200_f32_t *h, const char *name, int value);
int64_t spmm_b200_last_h2d_bytes_f32(const spmm_b200_f32_t *h); /* bytes of B the last spmm_b200_run_f32 uploaded: the rows between the first and the last touched column (all of B for the TF32 variant or with UPLOAD_WINDOW=0) */
int64_t spmm_b200_rows_f32(const spmm_b200_f32_t *h);
int64_t spmm_b200_nnz_f32(const spmm_b200_f32_t *h);
int64_t spmm_b200_launch_count_f32(const spmm_b200_f32_t *h);
double  spmm_b200_last_kernel_ms_f32(const spmm_b200_f32_t *h);
void    spmm_b200_destroy_f32(spmm_b200_f32_t *h);

/* ---- allocation-free SpMM.  No counterpart in the reference: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * C = A * B by the merge-path kernels on the CURRENT device, everything passed as DEVICE pointers: CSR arrays as for
 * spmm_b200_create (int32, 0-based, columns sorted per row), b_dev = row-major c x k, c_dev = row-major m x k (both
 * 16-byte aligned when k >= 2), k in {1, 2, 4, 8, 16, 32, 64, 128}; `stream` is a cudaStream_t (NULL = default).
 * Asynchronous.  Same arithmetic as a SPMM_B200_MERGE handle's spmm_b200_run_device (same bits).
 * The scratch of the kernels (tile coordinates, per-tile carries) lives in ONE static device array of
 * SPMM_B200_NOALLOC_SCRATCH_BYTES that every GPU which loads the library holds until exit; the call allocates and
 * frees nothing.  Limit: the scratch is about (12 + 8k) bytes per tile of T merge items, m + nnz items in all (T = 4096
 * for k <= 2, 1024 for k = 16 and k = 128, 2048 otherwise): m + nnz up to ~64 M at k = 32, ~8.3 M at k = 128, no
 * practical bound at k = 1.  Beyond it the call returns SPMM_B200_ESCRATCH (sizes: spmm_b200_last_cuda_error) and
 * launches nothing; such inputs take spmm_b200_create + spmm_b200_run_device.
 * All calls on a device SHARE the array and find the previous call's contents in it (what must start at zero is cleared
 * by the call): two calls that overlap on the GPU -- different streams, different threads -- race. */
#define SPMM_B200_NOALLOC_SCRATCH_BYTES (8u << 20)
int spmm_b200_spmm_merge_noalloc(const int32_t *row_ptr_dev, const int32_t *col_ind_dev, const double *values_dev,
                                 int64_t m, int64_t c, int64_t nnz, const double *b_dev, double *c_dev, int k, void *stream);

/* ---- introspection */
int64_t spmm_b200_rows(const spmm_b200_t *h);         /* rows held by this handle (shard)     */
int64_t spmm_b200_nnz(const spmm_b200_t *h);
int     spmm_b200_variant(const spmm_b200_t *h);      /* the variant AUTO resolved to         */
int64_t spmm_b200_bcsr_blocks(const spmm_b200_t *h);  /* 8x1 blocks stored (0 unless a BCSR variant); fill = nnz / (8*blocks) */
int     spmm_b200_last_rowband(const spmm_b200_t *h); /* row blocks of the last spmm_b200_run when it took the row-band host pipeline (banded matrix: B rows up, C rows down, block by block), else 0 */
int     spmm_b200_col_window(const spmm_b200_t *h, int64_t *first_col, int64_t *ncols); /* first touched column and last - first + 1 over the handle's nonzeros (0 columns: no nonzero): the rows of B this handle reads */
int     spmm_b200_last_transfer_bytes(const spmm_b200_t *h, int64_t *h2d_bytes, int64_t *d2h_bytes); /* bytes the last host-buffer call moved over PCIe: only the rows of B the matrix's nonzeros reference go up (UPLOAD_WINDOW) */
int     spmm_b200_kslice(const spmm_b200_t *h);       /* column-slice width the row-split path settled on for its last k (0: none / still measuring; k: one launch) */
double  spmm_b200_last_kernel_ms(const spmm_b200_t *h); /* CUDA-event time of the last run's kernels */
int64_t spmm_b200_launch_count(const spmm_b200_t *h); /* kernels launched by this handle so far */
/* replaces statistics_print_data()/statistics_print_labels() (spmv_kernel.h:20,30) */
int     spmm_b200_stats(const spmm_b200_t *h, char *buf, long buf_n);
int     spmm_b200_stats_labels(char *buf, long buf_n);
void    spmm_b200_destroy(spmm_b200_t *h);
const char *spmm_b200_strerror(int code);
const char *spmm_b200_last_cuda_error(void);
int     spmm_b200_device_count(void);

#ifdef __cplusplus
}
#endif
#endif /* SPMM_B200_H */
