"""spmm-research_b200 -- host-side mirror of the reference's SpMM plug-in interface.

The reference selects a kernel at link time: one translation unit supplies
`csr_to_format(row_ptr, col_ind, values, m, n, nnz, k)` and a `Matrix_Format`
subclass with `spmm(x, y, k)` (ref: benchmark_code/CPU/AMD/spmv_code_bench/
spmv_kernel.h:9-30).  The C++ twin of this file, csrc/spmm_kernel_b200.cpp, is that
translation unit for the reference's own harness; this module exposes the same
names, argument meaning and error behaviour to Python (tests, bench.py) through
ctypes over the C ABI in include/spmm_b200.h.

There is NO CPU fallback: if lib/libspmm_b200.so is missing, or no sm_100 device is
present, construction fails loudly.
"""
import ctypes as C
import os

import numpy as np

__all__ = ["csr_to_format", "csr_to_format_f32", "MatrixFormatF32", "coo_to_csr", "read_mtx", "read_smtx", "MatrixFormat", "partition", "shard_slices", "statistics_print_labels", "load_library",
           "SpmmB200Error", "ROWSPLIT", "MERGE", "BCSR", "AUTO", "BCSR_FMA", "B_COLMAJOR", "B_ROWMAJOR",
           "PART_ROWS", "PART_NNZ", "PART_MERGE", "PART_NNZ_CARRY", "LIB_PATH", "EXPORTS", "MultiGpuFormat", "csr_to_format_multi",
           "PipelineFormat", "csr_to_pipeline_format", "PipelineFormatF32", "csr_to_pipeline_format_f32", "SDDMM_REFERENCE", "SDDMM_ATTENTION", "SDDMM_ATTENTION_ORDERED"]

ROWSPLIT, MERGE, BCSR, AUTO, BCSR_FMA = 0, 1, 2, 3, 4
B_COLMAJOR, B_ROWMAJOR = 0, 1
PART_ROWS, PART_NNZ, PART_MERGE, PART_NNZ_CARRY = 0, 1, 2, 3

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "lib", "libspmm_b200.so")

# every symbol include/spmm_b200.h declares
EXPORTS = [
    "spmm_b200_create", "spmm_b200_create_shard", "spmm_b200_run", "spmm_b200_run_replicated", "spmm_b200_run_routed", "spmm_b200_spmm_merge_noalloc", "spmm_b200_set_b",
    "spmm_b200_register_host", "spmm_b200_unregister_host", "spmm_b200_set_option", "spmm_b200_last_timeline",
    "spmm_b200_create_multi", "spmm_b200_run_multi", "spmm_b200_set_b_multi", "spmm_b200_run_resident_multi",
    "spmm_b200_fetch_c_multi", "spmm_b200_register_host_multi", "spmm_b200_multi_ngpus", "spmm_b200_multi_shard",
    "spmm_b200_multi_first_row", "spmm_b200_multi_last_kernel_ms", "spmm_b200_multi_launch_count",
    "spmm_b200_multi_peer_access", "spmm_b200_multi_routes", "spmm_b200_destroy_multi",
    "spmm_b200_run_resident", "spmm_b200_fetch_c", "spmm_b200_run_device",
    "spmm_b200_transpose_b_device", "spmm_b200_partition", "spmm_b200_partition_host", "spmm_b200_coo_to_csr", "spmm_b200_rows", "spmm_b200_nnz",
    "spmm_b200_variant", "spmm_b200_bcsr_blocks", "spmm_b200_kslice", "spmm_b200_last_rowband", "spmm_b200_last_transfer_bytes", "spmm_b200_col_window", "spmm_b200_last_kernel_ms", "spmm_b200_launch_count", "spmm_b200_stats",
    "spmm_b200_stats_labels", "spmm_b200_destroy", "spmm_b200_strerror", "spmm_b200_last_cuda_error",
    "spmm_b200_create_f32", "spmm_b200_create_shard_f32", "spmm_b200_run_f32", "spmm_b200_run_device_f32",
    "spmm_b200_fetch_c_f32", "spmm_b200_variant_f32", "spmm_b200_set_option_f32", "spmm_b200_last_h2d_bytes_f32", "spmm_b200_rows_f32", "spmm_b200_nnz_f32", "spmm_b200_launch_count_f32",
    "spmm_b200_last_kernel_ms_f32", "spmm_b200_destroy_f32",
    "spmm_b200_device_count", "spmm_b200_read_mtx", "spmm_b200_read_smtx", "spmm_b200_free", "spmm_b200_check_accuracy",
    # include/spmm_b200_pipeline.h
    "spmm_b200_pipeline_create", "spmm_b200_pipeline_spmm", "spmm_b200_pipeline_sddmm", "spmm_b200_pipeline_softmax",
    "spmm_b200_pipeline_final", "spmm_b200_pipeline_set_weight", "spmm_b200_pipeline_attention",
    "spmm_b200_pipeline_device_ptr", "spmm_b200_pipeline_last_kernel_ms", "spmm_b200_pipeline_launch_count",
    "spmm_b200_pipeline_destroy", "spmm_b200_pipeline_invalidate", "spmm_b200_pipeline_set_option",
    "spmm_b200_pipeline_create_f32", "spmm_b200_pipeline_spmm_f32", "spmm_b200_pipeline_sddmm_f32",
    "spmm_b200_pipeline_softmax_f32", "spmm_b200_pipeline_final_f32", "spmm_b200_pipeline_launch_count_f32", "spmm_b200_pipeline_destroy_f32",
]

_lib = None
_vp, _i64, _int, _dbl = C.c_void_p, C.c_int64, C.c_int, C.c_double


# int (*spmm_b200_bcast_fn)(void *ctx, double *dev_ptr, size_t count, int parts, void *stream)
BCAST_FN = C.CFUNCTYPE(_int, _vp, _vp, C.c_size_t, _int, _vp)
XFER_FN = C.CFUNCTYPE(_int, _vp, _vp, C.c_size_t, _int, _int, _vp)


class SpmmB200Error(RuntimeError):
    """A C-ABI call returned a negative code.  The reference's convention is
    print + exit(1) (lib/debug.h error()); in Python that becomes an exception."""

    def __init__(self, code, where):
        lib = load_library()
        msg = lib.spmm_b200_strerror(code).decode()
        detail = lib.spmm_b200_last_cuda_error().decode()
        super().__init__(f"{where}: {msg} (code {code})" + (f" [{detail}]" if detail and code in (-2, -3, -4, -6, -10) else ""))
        self.code = code


def load_library():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} not built: run `make -C spmm-research_b200/csrc` "
                          "(there is no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    lib.spmm_b200_create.argtypes = [C.POINTER(_vp), _vp, _vp, _vp, _i64, _i64, _i64, _int, _int, _int, _int]
    lib.spmm_b200_create_shard.argtypes = lib.spmm_b200_create.argtypes + [_int, _int, C.POINTER(_i64), C.POINTER(_i64)]
    lib.spmm_b200_run.argtypes = [_vp, _vp, _vp, _int, _int]
    lib.spmm_b200_run_replicated.argtypes = [_vp, _vp, _vp, _int, _int, BCAST_FN, _vp, _int]
    lib.spmm_b200_run_routed.argtypes = [_vp, _vp, _vp, _int, _int, _int, _vp, _vp, _vp, _int, XFER_FN, _vp]
    lib.spmm_b200_set_b.argtypes = [_vp, _vp, _int]
    lib.spmm_b200_register_host.argtypes = [_vp, _vp, C.c_size_t]
    lib.spmm_b200_unregister_host.argtypes = [_vp, _vp]
    lib.spmm_b200_set_option.argtypes = [_vp, C.c_char_p, _int]
    lib.spmm_b200_last_timeline.argtypes = [_vp, _vp, _int]
    lib.spmm_b200_create_multi.argtypes = [C.POINTER(_vp), _vp, _vp, _vp, _i64, _i64, _i64, _int, _int, _vp, _int, _int]
    lib.spmm_b200_run_multi.argtypes = [_vp, _vp, _vp, _int, _int]
    lib.spmm_b200_set_b_multi.argtypes = [_vp, _vp, _int]
    lib.spmm_b200_run_resident_multi.argtypes = [_vp, _int]
    lib.spmm_b200_fetch_c_multi.argtypes = [_vp, _vp, _int]
    lib.spmm_b200_register_host_multi.argtypes = [_vp, _vp, C.c_size_t]
    lib.spmm_b200_multi_ngpus.argtypes = [_vp]
    lib.spmm_b200_multi_routes.argtypes = [_vp, _vp, _int]
    lib.spmm_b200_multi_shard.argtypes = [_vp, _int]; lib.spmm_b200_multi_shard.restype = _vp
    lib.spmm_b200_multi_first_row.argtypes = [_vp, _int]; lib.spmm_b200_multi_first_row.restype = _i64
    lib.spmm_b200_multi_last_kernel_ms.argtypes = [_vp]; lib.spmm_b200_multi_last_kernel_ms.restype = _dbl
    lib.spmm_b200_multi_launch_count.argtypes = [_vp]; lib.spmm_b200_multi_launch_count.restype = _i64
    lib.spmm_b200_multi_peer_access.argtypes = [_vp]
    lib.spmm_b200_destroy_multi.argtypes = [_vp]; lib.spmm_b200_destroy_multi.restype = None
    lib.spmm_b200_run_resident.argtypes = [_vp, _int]
    lib.spmm_b200_fetch_c.argtypes = [_vp, _vp, _int]
    lib.spmm_b200_run_device.argtypes = [_vp, _vp, _vp, _int, _vp]
    lib.spmm_b200_spmm_merge_noalloc.argtypes = [_vp, _vp, _vp, _i64, _i64, _i64, _vp, _vp, _int, _vp]
    lib.spmm_b200_transpose_b_device.argtypes = [_vp, _vp, _vp, _int, _vp]
    lib.spmm_b200_partition.argtypes = [_vp, _i64, _i64, _int, _int, _vp, _int]
    lib.spmm_b200_partition_host.argtypes = [_vp, _i64, _i64, _int, _int, _vp]
    lib.spmm_b200_coo_to_csr.argtypes = [_vp, _vp, _vp, _i64, _i64, _i64, _vp, _vp, _vp, _int, _int, _int]
    lib.spmm_b200_rows.argtypes = [_vp]; lib.spmm_b200_rows.restype = _i64
    lib.spmm_b200_nnz.argtypes = [_vp]; lib.spmm_b200_nnz.restype = _i64
    lib.spmm_b200_variant.argtypes = [_vp]
    lib.spmm_b200_bcsr_blocks.argtypes = [_vp]; lib.spmm_b200_bcsr_blocks.restype = _i64
    lib.spmm_b200_kslice.argtypes = [_vp]
    lib.spmm_b200_last_rowband.argtypes = [_vp]
    lib.spmm_b200_last_transfer_bytes.argtypes = [_vp, C.POINTER(_i64), C.POINTER(_i64)]
    lib.spmm_b200_col_window.argtypes = [_vp, C.POINTER(_i64), C.POINTER(_i64)]
    lib.spmm_b200_last_kernel_ms.argtypes = [_vp]; lib.spmm_b200_last_kernel_ms.restype = _dbl
    lib.spmm_b200_launch_count.argtypes = [_vp]; lib.spmm_b200_launch_count.restype = _i64
    lib.spmm_b200_stats.argtypes = [_vp, C.c_char_p, C.c_long]
    lib.spmm_b200_stats_labels.argtypes = [C.c_char_p, C.c_long]
    lib.spmm_b200_destroy.argtypes = [_vp]; lib.spmm_b200_destroy.restype = None
    lib.spmm_b200_strerror.argtypes = [_int]; lib.spmm_b200_strerror.restype = C.c_char_p
    lib.spmm_b200_last_cuda_error.restype = C.c_char_p
    lib.spmm_b200_device_count.restype = _int
    _pp = C.POINTER(_vp)
    lib.spmm_b200_read_mtx.argtypes = [C.c_char_p, _int, C.POINTER(_i64), C.POINTER(_i64), C.POINTER(_i64), _pp, _pp, _pp, C.c_char_p]
    lib.spmm_b200_read_smtx.argtypes = [C.c_char_p, C.c_uint64, C.POINTER(_i64), C.POINTER(_i64), C.POINTER(_i64), _pp, _pp, _pp]
    lib.spmm_b200_free.argtypes = [_vp]; lib.spmm_b200_free.restype = None
    lib.spmm_b200_check_accuracy.argtypes = [_vp, _vp, _vp, _int, _vp, _vp]
    lib.spmm_b200_create_f32.argtypes = [C.POINTER(_vp), _vp, _vp, _vp, _i64, _i64, _i64, _int, _int, _int, _int]
    lib.spmm_b200_variant_f32.argtypes = [_vp]
    lib.spmm_b200_set_option_f32.argtypes = [_vp, C.c_char_p, _int]
    lib.spmm_b200_create_shard_f32.argtypes = lib.spmm_b200_create_f32.argtypes + [_int, _int, C.POINTER(_i64), C.POINTER(_i64)]
    lib.spmm_b200_run_f32.argtypes = [_vp, _vp, _vp, _int, _int]
    lib.spmm_b200_run_device_f32.argtypes = [_vp, _vp, _vp, _int, _vp]
    lib.spmm_b200_fetch_c_f32.argtypes = [_vp, _vp, _int]
    lib.spmm_b200_rows_f32.argtypes = [_vp]; lib.spmm_b200_rows_f32.restype = _i64
    lib.spmm_b200_nnz_f32.argtypes = [_vp]; lib.spmm_b200_nnz_f32.restype = _i64
    lib.spmm_b200_launch_count_f32.argtypes = [_vp]; lib.spmm_b200_launch_count_f32.restype = _i64
    lib.spmm_b200_last_kernel_ms_f32.argtypes = [_vp]; lib.spmm_b200_last_kernel_ms_f32.restype = _dbl
    lib.spmm_b200_destroy_f32.argtypes = [_vp]; lib.spmm_b200_destroy_f32.restype = None
    lib.spmm_b200_pipeline_create.argtypes = [C.POINTER(_vp), _vp, _vp, _vp, _i64, _i64, _int, _int]
    lib.spmm_b200_pipeline_spmm.argtypes = [_vp, C.c_char, _i64, _i64, _int, _vp, _vp, _vp, _vp, _vp]
    lib.spmm_b200_pipeline_sddmm.argtypes = [_vp, _vp, _int]
    lib.spmm_b200_pipeline_softmax.argtypes = [_vp, _vp, _int]
    lib.spmm_b200_pipeline_final.argtypes = [_vp, _vp]
    lib.spmm_b200_pipeline_set_weight.argtypes = [_vp, C.c_char, _i64, _i64, _i64, _vp, _vp, _vp]
    lib.spmm_b200_pipeline_attention.argtypes = [_vp, _vp, _vp, _int, _int]
    lib.spmm_b200_pipeline_invalidate.argtypes = [_vp, C.c_char]
    lib.spmm_b200_pipeline_set_option.argtypes = [_vp, C.c_char_p, _int]
    lib.spmm_b200_pipeline_device_ptr.argtypes = [_vp, C.c_char]; lib.spmm_b200_pipeline_device_ptr.restype = _vp
    lib.spmm_b200_pipeline_last_kernel_ms.argtypes = [_vp]; lib.spmm_b200_pipeline_last_kernel_ms.restype = _dbl
    lib.spmm_b200_pipeline_launch_count.argtypes = [_vp]; lib.spmm_b200_pipeline_launch_count.restype = _i64
    lib.spmm_b200_pipeline_destroy.argtypes = [_vp]; lib.spmm_b200_pipeline_destroy.restype = None
    lib.spmm_b200_pipeline_create_f32.argtypes = lib.spmm_b200_pipeline_create.argtypes
    lib.spmm_b200_pipeline_spmm_f32.argtypes = lib.spmm_b200_pipeline_spmm.argtypes
    lib.spmm_b200_pipeline_sddmm_f32.argtypes = [_vp, _vp, _int]
    lib.spmm_b200_pipeline_final_f32.argtypes = [_vp, _vp]
    lib.spmm_b200_last_h2d_bytes_f32.argtypes = [_vp]; lib.spmm_b200_last_h2d_bytes_f32.restype = _i64
    lib.spmm_b200_pipeline_softmax_f32.argtypes = [_vp, _vp, _int]
    lib.spmm_b200_pipeline_launch_count_f32.argtypes = [_vp]; lib.spmm_b200_pipeline_launch_count_f32.restype = _i64
    lib.spmm_b200_pipeline_destroy_f32.argtypes = [_vp]; lib.spmm_b200_pipeline_destroy_f32.restype = None
    _lib = lib
    return lib


def _check(code, where):
    if code != 0:
        raise SpmmB200Error(code, where)


def _host(arr, dtype, name):
    if not isinstance(arr, np.ndarray) or arr.dtype != dtype or not arr.flags.c_contiguous:
        raise TypeError(f"{name} must be a C-contiguous numpy array of {np.dtype(dtype).name}")
    return arr.ctypes.data


class MatrixFormat:
    """Mirror of `struct Matrix_Format` (spmv_kernel.h:9-26): fields format_name, m, n,
    nnz, mem_footprint, csr_mem_footprint; methods spmm, statistics_start,
    statistics_print_data."""

    def __init__(self, handle, m, n, nnz, k, first_row=0, m_global=None):
        self._h = handle
        self.m, self.n, self.nnz, self.k = int(m), int(n), int(nnz), int(k)
        self.first_row = int(first_row)
        self.m_global = int(m if m_global is None else m_global)
        lib = load_library()
        self.variant = lib.spmm_b200_variant(handle)
        self.format_name = "B200_CSR_" + {ROWSPLIT: "ROWSPLIT", MERGE: "MERGE", BCSR: "BCSR_DMMA", BCSR_FMA: "BCSR_FMA"}.get(self.variant, "?")
        # spmv_kernel.h:24
        self.csr_mem_footprint = self.nnz * (8 + 4) + (self.m + 1) * 4
        self.mem_footprint = self.csr_mem_footprint

    # -- the hot call: Matrix_Format::spmm(x, y, k) with HOST buffers ------------------
    def spmm(self, x, y, k, copy_back=True):
        """y = A @ x.  x: float64[c*k] in the oracle layout x[n*c + j]; y: float64[m*k]
        row-major, fully overwritten (spmm_kernel_csr.cpp:88,93)."""
        k = int(k)
        if x.size != self.n * k:
            raise ValueError(f"x holds {x.size} values, expected c*k = {self.n * k}")
        if copy_back and y.size != self.m * k:
            raise ValueError(f"y holds {y.size} values, expected m*k = {self.m * k}")
        xp = _host(x, np.float64, "x")
        yp = _host(y, np.float64, "y") if copy_back else None
        _check(load_library().spmm_b200_run(self._h, xp, yp, k, 1 if copy_back else 0), "spmm")

    def spmm_replicated(self, x, y, k, bcast, copy_back=True, x_on_every_rank=False):
        """Multi-GPU spmm().  x is the host B on the root rank (None elsewhere), or -- with
        x_on_every_rank -- the same host B on every rank, of which each rank uploads only its part.
        `bcast(dev_ptr, count, parts, stream)` must enqueue the replication of `count` doubles at
        `dev_ptr` on the CUDA stream `stream` (parts == 1: broadcast from rank 0; parts == P: part r
        of ceil(count/P) doubles comes from rank r), see include/spmm_b200.h."""
        k = int(k)
        xp = _host(x, np.float64, "x") if x is not None else None
        yp = _host(y, np.float64, "y") if copy_back else None
        err = []

        def _cb(_ctx, ptr, count, parts, stream):
            try:
                bcast(ptr, count, parts, stream)
                return 0
            except Exception as e:      # an exception must not unwind through the C frames
                err.append(e)
                return 1
        cb = BCAST_FN(_cb) if bcast is not None else C.cast(None, BCAST_FN)
        rc = load_library().spmm_b200_run_replicated(self._h, xp, yp, k, 1 if copy_back else 0, cb, None, 1 if x_on_every_rank else 0)
        if err:
            raise err[0]
        _check(rc, "spmm_replicated")

    def spmm_routed(self, x, y_all, k, my_rank, downloader_rank, clients, xfer):
        """One process per GPU with routed downloads (spmm_b200_run_routed): y_all is the ONE m_global x k result every
        rank shares (shared memory, page-locked by every rank).  A rank whose downloader_rank is another rank sends its
        rows of C there; a downloader stores its own rows and those of `clients` = [(rank, first_row, rows), ...] into
        y_all.  `xfer(dev_ptr, count, peer_rank, is_send, stream)` enqueues the send / receive on `stream`."""
        err = []

        def _cb(_ctx, ptr, count, peer, is_send, stream):
            try:
                xfer(ptr, count, peer, bool(is_send), stream)
                return 0
            except Exception as e:      # an exception must not unwind through the C frames
                err.append(e)
                return 1
        cb = XFER_FN(_cb) if xfer is not None else C.cast(None, XFER_FN)
        ranks = np.array([c[0] for c in clients], np.int32)
        first = np.array([c[1] for c in clients], np.int64)
        rows = np.array([c[2] for c in clients], np.int64)
        rc = load_library().spmm_b200_run_routed(self._h, _host(x, np.float64, "x"), _host(y_all, np.float64, "y_all"), int(k),
                                                 int(my_rank), int(downloader_rank), ranks.ctypes.data, first.ctypes.data,
                                                 rows.ctypes.data, len(clients), cb, None)
        if err:
            raise err[0]
        _check(rc, "spmm_routed")

    ACCURACY_LABELS = ("max_rel_diff", "mae", "max_ae", "mse", "mape", "smape", "lnQ_error", "mlare", "gmare", "passed")

    def check_accuracy(self, x, y, k, return_gold=False):
        """Mirror of CheckAccuracy() (spmv_bench.cpp:121-206): gold y = A*x on the device in
        double-double and the reference's error metrics of y against it -> dict (+ gold)."""
        k = int(k)
        stats = np.zeros(10)
        gold = np.zeros(max(self.m * k, 1)) if return_gold else None
        _check(load_library().spmm_b200_check_accuracy(self._h, _host(x, np.float64, "x"), _host(y, np.float64, "y"), k,
                                                       stats.ctypes.data, gold.ctypes.data if return_gold else None),
               "check_accuracy")
        out = dict(zip(self.ACCURACY_LABELS, stats.tolist()))
        return (out, gold[: self.m * k]) if return_gold else out

    def register_host(self, arr):
        """Page-lock a host array for the lifetime of the handle (full PCIe rate; lets spmm() store C straight
        into it).  The array must stay alive until close()."""
        _check(load_library().spmm_b200_register_host(self._h, arr.ctypes.data, arr.nbytes), "spmm_b200_register_host")

    def set_option(self, name, value):
        """Change a tuning knob of this handle (the SPMM_B200_<NAME> variables are read once, at creation)."""
        _check(load_library().spmm_b200_set_option(self._h, name.encode(), int(value)), f"set_option({name})")

    def last_timeline(self):
        """Event times (ms since the first enqueue) of the last k-sliced host-buffer run:
        [{h2d, bcast, transpose, kernel_start, kernel_end, d2h}] per slice."""
        out = np.zeros(1 + 6 * 64)
        n = load_library().spmm_b200_last_timeline(self._h, out.ctypes.data, out.size)
        if n < 1:
            return []
        keys = ("h2d", "bcast", "transpose", "kernel_start", "kernel_end", "d2h")
        return [dict(zip(keys, out[1 + 6 * s: 7 + 6 * s].tolist())) for s in range(int(out[0]))]

    def set_b(self, x, k):
        _check(load_library().spmm_b200_set_b(self._h, _host(x, np.float64, "x"), int(k)), "set_b")

    def spmm_resident(self, k):
        _check(load_library().spmm_b200_run_resident(self._h, int(k)), "spmm_resident")

    def fetch_c(self, y, k):
        _check(load_library().spmm_b200_fetch_c(self._h, _host(y, np.float64, "y"), int(k)), "fetch_c")

    # -- device-pointer form for callers that own device memory / streams ---------------
    def spmm_device(self, b_ptr, c_ptr, k, stream=0):
        """b_ptr: device address of row-major B (c x k); c_ptr: device address of row-major
        C (m_local x k); stream: cudaStream_t as int (0 = default).  Asynchronous."""
        _check(load_library().spmm_b200_run_device(self._h, b_ptr, c_ptr, int(k), stream or None), "spmm_device")

    def transpose_b_device(self, x_ptr, b_ptr, k, stream=0):
        _check(load_library().spmm_b200_transpose_b_device(self._h, x_ptr, b_ptr, int(k), stream or None),
               "transpose_b_device")

    def bcsr_fill(self):
        """nnz / (8 * stored 8x1 blocks); None for the CSR variants."""
        nb = load_library().spmm_b200_bcsr_blocks(self._h)
        return self.nnz / (8.0 * nb) if nb > 0 else None

    def last_kernel_ms(self):
        return load_library().spmm_b200_last_kernel_ms(self._h)

    def launch_count(self):
        return load_library().spmm_b200_launch_count(self._h)

    def kslice(self):
        """Column-slice width the row-split path settled on (0: none / still measuring)."""
        return load_library().spmm_b200_kslice(self._h)

    def last_rowband(self):
        """Row blocks of the last host-buffer spmm() when it ran the row-band pipeline (banded matrix), else 0."""
        return load_library().spmm_b200_last_rowband(self._h)

    def col_window(self):
        """(first touched column, number of columns up to the last touched one): the rows of B this handle reads."""
        lo, n = _i64(0), _i64(0)
        _check(load_library().spmm_b200_col_window(self._h, C.byref(lo), C.byref(n)), "col_window")
        return int(lo.value), int(n.value)

    def last_transfer_bytes(self):
        """(host->device, device->host) bytes the last host-buffer spmm() copied: only the rows of B between the first
        and the last column the matrix touches are uploaded."""
        up, down = _i64(0), _i64(0)
        _check(load_library().spmm_b200_last_transfer_bytes(self._h, C.byref(up), C.byref(down)), "last_transfer_bytes")
        return int(up.value), int(down.value)

    def statistics_start(self):
        pass

    def statistics_print_data(self, buf_n=512):
        buf = C.create_string_buffer(buf_n)
        load_library().spmm_b200_stats(self._h, buf, buf_n)
        return buf.value.decode()

    def close(self):
        if self._h:
            load_library().spmm_b200_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class MatrixFormatF32:
    """fp32 twin of MatrixFormat: the reference's "_f" builds (ValueType = float, make.sh:98-108)."""

    def __init__(self, handle, m, n, nnz, k, first_row=0):
        self._h = handle
        self.m, self.n, self.nnz, self.k, self.first_row = int(m), int(n), int(nnz), int(k), int(first_row)
        self.variant = load_library().spmm_b200_variant_f32(handle)
        self.format_name = {ROWSPLIT: "B200_CSR_ROWSPLIT_F32", MERGE: "B200_CSR_MERGE_F32", BCSR: "B200_BLOCKED_TF32"}.get(self.variant, "B200_CSR_F32")
        self.csr_mem_footprint = self.nnz * (4 + 4) + (self.m + 1) * 4
        self.mem_footprint = self.csr_mem_footprint

    def spmm(self, x, y, k, copy_back=True):
        k = int(k)
        if x.size != self.n * k:
            raise ValueError(f"x holds {x.size} values, expected c*k = {self.n * k}")
        if copy_back and y.size != self.m * k:
            raise ValueError(f"y holds {y.size} values, expected m*k = {self.m * k}")
        _check(load_library().spmm_b200_run_f32(self._h, _host(x, np.float32, "x"), _host(y, np.float32, "y") if copy_back else None,
                                                k, 1 if copy_back else 0), "spmm(f32)")

    def last_h2d_bytes(self):
        """Bytes of B the last host-buffer spmm() uploaded (rows between the first and the last touched column)."""
        return int(load_library().spmm_b200_last_h2d_bytes_f32(self._h))

    def spmm_device(self, b_ptr, c_ptr, k, stream=0):
        _check(load_library().spmm_b200_run_device_f32(self._h, b_ptr, c_ptr, int(k), stream or None), "spmm_device(f32)")

    def set_option(self, name, value):
        _check(load_library().spmm_b200_set_option_f32(self._h, name.encode(), int(value)), f"set_option_f32({name})")

    def fetch_c(self, y, k):
        _check(load_library().spmm_b200_fetch_c_f32(self._h, _host(y, np.float32, "y"), int(k)), "fetch_c(f32)")

    def last_kernel_ms(self):
        return load_library().spmm_b200_last_kernel_ms_f32(self._h)

    def launch_count(self):
        return load_library().spmm_b200_launch_count_f32(self._h)

    def statistics_start(self):
        pass

    def close(self):
        if self._h:
            load_library().spmm_b200_destroy_f32(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


SDDMM_REFERENCE, SDDMM_ATTENTION, SDDMM_ATTENTION_ORDERED = 0, 1, 2


class PipelineFormat:
    """Mirror of the pipeline bench's `struct Matrix_Format` (pipeline_code_bench/sddmm_kernel.h:9-31):
    fields m, n, nnz; methods spmm(type, m, k, n, ia, ja, a, x, y), sddmm(y).  Built from the attention
    mask by csr_to_pipeline_format (csr_to_format(row_ptr, col_ind, values, m, nnz, n), sddmm_kernel.h:31).
    Host numpy arrays in and out; K, Q, V, y stay on the device between calls."""

    format_name = "B200_CSR_PIPELINE"

    def __init__(self, handle, m, nnz, n, keep):
        self._h = handle
        self.m, self.nnz, self.n = int(m), int(nnz), int(n)
        self._keep = {"mask": tuple(keep)}   # the C side remembers host pointers: keep the arrays alive

    def spmm(self, type, m, k, n, ia, ja, a, x, y):
        """y = CSR(ia, ja, a; m x k) * x (k x n row-major).  type 'K'|'Q'|'V' also keeps the result on the
        device for sddmm(); any other type (the harness passes 'final') is a plain product."""
        t = type.encode()[:1] if isinstance(type, str) else bytes(type)[:1]
        yp = None if y is None else _host(y, np.float64, "y")
        if y is not None and y.size < m * n:
            raise ValueError("y holds fewer than m*n values")
        if x.size < k * n:
            raise ValueError("x holds fewer than k*n values")
        if t in (b"K", b"Q", b"V"):
            self._keep[t] = (ia, ja, a)
        _check(load_library().spmm_b200_pipeline_spmm(self._h, t, m, k, n, _host(ia, np.int32, "ia"),
               _host(ja, np.int32, "ja") if len(ja) else None, _host(a, np.float64, "a") if len(a) else None,
               _host(x, np.float64, "x") if x.size else None, yp), "spmm_b200_pipeline_spmm")

    def sddmm(self, y=None, mode=SDDMM_REFERENCE):
        if y is not None and y.size < self.nnz:
            raise ValueError("y holds fewer than nnz values")
        _check(load_library().spmm_b200_pipeline_sddmm(self._h, None if y is None else _host(y, np.float64, "y"), mode),
               "spmm_b200_pipeline_sddmm")

    def softmax(self, y=None, rowwise=False):
        _check(load_library().spmm_b200_pipeline_softmax(self._h, None if y is None else _host(y, np.float64, "y"), int(bool(rowwise))),
               "spmm_b200_pipeline_softmax")

    def final(self, y_final):
        if y_final.size < self.m * self.n:
            raise ValueError("y_final holds fewer than m*n values")
        _check(load_library().spmm_b200_pipeline_final(self._h, _host(y_final, np.float64, "y_final")), "spmm_b200_pipeline_final")

    def set_weight(self, type, m, k, ia, ja, a):
        self._keep[type.encode()[:1]] = (ia, ja, a)
        _check(load_library().spmm_b200_pipeline_set_weight(self._h, type.encode()[:1], m, k, len(ja), _host(ia, np.int32, "ia"),
               _host(ja, np.int32, "ja") if len(ja) else None, _host(a, np.float64, "a") if len(a) else None),
               "spmm_b200_pipeline_set_weight")

    def invalidate(self, type="*"):
        """Forget the resident copy of weight 'K' | 'Q' | 'V' ('*': all): for callers that refill a weight in place."""
        _check(load_library().spmm_b200_pipeline_invalidate(self._h, type.encode()[:1]), "spmm_b200_pipeline_invalidate")

    def set_option(self, name, value):
        _check(load_library().spmm_b200_pipeline_set_option(self._h, name.encode(), int(value)), f"pipeline set_option({name})")

    def attention(self, x, y_final, mode=SDDMM_REFERENCE, softmax=0):
        """The whole layer, one upload of x and one download of y_final."""
        _check(load_library().spmm_b200_pipeline_attention(self._h, _host(x, np.float64, "x") if x.size else None,
               _host(y_final, np.float64, "y_final"), mode, softmax), "spmm_b200_pipeline_attention")

    def device_ptr(self, which):
        return load_library().spmm_b200_pipeline_device_ptr(self._h, which.encode()[:1])

    def last_kernel_ms(self):
        return load_library().spmm_b200_pipeline_last_kernel_ms(self._h)

    def launch_count(self):
        return load_library().spmm_b200_pipeline_launch_count(self._h)

    def close(self):
        if self._h:
            load_library().spmm_b200_pipeline_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def csr_to_pipeline_format(row_ptr, col_ind, values, m, nnz, n, *, device=0):
    """csr_to_format(row_ptr, col_ind, values, m, nnz, n) of the pipeline bench (sddmm_kernel.h:31): the format
    is built from the m x m attention mask; n is the dense width."""
    h = _vp()
    _check(load_library().spmm_b200_pipeline_create(C.byref(h), _host(row_ptr, np.int32, "row_ptr"),
           _host(col_ind, np.int32, "col_ind") if nnz else None, _host(values, np.float64, "values") if nnz else None,
           m, nnz, n, device), "spmm_b200_pipeline_create")
    return PipelineFormat(h, m, nnz, n, (row_ptr, col_ind, values))


class PipelineFormatF32:
    """PipelineFormat for the pipeline bench's "_f" build (float32 arrays): spmm, sddmm, final."""

    format_name = "B200_CSR_PIPELINE_F32"

    def __init__(self, handle, m, nnz, n, keep):
        self._h = handle
        self.m, self.nnz, self.n = int(m), int(nnz), int(n)
        self._keep = {"mask": tuple(keep)}

    def spmm(self, type, m, k, n, ia, ja, a, x, y):
        t = type.encode()[:1] if isinstance(type, str) else bytes(type)[:1]
        if y is not None and y.size < m * n:
            raise ValueError("y holds fewer than m*n values")
        if x.size < k * n:
            raise ValueError("x holds fewer than k*n values")
        if t in (b"K", b"Q", b"V"):
            self._keep[t] = (ia, ja, a)
        _check(load_library().spmm_b200_pipeline_spmm_f32(self._h, t, m, k, n, _host(ia, np.int32, "ia"),
               _host(ja, np.int32, "ja") if len(ja) else None, _host(a, np.float32, "a") if len(a) else None,
               _host(x, np.float32, "x") if x.size else None, None if y is None else _host(y, np.float32, "y")),
               "spmm_b200_pipeline_spmm_f32")

    def sddmm(self, y=None, mode=SDDMM_REFERENCE):
        if y is not None and y.size < self.nnz:
            raise ValueError("y holds fewer than nnz values")
        _check(load_library().spmm_b200_pipeline_sddmm_f32(self._h, None if y is None else _host(y, np.float32, "y"), mode),
               "spmm_b200_pipeline_sddmm_f32")

    def softmax(self, y=None, rowwise=True):
        """softmax of y in place on the device (fp32, sums in double); rowwise=False: over all nonzeros at once."""
        if y is not None and y.size < self.nnz:
            raise ValueError("y holds fewer than nnz values")
        _check(load_library().spmm_b200_pipeline_softmax_f32(self._h, None if y is None else _host(y, np.float32, "y"), 1 if rowwise else 0),
               "spmm_b200_pipeline_softmax_f32")

    def final(self, y_final):
        if y_final.size < self.m * self.n:
            raise ValueError("y_final holds fewer than m*n values")
        _check(load_library().spmm_b200_pipeline_final_f32(self._h, _host(y_final, np.float32, "y_final")),
               "spmm_b200_pipeline_final_f32")

    def launch_count(self):
        return load_library().spmm_b200_pipeline_launch_count_f32(self._h)

    def close(self):
        if self._h:
            load_library().spmm_b200_pipeline_destroy_f32(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def csr_to_pipeline_format_f32(row_ptr, col_ind, values, m, nnz, n, *, device=0):
    h = _vp()
    _check(load_library().spmm_b200_pipeline_create_f32(C.byref(h), _host(row_ptr, np.int32, "row_ptr"),
           _host(col_ind, np.int32, "col_ind") if nnz else None, _host(values, np.float32, "values") if nnz else None,
           m, nnz, n, device), "spmm_b200_pipeline_create_f32")
    return PipelineFormatF32(h, m, nnz, n, (row_ptr, col_ind, values))


NOALLOC_SCRATCH_BYTES = 8 << 20
ESCRATCH = -10


def spmm_merge_noalloc(row_ptr_dev, col_ind_dev, values_dev, m, c, nnz, b_dev, c_dev, k, stream=0):
    """Allocation-free merge-path SpMM on the current device (spmm_b200_spmm_merge_noalloc; no counterpart in the
    reference).  Every array is a DEVICE address (int); B row-major c x k, C row-major m x k.  Asynchronous.  Raises
    SpmmB200Error with code ESCRATCH when the input needs more than NOALLOC_SCRATCH_BYTES of scratch."""
    _check(load_library().spmm_b200_spmm_merge_noalloc(row_ptr_dev or None, col_ind_dev or None, values_dev or None, int(m), int(c), int(nnz),
                                                       b_dev or None, c_dev or None, int(k), stream or None), "spmm_merge_noalloc")


def csr_to_format_f32(row_ptr, col_ind, values, m, n, nnz, k, *, device=0, variant=AUTO, b_layout=B_COLMAJOR, rank=None, nranks=None):
    """csr_to_format() of an "_f" build: values float32.  variant: ROWSPLIT | MERGE | AUTO | BCSR (= the TF32
    tensor-core kernel for banded matrices, ~1e-3 relative accuracy; never picked by AUTO)."""
    lib = load_library()
    h = _vp()
    rp = _host(row_ptr, np.int32, "row_ptr")
    ci = _host(col_ind, np.int32, "col_ind") if nnz > 0 else None
    va = _host(values, np.float32, "values") if nnz > 0 else None
    if nranks is None:
        _check(lib.spmm_b200_create_f32(C.byref(h), rp, ci, va, m, n, nnz, k, device, variant, b_layout), "csr_to_format_f32")
        return MatrixFormatF32(h, m, n, nnz, k)
    r0, r1 = _i64(), _i64()
    _check(lib.spmm_b200_create_shard_f32(C.byref(h), rp, ci, va, m, n, nnz, k, device, variant, b_layout, int(rank), int(nranks),
                                          C.byref(r0), C.byref(r1)), "csr_to_format_f32(shard)")
    return MatrixFormatF32(h, lib.spmm_b200_rows_f32(h), n, lib.spmm_b200_nnz_f32(h), k, first_row=r0.value)


def csr_to_format(row_ptr, col_ind, values, m, n, nnz, k, *, device=0, variant=AUTO,
                  b_layout=B_COLMAJOR, rank=None, nranks=None):
    """Mirror of csr_to_format() (spmv_kernel.h:29).  With rank/nranks the handle holds
    only that rank's nnz-balanced row shard."""
    lib = load_library()
    h = _vp()
    rp = _host(row_ptr, np.int32, "row_ptr")
    ci = _host(col_ind, np.int32, "col_ind") if nnz > 0 else None
    va = _host(values, np.float64, "values") if nnz > 0 else None
    if row_ptr.size < m + 1:
        raise ValueError("row_ptr shorter than m+1")
    if nranks is None:
        _check(lib.spmm_b200_create(C.byref(h), rp, ci, va, m, n, nnz, k, device, variant, b_layout), "csr_to_format")
        return MatrixFormat(h, m, n, nnz, k)
    r0, r1 = _i64(), _i64()
    _check(lib.spmm_b200_create_shard(C.byref(h), rp, ci, va, m, n, nnz, k, device, variant, b_layout,
                                      int(rank), int(nranks), C.byref(r0), C.byref(r1)), "csr_to_format(shard)")
    return MatrixFormat(h, lib.spmm_b200_rows(h), n, lib.spmm_b200_nnz(h), k, first_row=r0.value, m_global=m)


class MultiGpuFormat:
    """Matrix_Format over `ngpus` devices driven by ONE process (include/spmm_b200.h, "_multi"): rows cut by the
    reference's nnz-balanced split, B replicated by a peer-memory gather kernel, C row-local in the caller's y."""

    def __init__(self, handle, m, n, nnz, k):
        self._h = handle
        self.m, self.n, self.nnz, self.k = int(m), int(n), int(nnz), int(k)
        lib = load_library()
        self.ngpus = lib.spmm_b200_multi_ngpus(handle)
        self.bounds = [lib.spmm_b200_multi_first_row(handle, g) for g in range(self.ngpus + 1)]
        self.peer_access = bool(lib.spmm_b200_multi_peer_access(handle))
        self.format_name = f"B200_CSR_x{self.ngpus}"
        self.csr_mem_footprint = self.nnz * (8 + 4) + (self.m + 1) * 4
        self.mem_footprint = self.csr_mem_footprint

    def spmm(self, x, y, k, copy_back=True):
        k = int(k)
        if x.size != self.n * k:
            raise ValueError(f"x holds {x.size} values, expected c*k = {self.n * k}")
        if copy_back and y.size != self.m * k:
            raise ValueError(f"y holds {y.size} values, expected m*k = {self.m * k}")
        _check(load_library().spmm_b200_run_multi(self._h, _host(x, np.float64, "x"), _host(y, np.float64, "y") if copy_back else None,
                                                  k, 1 if copy_back else 0), "spmm(multi)")

    def set_b(self, x, k):
        _check(load_library().spmm_b200_set_b_multi(self._h, _host(x, np.float64, "x"), int(k)), "set_b(multi)")

    def spmm_resident(self, k):
        _check(load_library().spmm_b200_run_resident_multi(self._h, int(k)), "spmm_resident(multi)")

    def fetch_c(self, y, k):
        _check(load_library().spmm_b200_fetch_c_multi(self._h, _host(y, np.float64, "y"), int(k)), "fetch_c(multi)")

    def register_host(self, arr):
        _check(load_library().spmm_b200_register_host_multi(self._h, arr.ctypes.data, arr.nbytes), "register_host(multi)")

    def set_option(self, name, value):
        lib = load_library()
        for g in range(self.ngpus):
            _check(lib.spmm_b200_set_option(lib.spmm_b200_multi_shard(self._h, g), name.encode(), int(value)), f"set_option({name})")

    def routes(self):
        """(route, probe): route[g] = device that downloads device g's rows of C; probe = measured GB/s of the candidate
        downloader sets {all, lower half, upper half, evens, odds} (zeros when nothing was measured)."""
        out = np.zeros(self.ngpus + 5)
        n = load_library().spmm_b200_multi_routes(self._h, out.ctypes.data, out.size)
        _check(min(n, 0), "multi_routes")
        return [int(v) for v in out[:self.ngpus]], out[self.ngpus:].tolist()

    def shard_variant(self, g):
        lib = load_library()
        return lib.spmm_b200_variant(lib.spmm_b200_multi_shard(self._h, g))

    def last_kernel_ms(self):
        return load_library().spmm_b200_multi_last_kernel_ms(self._h)

    def launch_count(self):
        return load_library().spmm_b200_multi_launch_count(self._h)

    def statistics_start(self):
        pass

    def close(self):
        if self._h:
            load_library().spmm_b200_destroy_multi(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def csr_to_format_multi(row_ptr, col_ind, values, m, n, nnz, k, ngpus, *, devices=None, variant=AUTO, b_layout=B_COLMAJOR):
    """csr_to_format() over `ngpus` devices of this box, one process (`int ngpus`)."""
    lib = load_library()
    h = _vp()
    dv = None
    if devices is not None:
        dv = (C.c_int * ngpus)(*[int(d) for d in devices])
    _check(lib.spmm_b200_create_multi(C.byref(h), _host(row_ptr, np.int32, "row_ptr"),
                                      _host(col_ind, np.int32, "col_ind") if nnz > 0 else None,
                                      _host(values, np.float64, "values") if nnz > 0 else None,
                                      m, n, nnz, k, int(ngpus), C.cast(dv, _vp) if dv is not None else None, variant, b_layout),
           "csr_to_format_multi")
    return MultiGpuFormat(h, m, n, nnz, k)


def partition(row_ptr, m, nnz, P, mode, device=0, host=False):
    """Bit-exact twins of the reference partitioners (include/spmm_b200.h), searched on the device
    (host=True: the same integer functions on the CPU, for shard planning).
    PART_NNZ -> int64[P+1] row starts; PART_ROWS -> int64[P,2]; PART_MERGE -> int64[P+1,2];
    PART_NNZ_CARRY -> int64[P,4] = (i_s, i_e, j_s, j_e) per worker (spmv_kernel_csr.cpp:68-80)."""
    n_out = {PART_NNZ: P + 1, PART_ROWS: 2 * P, PART_MERGE: 2 * (P + 1), PART_NNZ_CARRY: 4 * P}[mode]
    out = np.zeros(n_out, dtype=np.int64)
    rp = _host(row_ptr, np.int32, "row_ptr") if row_ptr is not None else None
    if host:
        _check(load_library().spmm_b200_partition_host(rp, m, nnz, P, mode, out.ctypes.data), "partition_host")
    else:
        _check(load_library().spmm_b200_partition(rp, m, nnz, P, mode, out.ctypes.data, device), "partition")
    if mode == PART_NNZ:
        return out
    return out.reshape(-1, 4) if mode == PART_NNZ_CARRY else out.reshape(-1, 2)


def shard_slices(row_ptr, m, nnz, nranks):
    """Row range and nnz range of every rank under the reference's nnz-balanced split
    (lib/parallel_util.h:141-165): [(r0, r1, j0, j1)] * nranks."""
    b = partition(row_ptr, m, nnz, nranks, PART_NNZ, host=True)
    out = []
    for r in range(nranks):
        r0, r1 = int(b[r]), int(max(b[r + 1], b[r]))
        out.append((r0, r1, int(row_ptr[r0]), int(row_ptr[r1])))
    return out


def coo_to_csr(R, C, V, m, n, nnz, sort_columns=1, transpose=0, device=0):
    """Mirror of coo_to_csr() (lib/storage_formats/csr/csr_gen.c:162-217), conversion on the device.
    Returns (row_ptr int32[m+1 | n+1], col_idx int32[nnz], values f64[nnz] | None)."""
    nrows = n if transpose else m
    row_ptr = np.zeros(nrows + 1, dtype=np.int32)
    col_idx = np.zeros(max(nnz, 1), dtype=np.int32)[:nnz]
    values = np.zeros(max(nnz, 1), dtype=np.float64)[:nnz] if V is not None else None
    rp = _host(R, np.int32, "R") if nnz > 0 else None
    cp = _host(C, np.int32, "C") if nnz > 0 else None
    vp = _host(V, np.float64, "V") if (V is not None and nnz > 0) else None
    _check(load_library().spmm_b200_coo_to_csr(
        rp, cp, vp, m, n, nnz, row_ptr.ctypes.data, col_idx.ctypes.data if nnz > 0 else None,
        values.ctypes.data if (values is not None and nnz > 0) else None, int(sort_columns), int(transpose), device),
        "coo_to_csr")
    return row_ptr, col_idx, values


def _take(ptr, n, dtype):
    """Copy a library-owned array into numpy and release it with spmm_b200_free."""
    n = int(n)
    ctype = {np.int32: C.c_int32, np.float64: C.c_double}[dtype]
    out = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(ctype)), shape=(max(n, 1),))[:n].astype(dtype, copy=True)
    load_library().spmm_b200_free(ptr)
    return out


def read_mtx(path, expand_symmetry=1):
    """Mirror of mtx_read(file, expand_symmetry, 1) + the harness's value conversion
    (matrix_market.c:249-322, spmv_bench.cpp:723-763).  Returns (m, n, R, C, V, field): 0-based COO, fp64 values."""
    m, n, nnz = _i64(), _i64(), _i64()
    R, Cc, V = _vp(), _vp(), _vp()
    field = C.create_string_buffer(16)
    _check(load_library().spmm_b200_read_mtx(os.fsencode(path), int(expand_symmetry), C.byref(m), C.byref(n), C.byref(nnz),
                                             C.byref(R), C.byref(Cc), C.byref(V), field), f"read_mtx({path})")
    return (m.value, n.value, _take(R, nnz.value, np.int32), _take(Cc, nnz.value, np.int32), _take(V, nnz.value, np.float64),
            field.value.decode())


def read_smtx(path, seed=0):
    """Mirror of smtx_read (dlcm_matrix.c:257-324): DLMC .smtx -> (m, n, row_ptr, col_ind, values); the file has no
    values, they are uniform [-1, 1) from `seed` (the reference seeds rand() with time(NULL))."""
    m, n, nnz = _i64(), _i64(), _i64()
    ia, ja, a = _vp(), _vp(), _vp()
    _check(load_library().spmm_b200_read_smtx(os.fsencode(path), int(seed), C.byref(m), C.byref(n), C.byref(nnz),
                                              C.byref(ia), C.byref(ja), C.byref(a)), f"read_smtx({path})")
    return m.value, n.value, _take(ia, m.value + 1, np.int32), _take(ja, nnz.value, np.int32), _take(a, nnz.value, np.float64)


def statistics_print_labels(buf_n=512):
    buf = C.create_string_buffer(buf_n)
    load_library().spmm_b200_stats_labels(buf, buf_n)
    return buf.value.decode()
