"""ctypes loaders for the checker libraries.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py (cpu_baseline leg,
--impl reference, and synthetic-input generation with the reference's own
generator) may import this module.  The product (spmm-research_b200/) never does.

  oracle()   -> oracle/liboracle.so        our C restatement (oracle/spmm_oracle.c)
  ref()      -> oracle/_ref/libspmm_ref.so the reference's compute_csr + partitioner macros
  merge_ref()-> oracle/_ref/libmerge_ref.so the reference's CPU merge-path code
  amg()      -> oracle/_ref/libamg_ref.so  the reference's artificial-matrix-generator
"""
import ctypes as C
import os
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
_f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
_cache = {}


def _load(path):
    if path not in _cache:
        if not os.path.exists(path):
            raise FileNotFoundError(
                f"{path} missing: run `make -C oracle oracle ref` (ref needs /root/reference)")
        _cache[path] = C.CDLL(path)
    return _cache[path]


def have_ref():
    return all(os.path.exists(os.path.join(HERE, "_ref", f))
               for f in ("libspmm_ref.so", "libamg_ref.so", "libmerge_ref.so"))


def oracle():
    lib = _load(os.path.join(HERE, "liboracle.so"))
    if not hasattr(lib, "_typed"):
        lib.oracle_spmm_csr.argtypes = [_i32p, _i32p, _f64p, C.c_long, C.c_long, _f64p, _f64p, C.c_int]
        lib.oracle_spmm_csr.restype = None
        lib.oracle_spmm_merge.argtypes = [_i32p, _i32p, _f64p, C.c_long, C.c_long, _f64p, _f64p, C.c_int, C.c_int]
        lib.oracle_spmm_merge.restype = None
        lib.oracle_balance_prefix_sums.argtypes = [_i32p, C.c_long, C.c_long, C.c_long, _i64p]
        lib.oracle_balance_iterations.argtypes = [C.c_long, C.c_long, C.c_long, _i64p]
        lib.oracle_merge_partition.argtypes = [_i32p, C.c_long, C.c_long, C.c_long, _i64p]
        lib.oracle_nnz_carry_partition.argtypes = [_i32p, C.c_long, C.c_long, C.c_long, _i64p]
        lib.oracle_nnz_carry_partition.restype = None
        lib.oracle_spmv_nnz_carry.argtypes = [_i32p, _i32p, _f64p, C.c_long, C.c_long, C.c_long, _f64p, _f64p]
        lib.oracle_spmv_nnz_carry.restype = None
        lib.oracle_coo_to_csr.argtypes = [_i32p, _i32p, C.c_void_p, C.c_long, C.c_long, C.c_long,
                                          _i32p, _i32p, C.c_void_p, C.c_int, C.c_int]
        lib.oracle_coo_to_csr.restype = None
        lib.oracle_fnv1a64.argtypes = [C.c_void_p, C.c_size_t]
        lib.oracle_fnv1a64.restype = C.c_uint64
        lib.oracle_num_threads.restype = C.c_int
        lib._typed = True
    return lib


def ref():
    lib = _load(os.path.join(HERE, "_ref", "libspmm_ref.so"))
    if not hasattr(lib, "_typed"):
        lib.ref_spmm_csr.argtypes = [_i32p, _i32p, _f64p, C.c_long, C.c_long, C.c_long, _f64p, _f64p, C.c_int]
        lib.ref_spmm_csr.restype = None
        lib.ref_spmm_csr_nocopy.argtypes = lib.ref_spmm_csr.argtypes
        lib.ref_spmm_csr_nocopy.restype = None
        lib.ref_balance_prefix_sums.argtypes = [_i32p, C.c_long, C.c_long, C.c_long, _i64p]
        lib.ref_balance_prefix_sums_pairs.argtypes = [_i32p, C.c_long, C.c_long, C.c_long, _i64p]
        lib.ref_balance_iterations.argtypes = [C.c_long, C.c_long, C.c_long, _i64p]
        lib.ref_num_threads.restype = C.c_int
        lib._typed = True
    return lib


_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")


def have_ref_f32():
    return os.path.exists(os.path.join(HERE, "_ref", "libspmm_ref_f.so"))


def spmm_ref_f32(ia, ja, a, m, c, x, k):
    """The reference's compute_csr in an "_f" build (ValueType = float)."""
    lib = _load(os.path.join(HERE, "_ref", "libspmm_ref_f.so"))
    lib.ref_spmm_csr.argtypes = [_i32p, _i32p, _f32p, C.c_long, C.c_long, C.c_long, _f32p, _f32p, C.c_int]
    lib.ref_spmm_csr.restype = None
    y = np.empty(m * k, dtype=np.float32)
    lib.ref_spmm_csr(ia, ja, a, m, c, len(ja), x, y, k)
    return y


def spmm_oracle_f32(ia, ja, a, m, c, x, k):
    lib = oracle()
    lib.oracle_spmm_csr_f32.argtypes = [_i32p, _i32p, _f32p, C.c_long, C.c_long, _f32p, _f32p, C.c_int]
    lib.oracle_spmm_csr_f32.restype = None
    y = np.empty(m * k, dtype=np.float32)
    lib.oracle_spmm_csr_f32(ia, ja, a, m, c, x, y, k)
    return y


def have_pbv_ref():
    return os.path.exists(os.path.join(HERE, "_ref", "libpbv_ref.so"))


def pbv_ref():
    """The reference's spmv_kernel_csr.cpp built as Custom_CSR_PBV (perfect nnz balance with row carry, a6)."""
    lib = _load(os.path.join(HERE, "_ref", "libpbv_ref.so"))
    if not hasattr(lib, "_typed"):
        lib.ref_pbv_partition.argtypes = [_i32p, _i32p, _f64p, C.c_long, C.c_long, C.c_long, C.c_int, _i64p]
        lib.ref_pbv_partition.restype = None
        lib.ref_pbv_spmv.argtypes = [_i32p, _i32p, _f64p, C.c_long, C.c_long, C.c_long, C.c_int, _f64p, _f64p]
        lib.ref_pbv_spmv.restype = None
        lib._typed = True
    return lib


def merge_ref():
    lib = _load(os.path.join(HERE, "_ref", "libmerge_ref.so"))
    if not hasattr(lib, "_typed"):
        lib.ref_merge_partition.argtypes = [_i32p, C.c_long, C.c_long, C.c_long, _i64p]
        lib.ref_spmv_merge.argtypes = [_i32p, _i32p, _f64p, C.c_long, C.c_long, _f64p, _f64p, C.c_int]
        lib.ref_spmv_merge.restype = None
        lib._typed = True
    return lib


def coo_ref():
    """The reference's coo_to_csr (csr_gen.c:162-217), instantiated by lib/aux/csr_converter_double.c."""
    lib = _load(os.path.join(HERE, "_ref", "libcoo_ref.so"))
    if not hasattr(lib, "_typed"):
        f = lib.coo_to_csr_csr_converter_double
        f.argtypes = [_i32p, _i32p, C.c_void_p, C.c_long, C.c_long, C.c_long, _i32p, _i32p, C.c_void_p, C.c_int, C.c_int]
        f.restype = None
        lib._typed = True
    return lib


def have_coo_ref():
    return os.path.exists(os.path.join(HERE, "_ref", "libcoo_ref.so"))


def _coo_call(fn, R, Cc, V, m, n, sort_columns, transpose):
    nnz = len(R)
    nrows = n if transpose else m
    row_ptr = np.zeros(nrows + 1, np.int32)
    col_idx = np.zeros(max(nnz, 1), np.int32)
    values = np.zeros(max(nnz, 1), np.float64) if V is not None else None
    Rb = np.ascontiguousarray(R, np.int32) if nnz else np.zeros(1, np.int32)
    Cb = np.ascontiguousarray(Cc, np.int32) if nnz else np.zeros(1, np.int32)
    Vb = np.ascontiguousarray(V, np.float64) if V is not None else None
    fn(Rb.copy(), Cb.copy(), Vb.ctypes.data if Vb is not None else None, m, n, nnz,
       row_ptr, col_idx, values.ctypes.data if values is not None else None, sort_columns, transpose)
    return row_ptr, col_idx[:nnz], (values[:nnz] if values is not None else None)


def coo_to_csr_oracle(R, Cc, V, m, n, sort_columns=1, transpose=0):
    return _coo_call(oracle().oracle_coo_to_csr, R, Cc, V, m, n, sort_columns, transpose)


def coo_to_csr_ref(R, Cc, V, m, n, sort_columns=1, transpose=0):
    return _coo_call(coo_ref().coo_to_csr_csr_converter_double, R, Cc, V, m, n, sort_columns, transpose)


class _MtxHead(C.Structure):
    # struct Matrix_Market (lib/storage_formats/matrix_market/matrix_market.h:29-46); struct DLCM_Matrix is laid out alike
    _fields_ = [("filename", C.c_char_p), ("format", C.c_char_p), ("field", C.c_char_p),
                ("symmetric", C.c_int), ("skew_symmetric", C.c_int),
                ("m", C.c_long), ("k", C.c_long), ("nnz", C.c_long), ("nnz_sym", C.c_long),
                ("R", C.POINTER(C.c_int)), ("C", C.POINTER(C.c_int)), ("V", C.c_void_p)]


def have_io_ref():
    return os.path.exists(os.path.join(HERE, "_ref", "libio_ref.so"))


def io_ref():
    lib = _load(os.path.join(HERE, "_ref", "libio_ref.so"))
    if not hasattr(lib, "_typed"):
        lib.mtx_read.argtypes = [C.c_char_p, C.c_long, C.c_long]
        lib.mtx_read.restype = C.POINTER(_MtxHead)
        lib.smtx_read.argtypes = [C.c_char_p, C.c_long, C.c_long]
        lib.smtx_read.restype = C.POINTER(_MtxHead)
        lib._typed = True
    return lib


def mtx_read_ref(path):
    """The reference's mtx_read(file, expand_symmetry=1, pattern_dummy_vals=1) followed by the value conversion of
    spmv_bench.cpp:730-763.  Returns (m, k, R, C, V as float64, field)."""
    h = io_ref().mtx_read(path.encode(), 1, 1).contents
    nnz, field = int(h.nnz), h.field.decode()
    R = np.ctypeslib.as_array(h.R, shape=(max(nnz, 1),))[:nnz].astype(np.int32, copy=True)
    Cc = np.ctypeslib.as_array(h.C, shape=(max(nnz, 1),))[:nnz].astype(np.int32, copy=True)
    if field == "integer":
        V = np.ctypeslib.as_array(C.cast(h.V, C.POINTER(C.c_int)), shape=(max(nnz, 1),))[:nnz].astype(np.float64)
    elif field == "complex":
        z = np.ctypeslib.as_array(C.cast(h.V, C.POINTER(C.c_double)), shape=(2 * max(nnz, 1),))[:2 * nnz]
        V = np.hypot(z[0::2], z[1::2])          # cabs()
    else:
        V = np.ctypeslib.as_array(C.cast(h.V, C.POINTER(C.c_double)), shape=(max(nnz, 1),))[:nnz].astype(np.float64, copy=True)
    return int(h.m), int(h.k), R, Cc, V, field


def smtx_read_ref(path):
    """The reference's smtx_read: (m, k, row_ptr, col_ind).  Its values are time-seeded rand(): not comparable."""
    h = io_ref().smtx_read(path.encode(), 0, 0).contents
    m, nnz = int(h.m), int(h.nnz)
    ia = np.ctypeslib.as_array(h.R, shape=(m + 1,)).astype(np.int32, copy=True)
    ja = np.ctypeslib.as_array(h.C, shape=(max(nnz, 1),))[:nnz].astype(np.int32, copy=True)
    return m, int(h.k), ia, ja


def check_accuracy_oracle(ia, ja, a, m, c, x, y, k):
    lib = oracle()
    lib.oracle_check_accuracy.argtypes = [_i32p, _i32p, _f64p, C.c_long, C.c_long, _f64p, _f64p, C.c_int, _f64p, _f64p]
    lib.oracle_check_accuracy.restype = None
    stats = np.zeros(10)
    gold = np.zeros(max(m * k, 1))
    lib.oracle_check_accuracy(ia, ja, a, m, c, x, y, k, stats, gold)
    return stats, gold[:m * k]


class _CsrHead(C.Structure):
    # leading fields of struct csr_matrix (artificial_matrix_generation.h:10-17)
    _fields_ = [("row_ptr", C.POINTER(C.c_int)), ("col_ind", C.POINTER(C.c_int)),
                ("values", C.POINTER(C.c_double)),
                ("nr_rows", C.c_uint), ("nr_cols", C.c_uint), ("nr_nzeros", C.c_uint)]


def amg():
    lib = _load(os.path.join(HERE, "_ref", "libamg_ref.so"))
    if not hasattr(lib, "_typed"):
        lib.artificial_matrix_generation.argtypes = [
            C.c_long, C.c_long, C.c_double, C.c_double, C.c_char_p, C.c_uint, C.c_char_p,
            C.c_double, C.c_double, C.c_double, C.c_double]
        lib.artificial_matrix_generation.restype = C.POINTER(_CsrHead)
        lib.free_csr_matrix.argtypes = [C.POINTER(_CsrHead)]
        lib._typed = True
    return lib


def generate(args):
    """Run the reference generator.  `args` is the harness's positional order
    (spmv_bench.cpp:851-862): rows cols avg std distribution placement bw skew
    neighbours cross_row_similarity seed -- a string or a sequence.
    Returns (row_ptr int32[m+1], col_ind int32[nnz], values f64[nnz], m, c)."""
    if isinstance(args, str):
        args = args.split()
    rows, cols, avg, std, distr, place, bw, skew, neigh, crs, seed = args[:11]
    lib = amg()
    p = lib.artificial_matrix_generation(
        int(rows), int(cols), float(avg), float(std), str(distr).encode(), int(seed),
        str(place).encode(), float(bw), float(skew), float(neigh), float(crs))
    h = p.contents
    m, c, nnz = int(h.nr_rows), int(h.nr_cols), int(h.nr_nzeros)
    ia = np.ctypeslib.as_array(h.row_ptr, shape=(m + 1,)).astype(np.int32, copy=True)
    ja = np.ctypeslib.as_array(h.col_ind, shape=(max(nnz, 1),))[:nnz].astype(np.int32, copy=True)
    a = np.ctypeslib.as_array(h.values, shape=(max(nnz, 1),))[:nnz].astype(np.float64, copy=True)
    lib.free_csr_matrix(p)
    return ia, ja, a, m, c


def fnv(arr):
    arr = np.ascontiguousarray(arr)
    return "%016x" % oracle().oracle_fnv1a64(arr.ctypes.data, arr.nbytes)


def spmm_oracle(ia, ja, a, m, c, x, k):
    y = np.empty(m * k, dtype=np.float64)
    oracle().oracle_spmm_csr(ia, ja, a, m, c, x, y, k)
    return y


def spmm_ref(ia, ja, a, m, c, x, k):
    y = np.empty(m * k, dtype=np.float64)
    ref().ref_spmm_csr(ia, ja, a, m, c, len(ja), x, y, k)
    return y


# ---- N4: sparse-attention pipeline (pipeline_code_bench) ---------------------------------------------------

def _pipeline_oracle():
    lib = oracle()
    if not hasattr(lib, "_typed_pipeline"):
        lib.oracle_csrmm_rowmajor.argtypes = [_i32p, _i32p, _f64p, C.c_long, C.c_long, _f64p, _f64p]
        lib.oracle_csrmm_rowmajor.restype = None
        lib.oracle_sddmm.argtypes = [_i32p, _i32p, _f64p, C.c_long, C.c_long, _f64p, _f64p, _f64p, C.c_int]
        lib.oracle_sddmm.restype = None
        lib.oracle_softmax.argtypes = [_f64p, C.c_long]
        lib.oracle_softmax.restype = None
        lib.oracle_softmax_rows.argtypes = [_i32p, C.c_long, _f64p]
        lib.oracle_softmax_rows.restype = None
        lib._typed_pipeline = True
    return lib


def csrmm_oracle(ia, ja, a, m, n, x):
    """y (m x n row-major) = CSR(ia, ja, a) * x (k x n row-major)."""
    y = np.empty(m * n, dtype=np.float64)
    _pipeline_oracle().oracle_csrmm_rowmajor(ia, ja, a, m, n, np.ascontiguousarray(x, dtype=np.float64).ravel(), y)
    return y


def sddmm_oracle(ia, ja, mask, m, n, Q, K, mode=0):
    y = np.empty(max(len(ja), 1), dtype=np.float64)
    _pipeline_oracle().oracle_sddmm(ia, ja, mask, m, n, np.ascontiguousarray(Q).ravel(), np.ascontiguousarray(K).ravel(), y, mode)
    return y[:len(ja)]


def softmax_oracle(y, ia=None):
    """ia None: the reference's softmax over the whole array; else row by row."""
    v = np.array(y, dtype=np.float64, copy=True)
    if ia is None:
        _pipeline_oracle().oracle_softmax(v, len(v))
    else:
        _pipeline_oracle().oracle_softmax_rows(ia, len(ia) - 1, v)
    return v


def pipeline_oracle(mask, weights, x, n, mode=0, softmax=0):
    """mask = (ia, ja, a, m); weights = {'K'|'Q'|'V': (ia, ja, a)}.  Returns dict K, Q, V, y, y_final."""
    mia, mja, ma, m = mask
    out = {t: csrmm_oracle(*weights[t], m, n, x) for t in "KQV"}
    y = sddmm_oracle(mia, mja, ma, m, n, out["Q"], out["K"], mode)
    if softmax:
        y = softmax_oracle(y, None if softmax == 1 else mia)
    out["y"] = y
    out["y_final"] = csrmm_oracle(mia, mja, np.ascontiguousarray(y), m, n, out["V"])
    return out


def have_pipeline_ref():
    return os.path.exists(os.path.join(HERE, "_ref", "libpipeline_ref.so"))


def pipeline_ref_lib():
    lib = _load(os.path.join(HERE, "_ref", "libpipeline_ref.so"))
    if not hasattr(lib, "_typed"):
        csr = [_i32p, _i32p, _f64p]
        lib.ref_pipeline.argtypes = csr + [C.c_long, C.c_long, C.c_int] + csr * 3 + [C.c_long] + [_f64p] * 6 + [C.c_int]
        lib.ref_pipeline.restype = None
        lib.ref_sddmm.argtypes = csr + [C.c_long, C.c_long, C.c_int, _f64p, _f64p, _f64p]
        lib.ref_sddmm.restype = None
        lib.ref_softmax.argtypes = [_f64p, C.c_int]
        lib.ref_softmax.restype = None
        lib._typed = True
    return lib


def pipeline_ref(mask, weights, x, n, kdim, softmax=0):
    """The reference's own csr_to_format / spmm('K','Q','V') / sddmm / spmm('final') sequence."""
    mia, mja, ma, m = mask
    nnz = len(mja)
    K, Q, V, yf = (np.zeros(m * n) for _ in range(4))
    y = np.zeros(max(nnz, 1))
    xx = np.ascontiguousarray(x, dtype=np.float64).ravel()
    pipeline_ref_lib().ref_pipeline(mia, mja, ma, m, nnz, n, *weights["K"], *weights["Q"], *weights["V"], kdim,
                                    xx, K, Q, V, y, yf, softmax)
    return {"K": K, "Q": Q, "V": V, "y": y[:nnz], "y_final": yf}


# ---- N4 in fp32 (the pipeline bench's "_f" build) -----------------------------------------------------------

_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")


def _pipeline_oracle_f32():
    lib = oracle()
    if not hasattr(lib, "_typed_pipeline_f32"):
        lib.oracle_csrmm_rowmajor_f32.argtypes = [_i32p, _i32p, _f32p, C.c_long, C.c_long, _f32p, _f32p]
        lib.oracle_csrmm_rowmajor_f32.restype = None
        lib.oracle_sddmm_f32.argtypes = [_i32p, _i32p, _f32p, C.c_long, C.c_long, _f32p, _f32p, _f32p, C.c_int]
        lib.oracle_sddmm_f32.restype = None
        lib._typed_pipeline_f32 = True
    return lib


def pipeline_oracle_f32(mask, weights, x, n, mode=0):
    """fp32 pipeline without softmax.  mask = (ia, ja, a float32, m); weights values float32; x float32."""
    mia, mja, ma, m = mask
    lib = _pipeline_oracle_f32()
    out = {}
    xx = np.ascontiguousarray(x, dtype=np.float32).ravel()
    for t in "KQV":
        ia, ja, a = weights[t]
        y = np.empty(m * n, dtype=np.float32)
        lib.oracle_csrmm_rowmajor_f32(ia, ja, np.ascontiguousarray(a, dtype=np.float32), m, n, xx, y)
        out[t] = y
    y = np.empty(max(len(mja), 1), dtype=np.float32)
    lib.oracle_sddmm_f32(mia, mja, np.ascontiguousarray(ma, dtype=np.float32), m, n, out["Q"], out["K"], y, mode)
    out["y"] = y[:len(mja)]
    yf = np.empty(m * n, dtype=np.float32)
    lib.oracle_csrmm_rowmajor_f32(mia, mja, np.ascontiguousarray(out["y"]), m, n, out["V"], yf)
    out["y_final"] = yf
    return out


def have_pipeline_ref_f32():
    return os.path.exists(os.path.join(HERE, "_ref", "libpipeline_ref_f.so"))


def pipeline_ref_f32(mask, weights, x, n, kdim):
    """The reference's own "_f" pipeline TU: csr_to_format / spmm('K','Q','V') / sddmm / spmm('final')."""
    lib = _load(os.path.join(HERE, "_ref", "libpipeline_ref_f.so"))
    if not hasattr(lib, "_typed"):
        csr = [_i32p, _i32p, _f32p]
        lib.ref_pipeline.argtypes = csr + [C.c_long, C.c_long, C.c_int] + csr * 3 + [C.c_long] + [_f32p] * 6 + [C.c_int]
        lib.ref_pipeline.restype = None
        lib._typed = True
    mia, mja, ma, m = mask
    nnz = len(mja)
    K, Q, V, yf = (np.zeros(m * n, np.float32) for _ in range(4))
    y = np.zeros(max(nnz, 1), np.float32)
    w = {t: (weights[t][0], weights[t][1], np.ascontiguousarray(weights[t][2], dtype=np.float32)) for t in "KQV"}
    lib.ref_pipeline(mia, mja, np.ascontiguousarray(ma, dtype=np.float32), m, nnz, n, *w["K"], *w["Q"], *w["V"], kdim,
                     np.ascontiguousarray(x, dtype=np.float32).ravel(), K, Q, V, y, yf, 0)
    return {"K": K, "Q": Q, "V": V, "y": y[:nnz], "y_final": yf}
