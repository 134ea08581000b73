"""N3: fp32 ("_f" builds, ValueType = float).  CPU: the oracle's fp32 restatement against the reference's own
compute_csr compiled with ValueType = float (oracle/_ref/libspmm_ref_f.so) and against committed outputs of it.
GPU: the fp32 kernels through the C ABI against the oracle.  Tolerance (stated): fp32 left-to-right sums on
positive operands, relative error <= 2e-6 per entry (bit-identical in practice: same order, same FMA contraction);
with mixed signs the comparison is absolute against sum |a||x| * 2e-6."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, random_csr

F32_TOL = 2e-6


def case(m, c, avg, seed, k, long_row=0):
    ia, ja, a = random_csr(m, c, avg, seed=seed, empty_frac=0.1, long_row=long_row)
    a = (np.abs(a) + 0.01).astype(np.float32)
    x = np.random.default_rng(seed + 100).random(c * k, dtype=np.float32)
    return ia, ja, a, x


def rel_err(y, y_ref):
    return float(np.max(np.abs(y.astype(np.float64) - y_ref) / np.maximum(np.abs(y_ref), 1e-30))) if y.size else 0.0


def test_oracle_f32_matches_reference_f_build(L):
    if not L.have_ref_f32():
        pytest.skip("oracle/_ref/libspmm_ref_f.so not built")
    for seed, k in ((1, 1), (2, 8), (3, 32), (4, 7)):
        ia, ja, a, x = case(3000, 2500, 9.0, seed, k, long_row=900)
        y_o = L.spmm_oracle_f32(ia, ja, a, 3000, 2500, x, k)
        y_r = L.spmm_ref_f32(ia, ja, a, 3000, 2500, x, k)
        assert np.array_equal(y_o, y_r)                     # same source order, same flags: bit-identical


def test_oracle_f32_matches_committed_reference_output(L):
    z = np.load(os.path.join(GOLDEN, "spmm_reference_outputs_f32.npz"))
    y_o = L.spmm_oracle_f32(z["ia"], z["ja"], z["a"], int(z["m"]), int(z["c"]), z["x"], int(z["k"]))
    assert rel_err(y_o, z["y"].astype(np.float64)) <= F32_TOL


@pytest.mark.gpu
@pytest.mark.parametrize("k", [1, 2, 4, 8, 16, 32, 64, 128, 256, 3, 20, 100])
def test_gpu_f32_all_k(pkg, L, k):
    m, c = 4000, 3500
    ia, ja, a, x = case(m, c, 9.0, k, k, long_row=1500)
    y_ref = L.spmm_oracle_f32(ia, ja, a, m, c, x, k).astype(np.float64)
    mf = pkg.csr_to_format_f32(ia, ja, a, m, c, len(ja), k, variant=pkg.ROWSPLIT)
    y = np.full(m * k, np.nan, np.float32); mf.spmm(x, y, k)
    assert mf.last_kernel_ms() > 0 and mf.launch_count() >= 2
    mf.close()
    assert not np.isnan(y).any()
    assert rel_err(y, y_ref) <= F32_TOL, k
    # row-major B gives the same result
    x_rm = np.ascontiguousarray(x.reshape(k, c).T).ravel()
    mf = pkg.csr_to_format_f32(ia, ja, a, m, c, len(ja), k, variant=pkg.ROWSPLIT, b_layout=pkg.B_ROWMAJOR)
    y2 = np.full(m * k, np.nan, np.float32); mf.spmm(x_rm, y2, k); mf.close()
    assert np.array_equal(y, y2)


@pytest.mark.gpu
def test_gpu_f32_edge_cases_and_shards(pkg, L):
    for m, c, ia, ja in ((0, 5, np.zeros(1, np.int32), np.zeros(0, np.int32)),
                         (7, 5, np.zeros(8, np.int32), np.zeros(0, np.int32))):
        mf = pkg.csr_to_format_f32(ia, ja, np.zeros(0, np.float32), m, c, 0, 32)
        y = np.full(m * 32, np.nan, np.float32); mf.spmm(np.ones(c * 32, np.float32), y, 32); mf.close()
        assert not np.isnan(y).any() and not y.any()
    m, c, k = 6000, 2000, 32
    ia, ja, a, x = case(m, c, 11.0, 4, k, long_row=1800)
    y_ref = L.spmm_oracle_f32(ia, ja, a, m, c, x, k).astype(np.float64)
    bounds = np.zeros(5, np.int64); L.oracle().oracle_balance_prefix_sums(ia, m, len(ja), 4, bounds)
    parts = []
    for r in range(4):
        mf = pkg.csr_to_format_f32(ia, ja, a, m, c, len(ja), k, rank=r, nranks=4)
        assert (mf.first_row, mf.first_row + mf.m) == (bounds[r], bounds[r + 1])
        y = np.empty(mf.m * k, np.float32); mf.spmm(x, y, k); mf.close(); parts.append(y)
    assert rel_err(np.concatenate(parts), y_ref) <= F32_TOL


@pytest.mark.gpu
def test_gpu_f32_generated_matrix(pkg, L):
    if not L.have_ref():
        pytest.skip("reference generator not built")
    ia, ja, a64, m, c = L.generate("200000 200000 20 10 gamma random 0.3 1000 0.5 0.5 14")
    a = a64.astype(np.float32)
    for k in (32, 128):
        x = np.random.default_rng(k).random(c * k, dtype=np.float32)
        y_ref = L.spmm_oracle_f32(ia, ja, a, m, c, x, k).astype(np.float64)
        mf = pkg.csr_to_format_f32(ia, ja, a, m, c, len(ja), k)
        y = np.empty(m * k, np.float32); mf.spmm(x, y, k); mf.close()
        # max row 20 040 against a mean of 20: AUTO takes merge-path, which re-associates the long rows' sums
        assert mf.format_name == "B200_CSR_MERGE_F32"
        assert rel_err(y, y_ref) <= 2e-5
        mf = pkg.csr_to_format_f32(ia, ja, a, m, c, len(ja), k, variant=pkg.ROWSPLIT)
        y = np.empty(m * k, np.float32); mf.spmm(x, y, k); mf.close()
        assert rel_err(y, y_ref) <= F32_TOL


@pytest.mark.gpu
def test_reference_harness_dropin_f_build(pkg):
    """The reference's unchanged harness compiled as an "_f" build (ValueType = float) with our plug-in TU."""
    import subprocess
    from conftest import ROOT
    exe = os.path.join(ROOT, "oracle", "_ref", "spmm_b200_f.exe")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/spmm_b200_f.exe not built")
    env = dict(os.environ, USE_PROCESSES="0", USE_ARTIFICIAL_MATRICES="1", USE_DLCM_MATRICES="0", NUM_COLS="32",
               COOLDOWN="0", RAPL_REGISTERS="0", OMP_NUM_THREADS="4")
    r = subprocess.run([exe] + "100000 100000 10 2 normal random 1 0 0 0 14".split(), env=env,
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:]
    row = r.stderr.strip().splitlines()[-1].split(",")
    assert row[0] == "synthetic" and row[6] == "999475" and row[23] == "B200_CSR_ROWSPLIT_F32"
    assert float(row[25]) > 0


F32_MERGE_TOL = 2e-5     # merge-path re-associates a row's sum at the carries: fp32, rows of up to 1500 positive terms


@pytest.mark.gpu
@pytest.mark.parametrize("k", [1, 2, 4, 8, 16, 32, 64, 128, 20])
def test_gpu_f32_merge_path(pkg, L, k):
    """fp32 merge-path (a skewed matrix in an "_f" build serialised on one lane group): the merge kernel of
    spmm_merge.cuh instantiated for float.  AUTO takes it for the skewed matrix; k without an instantiation (20) falls
    back to row-split."""
    m, c = 4000, 3500
    ia, ja, a, x = case(m, c, 9.0, 50 + k, k, long_row=1500)
    y_ref = L.spmm_oracle_f32(ia, ja, a, m, c, x, k).astype(np.float64)
    for variant in (pkg.MERGE, pkg.AUTO):
        mf = pkg.csr_to_format_f32(ia, ja, a, m, c, len(ja), k, variant=variant)
        assert mf.format_name == "B200_CSR_MERGE_F32"
        y = np.full(m * k, np.nan, np.float32); mf.spmm(x, y, k); mf.close()
        assert not np.isnan(y).any()
        assert rel_err(y, y_ref) <= F32_MERGE_TOL, (k, variant)
    # ragged shapes: no rows, no nonzeros, one huge row
    for ia2, ja2, a2, m2, c2 in ((np.zeros(1, np.int32), np.zeros(0, np.int32), np.zeros(0, np.float32), 0, 5),
                                 (np.zeros(8, np.int32), np.zeros(0, np.int32), np.zeros(0, np.float32), 7, 5),
                                 (np.array([0, 3000], np.int32), np.arange(3000, dtype=np.int32), np.ones(3000, np.float32), 1, 3000)):
        x2 = np.random.default_rng(1).random(c2 * k, dtype=np.float32)
        mf = pkg.csr_to_format_f32(ia2, ja2, a2, m2, c2, len(ja2), k, variant=pkg.MERGE)
        y = np.full(m2 * k, np.nan, np.float32); mf.spmm(x2, y, k); mf.close()
        y_ref2 = L.spmm_oracle_f32(ia2, ja2, a2, m2, c2, x2, k).astype(np.float64)
        assert rel_err(y, y_ref2) <= 1e-4 and not np.isnan(y).any()


TF32_TOL = 4e-3          # TF32 operands (10 mantissa bits: A rounded to nearest, B truncated by the tensor core), fp32 accumulate


def banded(m, c, band, per_row, seed, empty_rows=()):
    """rows with `per_row` nonzeros inside a band of `band` columns around the diagonal (cfg4's structure, small)"""
    rng = np.random.default_rng(seed)
    cols = []
    for i in range(m):
        if i in empty_rows:
            cols.append(np.zeros(0, np.int64))
            continue
        centre = i * c // max(m, 1)
        lo = max(0, min(c - band, centre - band // 2))
        cols.append(lo + np.sort(rng.choice(band, min(per_row, band), replace=False)))
    ia = np.zeros(m + 1, np.int32); ia[1:] = np.cumsum([len(x) for x in cols])
    ja = np.concatenate(cols).astype(np.int32) if m else np.zeros(0, np.int32)
    a = (rng.random(len(ja)) + 0.05).astype(np.float32)
    return ia, ja, a


@pytest.mark.gpu
@pytest.mark.parametrize("k", [32, 64, 128])
def test_gpu_tf32_blocked_kernel(pkg, L, k):
    """N3: the TF32 blocked kernel (tcgen05.mma kind::tf32, accumulators in TMEM) on banded matrices, against the fp32
    reference restatement with the stated TF32 tolerance.  Shapes: m not a multiple of the 128-row slab, empty rows, a
    slab with no nonzero at all, windows that touch the last column, a window far wider than a stage (chunk skipping)."""
    cases = [banded(5003, 4000, 200, 40, 1, empty_rows=set(range(256, 384)) | {5, 77, 5002}),
             banded(300, 300, 64, 20, 2),
             banded(128, 5000, 48, 48, 3)]
    ia, ja, a = random_csr(700, 20000, 6.0, seed=4)                    # not banded: windows of ~20000 columns, mostly empty chunks
    cases.append((ia, ja, (np.abs(a) + 0.05).astype(np.float32)))
    for ia, ja, a in cases:
        m = len(ia) - 1
        c = int(ja.max()) + 1
        x = np.random.default_rng(k).random(c * k, dtype=np.float32)
        y_ref = L.spmm_oracle_f32(ia, ja, a, m, c, x, k).astype(np.float64)
        mf = pkg.csr_to_format_f32(ia, ja, a, m, c, len(ja), k, variant=pkg.BCSR)
        assert mf.format_name == "B200_BLOCKED_TF32"
        for rep in range(2):
            y = np.full(m * k, np.nan, np.float32); mf.spmm(x, y, k)
            assert not np.isnan(y).any()
            assert rel_err(y, y_ref) <= TF32_TOL, (k, m, rep)
        mf.close()
    # AUTO never rounds to TF32; k without an instantiation runs the fp32 row-split kernel (bit-exact again)
    ia, ja, a = cases[1]
    x = np.random.default_rng(0).random(300 * 16, dtype=np.float32)
    mf = pkg.csr_to_format_f32(ia, ja, a, 300, 300, len(ja), 16, variant=pkg.BCSR)
    y = np.full(300 * 16, np.nan, np.float32); mf.spmm(x, y, 16); mf.close()
    assert rel_err(y, L.spmm_oracle_f32(ia, ja, a, 300, 300, x, 16).astype(np.float64)) <= F32_TOL
    assert pkg.csr_to_format_f32(ia, ja, a, 300, 300, len(ja), 32).format_name == "B200_CSR_ROWSPLIT_F32"


@pytest.mark.gpu
@pytest.mark.parametrize("variant,k,layout", [("ROWSPLIT", 32, "col"), ("ROWSPLIT", 20, "col"), ("ROWSPLIT", 1, "col"), ("ROWSPLIT", 64, "row"),
                                              ("MERGE", 32, "col"), ("MERGE", 8, "row"), ("BCSR", 32, "col")])
def test_gpu_f32_upload_window(pkg, L, variant, k, layout):
    """The fp32 host call uploads only rows [first, last touched column] of B (the fp64 rule): what lies
    outside may be NaN; the device copy is poisoned with an all-NaN operand first so that a read outside the window
    shows.  The TF32 blocked kernel reads whole chunks of B and keeps the full upload."""
    m, c = 3000, 9000
    lo = 4100
    ia, ja, a, _ = case(m, 700, 11.0, 300 + k, k, long_row=600 if variant == "MERGE" else 0)
    ja = (ja + lo).astype(np.int32)
    x = np.random.default_rng(5).random(c * k, dtype=np.float32)
    y_ref = L.spmm_oracle_f32(ia, ja, a, m, c, x, k).astype(np.float64)
    span = int(ja.max()) - int(ja.min()) + 1
    full = variant == "BCSR"
    xp = x.reshape(k, c).copy()
    if not full:
        xp[:, :int(ja.min())] = np.nan; xp[:, int(ja.max()) + 1:] = np.nan
    xp = (np.ascontiguousarray(xp.T) if layout == "row" else xp).ravel()
    mf = pkg.csr_to_format_f32(ia, ja, a, m, c, len(ja), k, variant=getattr(pkg, variant),
                               b_layout=pkg.B_ROWMAJOR if layout == "row" else pkg.B_COLMAJOR)
    mf.set_option("UPLOAD_WINDOW", 0)
    mf.spmm(np.full(c * k, np.nan, np.float32), np.empty(m * k, np.float32), k)       # poison d_B / d_x
    assert mf.last_h2d_bytes() == c * k * 4
    mf.set_option("UPLOAD_WINDOW", 1)
    for _ in range(2):
        y = np.full(m * k, np.nan, np.float32); mf.spmm(xp, y, k)
        assert np.isfinite(y).all(), (variant, k)
        assert rel_err(y, y_ref) <= (2e-3 if full else F32_TOL), (variant, k)
        assert mf.last_h2d_bytes() == (c if full else span) * k * 4
    mf.close()
