"""GPU parity tests proper: every call goes through the C ABI (libspmm_b200.so) and is
compared with the oracle / the committed reference outputs.  fp64 tolerance 1e-12 relative
(north_star), integer outputs exact."""
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT, FP64_TOL, dense_b, max_rel_err, random_csr

pytestmark = pytest.mark.gpu

KS = [1, 2, 4, 8, 16, 32, 64, 128, 3, 20, 100]


def variants(pkg):
    names = os.environ.get("SPMM_B200_TEST_VARIANTS", "rowsplit,merge,auto,bcsr,bcsr_fma").split(",")
    table = {"rowsplit": pkg.ROWSPLIT, "merge": pkg.MERGE, "bcsr": pkg.BCSR, "auto": pkg.AUTO, "bcsr_fma": pkg.BCSR_FMA}
    return [table[n] for n in names if n]


def run_spmm(pkg, ia, ja, a, m, c, x, k, variant, **kw):
    mf = pkg.csr_to_format(ia, ja, a, m, c, len(ja), k, variant=variant, **kw)
    y = np.full(mf.m * k, np.nan)
    mf.spmm(x, y, k)
    info = (mf.format_name, mf.first_row, mf.m)
    mf.close()
    return y, info


def test_golden_reference_outputs(pkg, golden_cases):
    """CUDA path vs outputs of the reference's own compute_csr (tests/golden, committed)."""
    for g in golden_cases:
        m, c = len(g["ia"]) - 1, int(g["args"].split()[1])
        x = dense_b(c, g["k"], g["seed"])
        for v in variants(pkg):
            y, info = run_spmm(pkg, g["ia"], g["ja"], g["a"], m, c, x, g["k"], v)
            assert max_rel_err(y, g["y"]) <= FP64_TOL, (g["args"], g["k"], info)


@pytest.mark.parametrize("k", KS)
def test_all_k_against_oracle(pkg, L, k):
    ia, ja, a = random_csr(4000, 3500, 9.0, seed=k, empty_frac=0.1, long_row=1500)
    a = np.abs(a) + 0.01      # positive values: relative error is meaningful for every entry
    m, c = 4000, 3500
    x = dense_b(c, k, 5)
    y_ref = L.spmm_oracle(ia, ja, a, m, c, x, k)
    for v in variants(pkg):
        y, info = run_spmm(pkg, ia, ja, a, m, c, x, k, v)
        assert max_rel_err(y, y_ref) <= FP64_TOL, (k, info)


def test_row_major_b_layout(pkg, L):
    ia, ja, a = random_csr(1000, 800, 7.0, seed=1)
    k = 32
    x = dense_b(800, k, 2)                               # oracle layout [k][c]
    x_rm = np.ascontiguousarray(x.reshape(k, 800).T).ravel()   # [c][k]
    y_ref = L.spmm_oracle(ia, ja, a, 1000, 800, x, k)
    y, _ = run_spmm(pkg, ia, ja, a, 1000, 800, x_rm, k, pkg.AUTO, b_layout=pkg.B_ROWMAJOR)
    assert np.allclose(y, y_ref, rtol=1e-12, atol=1e-13)


@pytest.mark.parametrize("case", ["m0", "nnz0", "one_row", "one_col", "all_empty_but_last", "dense_row"])
def test_edge_cases(pkg, L, case):
    if case == "m0":
        ia, ja, a, m, c = np.zeros(1, np.int32), np.zeros(0, np.int32), np.zeros(0), 0, 5
    elif case == "nnz0":
        ia, ja, a, m, c = np.zeros(8, np.int32), np.zeros(0, np.int32), np.zeros(0), 7, 5
    elif case == "one_row":
        ia, ja, a = random_csr(1, 300, 100.0, seed=3); m, c = 1, 300
    elif case == "one_col":
        m, c = 50, 1
        ia = np.arange(m + 1, dtype=np.int32); ja = np.zeros(m, np.int32); a = np.arange(1, m + 1, dtype=np.float64)
    elif case == "all_empty_but_last":
        m, c = 3000, 64
        ia = np.zeros(m + 1, np.int32); ia[-1] = 64; ja = np.arange(64, dtype=np.int32); a = np.ones(64)
    else:
        m, c = 33, 5000
        ia = np.zeros(m + 1, np.int32); ia[17:] = 5000; ja = np.arange(5000, dtype=np.int32)
        a = np.random.default_rng(0).random(5000)
    for k in (1, 8, 32, 7):
        x = dense_b(c, k, 9)
        y_ref = L.spmm_oracle(ia, ja, a, m, c, x, k)
        for v in variants(pkg):
            y, info = run_spmm(pkg, ia, ja, a, m, c, x, k, v)
            assert y.shape == y_ref.shape
            assert not np.isnan(y).any(), (case, k, info)      # every C entry is written (beta = 0)
            assert max_rel_err(y, y_ref) <= FP64_TOL, (case, k, info)


def test_partition_device_bit_exact(pkg, L, pins, golden_cases):
    """Device partitioners == reference vectors (committed) and == oracle on ragged inputs."""
    for g in golden_cases[:4:3]:
        p = pins[g["args"]]
        ia = g["ia"]; m = len(ia) - 1; nnz = int(ia[-1])
        assert pkg.partition(ia, m, nnz, 8, pkg.PART_NNZ).tolist() == p["nnz_split_P8"]
        assert pkg.partition(ia, m, nnz, 3, pkg.PART_NNZ).tolist() == p["nnz_split_P3"]
        assert pkg.partition(None, m, nnz, 8, pkg.PART_ROWS).ravel().tolist() == p["rows_split_P8"]
        assert pkg.partition(ia, m, nnz, 8, pkg.PART_MERGE).ravel().tolist() == p["merge_P8"]
    for seed in range(5):
        m = 1 + 977 * seed
        ia, ja, a = random_csr(m, 500, 4.0, seed, empty_frac=0.4, long_row=(seed % 2) * 400)
        nnz = len(ja)
        for P in (1, 2, 7, 148, 5000):
            o = np.zeros(P + 1, np.int64); L.oracle().oracle_balance_prefix_sums(ia, m, nnz, P, o)
            assert np.array_equal(pkg.partition(ia, m, nnz, P, pkg.PART_NNZ), o)
            o = np.zeros(2 * P, np.int64); L.oracle().oracle_balance_iterations(0, m, P, o)
            assert np.array_equal(pkg.partition(None, m, nnz, P, pkg.PART_ROWS).ravel(), o)
            o = np.zeros(2 * P + 2, np.int64); L.oracle().oracle_merge_partition(ia, m, nnz, P, o)
            assert np.array_equal(pkg.partition(ia, m, nnz, P, pkg.PART_MERGE).ravel(), o)


def test_generated_configs_against_oracle(pkg, L, pins):
    """configs[0] (k=1) and the skewed / banded pinned matrices at k=32, with partition pins."""
    if not L.have_ref():
        pytest.skip("reference generator not built")
    for args, k in (("100000 100000 10 2 normal random 1 0 0 0 14", 1),
                    ("200000 200000 20 10 gamma random 0.3 1000 0.5 0.5 14", 32),
                    ("50000 50000 40 5 normal diagonal 0.002 0 1.9 0.9 14", 32),
                    ("50000 50000 40 5 normal diagonal 0.002 0 1.9 0.9 14", 128)):
        ia, ja, a, m, c = L.generate(args)
        nnz = len(ja)
        assert pkg.partition(ia, m, nnz, 8, pkg.PART_NNZ).tolist() == pins[args]["nnz_split_P8"]
        assert pkg.partition(ia, m, nnz, 8, pkg.PART_MERGE).ravel().tolist() == pins[args]["merge_P8"]
        x = dense_b(c, k, 77)
        y_ref = L.spmm_oracle(ia, ja, a, m, c, x, k)
        for v in variants(pkg):
            y, info = run_spmm(pkg, ia, ja, a, m, c, x, k, v)
            assert max_rel_err(y, y_ref) <= FP64_TOL, (args, k, info)


def test_bcsr_block_count_and_banded_parity(pkg, L):
    """CSR -> BCSR(8x1) on the device: the number of stored blocks equals the number of distinct
    (row // 8, col) pairs, and both blocked kernels reproduce the oracle on a banded matrix with
    similar neighbouring rows (the structure the blocked path is for) at k = 32 and 128."""
    rng = np.random.default_rng(3)
    m, c = 5003, 4000                         # m not a multiple of 8: ragged last block row
    cols = []
    prev = np.sort(rng.choice(200, 40, replace=False))
    for i in range(m):
        keep = prev[rng.random(prev.size) < 0.9]
        lo = max(0, min(c - 200, i * c // m - 100))
        new = lo + rng.choice(200, 6, replace=False)
        cur = np.unique(np.concatenate([keep, new]))
        cur = cur[(cur >= lo) & (cur < lo + 200)]
        cols.append(cur)
        prev = cur
    ia = np.zeros(m + 1, np.int32); ia[1:] = np.cumsum([len(x) for x in cols])
    ja = np.concatenate(cols).astype(np.int32)
    a = rng.random(len(ja)) + 0.1
    rows = np.repeat(np.arange(m, dtype=np.int64), np.diff(ia))
    nblocks = len(np.unique((rows // 8) * c + ja))
    for k in (32, 128):
        x = dense_b(c, k, 11)
        y_ref = L.spmm_oracle(ia, ja, a, m, c, x, k)
        for v in (pkg.BCSR, pkg.BCSR_FMA):
            mf = pkg.csr_to_format(ia, ja, a, m, c, len(ja), k, variant=v)
            assert abs(mf.bcsr_fill() - len(ja) / (8.0 * nblocks)) < 1e-12
            y = np.full(m * k, np.nan); mf.spmm(x, y, k); mf.close()
            assert max_rel_err(y, y_ref) <= FP64_TOL, (k, v)


def test_shards_cover_rows_and_match_reference_split(pkg, L):
    """nnz-balanced row shards (multi-GPU layout) emulated on one device: shard
    boundaries == reference split, and the stacked row-local C's == the full result."""
    ia, ja, a = random_csr(6000, 2000, 11.0, seed=4, empty_frac=0.05, long_row=1800)
    m, c, nnz, k = 6000, 2000, len(ja), 32
    x = dense_b(c, k, 8)
    y_ref = L.spmm_oracle(ia, ja, a, m, c, x, k)
    for P in (2, 4, 8):
        bounds = np.zeros(P + 1, np.int64); L.oracle().oracle_balance_prefix_sums(ia, m, nnz, P, bounds)
        parts = []
        for r in range(P):
            y, (name, first, rows) = run_spmm(pkg, ia, ja, a, m, c, x, k, pkg.AUTO, rank=r, nranks=P)
            assert (first, first + rows) == (bounds[r], bounds[r + 1])
            parts.append(y)
        assert max_rel_err(np.concatenate(parts), y_ref) <= FP64_TOL


def test_full_size_properties(pkg, L):
    """BASELINE configs[1] at full size (1M x 1M, avg 50, k=32): properties that need no CPU
    SpMM -- (i) with B = 1 every C entry equals the row sum of A (numpy reduceat);
    (ii) linearity A(x1 + 2 x2) = A x1 + 2 A x2; (iii) a sampled set of rows against the oracle."""
    if not L.have_ref():
        pytest.skip("reference generator not built")
    ia, ja, a, m, c = L.generate("1000000 1000000 50 16.6667 normal random 0.05 0 0.05 0.05 14")
    nnz, k = len(ja), 32
    assert nnz == 49994220 or nnz > 4.9e7
    mf = pkg.csr_to_format(ia, ja, a, m, c, nnz, k)
    y1 = np.empty(m * k); mf.spmm(np.ones(c * k), y1, k)
    starts = ia[:-1].astype(np.int64)
    rowsum = np.add.reduceat(np.concatenate([a, [0.0]]), np.minimum(starts, nnz))
    rowsum[ia[1:] == ia[:-1]] = 0.0
    assert max_rel_err(y1.reshape(m, k)[:, 0], rowsum) <= FP64_TOL
    assert max_rel_err(y1.reshape(m, k)[:, 31], rowsum) <= FP64_TOL
    x1, x2 = dense_b(c, k, 1), dense_b(c, k, 2)
    ya = np.empty(m * k); mf.spmm(x1, ya, k)
    yb = np.empty(m * k); mf.spmm(x2, yb, k)
    yc = np.empty(m * k); mf.spmm(x1 + 2.0 * x2, yc, k)
    assert max_rel_err(yc, ya + 2.0 * yb) <= 1e-11
    mf.close()
    # sampled rows vs the oracle (sub-matrix of 2000 rows)
    rows = np.random.default_rng(0).choice(m, 2000, replace=False); rows.sort()
    sub_ia = np.zeros(2001, np.int32)
    sub_ia[1:] = np.cumsum(ia[rows + 1] - ia[rows])
    idx = np.concatenate([np.arange(ia[r], ia[r + 1]) for r in rows])
    y_ref = L.spmm_oracle(sub_ia, ja[idx].copy(), a[idx].copy(), 2000, c, x1, k)
    assert max_rel_err(ya.reshape(m, k)[rows].ravel(), y_ref) <= FP64_TOL


def test_reference_harness_dropin(pkg):
    """The reference's UNCHANGED harness (spmv_bench.cpp) linked with our plug-in TU runs on the
    GPU and emits the reference's CSV row on stderr with our format name."""
    exe = os.path.join(ROOT, "oracle", "_ref", "spmm_b200_d.exe")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/spmm_b200_d.exe not built")
    env = dict(os.environ, USE_PROCESSES="0", USE_ARTIFICIAL_MATRICES="1", USE_DLCM_MATRICES="0", NUM_COLS="32",
               COOLDOWN="0", RAPL_REGISTERS="0", OMP_NUM_THREADS="4", SPMM_B200_COPYBACK="0")
    r = subprocess.run([exe] + "100000 100000 10 2 normal random 1 0 0 0 14".split(), env=env,
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:]
    row = r.stderr.strip().splitlines()[-1].split(",")
    assert row[0] == "synthetic" and row[6] == "999475" and row[23].startswith("B200_CSR")
    assert float(row[25]) > 0          # gflops column


@pytest.mark.parametrize("layout", ["col", "row"])
def test_k_sliced_host_pipeline(pkg, L, layout, monkeypatch):
    """spmm() with HOST buffers cuts k into 16-column slices (upload | kernel | download overlap,
    DESIGN 5.3).  Forced on for a small matrix: same result as the oracle for every variant, for
    copy_back=False + fetch_c, and the one-shot path still works on the same handle afterwards."""
    monkeypatch.setenv("SPMM_B200_PIPELINE_MIN_KB", "0")
    ia, ja, a = random_csr(3001, 2500, 9.0, seed=21, empty_frac=0.1, long_row=1200)
    a = np.abs(a) + 0.01
    m, c = 3001, 2500
    for k in (32, 48, 128):
        x = dense_b(c, k, 5)
        y_ref = L.spmm_oracle(ia, ja, a, m, c, x, k)
        xin = x if layout == "col" else np.ascontiguousarray(x.reshape(k, c).T).ravel()
        lay = pkg.B_COLMAJOR if layout == "col" else pkg.B_ROWMAJOR
        for v in variants(pkg):
            mf = pkg.csr_to_format(ia, ja, a, m, c, len(ja), k, variant=v, b_layout=lay)
            y = np.full(m * k, np.nan); mf.spmm(xin, y, k)
            assert max_rel_err(y, y_ref) <= FP64_TOL, (k, v, layout)
            assert mf.last_kernel_ms() > 0
            y2 = np.full(m * k, np.nan); mf.spmm(xin, None, k, copy_back=False); mf.fetch_c(y2, k)
            assert np.array_equal(y, y2)
            mf.set_option("PIPELINE_COLS", 0)                            # one-shot path on the same handle
            y3 = np.full(m * k, np.nan); mf.spmm(xin, y3, k)
            mf.set_option("PIPELINE_COLS", -1)
            assert max_rel_err(y3, y_ref) <= FP64_TOL
            mf.close()


def test_direct_store_into_page_locked_y(pkg, L, monkeypatch):
    """With a page-locked (device-mapped) y the sliced row-split kernels store C straight into host memory
    (no staging copy).  Same bits as the staged download; fetch_c has nothing to fetch afterwards;
    untouched guard elements around y stay untouched."""
    monkeypatch.setenv("SPMM_B200_PIPELINE_MIN_KB", "0")
    m, c = 4003, 3000
    ia, ja, a = random_csr(m, c, 11.0, seed=33, empty_frac=0.05)
    a = np.abs(a) + 0.01                      # no cancellation: the element-wise 1e-12 bound applies
    for k in (32, 64, 128):
        x = dense_b(c, k, 6)
        y_ref = L.spmm_oracle(ia, ja, a, m, c, x, k)
        mf = pkg.csr_to_format(ia, ja, a, m, c, len(ja), k, variant=pkg.ROWSPLIT)
        guard = 64
        buf = np.full(m * k + 2 * guard, -7.0)
        mf.register_host(buf)
        y = buf[guard:guard + m * k]
        before = mf.launch_count()
        mf.spmm(x, y, k)
        assert max_rel_err(y, y_ref) <= FP64_TOL, k
        assert (buf[:guard] == -7.0).all() and (buf[-guard:] == -7.0).all()
        assert mf.launch_count() - before >= 2                      # at least two slices ran
        with pytest.raises(pkg.SpmmB200Error):
            mf.fetch_c(np.empty(m * k), k)
        mf.set_option("DIRECT_STORE", 0)
        y2 = np.full(m * k, np.nan); mf.spmm(x, y2, k)
        mf.set_option("DIRECT_STORE", 1)
        assert np.array_equal(y2, y)                                 # staged download: the same bits
        y3 = np.full(m * k, np.nan); mf.fetch_c(y3, k)
        assert np.array_equal(y3, y)
        mf.close()


def test_kslice_measurement_keeps_results_identical(pkg, L):
    """L2-aware k slicing: a matrix whose touched B window (column span x 8k bytes) exceeds
    192 MB makes the handle time the candidate slice widths on its first 8 device calls.  Every one of those
    calls -- whatever width it runs -- must produce the oracle's result, and a width must have been chosen."""
    m, c, k = 3000, 210000, 128                     # span ~210k columns x 1 KB = 215 MB
    rng = np.random.default_rng(12)
    deg = rng.integers(1, 9, m)
    ia = np.zeros(m + 1, np.int32); ia[1:] = np.cumsum(deg)
    ja = np.concatenate([np.sort(rng.choice(c, d, replace=False)) for d in deg]).astype(np.int32)
    ja[0], ja[-1] = 0, c - 1                        # pin the span
    ja[: deg[0]].sort(); ja[-deg[-1]:].sort()
    a = rng.random(len(ja)) + 0.05
    x = dense_b(c, k, 4)
    y_ref = L.spmm_oracle(ia, ja, a, m, c, x, k)
    mf = pkg.csr_to_format(ia, ja, a, m, c, len(ja), k, variant=pkg.ROWSPLIT)
    mf.set_b(x, k)
    assert mf.kslice() == 0
    for call in range(10):
        mf.spmm_resident(k)
        y = np.full(m * k, np.nan); mf.fetch_c(y, k)
        assert max_rel_err(y, y_ref) <= FP64_TOL, call
    assert mf.kslice() in (16, 32, 64, 128)
    mf.close()


@pytest.mark.parametrize("args,k,expect", [
    ("2000000 2000000 20 20 gamma random 0.3 10000 0.5 0.5 14", 32, "MERGE"),              # cfg3a: one 200 040-nnz row
    ("1000000 1000000 50 16.6667 normal diagonal 0.0002 0 1.9 0.95 14", 128, "BCSR_DMMA"),   # cfg4: banded, blocked path
    ("8388607 8388607 5 1.6667 normal random 0.05 0 0.05 0.05 14", 8, "ROWSPLIT"),           # cfg5-large matrix at k=8 (host arrays stay < 1 GB each)
])
def test_full_size_properties_other_configs(pkg, L, pins, args, k, expect):
    """BASELINE configs[2..4] at full size through size-independent properties: generator pins (the input is the
    reference's matrix), AUTO's choice of kernel, row sums with B = 1, linearity, and sampled rows vs the oracle."""
    if not L.have_ref():
        pytest.skip("reference generator not built")
    ia, ja, a, m, c = L.generate(args)
    nnz = len(ja)
    if args in pins:
        assert L.fnv(ia) == pins[args]["fnv_row_ptr"] and L.fnv(ja) == pins[args]["fnv_col_ind"]
    mf = pkg.csr_to_format(ia, ja, a, m, c, nnz, k)
    assert mf.format_name.endswith(expect), mf.format_name
    y1 = np.empty(m * k); mf.spmm(np.ones(c * k), y1, k)
    starts = ia[:-1].astype(np.int64)
    rowsum = np.add.reduceat(np.concatenate([a, [0.0]]), np.minimum(starts, nnz))
    rowsum[ia[1:] == ia[:-1]] = 0.0
    # merge-path and the blocked kernels reorder the sum of a row: tolerance against sum |a|
    scale = np.maximum(np.abs(rowsum), 1e-300)
    for col in (0, k - 1):
        assert np.max(np.abs(y1.reshape(m, k)[:, col] - rowsum) / scale) <= 1e-11
    x1, x2 = dense_b(c, k, 1), dense_b(c, k, 2)
    ya = np.empty(m * k); mf.spmm(x1, ya, k)
    yb = np.empty(m * k); mf.spmm(x2, yb, k)
    yc = np.empty(m * k); mf.spmm(x1 + 2.0 * x2, yc, k)
    assert max_rel_err(yc, ya + 2.0 * yb) <= 1e-10
    mf.close()
    rows = np.random.default_rng(0).choice(m, 1500, replace=False); rows.sort()
    if expect == "MERGE":
        rows[0] = int(np.argmax(np.diff(ia)))        # include the monster row
        rows.sort()
    sub_ia = np.zeros(len(rows) + 1, np.int32)
    sub_ia[1:] = np.cumsum(ia[rows + 1] - ia[rows])
    idx = np.concatenate([np.arange(ia[r], ia[r + 1]) for r in rows])
    y_ref = L.spmm_oracle(sub_ia, ja[idx].copy(), a[idx].copy(), len(rows), c, x1, k)
    assert max_rel_err(ya.reshape(m, k)[rows].ravel(), y_ref) <= 1e-11
