"""N2 (on-disk formats) and N1b (CheckAccuracy): the product's Matrix Market / DLMC readers against the
arrays the REFERENCE's readers returned for the committed fixtures (tests/golden/io_reference.npz, made by
tests/golden/make_io_golden.py) and, where oracle/_ref/libio_ref.so is present, against the reference run
live; the oracle's CheckAccuracy restatement against the line the reference harness printed.
Readers are host code: no GPU needed.  Integers and values compare EXACTLY (same strtol / strtod)."""
import os

import numpy as np
import pytest

from conftest import GOLDEN

IO = os.path.join(GOLDEN, "io")
MTX = ["real_general", "real_symmetric", "real_skew", "integer_general", "pattern_symmetric", "complex_hermitian", "no_banner"]


@pytest.fixture(scope="module")
def ref_io():
    return np.load(os.path.join(GOLDEN, "io_reference.npz"))


@pytest.mark.parametrize("name", MTX)
def test_read_mtx_matches_reference_arrays(pkg, ref_io, name):
    m, n, R, C, V, field = pkg.read_mtx(os.path.join(IO, name + ".mtx"))
    key = name + "_mtx"
    assert [m, n] == ref_io[key + "_shape"].tolist()
    assert field == str(ref_io[key + "_field"])
    assert np.array_equal(R, ref_io[key + "_R"])
    assert np.array_equal(C, ref_io[key + "_C"])
    assert np.array_equal(V, ref_io[key + "_V"])          # bit-exact values
    assert R.dtype == np.int32 and V.dtype == np.float64


@pytest.mark.parametrize("name", MTX)
def test_read_mtx_matches_reference_live(pkg, L, name):
    if not L.have_io_ref():
        pytest.skip("oracle/_ref/libio_ref.so not built")
    path = os.path.join(IO, name + ".mtx")
    m, n, R, C, V, field = pkg.read_mtx(path)
    rm, rn, rR, rC, rV, rfield = L.mtx_read_ref(path)
    assert (m, n, field) == (rm, rn, rfield)
    assert np.array_equal(R, rR) and np.array_equal(C, rC) and np.array_equal(V, rV)


def test_read_mtx_without_symmetry_expansion(pkg):
    m, n, R, C, V, _ = pkg.read_mtx(os.path.join(IO, "real_symmetric.mtx"), expand_symmetry=0)
    assert len(R) == 200 and np.all(C <= R)                  # the file's lower triangle only


def test_read_smtx_matches_reference(pkg, L, ref_io):
    path = os.path.join(IO, "dlmc_small.smtx")
    m, n, ia, ja, a = pkg.read_smtx(path, seed=3)
    assert [m, n] == ref_io["dlmc_small_smtx_shape"].tolist()
    assert np.array_equal(ia, ref_io["dlmc_small_smtx_ia"])
    assert np.array_equal(ja, ref_io["dlmc_small_smtx_ja"])
    assert len(a) == len(ja) and np.all(a >= -1.0) and np.all(a < 1.0) and np.unique(a).size == a.size
    _, _, _, _, a2 = pkg.read_smtx(path, seed=3)
    _, _, _, _, a3 = pkg.read_smtx(path, seed=4)
    assert np.array_equal(a, a2) and not np.array_equal(a, a3)       # seeded, reproducible
    if L.have_io_ref():
        rm, rn, ria, rja = L.smtx_read_ref(path)
        assert (m, n) == (rm, rn) and np.array_equal(ia, ria) and np.array_equal(ja, rja)


def test_reader_error_codes(pkg, tmp_path):
    with pytest.raises(pkg.SpmmB200Error) as e:
        pkg.read_mtx(str(tmp_path / "missing.mtx"))
    assert e.value.code == -8                                 # SPMM_B200_EIO
    bad = tmp_path / "bad.mtx"
    bad.write_text("%%MatrixMarket matrix coordinate real general\n3 3 2\n1 1 1.0\n")      # one entry short
    with pytest.raises(pkg.SpmmB200Error) as e:
        pkg.read_mtx(str(bad))
    assert e.value.code == -9                                 # SPMM_B200_EFORMAT
    bad.write_text("%%MatrixMarket matrix coordinate real general\n3 3 1\n4 1 1.0\n")      # row out of range
    with pytest.raises(pkg.SpmmB200Error) as e:
        pkg.read_mtx(str(bad))
    assert e.value.code == -9
    arr = tmp_path / "arr.mtx"
    arr.write_text("%%MatrixMarket matrix array real general\n2 1\n1.0\n2.0\n")
    with pytest.raises(pkg.SpmmB200Error) as e:
        pkg.read_mtx(str(arr))
    assert e.value.code == -6                                 # SPMM_B200_ENOSUP
    empty = tmp_path / "e.smtx"
    empty.write_text("2, 2, 1\n0 1\n")                        # offsets line too short
    with pytest.raises(pkg.SpmmB200Error):
        pkg.read_smtx(str(empty))


def test_oracle_check_accuracy_pinned_to_reference_harness(pkg, L, ref_io):
    """oracle_check_accuracy == the 'errors spmv:' line the reference harness printed for
    real_symmetric.mtx at k=4 (reference reader -> coo_to_csr -> compute_csr -> CheckAccuracy); the line
    carries 6 significant digits."""
    m, n, R, C, V, _ = pkg.read_mtx(os.path.join(IO, "real_symmetric.mtx"))
    ia, ja, a = L.coo_to_csr_oracle(R, C, V, m, n)
    k = 4
    x = np.ones(n * k)                                        # the harness's x (spmv_bench.cpp:899-903)
    y = L.spmm_oracle(ia, ja, a, m, n, x, k)
    stats, gold = L.check_accuracy_oracle(ia, ja, a, m, n, x, y, k)
    want = ref_io["harness_errors_real_symmetric_k4"]       # mae, max_ae, mse, mape, smape, lnQ_error, mlare, gmare
    assert np.allclose(stats[1:9], want, rtol=2e-5, atol=0), (stats[1:9], want)
    assert stats[9] == 1.0 and stats[0] <= 1e-10
    assert np.max(np.abs(gold - y)) <= 1e-14
