"""Real-matrix path end to end on the GPU (SURVEY 8(f) N1 + N2): file -> read_mtx / read_smtx -> device
coo_to_csr -> csr_to_format -> spmm -> check_accuracy, every stage against the oracle; and the reference's
UNCHANGED harness with our plug-in TU on a .mtx file (its own CheckAccuracy judges our y)."""
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import ROOT, GOLDEN, FP64_TOL, dense_b, max_rel_err, random_csr

pytestmark = pytest.mark.gpu
IO = os.path.join(GOLDEN, "io")


@pytest.mark.parametrize("name", ["real_general", "real_symmetric", "real_skew", "pattern_symmetric", "complex_hermitian"])
def test_mtx_file_to_result(pkg, L, name):
    m, n, R, C, V, _ = pkg.read_mtx(os.path.join(IO, name + ".mtx"))
    ia, ja, a = pkg.coo_to_csr(R, C, V, m, n, len(R))
    ia_o, ja_o, a_o = L.coo_to_csr_oracle(R, C, V, m, n)
    assert np.array_equal(ia, ia_o) and np.array_equal(ja, ja_o) and np.array_equal(a, a_o)
    for k in (1, 4, 32):
        x = dense_b(n, k, 3)
        mf = pkg.csr_to_format(ia, ja, a, m, n, len(ja), k)
        y = np.full(m * k, np.nan); mf.spmm(x, y, k)
        y_ref = L.spmm_oracle(ia, ja, a, m, n, x, k)
        assert np.allclose(y, y_ref, rtol=1e-12, atol=1e-13)
        stats, gold = mf.check_accuracy(x, y, k, return_gold=True)
        o_stats, o_gold = L.check_accuracy_oracle(ia, ja, a, m, n, x, y, k)
        mf.close()
        assert np.array_equal(gold, o_gold)                    # double-double gold == fp128 Kahan gold, rounded to fp64
        got = np.array([stats[l] for l in pkg.MatrixFormat.ACCURACY_LABELS])
        assert got[9] == o_stats[9]
        assert np.allclose(got[:6], o_stats[:6], rtol=1e-9, atol=1e-300), (got, o_stats)
        assert abs(got[6] - o_stats[6]) < 1e-15                 # lnQ_error: a sum of cancelling log10 differences
        # mlare / gmare amplify the last bits of lnQ_error (log10|10^e - 1|): compare loosely
        if np.isfinite(o_stats[7]) and np.isfinite(got[7]) and abs(o_stats[6]) > 1e-14:
            assert abs(got[7] - o_stats[7]) < 1e-3 * max(1.0, abs(o_stats[7]))


def test_check_accuracy_flags_a_wrong_result(pkg, L):
    ia, ja, a = random_csr(2000, 1500, 12.0, seed=5, empty_frac=0.05)
    a = np.abs(a) + 0.05
    m, c, k = 2000, 1500, 8
    x = dense_b(c, k, 1)
    mf = pkg.csr_to_format(ia, ja, a, m, c, len(ja), k)
    y = np.empty(m * k); mf.spmm(x, y, k)
    good = mf.check_accuracy(x, y, k)
    assert good["passed"] == 1.0 and good["max_rel_diff"] < 1e-13 and good["max_ae"] < 1e-12
    y_bad = y.copy(); i = int(np.argmax(y)); y_bad[i] *= 1.0 + 1e-6
    bad = mf.check_accuracy(x, y_bad, k)
    o_bad, _ = L.check_accuracy_oracle(ia, ja, a, m, c, x, y_bad, k)
    mf.close()
    assert bad["passed"] == 0.0 and abs(bad["max_rel_diff"] - 1e-6) < 1e-9
    assert np.isclose(bad["max_rel_diff"], o_bad[0], rtol=1e-9) and np.isclose(bad["max_ae"], o_bad[2], rtol=1e-12)
    assert np.isclose(bad["mse"], o_bad[3], rtol=1e-9) and np.isclose(bad["mape"], o_bad[4], rtol=1e-9)


def test_smtx_file_to_result(pkg, L):
    m, n, ia, ja, a = pkg.read_smtx(os.path.join(IO, "dlmc_small.smtx"), seed=11)
    k = 32
    x = dense_b(n, k, 2)
    mf = pkg.csr_to_format(ia, ja, a, m, n, len(ja), k)
    y = np.full(m * k, np.nan); mf.spmm(x, y, k); mf.close()
    y_ref = L.spmm_oracle(ia, ja, a, m, n, x, k)
    assert np.allclose(y, y_ref, rtol=1e-12, atol=1e-14)


def test_reference_harness_dropin_on_mtx_file(pkg):
    """The reference harness + our TU on a Matrix Market file: the REFERENCE's reader, coo_to_csr and
    CheckAccuracy surround our csr_to_format / spmm; its 'errors spmv:' line must report fp64-level errors
    and no 'Test failed'."""
    exe = os.path.join(ROOT, "oracle", "_ref", "spmm_b200_d.exe")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/spmm_b200_d.exe not built")
    env = dict(os.environ, USE_PROCESSES="0", USE_ARTIFICIAL_MATRICES="0", USE_DLCM_MATRICES="0", NUM_COLS="4",
               COOLDOWN="0", RAPL_REGISTERS="0", OMP_NUM_THREADS="4")
    r = subprocess.run([exe, os.path.join(IO, "real_symmetric.mtx")], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:]
    assert "Test failed" not in r.stdout
    line = [l for l in r.stdout.splitlines() if l.startswith("errors spmv:")][-1]
    mae = float(re.search(r"mae=([^,\s]+)", line).group(1))
    max_ae = float(re.search(r"max_ae=([^,\s]+)", line).group(1))
    assert mae < 1e-14 and max_ae < 1e-13, line
    row = r.stderr.strip().splitlines()[-1].split(",")          # the real-matrix CSV row: name, threads, k, m, n, nnz, time, gflops ...
    assert row[3:6] == ["50", "50", "395"] and float(row[7]) > 0
