"""N4: the sparse-attention pipeline (pipeline_code_bench: spmm('K'|'Q'|'V') -> sddmm -> [softmax] -> spmm('final')).

CPU part: the oracle restatement (oracle/spmm_oracle.c: oracle_csrmm_rowmajor, oracle_sddmm, oracle_softmax)
against the golden outputs of the reference's own TU (tests/golden/pipeline_reference.npz) and, where the
compiled reference is present, against it directly.  GPU part: the C ABI of include/spmm_b200_pipeline.h
against the oracle and the same golden vectors.  fp64 tolerance where the order of a sum differs (softmax):
1e-12 relative; everything else is compared bit for bit."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, random_csr
from pipeline_util import band_and_random_mask, pipeline_case

RTOL = 1e-12


def golden():
    z = np.load(os.path.join(GOLDEN, "pipeline_reference.npz"))
    for i in range(int(z["ncases"])):
        seed, m, kdim, n, band, sm = (int(v) for v in z[f"case{i}_params"])
        yield dict(i=i, seed=seed, m=m, kdim=kdim, n=n, band=band, softmax=sm, fill=float(z[f"case{i}_fill"]),
                   mask_values=str(z[f"case{i}_mask_values"]), mask_nnz=int(z[f"case{i}_mask_nnz"]),
                   out={k: z[f"case{i}_{k}"] for k in ("K", "Q", "V", "y", "y_final")})


def case_inputs(g):
    mask, W, x = pipeline_case(g["seed"], g["m"], g["kdim"], g["n"], g["band"], g["fill"], mask_values=g["mask_values"])
    assert len(mask[1]) == g["mask_nnz"]
    return mask, W, x


# ---------------------------------------------------------------- CPU: oracle pinned to the reference

def test_oracle_matches_reference_golden(L):
    for g in golden():
        mask, W, x = case_inputs(g)
        o = L.pipeline_oracle(mask, W, x, g["n"], mode=0, softmax=g["softmax"])
        for k, ref in g["out"].items():
            assert np.array_equal(o[k], ref), (g["i"], k)


def test_oracle_matches_compiled_reference(L):
    if not L.have_pipeline_ref():
        pytest.skip("oracle/_ref/libpipeline_ref.so not built")
    for seed, m, kdim, n in ((11, 70, 50, 8), (12, 130, 200, 48), (13, 33, 40, 5)):
        mask, W, x = pipeline_case(seed, m, kdim, n, 3, 0.12, mask_values="random", empty_frac=0.2)
        for sm in (0, 1):
            o = L.pipeline_oracle(mask, W, x, n, 0, sm)
            r = L.pipeline_ref(mask, W, x, n, kdim, sm)
            for k in o:
                assert np.array_equal(o[k], r[k]), (seed, sm, k)


def test_oracle_reference_mode_reads_k_at_the_output_row(L):
    """sddmm_taco_naive.cpp:128: both factors at row i -- every nonzero of a row gets the same dot product."""
    mask, W, x = pipeline_case(21, 60, 40, 16, 3, 0.15)
    mia, mja, ma, m = mask
    Q = np.random.default_rng(1).standard_normal((m, 16))
    K = np.random.default_rng(2).standard_normal((m, 16))
    y = L.sddmm_oracle(mia, mja, ma, m, 16, Q, K, 0)
    rows = np.repeat(np.arange(m), np.diff(mia))
    np.testing.assert_allclose(y, (Q * K).sum(1)[rows], rtol=1e-13)
    y1 = L.sddmm_oracle(mia, mja, ma, m, 16, Q, K, 1)
    np.testing.assert_allclose(y1, np.einsum("pd,pd->p", Q[rows], K[mja]), rtol=1e-12, atol=1e-14)


def test_oracle_softmax(L):
    v = np.random.default_rng(3).standard_normal(1000) * 5
    s = L.softmax_oracle(v)
    e = np.exp(v - v.max())
    np.testing.assert_allclose(s, e / e.sum(), rtol=1e-13)
    ia = np.array([0, 3, 3, 10, 1000], np.int32)
    r = L.softmax_oracle(v, ia)
    for a, b in zip(ia[:-1], ia[1:]):
        if b > a:
            assert abs(r[a:b].sum() - 1) < 1e-12


def test_mask_generator_shape():
    ia, ja, a = band_and_random_mask(64, 4, 0.2, 0)
    assert ia[-1] == len(ja) == int(0.2 * 64 * 64) and (a == 1.0).all()
    rows = np.repeat(np.arange(64), np.diff(ia))
    assert (ja[np.abs(rows - ja) >= 4] < rows[np.abs(rows - ja) >= 4]).all()      # random part: lower triangle only
    for i in range(64):
        band = set(range(max(0, i - 3), min(63, i + 3) + 1))
        assert band <= set(ja[ia[i]:ia[i + 1]])


# ---------------------------------------------------------------- GPU: the C ABI against the oracle

def run_steps(pkg, mask, W, x, n, kdim, mode=0, softmax=0):
    mia, mja, ma, m = mask
    MF = pkg.csr_to_pipeline_format(mia, mja, ma, m, len(mja), n)
    out = {}
    for t in "KQV":
        out[t] = np.full(m * n, np.nan)
        MF.spmm(t, m, kdim, n, *W[t], x, out[t])
    y = np.full(max(len(mja), 1), np.nan)
    MF.sddmm(y, mode)
    if softmax:
        MF.softmax(y, rowwise=(softmax == 2))
    out["y"] = y[:len(mja)]
    out["y_final"] = np.full(m * n, np.nan)
    MF.final(out["y_final"])
    # the harness's own last call: spmm('final', m, m, n, mask_ia, mask_ja, y, V, y_final) with HOST y and V
    yf2 = np.full(m * n, np.nan)
    MF.spmm("final", m, m, n, mia, mja, np.ascontiguousarray(y[:len(mja)]) if len(mja) else np.zeros(0), out["V"], yf2)
    assert np.array_equal(yf2, out["y_final"])
    assert MF.launch_count() > 0
    MF.close()
    return out


@pytest.mark.gpu
def test_gpu_pipeline_matches_reference_golden(pkg):
    for g in golden():
        mask, W, x = case_inputs(g)
        o = run_steps(pkg, mask, W, x, g["n"], g["kdim"], 0, g["softmax"])
        for k, ref in g["out"].items():
            if g["softmax"] and k in ("y", "y_final"):
                # the softmax sums run in a different order; y_final sums signed terms, so its bound is norm-wise
                np.testing.assert_allclose(o[k], ref, rtol=RTOL, atol=0 if k == "y" else RTOL * np.abs(ref).max(),
                                           err_msg=f"case {g['i']} {k}")
            else:
                assert np.array_equal(o[k], ref), (g["i"], k)


def assert_close(got, ref, scale=None, msg=""):
    """fp64 tolerance of the path: 1e-12, norm-wise (sums of signed terms)."""
    bound = RTOL * (np.abs(ref).max() if scale is None else scale)
    assert np.abs(got - ref).max() <= bound, (msg, float(np.abs(got - ref).max()), bound)


@pytest.mark.gpu
@pytest.mark.parametrize("n", [1, 2, 4, 7, 16, 32, 64, 96, 128])
@pytest.mark.parametrize("mode", [0, 1, 2])
def test_gpu_pipeline_matches_oracle(pkg, L, n, mode):
    """mode 0 (the reference's SDDMM) and mode 2 (ordered attention SDDMM) bit for bit; mode 1 sums in tree order
    for n = 16, 32, 64, 128 and is held to the path's fp64 tolerance.  The SpMV-shaped projections (n <= 2) run on
    the small-k kernel, whose lanes split a row: tolerance there too."""
    mask, W, x = pipeline_case(100 + n, 300, 220, n, 5, 0.06, mask_values="random", empty_frac=0.1)
    o = run_steps(pkg, mask, W, x, n, 220, mode, 0)
    r = L.pipeline_oracle(mask, W, x, n, min(mode, 1), 0)
    exact = n > 2 and (mode != 1 or n not in (16, 32, 64, 128))
    for k in r:
        if exact:
            assert np.array_equal(o[k], r[k]), (n, mode, k)
        else:
            assert_close(o[k], r[k], msg=(n, mode, k))


@pytest.mark.gpu
@pytest.mark.parametrize("softmax", [1, 2])
def test_gpu_softmax(pkg, L, softmax):
    mask, W, x = pipeline_case(77, 400, 100, 32, 6, 0.05)
    o = run_steps(pkg, mask, W, x, 32, 100, 1, softmax)
    r = L.pipeline_oracle(mask, W, x, 32, 1, softmax)
    np.testing.assert_allclose(o["y"], r["y"], rtol=RTOL, atol=0)
    np.testing.assert_allclose(o["y_final"], r["y_final"], rtol=RTOL, atol=RTOL * np.abs(r["y_final"]).max())
    if softmax == 1:
        assert abs(o["y"].sum() - 1) < 1e-12
    else:
        mia = mask[0]
        sums = np.add.reduceat(o["y"], mia[:-1][np.diff(mia) > 0])
        assert np.abs(sums - 1).max() < 1e-12


@pytest.mark.gpu
@pytest.mark.parametrize("mode,softmax", [(0, 0), (1, 0), (2, 0), (1, 2)])
def test_gpu_attention_one_call_equals_the_steps(pkg, mode, softmax):
    n, kdim = 64, 500
    mask, W, x = pipeline_case(5, 700, kdim, n, 8, 0.04)
    steps = run_steps(pkg, mask, W, x, n, kdim, mode, softmax)
    mia, mja, ma, m = mask
    MF = pkg.csr_to_pipeline_format(mia, mja, ma, m, len(mja), n)
    for t in "KQV":
        MF.set_weight(t, m, kdim, *W[t])
    yf = np.full(m * n, np.nan)
    before = MF.launch_count()
    MF.attention(x, yf, mode, softmax)
    if (mode, softmax) == (1, 2):
        # one fused kernel (SDDMM + row softmax + SpMM, online softmax): rounding differs, y is not materialised
        assert_close(yf, steps["y_final"], msg="fused attention")
        assert MF.launch_count() - before == 4 and not MF.device_ptr("y")
    else:
        assert np.array_equal(yf, steps["y_final"])
        assert MF.launch_count() - before >= 5 and MF.device_ptr("y")
    assert MF.last_kernel_ms() > 0 and MF.device_ptr("f") and MF.device_ptr("K")
    MF.close()


@pytest.mark.gpu
@pytest.mark.parametrize("n", [32, 64, 128])
def test_gpu_fused_attention_matches_oracle(pkg, L, n):
    """attention(mode 1, row-wise softmax) = attention_fused_kernel: against the oracle's three steps, with empty
    mask rows, rows shorter and longer than one 32-nonzero batch, and non-unit mask values."""
    m, kdim = 500, 260
    mask, W, x = pipeline_case(40 + n, m, kdim, n, 3, 0.11, mask_values="random", empty_frac=0.05)
    mia, mja, ma, _ = mask
    cut = int(mia[m - 40])
    mia = mia.copy(); mia[m - 40:] = cut                      # the last 40 rows empty
    mask = (mia, mja[:cut].copy(), ma[:cut].copy(), m)
    r = L.pipeline_oracle(mask, W, x, n, 1, 2)
    MF = pkg.csr_to_pipeline_format(mask[0], mask[1], mask[2], m, cut, n)
    for t in "KQV":
        MF.set_weight(t, m, kdim, *W[t])
    yf = np.full(m * n, np.nan)
    MF.attention(x, yf, 1, 2)
    assert_close(yf, r["y_final"], msg=f"fused n={n}")
    assert (yf.reshape(m, n)[m - 40:] == 0).all()
    MF.set_option("ATTENTION_FUSED", 0)
    yf2 = np.full(m * n, np.nan)
    MF.attention(x, yf2, 1, 2)
    MF.set_option("ATTENTION_FUSED", 1)
    assert_close(yf2, r["y_final"], msg=f"unfused n={n}")
    MF.close()


@pytest.mark.gpu
def test_gpu_pipeline_edges(pkg, L):
    n = 8
    # order errors: sddmm before K and Q, final before V
    mask, W, x = pipeline_case(9, 40, 30, n, 2, 0.1)
    mia, mja, ma, m = mask
    MF = pkg.csr_to_pipeline_format(mia, mja, ma, m, len(mja), n)
    with pytest.raises(pkg.SpmmB200Error) as e:
        MF.sddmm(None)
    assert e.value.code == -1
    MF.spmm("K", m, 30, n, *W["K"], x, None)
    with pytest.raises(pkg.SpmmB200Error):
        MF.sddmm(None)
    MF.spmm("Q", m, 30, n, *W["Q"], x, None)
    MF.sddmm(None)
    with pytest.raises(pkg.SpmmB200Error):
        MF.final(np.zeros(m * n))
    with pytest.raises(pkg.SpmmB200Error):                       # a weight with the wrong number of rows
        MF.spmm("V", m - 1, 30, n, *random_csr(m - 1, 30, 2.0, 1), x, None)
    with pytest.raises(pkg.SpmmB200Error):                       # dense width differs from the format's
        MF.spmm("V", m, 30, n + 1, *W["V"], np.zeros(30 * (n + 1)), None)
    MF.close()
    # a mask with empty rows and a mask with no nonzeros at all
    ia = np.zeros(m + 1, np.int32)
    MF = pkg.csr_to_pipeline_format(ia, np.zeros(0, np.int32), np.zeros(0), m, 0, n)
    for t in "KQV":
        MF.spmm(t, m, 30, n, *W[t], x, None)
    MF.sddmm(None)
    MF.softmax(None)
    yf = np.full(m * n, np.nan)
    MF.final(yf)
    assert (yf == 0).all()
    MF.close()
    mia2 = mia.copy()
    cut = int(mia[m // 2])
    mia2[m // 2:] = cut                                         # rows m/2.. empty
    o = run_steps(pkg, (mia2, mja[:cut].copy(), ma[:cut].copy(), m), W, x, n, 30, 1, 0)
    r = L.pipeline_oracle((mia2, mja[:cut].copy(), ma[:cut].copy(), m), W, x, n, 1, 0)
    for k in r:
        assert np.array_equal(o[k], r[k]), k
    # a plain product through the "other type" path on a matrix that is not the mask
    ia3, ja3, a3 = random_csr(55, 30, 3.0, 4, empty_frac=0.2)
    MF = pkg.csr_to_pipeline_format(mia, mja, ma, m, len(mja), n)
    y3 = np.full(55 * n, np.nan)
    MF.spmm("x", 55, 30, n, ia3, ja3, a3, x, y3)
    assert np.array_equal(y3, L.csrmm_oracle(ia3, ja3, a3, 55, n, x))
    MF.close()


@pytest.mark.gpu
def test_gpu_pipeline_full_size_properties(pkg, L):
    """Sequence length 8192, width 64, band 16 + random fill to 5 % (the reference's SPARSITY=0.95 setting,
    sddmm_mask.h:29-31): rows checked against the oracle, linearity in the mask values, softmax rows sum to 1."""
    m, kdim, n = 8192, 4096, 64
    mask, W, x = pipeline_case(3, m, kdim, n, 16, 0.05, avg_w=16.0)
    mia, mja, ma, _ = mask
    MF = pkg.csr_to_pipeline_format(mia, mja, ma, m, len(mja), n)
    for t in "KQV":
        MF.set_weight(t, m, kdim, *W[t])
    yf = np.empty(m * n)
    MF.attention(x, yf, 2, 0)
    r = L.pipeline_oracle(mask, W, x, n, 1, 0)
    assert np.array_equal(yf, r["y_final"])
    MF.attention(x, yf, 1, 0)                                   # tree-ordered SDDMM
    assert_close(yf, r["y_final"], msg="fast attention mode")
    y = np.empty(len(mja))
    MF.sddmm(y, 1)
    assert_close(y, r["y"], msg="fast sddmm")
    MF.attention(x, yf, 2, 0)
    MF.close()
    # linearity: doubling the mask values doubles y and y_final exactly (power of two)
    MF2 = pkg.csr_to_pipeline_format(mia, mja, ma * 2.0, m, len(mja), n)
    for t in "KQV":
        MF2.set_weight(t, m, kdim, *W[t])
    yf2 = np.empty(m * n)
    MF2.attention(x, yf2, 2, 0)
    assert np.array_equal(yf2, 2.0 * yf)
    y = np.empty(len(mja))
    MF2.softmax(y, rowwise=True)
    sums = np.add.reduceat(y, mia[:-1])
    assert np.abs(sums - 1).max() < 1e-12
    MF2.close()


# ---------------------------------------------------------------- the reference's own pipeline harness

HARNESS_ENV = dict(USE_PROCESSES="0", USE_ARTIFICIAL_MATRICES="0", USE_DLCM_MATRICES="1", COOLDOWN="0", RAPL_REGISTERS="0",
                   OMP_NUM_THREADS="4", BAND_SIZE="8", SPARSE_ATTENTION_TYPE="band_and_random",
                   SDDMM_SPARSIFICATION_TYPE="none", SPARSITY="0.05", L_SPARSITY="0.1")


def run_pipeline_harness(exe, tmp_path, m, kdim, n, extra_env=None):
    """Runs pipeline_code_bench/sddmm_bench.cpp's main (built by `make -C oracle pipeline_harness`) on three DLMC
    weight files; returns (stdout, fields of the CSV line it writes to stderr)."""
    import subprocess
    files = []
    for i, t in enumerate("kqv"):
        ia, ja, _ = random_csr(m, kdim, 8.0, 50 + i)
        path = os.path.join(str(tmp_path), f"w_{t}.smtx")
        with open(path, "w") as f:
            f.write(f"{m}, {kdim}, {len(ja)}\n" + " ".join(map(str, ia)) + "\n" + " ".join(map(str, ja)) + "\n")
        files.append(path)
    env = dict(os.environ, NUM_COLS=str(n), **HARNESS_ENV)
    env.update(extra_env or {})
    r = subprocess.run([exe] + files, env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "failed" not in r.stdout
    line = [ln for ln in r.stderr.splitlines() if ln.startswith(files[0])]
    assert len(line) == 1, r.stderr
    return r.stdout, line[0].split(",")


def harness_mask_nnz(m, band=8, fill=0.05):
    """band_and_random (sddmm_mask.h:47-76): the band |i-j| < band first, then random fill up to fill*m*m."""
    return max(m * (2 * band - 1) - band * (band - 1), int(fill * m * m))


def test_reference_pipeline_harness_runs_on_cpu(tmp_path):
    exe = os.path.join(os.path.dirname(GOLDEN), "..", "oracle", "_ref", "sddmm_taco_naive_d.exe")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/sddmm_taco_naive_d.exe not built (make -C oracle pipeline_harness)")
    out, f = run_pipeline_harness(exe, tmp_path, 256, 192, 16)
    assert (int(f[2]), int(f[3]), int(f[4])) == (16, 256, 192)           # input_columns, csr_m, csr_k
    assert (int(f[-3]), int(f[-2]), int(f[-1])) == (256, 16, harness_mask_nnz(256))   # MF->m, MF->n, MF->nnz (the mask)


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["0", "1"])
def test_gpu_reference_pipeline_harness_runs_our_plugin(tmp_path, mode):
    """The UNCHANGED reference harness objects linked with csrc/sddmm_kernel_b200.cpp: csr_to_format(mask),
    spmm('K'|'Q'|'V'), sddmm, spmm('final') all land in the C ABI."""
    exe = os.path.join(os.path.dirname(GOLDEN), "..", "oracle", "_ref", "sddmm_b200_d.exe")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/sddmm_b200_d.exe not built (make -C oracle pipeline_harness)")
    out, f = run_pipeline_harness(exe, tmp_path, 2048, 1024, 64, {"SPMM_B200_SDDMM_MODE": mode})
    assert (int(f[2]), int(f[3]), int(f[4])) == (64, 2048, 1024)
    assert (int(f[-3]), int(f[-2]), int(f[-1])) == (2048, 64, harness_mask_nnz(2048))
    assert float(f[17]) > 0                                               # gflops of the whole pipeline


def test_oracle_sddmm_property_random_shapes(L):
    """Both SDDMM modes against a dense numpy evaluation over random shapes, including empty rows and n = 1."""
    from hypothesis import given, settings, strategies as st

    @settings(max_examples=25, deadline=None)
    @given(st.integers(1, 40), st.integers(1, 33), st.integers(0, 2 ** 16))
    def check(m, n, seed):
        rng = np.random.default_rng(seed)
        ia, ja, a = random_csr(m, m, 3.0, seed, empty_frac=0.3)
        Q = rng.standard_normal((m, n))
        K = rng.standard_normal((m, n))
        rows = np.repeat(np.arange(m), np.diff(ia))
        for mode in (0, 1):
            y = L.sddmm_oracle(ia, ja, a, m, n, Q, K, mode)
            kr = K[rows] if mode == 0 else K[ja]
            ref = np.einsum("pd,pd->p", Q[rows], kr) * a
            scale = (np.abs(Q[rows]) * np.abs(kr)).sum(1) * np.abs(a) + 1e-300
            assert (np.abs(y - ref) <= 1e-13 * scale * n).all()

    check()


def test_oracle_softmax_rows_with_empty_rows(L):
    ia = np.array([0, 0, 4, 4, 9, 9], np.int32)
    v = np.random.default_rng(5).standard_normal(9) * 3
    r = L.softmax_oracle(v, ia)
    assert abs(r[0:4].sum() - 1) < 1e-14 and abs(r[4:9].sum() - 1) < 1e-14
    assert (r > 0).all()
