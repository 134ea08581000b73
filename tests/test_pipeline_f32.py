"""N4 in fp32: the pipeline bench's "_f" build (ValueType = float).  CPU: the fp32 oracle restatement against golden
outputs of the reference's own "_f" TU (tests/golden/pipeline_reference_f32.npz) and against the compiled TU.
GPU: spmm_b200_pipeline_*_f32 against the oracle and the golden outputs.  fp32 tolerance of the path: 2e-6,
norm-wise (tests/test_f32.py); the SDDMM is checked bit for bit on the GPU's own Q and K."""
import os

import numpy as np
import pytest

from conftest import GOLDEN
from pipeline_util import pipeline_case

TOL32 = 2e-6


def f32_case(seed, m, kdim, n, band, fill, mask_values):
    mask, W, x = pipeline_case(seed, m, kdim, n, band, fill, mask_values=mask_values)
    mask = (mask[0], mask[1], mask[2].astype(np.float32), mask[3])
    W = {t: (W[t][0], W[t][1], W[t][2].astype(np.float32)) for t in "KQV"}
    return mask, W, x.astype(np.float32)


def golden32():
    z = np.load(os.path.join(GOLDEN, "pipeline_reference_f32.npz"))
    for i in range(int(z["ncases"])):
        seed, m, kdim, n, band, _ = (int(v) for v in z[f"case{i}_params"])
        yield dict(i=i, seed=seed, m=m, kdim=kdim, n=n, band=band, fill=float(z[f"case{i}_fill"]),
                   mask_values=str(z[f"case{i}_mask_values"]),
                   out={k: z[f"case{i}_{k}"] for k in ("K", "Q", "V", "y", "y_final")})


def close32(got, ref, msg=""):
    bound = TOL32 * max(float(np.abs(ref).max()), 1e-30)
    assert float(np.abs(got.astype(np.float64) - ref.astype(np.float64)).max()) <= bound, msg


def test_oracle_f32_matches_reference_golden(L):
    for g in golden32():
        mask, W, x = f32_case(g["seed"], g["m"], g["kdim"], g["n"], g["band"], g["fill"], g["mask_values"])
        o = L.pipeline_oracle_f32(mask, W, x, g["n"], 0)
        for k, ref in g["out"].items():
            assert np.array_equal(o[k], ref), (g["i"], k)


def test_oracle_f32_matches_compiled_reference(L):
    if not L.have_pipeline_ref_f32():
        pytest.skip("oracle/_ref/libpipeline_ref_f.so not built")
    for seed, m, kdim, n in ((11, 70, 50, 8), (12, 130, 200, 48), (13, 33, 40, 5)):
        mask, W, x = f32_case(seed, m, kdim, n, 3, 0.12, "random")
        o = L.pipeline_oracle_f32(mask, W, x, n, 0)
        r = L.pipeline_ref_f32(mask, W, x, n, kdim)
        for k in o:
            assert np.array_equal(o[k], r[k]), (seed, k)


def run_steps32(pkg, L, mask, W, x, n, kdim, mode):
    mia, mja, ma, m = mask
    MF = pkg.csr_to_pipeline_format_f32(mia, mja, ma, m, len(mja), n)
    out = {}
    for t in "KQV":
        out[t] = np.full(m * n, np.nan, np.float32)
        MF.spmm(t, m, kdim, n, *W[t], x, out[t])
    y = np.full(max(len(mja), 1), np.nan, np.float32)
    MF.sddmm(y, mode)
    out["y"] = y[:len(mja)]
    # the SDDMM itself, bit for bit, on the Q and K the GPU produced
    lib = L._pipeline_oracle_f32()
    y_chk = np.empty(max(len(mja), 1), np.float32)
    lib.oracle_sddmm_f32(mia, mja, ma, m, n, out["Q"], out["K"], y_chk, 1 if mode else 0)
    assert np.array_equal(out["y"], y_chk[:len(mja)]), "sddmm bits"
    out["y_final"] = np.full(m * n, np.nan, np.float32)
    MF.final(out["y_final"])
    yf2 = np.full(m * n, np.nan, np.float32)
    MF.spmm("final", m, m, n, mia, mja, np.ascontiguousarray(out["y"]), out["V"], yf2)
    assert np.array_equal(yf2, out["y_final"])
    assert MF.launch_count() > 0
    MF.close()
    return out


@pytest.mark.gpu
def test_gpu_pipeline_f32_matches_reference_golden(pkg, L):
    for g in golden32():
        mask, W, x = f32_case(g["seed"], g["m"], g["kdim"], g["n"], g["band"], g["fill"], g["mask_values"])
        o = run_steps32(pkg, L, mask, W, x, g["n"], g["kdim"], 0)
        for k, ref in g["out"].items():
            close32(o[k], ref, (g["i"], k))


@pytest.mark.gpu
@pytest.mark.parametrize("n", [4, 7, 32, 64, 128])
@pytest.mark.parametrize("mode", [0, 1, 2])
def test_gpu_pipeline_f32_matches_oracle(pkg, L, n, mode):
    mask, W, x = f32_case(200 + n, 300, 220, n, 5, 0.06, "random")
    o = run_steps32(pkg, L, mask, W, x, n, 220, mode)
    r = L.pipeline_oracle_f32(mask, W, x, n, 1 if mode else 0)
    for k in r:
        close32(o[k], r[k], (n, mode, k))


@pytest.mark.gpu
def test_gpu_pipeline_f32_errors(pkg):
    mask, W, x = f32_case(9, 40, 30, 8, 2, 0.1, "ones")
    mia, mja, ma, m = mask
    MF = pkg.csr_to_pipeline_format_f32(mia, mja, ma, m, len(mja), 8)
    with pytest.raises(pkg.SpmmB200Error):
        MF.sddmm(None)
    MF.spmm("K", m, 30, 8, *W["K"], x, None)
    MF.spmm("Q", m, 30, 8, *W["Q"], x, None)
    MF.sddmm(None)
    with pytest.raises(pkg.SpmmB200Error):
        MF.final(np.zeros(m * 8, np.float32))
    MF.close()


@pytest.mark.gpu
@pytest.mark.parametrize("rowwise", [True, False])
def test_gpu_pipeline_f32_softmax(pkg, L, rowwise):
    """fp32 softmax of y on the device (row-wise = what attention uses; global = the reference's commented-out host
    softmax(y)), against a float64 softmax of the GPU's own y, then the final product on the normalised y."""
    n, kdim = 32, 220
    mask, W, x = f32_case(77, 600, kdim, n, 9, 0.08, "random")
    mia, mja, ma, m = mask
    MF = pkg.csr_to_pipeline_format_f32(mia, mja, ma, m, len(mja), n)
    with pytest.raises(pkg.SpmmB200Error):
        MF.softmax()                                         # no y yet
    V = np.full(m * n, np.nan, np.float32)
    for t, out in (("K", None), ("Q", None), ("V", V)):
        MF.spmm(t, m, kdim, n, *W[t], x, out)
    y = np.full(len(mja), np.nan, np.float32)
    MF.sddmm(y, 1)
    y = (y * np.float32(0.05)).astype(np.float32)            # keep exp() away from a one-hot result
    MF.spmm("final", m, m, n, mia, mja, y, V, np.empty(m * n, np.float32))      # puts the scaled y on the device
    p = np.full(len(mja), np.nan, np.float32)
    MF.softmax(p, rowwise=rowwise)
    y64 = y.astype(np.float64)
    ref = np.empty_like(y64)
    if rowwise:
        for i in range(m):
            seg = y64[mia[i]:mia[i + 1]]
            if seg.size:
                e = np.exp(seg - seg.max()); ref[mia[i]:mia[i + 1]] = e / e.sum()
        sums = np.add.reduceat(p.astype(np.float64), mia[:-1][np.diff(mia) > 0])
        assert np.allclose(sums, 1.0, atol=1e-5)
    else:
        e = np.exp(y64 - y64.max()); ref = e / e.sum()
        assert abs(float(p.astype(np.float64).sum()) - 1.0) <= 1e-5
    assert float(np.abs(p - ref).max()) <= 4e-6 * float(ref.max())
    yf = np.full(m * n, np.nan, np.float32)
    MF.final(yf)
    import scipy.sparse as sp
    want = sp.csr_matrix((p.astype(np.float64), mja, mia), shape=(m, m)) @ V.astype(np.float64).reshape(m, n)
    close32(yf.reshape(m, n), want, "final after softmax")
    MF.close()


@pytest.mark.gpu
def test_gpu_reference_pipeline_harness_f_build(tmp_path):
    """The unchanged pipeline harness compiled with the reference's float defines + our plug-in TU."""
    from test_pipeline import run_pipeline_harness, harness_mask_nnz
    exe = os.path.join(os.path.dirname(GOLDEN), "..", "oracle", "_ref", "sddmm_b200_f.exe")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/sddmm_b200_f.exe not built (make -C oracle pipeline_harness)")
    out, f = run_pipeline_harness(exe, tmp_path, 2048, 1024, 64)
    assert (int(f[2]), int(f[3]), int(f[4])) == (64, 2048, 1024)
    assert (int(f[-3]), int(f[-2]), int(f[-1])) == (2048, 64, harness_mask_nnz(2048))
