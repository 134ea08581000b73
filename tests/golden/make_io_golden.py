#!/usr/bin/env python
"""Writes the small Matrix Market / DLMC fixtures under tests/golden/io/ and the arrays the REFERENCE's
own readers (oracle/_ref/libio_ref.so = mtx_read / smtx_read compiled from /root/reference) return for
them -> tests/golden/io_reference.npz.  Also runs the reference HARNESS (spmm_csr_naive_d.exe) on one
.mtx file and stores the "errors spmv:" line its CheckAccuracy prints -> pins oracle_check_accuracy.
Run in the build container (needs /root/reference): python tests/golden/make_io_golden.py"""
import os
import re
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import loader as L   # noqa: E402

IO = os.path.join(HERE, "io")


def write(name, text):
    with open(os.path.join(IO, name), "w") as f:
        f.write(text)


def coords(rng, m, n, nnz, lower=False, strict=False):
    seen = set()
    while len(seen) < nnz:
        r, c = int(rng.integers(1, m + 1)), int(rng.integers(1, n + 1))
        if lower and c > r:
            r, c = c, r
        if strict and r == c:
            continue
        seen.add((r, c))
    out = list(seen)
    rng.shuffle(out)
    return out


def main():
    os.makedirs(IO, exist_ok=True)
    rng = np.random.default_rng(7)
    # real general, with comments and awkward number spellings
    ents = coords(rng, 40, 30, 150)
    vals = rng.standard_normal(len(ents)) * 10.0 ** rng.integers(-8, 8, len(ents))
    body = "".join(f"{r} {c} {v:.17g}\n" for (r, c), v in zip(ents, vals))
    write("real_general.mtx", "%%MatrixMarket matrix coordinate real general\n% a comment\n%another\n40 30 150\n" + body)
    # real symmetric (lower triangle + diagonal)
    ents = coords(rng, 50, 50, 200, lower=True)
    vals = rng.random(len(ents))
    write("real_symmetric.mtx", "%%MatrixMarket matrix coordinate real symmetric\n50 50 200\n" +
          "".join(f"{r} {c} {float(v)!r}\n" for (r, c), v in zip(ents, vals)))
    # real skew-symmetric (strictly lower)
    ents = coords(rng, 30, 30, 90, lower=True, strict=True)
    vals = rng.random(len(ents)) - 0.5
    write("real_skew.mtx", "%%MatrixMarket matrix coordinate real skew-symmetric\n30 30 90\n" +
          "".join(f"{r}\t{c}\t{v:.6e}\n" for (r, c), v in zip(ents, vals)))
    # integer general
    ents = coords(rng, 25, 60, 120)
    vals = rng.integers(-1000, 1000, len(ents))
    write("integer_general.mtx", "%%MatrixMarket matrix coordinate integer general\n25 60 120\n" +
          "".join(f"{r} {c} {v}\n" for (r, c), v in zip(ents, vals)))
    # pattern symmetric
    ents = coords(rng, 64, 64, 300, lower=True)
    write("pattern_symmetric.mtx", "%%MatrixMarket matrix coordinate pattern symmetric\n64 64 300\n" +
          "".join(f"{r} {c}\n" for (r, c) in ents))
    # complex Hermitian
    ents = coords(rng, 20, 20, 70, lower=True)
    vals = rng.standard_normal((len(ents), 2))
    write("complex_hermitian.mtx", "%%MatrixMarket matrix coordinate complex Hermitian\n20 20 70\n" +
          "".join(f"{r} {c} {v[0]:.15g} {v[1]:.15g}\n" for (r, c), v in zip(ents, vals)))
    # no banner: the reference treats it as coordinate real general (matrix_market.c:170-177)
    ents = coords(rng, 10, 12, 30)
    vals = rng.random(len(ents))
    write("no_banner.mtx", "10 12 30\n" + "".join(f"{r} {c} {v:.12f}\n" for (r, c), v in zip(ents, vals)))
    # DLMC .smtx: "m, k, nnz" / row offsets / column indices
    m, k = 48, 96
    deg = rng.integers(0, 12, m); deg[5] = 0
    ia = np.concatenate([[0], np.cumsum(deg)])
    ja = np.concatenate([np.sort(rng.choice(k, d, replace=False)) for d in deg]).astype(int)
    write("dlmc_small.smtx", f"{m}, {k}, {ia[-1]}\n" + " ".join(map(str, ia)) + "\n" + " ".join(map(str, ja)) + "\n")

    out = {}
    for f in sorted(os.listdir(IO)):
        p = os.path.join(IO, f)
        key = f.replace(".", "_")
        if f.endswith(".mtx"):
            mm, kk, R, C, V, field = L.mtx_read_ref(p)
            out[key + "_shape"] = np.array([mm, kk]); out[key + "_R"] = R; out[key + "_C"] = C; out[key + "_V"] = V
            out[key + "_field"] = np.array(field)
        else:
            mm, kk, ia_r, ja_r = L.smtx_read_ref(p)
            out[key + "_shape"] = np.array([mm, kk]); out[key + "_ia"] = ia_r; out[key + "_ja"] = ja_r

    # the reference harness end to end on a .mtx file: reader -> coo_to_csr -> compute_csr -> CheckAccuracy
    exe = os.path.join(ROOT, "oracle", "_ref", "spmm_csr_naive_d.exe")
    env = dict(os.environ, USE_PROCESSES="0", USE_ARTIFICIAL_MATRICES="0", USE_DLCM_MATRICES="0", NUM_COLS="4",
               COOLDOWN="0", RAPL_REGISTERS="0", OMP_NUM_THREADS="4")
    r = subprocess.run([exe, os.path.join(IO, "real_symmetric.mtx")], env=env, capture_output=True, text=True, timeout=300)
    line = [l for l in r.stdout.splitlines() if l.startswith("errors spmv:")][-1]
    print(line)
    names = ["mae", "max_ae", "mse", "mape", "smape", "lnQ_error", "mlare", "gmare"]
    vals = [float(re.search(name + r"=([^,\s]+)", line).group(1)) for name in names]
    out["harness_errors_real_symmetric_k4"] = np.array(vals)
    out["harness_errors_line"] = np.array(line)
    np.savez_compressed(os.path.join(HERE, "io_reference.npz"), **out)
    print("wrote", len(out), "arrays")


if __name__ == "__main__":
    main()
