#!/usr/bin/env python
"""Small structured cases for the TF32 blocked kernel: identity, shifted diagonals, a dense band -- prints where the
result differs from the fp32 oracle (which rows / columns, and which B entry the wrong value equals)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g  # noqa: E402
from oracle import loader as L  # noqa: E402


def csr_from_rows(rows, m):
    ia = np.zeros(m + 1, np.int32)
    ia[1:] = np.cumsum([len(r) for r in rows])
    ja = np.concatenate([np.asarray(r, np.int32) for r in rows]) if m else np.zeros(0, np.int32)
    return ia, ja.astype(np.int32)


def report(name, pkg, ia, ja, a, m, c, k):
    x = (np.arange(c * k, dtype=np.float32) % 251.0) + 1.0             # integers: exact in TF32
    y_ref = L.spmm_oracle_f32(ia, ja, a, m, c, x, k).reshape(m, k)
    mf = pkg.csr_to_format_f32(ia, ja, a, m, c, len(ja), k, variant=pkg.BCSR)
    y = np.full(m * k, np.nan, np.float32)
    mf.spmm(x, y, k)
    mf.close()
    y = y.reshape(m, k)
    bad = np.argwhere(y != y_ref)
    print(f"{name}: m={m} c={c} k={k} kernel={mf.format_name} mismatches {len(bad)} / {m * k}; zeros {int((y == 0).sum())}; nan {int(np.isnan(y).sum())}")
    for (i, n) in bad[:6]:
        print(f"   y[{i},{n}] = {y[i, n]}  ref {y_ref[i, n]}")
    if len(bad):
        rows = np.unique(bad[:, 0]); cols = np.unique(bad[:, 1])
        print(f"   bad rows: {rows[:12]}... ({len(rows)})   bad cols: {cols[:12]}... ({len(cols)})")


def main():
    pkg = g.load_package()
    for k in (32, 128):
        m = c = 256
        rows = [[i] for i in range(m)]
        ia, ja = csr_from_rows(rows, m)
        report("identity", pkg, ia, ja, np.ones(len(ja), np.float32), m, c, k)
        rows = [[min(i + 5, c - 1)] for i in range(m)]
        ia, ja = csr_from_rows(rows, m)
        report("shift+5", pkg, ia, ja, np.full(len(ja), 2.0, np.float32), m, c, k)
        rows = [list(range(max(0, i - 3), min(c, i + 4))) for i in range(m)]
        ia, ja = csr_from_rows(rows, m)
        report("band7", pkg, ia, ja, (np.arange(len(ja)) % 5 + 1).astype(np.float32), m, c, k)
        m, c = 128, 40
        rows = [list(range(c)) for i in range(m)]
        ia, ja = csr_from_rows(rows, m)
        report("dense128x40", pkg, ia, ja, (np.arange(len(ja)) % 3 + 1).astype(np.float32), m, c, k)


if __name__ == "__main__":
    main()
