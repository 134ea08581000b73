#!/usr/bin/env python
"""Golden vectors for SURVEY 8(a) row a6 ("perfect nnz balance with row carry"), produced by the REFERENCE itself:
benchmark_code/CPU/AMD/spmv_code_bench/spmv_kernel_csr.cpp built as Custom_CSR_PBV (oracle/_ref/libpbv_ref.so,
oracle/ref_shim_pbv.cpp).  Run in the build container (needs /root/reference):  python tests/golden/make_pbv_golden.py
Writes tests/golden/pbv_pins.json: per generated matrix and worker count P the reference's
(thread_i_s, thread_i_e, thread_j_s, thread_j_e) and the FNV-1a-64 of its SpMV result for a seeded x."""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle import loader as L  # noqa: E402
from make_golden import dense_b  # noqa: E402

CASES = ["100000 100000 10 2 normal random 1 0 0 0 14",
         "200000 200000 20 10 gamma random 0.3 1000 0.5 0.5 14",
         "3000 3000 8 3 normal random 0.5 0 0.1 0.1 7",
         "2000 5000 5 4 gamma random 1 100 0 0 3"]


def main():
    out = {}
    for args in CASES:
        ia, ja, a, m, c = L.generate(args)
        nnz = len(ja)
        x = dense_b(c, 1, seed=77)
        entry = {"m": m, "nnz": nnz}
        for P in (3, 8, 148):
            part = np.zeros(4 * P, np.int64)
            L.pbv_ref().ref_pbv_partition(ia, ja, a, m, c, nnz, P, part)
            y = np.full(m, np.nan)
            L.pbv_ref().ref_pbv_spmv(ia, ja, a, m, c, nnz, P, x, y)
            entry[f"P{P}"] = part.tolist() if P <= 8 else L.fnv(part)
            entry[f"fnv_y_P{P}"] = L.fnv(y)
        out[args] = entry
        print(args, "->", entry["P3"])
    with open(os.path.join(HERE, "pbv_pins.json"), "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
