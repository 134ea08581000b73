#!/usr/bin/env python
"""L2-aware k slicing probe: kernel ms of the row-split CSR SpMM with SPMM_B200_KSLICE in {0, ...}
on matrices whose touched B window x 8k bytes exceeds the 126 MB L2.  B, A resident; CUDA events
inside the library.  Also checks that the sliced result is bit-identical to the unsliced one."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g          # noqa: E402
from oracle import loader as L       # noqa: E402  (input generation only)

CASES = [
    ("8388607 8388607 5 1.6667 normal random 0.05 0 0.05 0.05 14", 128, [0, 64, 32, 16]),
    ("8388607 8388607 5 1.6667 normal random 0.3 0 0.05 0.05 14", 128, [0, 64, 32, 16]),
    ("3517803 3517803 10 3.3333 normal random 0.3 0 0.05 0.05 14", 128, [0, 64, 32, 16]),
    ("1000000 1000000 50 16.6667 normal random 1 0 0.05 0.05 14", 32, [0, 16, 8]),
    ("1000000 1000000 50 16.6667 normal random 0.3 0 0.05 0.05 14", 32, [0, 16, 8]),
    ("1000000 1000000 50 16.6667 normal random 0.05 0 0.05 0.05 14", 128, [0, 64, 32, 16]),
    ("1787736 1787736 20 6.6667 normal random 0.3 0 0.05 0.05 14", 128, [0, 64, 32, 16]),
]


def main():
    pkg = g.load_package()
    print("| matrix | k | window MB | KSLICE | ms | GFLOP/s | same bits |")
    print("|---|---|---|---|---|---|---|")
    for args, k, slices in CASES:
        ia, ja, a, m, c = L.generate(args)
        nnz = len(ja)
        span = int(ja.max()) - int(ja.min()) + 1
        x = np.random.default_rng(1).random(c * k)
        mf = pkg.csr_to_format(ia, ja, a, m, c, nnz, k, variant=pkg.ROWSPLIT)
        mf.set_b(x, k)
        y0 = None
        for w in slices:
            os.environ["SPMM_B200_KSLICE"] = str(w)
            for _ in range(3):
                mf.spmm_resident(k)
            ts = []
            for _ in range(10):
                mf.spmm_resident(k)
                ts.append(mf.last_kernel_ms())
            ms = float(np.median(ts))
            y = np.empty(m * k); mf.fetch_c(y, k)
            same = "-" if y0 is None else str(bool(np.array_equal(y, y0)))
            if y0 is None:
                y0 = y
            print(f"| {args.split()[0]} avg {args.split()[2]} bw {args.split()[6]} | {k} | {span * 8 * k / 1e6:.0f} | {w} | {ms:.4f} | "
                  f"{2.0 * nnz * k / ms / 1e6:.0f} | {same} |", flush=True)
        mf.close()
        del y0


if __name__ == "__main__":
    main()
