#!/usr/bin/env python
"""Kernel-variant probe: ms of the row-split family under env knobs on cfg2-like matrices (B, A resident)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g
from oracle import loader as L

CASES = [("1000000 1000000 50 16.6667 normal random 0.05 0 0.05 0.05 14", [32, 64, 128, 16]),
         ("1000000 1000000 50 16.6667 normal random 1 0 0.05 0.05 14", [32]),
         ("8388607 8388607 5 1.6667 normal random 0.05 0 0.05 0.05 14", [32, 128]),
         ("2000000 2000000 20 40 gamma random 0.3 0 0.5 0.5 14", [32])]
GVU = {32: [842, 844, 848], 64: [882, 884, 1642, 1644], 128: [1682, 1684, 3242, 3244], 16: [824, 828, 444]}
CPS = [None, "16", "64"]
only = os.environ.get("PROBE_KNOBS")


def main():
    pkg = g.load_package()
    print("| matrix | k | knobs | ms | GFLOP/s | max rel diff vs default |")
    print("|---|---|---|---|---|---|")
    for args, ks in CASES:
        ia, ja, a, m, c = L.generate(args)
        nnz = len(ja)
        for k in ks:
            x = np.random.default_rng(1).random(c * k)
            y0 = None
            knob_list = [{}] + [dict(SPMM_B200_GVU=str(gv), **({"SPMM_B200_CTAS_PER_SM": cp} if cp else {}))
                                for gv in GVU[k] for cp in CPS]
            for knobs in knob_list:
                for kk in ("SPMM_B200_GVU", "SPMM_B200_VEC", "SPMM_B200_CTAS_PER_SM"):
                    os.environ.pop(kk, None)
                os.environ.update(knobs)
                mf = pkg.csr_to_format(ia, ja, a, m, c, nnz, k, variant=pkg.ROWSPLIT)
                mf.set_b(x, k)
                for _ in range(3):
                    mf.spmm_resident(k)
                ts = []
                for _ in range(10):
                    mf.spmm_resident(k); ts.append(mf.last_kernel_ms())
                ms = float(np.median(ts))
                y = np.empty(m * k); mf.fetch_c(y, k); mf.close()
                d = "-" if y0 is None else "%.2e" % float(np.max(np.abs(y - y0) / np.maximum(np.abs(y0), 1e-300)))
                if y0 is None:
                    y0 = y
                f = args.split()
                print(f"| {f[0]} avg {f[2]} bw {f[6]} {f[4]} | {k} | {knobs or 'default'} | {ms:.4f} | {2.0 * nnz * k / ms / 1e6:.0f} | {d} |", flush=True)


if __name__ == "__main__":
    main()
