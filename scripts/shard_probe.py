"""Kernel time of ONE rank's shard (rank 0 of 8, cfg2) on one GPU against the CTA count of the work list."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g
from oracle import loader as L
pkg = g.load_package()
ia, ja, a, m, c = L.generate("1000000 1000000 50 16.6667 normal random 0.05 0 0.05 0.05 14")
nnz = len(ja); k = 32
x = np.random.default_rng(1).random(c * k)
for nr in (8, 4, 1):
    for cps in ("2", "4", "8", "16", "32", "64"):
        os.environ["SPMM_B200_CTAS_PER_SM"] = cps
        mf = pkg.csr_to_format(ia, ja, a, m, c, nnz, k, variant=pkg.ROWSPLIT, rank=0, nranks=nr)
        mf.set_b(x, k)
        for _ in range(5): mf.spmm_resident(k)
        ts = []
        for _ in range(20):
            mf.spmm_resident(k); ts.append(mf.last_kernel_ms())
        print(f"shard 0/{nr} rows {mf.m} nnz {mf.nnz} CTAs/SM {cps}: median {np.median(ts):.4f} ms  min {np.min(ts):.4f}", flush=True)
        mf.close()
