"""Times csr_to_format (handle construction) for a small and a large matrix, per variant."""
import os, sys, time
import numpy as np
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as g
from conftest import random_csr
pkg = g.load_package()
for m, avg in ((8192, 64.0), (200000, 50.0)):
    ia, ja, a = random_csr(m, m, avg, 1)
    for variant, name in ((pkg.ROWSPLIT, "rowsplit"), (pkg.AUTO, "auto"), (pkg.MERGE, "merge")):
        ts = []
        for _ in range(6):
            t0 = time.perf_counter()
            mf = pkg.csr_to_format(ia, ja, a, m, m, len(ja), 64, variant=variant, b_layout=pkg.B_ROWMAJOR)
            ts.append((time.perf_counter() - t0) * 1e3)
            mf.close()
        print(f"m={m} nnz={len(ja)} {name}: create ms {['%.2f' % t for t in ts]}")
