#!/usr/bin/env python
"""Host-buffer call (spmm_b200_run: upload | kernel | download pipeline) of one workload under a list of handle
options: wall time per call and the event-stamped timeline of the last call.  profiles/r2_e2e_probe.md."""
import argparse
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g  # noqa: E402
from bench import WORKLOADS  # noqa: E402
from oracle import loader as L  # noqa: E402


def main():
    import torch
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="cfg2")
    ap.add_argument("--k", type=int, default=None)
    ap.add_argument("--set", action="append", default=[])
    ap.add_argument("--multi", type=int, default=0, help="drive this many GPUs from this one process (spmm_b200_*_multi)")
    args = ap.parse_args()
    pkg = g.load_package()
    gen, k = WORKLOADS[args.workload]
    k = args.k or k
    ia, ja, a, m, c = L.generate(gen)
    nnz = len(ja)
    x = torch.empty(c * k, dtype=torch.float64, pin_memory=True)
    x.numpy()[:] = np.random.default_rng(42).random(c * k)
    y = torch.empty(m * k, dtype=torch.float64, pin_memory=True)
    if args.multi:
        mf = pkg.csr_to_format_multi(ia, ja, a, m, c, nnz, k, args.multi)
        print(f"# {args.workload} k={k} multi={args.multi} peer_access={mf.peer_access} bounds={mf.bounds}")
    else:
        mf = pkg.csr_to_format(ia, ja, a, m, c, nnz, k)
        print(f"# {args.workload} k={k} kernel={mf.format_name}")
    y0 = None
    for conf in (args.set or [""]):
        for kv in conf.split(","):
            if kv:
                name, val = kv.split("=")
                mf.set_option(name, int(val))
        for _ in range(3):
            mf.spmm(x.numpy(), y.numpy(), k)
        ts = []
        for _ in range(10):
            t0 = time.perf_counter()
            mf.spmm(x.numpy(), y.numpy(), k)
            ts.append((time.perf_counter() - t0) * 1e3)
        same = ""
        if y0 is None:
            y0 = y.numpy().copy()
        else:
            same = "identical" if np.array_equal(y.numpy(), y0) else "DIFFERS"
        band = "" if args.multi else f"rowband {mf.last_rowband()}  "
        print(f"{conf or '(defaults)':36s} median {np.median(ts):.3f} ms  min {np.min(ts):.3f} ms  kernels {mf.last_kernel_ms():.3f} ms  {band}{same}", flush=True)
        if args.multi:
            route, probe = mf.routes()
            print(f"    download routes {route}  probe GB/s [all, lower, upper, even, odd] {[round(v, 1) for v in probe]}", flush=True)
        if not args.multi:
            for s, t in enumerate(mf.last_timeline()):
                print(f"    slice {s}: " + "  ".join(f"{kk} {v:.3f}" for kk, v in t.items()), flush=True)
    mf.close()


if __name__ == "__main__":
    main()
