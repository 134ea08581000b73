#!/usr/bin/env python
"""Kernel time of one workload (or of one rank's shard of it) under a list of handle options.

  python scripts/knob_probe.py --workload cfg2 [--k 32] [--shard R/N] [--variant rowsplit] \
         --set TICKET=0 --set TICKET=8 --set TICKET=8,STAGE=1 ...

Device-resident B and C, CUDA events around each call (spmm_b200_last_kernel_ms), 5 warm-ups + 30 timed calls,
median / min.  Every configuration's result is compared with the first one's (bit-identical for the row-split family).
Used for profiles/r2_rowsplit_stage_probe.md and r2_shard_probe.md."""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g  # noqa: E402
from bench import WORKLOADS  # noqa: E402
from oracle import loader as L  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="cfg2")
    ap.add_argument("--k", type=int, default=None)
    ap.add_argument("--shard", default=None, help="R/N: rank R of N nnz-balanced row shards")
    ap.add_argument("--variant", default="rowsplit")
    ap.add_argument("--set", action="append", default=[], help="comma-separated NAME=VALUE handle options of one configuration")
    ap.add_argument("--calls", type=int, default=30)
    ap.add_argument("--gen", default=None, help="generator arguments instead of a named workload (then --k is required)")
    args = ap.parse_args()
    pkg = g.load_package()
    gen, k = (args.gen, args.k) if args.gen else WORKLOADS[args.workload]
    k = args.k or k
    ia, ja, a, m, c = L.generate(gen)
    nnz = len(ja)
    kw = {}
    if args.shard:
        r, n = args.shard.split("/")
        kw = dict(rank=int(r), nranks=int(n))
    variant = {"auto": pkg.AUTO, "rowsplit": pkg.ROWSPLIT, "merge": pkg.MERGE, "bcsr": pkg.BCSR}[args.variant]
    x = np.random.default_rng(1).random(c * k)
    mf = pkg.csr_to_format(ia, ja, a, m, c, nnz, k, variant=variant, **kw)
    mf.set_b(x, k)
    print(f"# {args.workload} k={k} shard={args.shard} rows={mf.m} nnz={mf.nnz} kernel={mf.format_name}", flush=True)
    y0 = None
    for conf in (args.set or [""]):
        opts = dict(kv.split("=") for kv in conf.split(",") if kv)
        for name, val in opts.items():
            mf.set_option(name, int(val))
        for _ in range(5):
            mf.spmm_resident(k)
        ts = []
        for _ in range(args.calls):
            mf.spmm_resident(k)
            ts.append(mf.last_kernel_ms())
        y = np.empty(mf.m * k)
        mf.fetch_c(y, k)
        same = "ref" if y0 is None else ("identical" if np.array_equal(y, y0) else f"DIFFERS max {np.max(np.abs(y - y0)):.3e}")
        if y0 is None:
            y0 = y
        gf = 2.0 * mf.nnz * k / (np.median(ts) * 1e6)
        print(f"{conf or '(defaults)':40s} median {np.median(ts):.4f} ms  min {np.min(ts):.4f} ms  {gf:8.0f} GFLOP/s  {same}", flush=True)
    mf.close()


if __name__ == "__main__":
    main()
