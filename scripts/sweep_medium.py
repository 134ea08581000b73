#!/usr/bin/env python
"""BASELINE configs[4]: parameter sweep over a subset of synthetic_matrices_medium_dataset.txt
(bench/medium_subset.txt) at k in {1, 8, 32, 128} on ONE B200, kernel time by CUDA events inside
the library (spmm_b200_last_kernel_ms), B and A resident.  Every result is compared -- whole output -- with the
reference's own CPU kernel (compute_csr from oracle/_ref, all host threads; its one call is also timed: the "CPU
GFLOP/s" column).  Prints a markdown table and writes gpurun_out/sweep_medium.json.
Matrices come from the reference generator (oracle/_ref).  SWEEP_CPU=0 skips the CPU leg."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g          # noqa: E402
from oracle import loader as L       # noqa: E402  (input generation only)

PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(
    os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0


def main():
    pkg = g.load_package()
    ks = [int(x) for x in os.environ.get("SWEEP_KS", "1,8,32,128").split(",")]
    variants = os.environ.get("SWEEP_VARIANTS", "auto").split(",")
    vmap = {"auto": pkg.AUTO, "rowsplit": pkg.ROWSPLIT, "merge": pkg.MERGE, "bcsr": pkg.BCSR}
    lines = [l.strip() for l in open(os.path.join(ROOT, "bench", "medium_subset.txt")) if l.strip() and l[0] != "#"]
    limit = int(os.environ.get("SWEEP_LIMIT", "0"))
    if limit:
        lines = lines[:limit]
    rows = []
    do_cpu = os.environ.get("SWEEP_CPU", "1") != "0"
    ncores = len(os.sched_getaffinity(0))
    print("| rows | avg | bw | skew | nnz | k | variant | ms | GFLOP/s | alg. GB/s | frac of %.0f GB/s | max rel err vs compute_csr | CPU GFLOP/s (%d thr) |" % (PEAK, ncores))
    print("|---|---|---|---|---|---|---|---|---|---|---|---|---|")
    for args in lines:
        f = args.split()
        t0 = time.time()
        ia, ja, a, m, c = L.generate(args)
        nnz = len(ja)
        gen_s = time.time() - t0
        for k in ks:
            if (m + c) * k * 8 > 24e9:
                continue
            x = np.random.default_rng(1).random(c * k)
            y_ref, cpu_gf = None, None
            if do_cpu:
                t0 = time.perf_counter()
                y_ref = L.spmm_ref(ia, ja, a, m, c, x, k) if L.have_ref() else L.spmm_oracle(ia, ja, a, m, c, x, k)
                cpu_gf = 2.0 * nnz * k / (time.perf_counter() - t0) / 1e9
            for vname in variants:
                mf = pkg.csr_to_format(ia, ja, a, m, c, nnz, k, variant=vmap[vname])
                mf.set_b(x, k)
                for _ in range(9):          # 8 = the k-slice measurement calls of a large-window matrix
                    mf.spmm_resident(k)
                ts = []
                for _ in range(10):
                    mf.spmm_resident(k)
                    ts.append(mf.last_kernel_ms())
                ms = float(np.median(ts))
                name = mf.format_name
                ksl = mf.kslice()
                err = None
                if y_ref is not None:
                    y = np.empty(m * k)
                    mf.fetch_c(y, k)
                    err = 0.0
                    for s0 in range(0, y.size, 1 << 26):
                        yy, rr = y[s0:s0 + (1 << 26)], y_ref[s0:s0 + (1 << 26)]
                        err = max(err, float(np.max(np.abs(yy - rr) / np.maximum(np.abs(rr), 1e-300))))
                    del y
                mf.close()
                byts = 12.0 * nnz + 4.0 * (m + 1) + 8.0 * k * c + 8.0 * k * m
                gf = 2.0 * nnz * k / (ms * 1e-3) / 1e9
                gbs = byts / (ms * 1e-3) / 1e9
                rows.append(dict(args=args, m=m, nnz=nnz, k=k, variant=name, kslice=ksl, ms=ms, gflops=gf, alg_gbs=gbs,
                                 frac=gbs / PEAK, gen_s=gen_s, max_rel_err=err, cpu_gflops=cpu_gf, cpu_threads=ncores))
                print(f"| {m} | {f[2]} | {f[6]} | {f[7]} | {nnz} | {k} | {name.replace('B200_CSR_', '')}{'/w' + str(ksl) if ksl and ksl != k else ''} | {ms:.4f} | {gf:.0f} | {gbs:.0f} | {gbs / PEAK:.3f} | "
                      f"{'-' if err is None else format(err, '.1e')} | {'-' if cpu_gf is None else format(cpu_gf, '.1f')} |", flush=True)
                if err is not None and err > 1e-12:
                    raise SystemExit(f"sweep: {name} k={k} on '{args}' differs from the reference: {err}")
            del y_ref
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(rows, open(os.path.join(ROOT, "gpurun_out", "sweep_medium.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
