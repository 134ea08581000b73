#!/usr/bin/env python
"""fp32 ("_f") kernels on one workload: row-split, merge-path and the TF32 blocked tensor-core kernel (tcgen05 / TMEM).
Kernel time (CUDA events, device-resident operands), GFLOP/s, fraction of the fp32 HBM roofline
(8*nnz + 4*(m+1) + 4*k*c + 4*k*m bytes) and the error against the fp32 oracle on the first rows.
profiles/r2_tf32_probe.md."""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g  # noqa: E402
from bench import WORKLOADS, measured_peak_gbs  # noqa: E402
from oracle import loader as L  # noqa: E402


def main():
    import torch
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="cfg4")
    ap.add_argument("--ks", default="32,64,128")
    ap.add_argument("--variants", default="rowsplit,merge,bcsr")
    ap.add_argument("--set", default="", help="comma-separated NAME=VALUE handle options (e.g. TF32_STAGES=3)")
    args = ap.parse_args()
    pkg = g.load_package()
    gen, _ = WORKLOADS[args.workload]
    ia, ja, a, m, c = L.generate(gen)
    a = a.astype(np.float32)
    nnz = len(ja)
    peak, _ = measured_peak_gbs()
    check_rows = min(m, 20000)
    for k in [int(v) for v in args.ks.split(",")]:
        x = np.random.default_rng(k).random(c * k, dtype=np.float32)
        b_rm = torch.from_numpy(np.ascontiguousarray(x.reshape(k, c).T)).cuda()
        y_ref = L.spmm_oracle_f32(ia[: check_rows + 1].copy(), ja, a, check_rows, c, x, k).astype(np.float64)
        bytes_ = 8.0 * nnz + 4.0 * (m + 1) + 4.0 * k * c + 4.0 * k * m
        for vname in args.variants.split(","):
            variant = {"rowsplit": pkg.ROWSPLIT, "merge": pkg.MERGE, "bcsr": pkg.BCSR, "auto": pkg.AUTO}[vname]
            mf = pkg.csr_to_format_f32(ia, ja, a, m, c, nnz, k, variant=variant)
            for kv in args.set.split(","):
                if kv:
                    mf.set_option(kv.split("=")[0], int(kv.split("=")[1]))
            c_dev = torch.empty(m * k, dtype=torch.float32, device="cuda")
            ts = []
            for i in range(25):
                mf.spmm_device(b_rm.data_ptr(), c_dev.data_ptr(), k, torch.cuda.current_stream().cuda_stream)
                torch.cuda.synchronize()
                if i >= 5:
                    ts.append(mf.last_kernel_ms())
            y = c_dev[: check_rows * k].cpu().numpy().astype(np.float64)
            err = float(np.max(np.abs(y - y_ref) / np.maximum(np.abs(y_ref), 1e-30)))
            t = float(np.median(ts))
            print(json.dumps({"workload": args.workload, "k": k, "kernel": mf.format_name + (" " + args.set if args.set else ""), "ms": round(t, 4),
                              "gflops": round(2.0 * nnz * k / t / 1e6, 1), "frac_hbm": round(bytes_ / (t * 1e-3) / 1e9 / peak, 4),
                              "max_rel_err_vs_fp32_oracle": err, "rows_checked": check_rows}), flush=True)
            mf.close()


if __name__ == "__main__":
    main()
