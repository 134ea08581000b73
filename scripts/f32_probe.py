#!/usr/bin/env python
"""fp32 row-split kernel times on the BASELINE matrices (A, B resident; CUDA events via torch on the default stream)."""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g
from oracle import loader as L

CASES = [("cfg2", "1000000 1000000 50 16.6667 normal random 0.05 0 0.05 0.05 14", [32, 64, 128]),
         ("cfg4", "1000000 1000000 50 16.6667 normal diagonal 0.0002 0 1.9 0.95 14", [128]),
         ("cfg5_large", "8388607 8388607 5 1.6667 normal random 0.05 0 0.05 0.05 14", [32, 128])]
PEAK = 6552.0


def main():
    pkg = g.load_package()
    print("| matrix | k | fp32 ms | GFLOP/s | alg. GB/s (8 nnz + 4(m+1) + 4kc + 4km) | frac of %.0f |" % PEAK)
    print("|---|---|---|---|---|---|")
    only = os.environ.get("F32_ONLY")
    for name, args, ks in CASES:
        if only and name != only:
            continue
        ia, ja, a, m, c = L.generate(args)
        a = a.astype(np.float32); nnz = len(ja)
        for k in ks:
            mf = pkg.csr_to_format_f32(ia, ja, a, m, c, nnz, k)
            B = torch.rand(c * k, dtype=torch.float32, device="cuda")
            Cd = torch.empty(m * k, dtype=torch.float32, device="cuda")
            s = torch.cuda.current_stream()
            for _ in range(3):
                mf.spmm_device(B.data_ptr(), Cd.data_ptr(), k, s.cuda_stream)
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(11)]
            torch.cuda.synchronize()
            ev[0].record(s)
            for i in range(10):
                mf.spmm_device(B.data_ptr(), Cd.data_ptr(), k, s.cuda_stream); ev[i + 1].record(s)
            torch.cuda.synchronize()
            ms = float(np.median([ev[i].elapsed_time(ev[i + 1]) for i in range(10)]))
            mf.close()
            byts = 8.0 * nnz + 4.0 * (m + 1) + 4.0 * k * c + 4.0 * k * m
            print(f"| {name} | {k} | {ms:.4f} | {2.0 * nnz * k / ms / 1e6:.0f} | {byts / ms / 1e6:.0f} | {byts / ms / 1e6 / PEAK:.3f} |", flush=True)


if __name__ == "__main__":
    main()
