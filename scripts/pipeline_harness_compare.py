"""Runs the reference's pipeline harness (pipeline_code_bench/sddmm_bench.cpp, built by `make -C oracle
pipeline_harness`) twice on the same DLMC weight files: linked with the reference's own kernel TU
(sddmm_taco_naive_d.exe, CPU, OpenMP) and linked with our plug-in TU (sddmm_b200_d.exe, one B200), and prints the
CSV line each writes to stderr.  The mask is drawn by the harness itself (time-seeded), so the two runs see
different masks of the same size and fill.
   python scripts/pipeline_harness_compare.py [length] [n] [threads]"""
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import random_csr          # noqa: E402

length = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
n = int(sys.argv[2]) if len(sys.argv) > 2 else 64
threads = sys.argv[3] if len(sys.argv) > 3 else str(os.cpu_count())
tmp = tempfile.mkdtemp()
files = []
for i, t in enumerate("kqv"):
    ia, ja, _ = random_csr(length, length, 64.0, 50 + i)
    path = os.path.join(tmp, f"w_{t}.smtx")
    with open(path, "w") as f:
        f.write(f"{length}, {length}, {len(ja)}\n" + " ".join(map(str, ia)) + "\n" + " ".join(map(str, ja)) + "\n")
    files.append(path)
env = dict(os.environ, USE_PROCESSES="0", USE_ARTIFICIAL_MATRICES="0", USE_DLCM_MATRICES="1", COOLDOWN="0",
           RAPL_REGISTERS="0", OMP_NUM_THREADS=threads, BAND_SIZE="16", SPARSE_ATTENTION_TYPE="band_and_random",
           SDDMM_SPARSIFICATION_TYPE="none", SPARSITY="0.05", L_SPARSITY="0.1", NUM_COLS=str(n))
labels = ("matrix_name,num_threads,input_columns,csr_m,csr_k,csr_nnz,time_spmm_K,time_spmm_Q,time_spmm_V,time_sddmm,"
          "time_final_spmm,time,gflops_spmm_K,gflops_spmm_Q,gflops_spmm_V,gflops_sddmm,gflops_final_spmm,gflops,"
          "csr_mem_footprint,m,n,nnz").split(",")
print(f"# pipeline harness, length {length}, n {n}, OMP threads {threads}")
for exe, extra in (("sddmm_taco_naive_d.exe", {}), ("sddmm_b200_d.exe", {"SPMM_B200_SDDMM_MODE": "0"}),
                   ("sddmm_b200_d.exe", {"SPMM_B200_SDDMM_MODE": "1"})):
    path = os.path.join(ROOT, "oracle", "_ref", exe)
    if not os.path.exists(path):
        print(exe, "not built")
        continue
    r = subprocess.run([path] + files, env=dict(env, **extra), capture_output=True, text=True, timeout=900)
    line = [ln for ln in r.stderr.splitlines() if ln.startswith(files[0])]
    if r.returncode != 0 or not line:
        print(exe, "FAILED rc", r.returncode, r.stdout[-500:], r.stderr[-500:])
        continue
    f = dict(zip(labels, line[0].split(",")))
    print(f"{exe} {extra}: " + ", ".join(f"{k}={f[k]}" for k in labels[6:18]) + f", mask nnz={f['nnz']}")
