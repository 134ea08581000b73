#!/bin/bash
# ncu evidence for the N4 pipeline: launch list of one attention() sequence + full capture of the SDDMM group kernel
mkdir -p gpurun_out
cat > /tmp/pipe_once.py <<'PY'
import os, sys
sys.path.insert(0, os.environ.get("GRAFT_REPO_ROOT", "/root/repo")); sys.path.insert(0, os.path.join(os.environ.get("GRAFT_REPO_ROOT", "/root/repo"), "tests"))
import numpy as np
import __graft_entry__ as g
from pipeline_util import pipeline_case
pkg = g.load_package()
m = kdim = 8192; n = 64
mask, W, x = pipeline_case(3, m, kdim, n, 16, 0.05, avg_w=64.0)
mia, mja, ma, _ = mask
MF = pkg.csr_to_pipeline_format(mia, mja, ma, m, len(mja), n)
for t in "KQV":
    MF.set_weight(t, m, kdim, *W[t])
yf = np.empty(m * n)
for mode in (0, 1, 2):
    for _ in range(3):
        MF.attention(x, yf, mode, 2)
print("ok", MF.launch_count())
PY
timeout 300 python /tmp/pipe_once.py > gpurun_out/r47_plain.log 2>&1 || { tail -5 gpurun_out/r47_plain.log; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r47_launches.csv python /tmp/pipe_once.py > gpurun_out/r47_ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sddmm_attention_group -s 2 -c 1 -o gpurun_out/r47_sddmm_group -f python /tmp/pipe_once.py > gpurun_out/r47_ncu2.log 2>&1
ls -la gpurun_out/r47*; tail -3 gpurun_out/r47_ncu2.log
