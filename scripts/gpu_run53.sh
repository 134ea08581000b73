#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_pipeline.py -m gpu -q > gpurun_out/r53_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r53_pytest.log
tail -4 gpurun_out/r53_pytest.log
cat > /tmp/pipe_fused.py <<'PY'
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
for _ in range(4):
    MF.attention(x, yf, 1, 2)
print("ok", MF.launch_count(), MF.last_kernel_ms())
PY
timeout 300 python /tmp/pipe_fused.py > gpurun_out/r53_plain.log 2>&1 || { tail -5 gpurun_out/r53_plain.log; exit 1; }
cat gpurun_out/r53_plain.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:attention_fused -s 2 -c 1 -o gpurun_out/r53_attention_fused -f python /tmp/pipe_fused.py > gpurun_out/r53_ncu.log 2>&1
tail -2 gpurun_out/r53_ncu.log
