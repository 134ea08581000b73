#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "direct_store or k_sliced or harness" > gpurun_out/r56_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r56_pytest.log
tail -8 gpurun_out/r56_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r56_bench_direct.log 2>&1; python - <<'PY'
import json
for f in ("gpurun_out/r56_bench_direct.log",):
    for line in open(f):
        if line.startswith("{"):
            d = json.loads(line); print(f, "ms_per_step", d["ms_per_step"], "e2e", d["e2e"])
PY
SPMM_B200_DIRECT_STORE=0 timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r56_bench_staged.log 2>&1; python - <<'PY'
import json
for line in open("gpurun_out/r56_bench_staged.log"):
    if line.startswith("{"):
        d = json.loads(line); print("staged: ms_per_step", d["ms_per_step"], "e2e", d["e2e"])
PY
