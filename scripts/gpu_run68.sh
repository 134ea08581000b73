#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/r68_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r68_pytest.log
tail -3 gpurun_out/r68_pytest.log
timeout 300 python bench.py > gpurun_out/r68_bench.log 2>&1; python - <<'PY'
import json
for line in open("gpurun_out/r68_bench.log"):
    if line.startswith("{"):
        d = json.loads(line); print("ms", d["ms_per_step"], "value", d["value"], "frac", d["roofline"]["frac"], "e2e", d["e2e"], "launches", d["gpu_launches"])
PY
