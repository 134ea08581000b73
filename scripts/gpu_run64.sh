#!/bin/bash
# full verification: GPU suite, smoke, both bench arms
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r64_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r64_pytest.log
tail -5 gpurun_out/r64_pytest.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/r64_smoke.log 2>&1; tail -2 gpurun_out/r64_smoke.log
timeout 600 python bench.py > gpurun_out/r64_bench.log 2>&1; tail -1 gpurun_out/r64_bench.log | cut -c1-300
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r64_bench_ref.log 2>&1; tail -1 gpurun_out/r64_bench_ref.log | cut -c1-300
python - <<'PY'
import json
for f in ("gpurun_out/r64_bench.log", "gpurun_out/r64_bench_ref.log"):
    for line in open(f):
        if line.startswith("{"):
            d = json.loads(line); print(f, d.get("impl"), "value", round(d["value"], 1), "ms", round(d["ms_per_step"], 3), "e2e", d["e2e"]["value"], "frac", d.get("roofline", {}).get("frac"), "launches", d.get("gpu_launches"))
PY
