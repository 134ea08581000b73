#!/bin/bash
mkdir -p gpurun_out
for w in 8 16; do
  SPMM_B200_PIPELINE_COLS=$w timeout 600 python bench.py --steps 10 --warmup 5 > gpurun_out/r59_bench_w$w.log 2>&1
  python - "$w" <<'PY'
import json, sys
for line in open(f"gpurun_out/r59_bench_w{sys.argv[1]}.log"):
    if line.startswith("{"):
        d = json.loads(line); print("direct, slice width", sys.argv[1], "e2e ms", d["e2e"]["ms_per_step"], "GFLOP/s", d["e2e"]["value"])
PY
done
