#!/bin/bash
# Round 2, GPU call 22: last sanity on one GPU after the routing / no-alloc commits: smoke, both bench arms
mkdir -p gpurun_out
O=gpurun_out
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > $O/r2o_smoke.log 2>&1; tail -2 $O/r2o_smoke.log
timeout 600 python bench.py --impl reference > $O/r2o_bench_reference.log 2>&1; tail -c 400 $O/r2o_bench_reference.log; echo
timeout 600 python bench.py > $O/r2o_bench_default.log 2>&1; python - <<PY
import json
for ln in open("$O/r2o_bench_default.log"):
    if ln.startswith("{"):
        d = json.loads(ln); e = d["e2e"]
        print("default", "ms", round(d["ms_per_step"], 4), "GF", round(d["value"]), "frac", round(d["roofline"]["frac"], 4), "e2e_ms", round(e["ms_per_step"], 3), e["pipeline"], "parity", d["parity"], "cpu", d["cpu_baseline"]["value"], "launches", d["gpu_launches"], d["clocks"])
PY
